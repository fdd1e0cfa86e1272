"""GPU parity: the CUDA voxelizer against the CPU oracle (bit-exact) and the reference's golden digests."""
import numpy as np
import pytest
import torch

from oracle import cpu_ops
from gndnet_b200 import synth
from tests import cases

pytestmark = pytest.mark.gpu


def _shifted(cfg, raw):
    p = raw.copy()
    cpu_ops.shift_cloud(p, cfg.lidar_height)
    return p


@pytest.mark.parametrize("name", list(cases.CASES))
def test_points_to_voxel_dropin_bit_exact(name, cuda):
    """numpy in / numpy out, same call as predict_ground.py:114."""
    from gndnet_b200.point_cloud_ops import points_to_voxel
    cfg, raw, gold = cases.load_case(name)
    pts = _shifted(cfg, raw)
    keep = pts.copy()
    v, c, n = points_to_voxel(pts, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
    assert pts.tobytes() == keep.tobytes(), "points_to_voxel must not mutate its input"
    assert v.dtype == np.float32 and c.dtype == np.int32 and n.dtype == np.int32
    ov, oc, on = cpu_ops.points_to_voxel(pts, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
    assert v.shape == ov.shape and len(n) == gold["M"]
    assert c.tobytes() == oc.tobytes(), "coors differ"
    assert n.tobytes() == on.tobytes(), "num_points differ"
    assert v.tobytes() == ov.tobytes(), "voxels differ"
    assert synth.digest(v) == gold["voxels_digest"] and synth.digest(c) == gold["coors_digest"]
    assert synth.digest(n) == gold["num_points_digest"]


def test_points_to_voxel_float64_is_cast_with_a_warning(cuda):
    """predict_ground.py:129-131 builds a float64 cloud for short .npy files and the reference then computes in float64;
    here such a cloud is voxelized in float32, loudly, and the pillars come back in the caller's dtype."""
    from gndnet_b200.point_cloud_ops import points_to_voxel
    cfg, raw, _ = cases.load_case("small_kitti")
    pts = _shifted(cfg, raw)
    v32, c32, n32 = points_to_voxel(pts, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
    with pytest.warns(UserWarning, match="float64"):
        v64, c64, n64 = points_to_voxel(pts.astype(np.float64), cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True,
                                        cfg.max_voxels)
    assert v64.dtype == np.float64 and np.array_equal(v64, v32.astype(np.float64))
    assert c64.tobytes() == c32.tobytes() and n64.tobytes() == n32.tobytes()
    with pytest.raises(TypeError):
        points_to_voxel(pts.astype(np.float16), cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)


@pytest.mark.parametrize("name", ["small_kitti", "boundary_kitti"])
def test_points_to_voxel_non_reversed(name, cuda):
    from gndnet_b200.point_cloud_ops import points_to_voxel
    cfg, raw, _ = cases.load_case(name)
    pts = _shifted(cfg, raw)
    a = points_to_voxel(torch.from_numpy(pts).cuda(), cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, False, cfg.max_voxels)
    b = cpu_ops.points_to_voxel(pts, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, False, cfg.max_voxels)
    assert all(x.is_cuda for x in a)
    assert all(x.cpu().numpy().tobytes() == y.tobytes() for x, y in zip(a, b))


def test_fused_batched_ragged_with_z_shift(cuda):
    """The fused path: raw scans, z-shift inside the kernel, ragged batch via counts, pillar lists instead
    of the dense tensor.  Every scan must equal the oracle run on that scan alone."""
    from gndnet_b200.point_cloud_ops import voxelize_batch
    names = ["hdl64_s0", "uniform_s0", "small_kitti", "empty", "boundary_kitti", "real_000000", "single_cell"]
    loaded = [cases.load_case(n) for n in names]
    cfg = loaded[0][0]
    nmax = max(len(r) for _, r, _ in loaded)
    batch = np.full((len(names), nmax, 4), 1e9, dtype=np.float32)          # padding rows are out of range anyway
    counts = np.array([len(r) for _, r, _ in loaded], dtype=np.int32)
    for b, (_, r, _) in enumerate(loaded):
        batch[b, :len(r)] = r
    out = voxelize_batch(torch.from_numpy(batch).cuda(), cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel,
                         cfg.max_voxels, True, counts=torch.from_numpy(counts).cuda(), z_shift=cfg.lidar_height,
                         want_slot=True)
    out = {k: v.cpu().numpy() for k, v in out.items()}
    for b, (name, (_, raw, gold)) in enumerate(zip(names, loaded)):
        pts = _shifted(cfg, raw)
        _, oc, on, oslot = cpu_ops.points_to_voxel(pts, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True,
                                                   cfg.max_voxels, return_point_slot=True)
        m = int(out["voxel_count"][b])
        assert m == len(on) == gold["M"], name
        assert np.array_equal(out["coors"][b, :m], oc), name
        assert np.array_equal(out["num_points"][b, :m], on), name
        assert np.array_equal(out["point_slot"][b, :len(raw)].astype(np.int64), oslot), name
        assert (out["point_slot"][b, len(raw):] == -1).all()
        off = out["point_offsets"][b, :m + 1]
        assert off[0] == 0 and np.array_equal(np.diff(off), on), name
        # pillar lists: point indices of each voxel in input order
        order = np.argsort(oslot[oslot >= 0], kind="stable")
        expect = np.nonzero(oslot >= 0)[0][order]
        assert np.array_equal(out["point_list"][b, :off[m]], expect.astype(np.int32)), name


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_random_grids_and_limits(seed, cuda):
    """Random clouds over random grids / max_points / max_voxels, duplicates and shuffles included."""
    from gndnet_b200.point_cloud_ops import points_to_voxel
    rng = np.random.default_rng(100 + seed)
    for _ in range(6):
        n = int(rng.integers(1, 40000))
        f = int(rng.choice([3, 4, 5]))
        vs = [float(rng.choice([0.25, 0.4, 1.0, 2.5])), float(rng.choice([0.3, 0.5, 1.0])), float(rng.choice([2.0, 4.0, 8.0]))]
        lo = [-float(rng.integers(5, 40)), -float(rng.integers(5, 40)), -4.0]
        hi = [lo[0] + vs[0] * int(rng.integers(3, 150)), lo[1] + vs[1] * int(rng.integers(3, 150)), 4.0]
        pts = (rng.random((n, f)) * 2 - 1).astype(np.float32) * np.array([60, 60, 5] + [1] * (f - 3), np.float32)
        pts[rng.integers(0, n, n // 10)] = pts[rng.integers(0, n, n // 10)]       # duplicates
        mp, mv = int(rng.choice([1, 3, 35, 100])), int(rng.choice([7, 100, 2000, 20000]))
        got = points_to_voxel(pts, vs, lo + hi, mp, True, mv)
        exp = cpu_ops.points_to_voxel(pts, vs, lo + hi, mp, True, mv)
        for g, e in zip(got, exp):
            assert g.shape == e.shape and g.tobytes() == e.tobytes()


def test_nan_points_are_dropped(cuda):
    from gndnet_b200.point_cloud_ops import points_to_voxel
    cfg, raw, _ = cases.load_case("small_kitti")
    pts = _shifted(cfg, raw)
    bad = pts.copy()
    bad[5, 0] = np.nan
    bad[17, 2] = np.nan
    got = points_to_voxel(bad, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
    exp = cpu_ops.points_to_voxel(np.delete(pts, [5, 17], axis=0), cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
    for g, e in zip(got, exp):
        assert g.tobytes() == e.tobytes()


def test_maximum_sizes(cuda):
    """The documented envelope: 1.6 M points per scan and 100 000 cells are served bit-exactly (the latter by the
    one-ranking-warp kernel, grids > 11 264 cells), one more of either is refused loudly -- never a silent fallback."""
    from gndnet_b200 import _lib
    from gndnet_b200.point_cloud_ops import points_to_voxel, voxelize_batch
    rng = np.random.default_rng(77)
    # 1.6 M points on the kitti grid
    n = 1_600_000
    pts = (rng.random((n, 4)) * np.array([110, 110, 9, 1]) - np.array([52, 55, 4.5, 0])).astype(np.float32)
    got = points_to_voxel(torch.from_numpy(pts).cuda(), [1, 1, 8], [-47, -50, -4, 53, 50, 4], 100, True, 10000)
    exp = cpu_ops.points_to_voxel(pts, [1, 1, 8], [-47, -50, -4, 53, 50, 4], 100, True, 10000)
    for g, e in zip(got, exp):
        assert g.cpu().numpy().tobytes() == e.tobytes()
    with pytest.raises(_lib.GndError) as ei:
        voxelize_batch(torch.zeros((1, n + 1, 4), device="cuda"), [1, 1, 8], [-47, -50, -4, 53, 50, 4], 100, 10000)
    assert ei.value.code == -4
    # 316 x 316 = 99 856 cells (serial ranking kernel), with max_voxels overflow and max_points truncation
    n = 300_000
    pts = (rng.random((n, 3)) * np.array([170, 170, 9]) - np.array([85, 85, 4.5])).astype(np.float32)
    pts[: n // 4] = pts[n // 4: n // 2] * np.float32(0.05)            # a dense core: cells far beyond max_points
    vs, cr = [0.5, 0.5, 8], [-79, -79, -4, 79, 79, 4]
    for mv in (99856, 5000):
        got = points_to_voxel(pts, vs, cr, 35, True, mv)
        exp = cpu_ops.points_to_voxel(pts, vs, cr, 35, True, mv)
        for g, e in zip(got, exp):
            assert g.shape == e.shape and g.tobytes() == e.tobytes()
    with pytest.raises(_lib.GndError) as ei:
        points_to_voxel(pts, [0.5, 0.5, 8], [-79.5, -79, -4, 79, 79, 4], 35, True, 20000)     # 317 x 316 cells
    assert ei.value.code == -4


LEAN_CASES = ["hdl64_s0", "uniform_s0", "real_000000", "boundary_kitti", "single_cell", "overflow_64", "overflow_1000",
              "overflow_5000", "maxpts_5", "small_kitti", "empty", "all_outside", "one_point", "boundary_camera",
              "small_camera", "camera_1m"]


@pytest.mark.parametrize("name", LEAN_CASES)
def test_fused_set_exact_voxelizer_equals_order_exact(name, cuda):
    """gnd_infer voxelizes with the set-exact kernels (no per-point ranking).  Against the order-exact kernels on the
    same scans: identical coors / num_points / voxel_count / offsets, identical SET of kept points in every cell, and --
    the pillar encoder being order-independent inside a pillar -- bit-identical height maps and labels.  And against the
    oracle: coors / num_points equal the reference's."""
    from tests import helpers
    cfgname, gen, ov = cases.CASES[name]
    cfg, raw, gold = cases.load_case(name)
    cfg.batch_size = 1
    model, sd = helpers.load_model(cfg, cfgname, "mixed", cuda)
    pts = torch.from_numpy(raw).cuda().unsqueeze(0).contiguous()
    if raw.shape[0] == 0:
        pts = torch.zeros((1, 4, cfg.input_features), device="cuda")
    counts = torch.tensor([raw.shape[0]], dtype=torch.int32, device="cuda")
    out = {}
    with torch.no_grad():
        for path in (1, 0, 0):                      # exact, set-exact, set-exact again (self-resetting state)
            model.set_option("vox_path", 2 if path == 0 else 1)
            h, l = model.infer(pts, threshold=0.08, counts=counts, labels="int8")
            pil = {k: v.clone() for k, v in model.pillars(1).items()}
            out.setdefault(path, []).append((h.clone(), l.clone(), pil))
    (he, le, pe), = out[1]
    for hl, ll, pl in out[0]:
        m = int(pe["voxel_count"][0])
        assert int(pl["voxel_count"][0]) == m == gold["M"]
        assert torch.equal(pl["coors"][0, :m], pe["coors"][0, :m])
        assert torch.equal(pl["num_points"][0, :m], pe["num_points"][0, :m])
        assert int(pl["point_offsets"][0, m]) == int(pe["point_offsets"][0, m])      # total kept points
        assert torch.equal(pl["cell_offsets"], pe["cell_offsets"])
        kept = int(pe["cell_offsets"][0, -1])
        assert kept == gold["kept"]
        # same set of points per cell: sort the rows of every cell (cells are contiguous runs of cell_points)
        offs = pe["cell_offsets"][0].long()
        seg = torch.repeat_interleave(torch.arange(offs.numel() - 1, device="cuda"), offs[1:] - offs[:-1])

        def canon(rows):
            rows = rows[:kept]
            order = torch.arange(kept, device="cuda")
            for col in (3, 2, 1, 0):                  # lexicographic by successive stable sorts, then by cell
                order = order[torch.argsort(rows[order, col], stable=True)]
            order = order[torch.argsort(seg[order], stable=True)]
            return rows[order]
        assert torch.equal(canon(pl["cell_points"][0]), canon(pe["cell_points"][0])), "kept point sets differ"
        assert torch.equal(hl, he) and torch.equal(ll, le), "set-exact voxelization changed the result"
    assert synth.digest(pe["coors"][0, :gold["M"]].cpu().numpy()) == gold["coors_digest"]
    assert synth.digest(pe["num_points"][0, :gold["M"]].cpu().numpy()) == gold["num_points_digest"]


def test_fused_set_exact_voxelizer_on_a_large_grid(cuda):
    """A 200 x 200 grid (40 000 cells) takes the multi-CTA per-cell kernel of the set-exact path (the shipped grids
    take the one-CTA-per-scan kernel): same checks as above, plus coors / num_points against the oracle, with and
    without the max_voxels break."""
    from tests import helpers
    raw = synth.hdl64_scan(21)
    for mv in (40000, 3000):
        cfg = synth.make_cfg("kitti", voxel_size=[0.5, 0.5, 8], max_voxels=mv, batch_size=1)
        model, sd = helpers.load_model(cfg, "kitti", "mixed", cuda)
        pts = torch.from_numpy(raw).cuda().unsqueeze(0).contiguous()
        res = {}
        with torch.no_grad():
            for path in (1, 0, 0):
                model.set_option("vox_path", 2 if path == 0 else 1)
                h, l = model.infer(pts, threshold=0.08, labels="int8")
                pil = {k: v.clone() for k, v in model.pillars(1).items()}
                res.setdefault(path, []).append((h.clone(), l.clone(), pil))
        (he, le, pe), = res[1]
        p = _shifted(cfg, raw)
        ov, oc, on = cpu_ops.points_to_voxel(p, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
        m = len(on)
        for hl, ll, pl in res[0] + res[1]:
            assert int(pl["voxel_count"][0]) == m
            assert np.array_equal(pl["coors"][0, :m].cpu().numpy(), oc) and np.array_equal(pl["num_points"][0, :m].cpu().numpy(), on)
            assert torch.equal(pl["cell_offsets"], pe["cell_offsets"])
            assert torch.equal(hl, he) and torch.equal(ll, le)
        model.release()


@pytest.mark.parametrize("max_voxels", [2500, 700])
def test_fused_set_exact_voxelizer_dense_cloud_chunked_counts(max_voxels, cuda):
    """Dense clouds (> 32 points per cell: the camera frames) keep the per-cell counts per CHUNK of the index range, so
    that only the chunk holding a cell's max_points-th point is searched.  Two different 300 k-point frames in one
    batch (every cell over-full, max_points truncation everywhere), with and without the max_voxels break: coors /
    num_points against the oracle, kept sets / heights / labels against the order-exact kernels, twice (self-resetting
    state)."""
    from tests import helpers
    cfg = synth.make_cfg("camera", max_voxels=max_voxels, batch_size=2)
    model, sd = helpers.load_model(cfg, "camera", "mixed", cuda)
    raws = [synth.camera_scan(31)[:300000].copy(), synth.camera_scan(32)[100000:400000].copy()]
    pts = torch.from_numpy(np.stack(raws)).cuda().contiguous()
    assert pts.shape[1] > 32 * 2500
    res = {}
    with torch.no_grad():
        for path in (1, 2, 2):
            model.set_option("vox_path", path)
            h, l = model.infer(pts, threshold=0.08, labels="int8")
            pil = {k: v.clone() for k, v in model.pillars(2).items()}
            res.setdefault(path, []).append((h.clone(), l.clone(), pil))
    (he, le, pe), = res[1]
    for b, raw in enumerate(raws):
        p = _shifted(cfg, raw)
        ov, oc, on = cpu_ops.points_to_voxel(p, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
        m = len(on)
        assert m == min(max_voxels, 2500) or m < max_voxels
        for hl, ll, pl in res[2] + res[1]:
            assert int(pl["voxel_count"][b]) == m
            assert np.array_equal(pl["coors"][b, :m].cpu().numpy(), oc) and np.array_equal(pl["num_points"][b, :m].cpu().numpy(), on)
    for hl, ll, pl in res[2]:
        assert torch.equal(pl["cell_offsets"], pe["cell_offsets"])
        for b in range(2):
            kept = int(pe["cell_offsets"][b, -1])
            offs = pe["cell_offsets"][b].long()
            seg = torch.repeat_interleave(torch.arange(offs.numel() - 1, device="cuda"), offs[1:] - offs[:-1])

            def canon(rows):
                rows = rows[:kept]
                order = torch.arange(kept, device="cuda")
                for col in (3, 2, 1, 0):
                    order = order[torch.argsort(rows[order, col], stable=True)]
                return rows[order[torch.argsort(seg[order], stable=True)]]
            assert torch.equal(canon(pl["cell_points"][b]), canon(pe["cell_points"][b])), "kept point sets differ"
        assert torch.equal(hl, he) and torch.equal(ll, le)
    model.release()
