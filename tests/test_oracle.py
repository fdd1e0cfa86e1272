"""CPU: the oracle (C restatement + torch-fp32 model restatement) against the golden fixtures that
oracle/ref_harness.py wrote while running the unmodified reference next to it."""
import numpy as np
import pytest

from oracle import cpu_ops, model_ref
from gndnet_b200 import synth
from tests import cases


@pytest.mark.parametrize("name", [n for n in cases.CASES if n != "camera_1m"])
def test_voxelizer_and_segment_match_reference_goldens(name):
    cfg, raw, gold = cases.load_case(name)
    assert synth.digest(cases.CASES[name][1]()) == gold["input_digest"], "synthetic input is not reproducible"
    pts = raw.copy()
    cpu_ops.shift_cloud(pts, cfg.lidar_height)
    v, c, n, slot = cpu_ops.points_to_voxel(pts, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True,
                                            cfg.max_voxels, return_point_slot=True)
    assert len(n) == gold["M"]
    assert synth.digest(v) == gold["voxels_digest"]
    assert synth.digest(c) == gold["coors_digest"]
    assert synth.digest(n) == gold["num_points_digest"]
    assert synth.digest(slot) == gold["slot_digest"]
    pred = cases.seg_test_map(cfg)
    for thr, rec in gold["segment"].items():
        if thr.startswith("noground"):
            a, b = pts[:, :3].copy(), pts[:, :3].copy()
            lab, obs = cpu_ops.segment_cloud_noground(a, b, np.asarray(cfg.grid_range), cfg.voxel_size[0], pred.T, 0.16)
            assert len(obs) == rec["n_obs"] and synth.digest(obs) == rec["digest"]
        else:
            lab = cpu_ops.segment_cloud(pts.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], pred.T, float(thr))
            assert synth.digest(lab.astype(np.int8)) == rec["digest"]
            assert [int((lab == k).sum()) for k in (-1, 0, 1)] == rec["counts"]


def test_small_case_arrays():
    with np.load(cases.GOLDEN + "/small_cases.npz") as z:
        for name in ("boundary_kitti", "small_kitti", "one_point", "boundary_camera", "small_camera"):
            cfg, raw, _ = cases.load_case(name)
            assert raw.tobytes() == z[name + ".points"][:, :cfg.input_features].tobytes()
            pts = raw.copy()
            cpu_ops.shift_cloud(pts, cfg.lidar_height)
            _, c, n, slot = cpu_ops.points_to_voxel(pts, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True,
                                                    cfg.max_voxels, return_point_slot=True)
            assert np.array_equal(c, z[name + ".coors"]) and np.array_equal(n, z[name + ".num_points"])
            assert np.array_equal(slot, z[name + ".slot"])


def test_model_restatement_matches_reference_heights():
    """camera + use_norm cases (small); the big kitti cases run in the GPU suite's oracle leg."""
    with np.load(cases.GOLDEN + "/model_heights.npz") as z:
        for key, scans, cfgname, ov in (("camera__small_camera", ["small_camera"], "camera", {}),
                                        ("kitti__small_kitti_norm", ["small_kitti"], "kitti", dict(use_norm=True))):
            cfg = synth.make_cfg(cfgname, batch_size=len(scans), **ov)
            sd = synth.make_state_dict(cfg, 0, synth.load_bn_stats(cfgname))
            vs, cs, ns = [], [], []
            for b, s in enumerate(scans):
                _, raw, _ = cases.load_case(s)
                p = raw.copy()
                cpu_ops.shift_cloud(p, cfg.lidar_height)
                v, c, n = cpu_ops.points_to_voxel(p, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
                vs.append(v); ns.append(n.astype(np.float32))
                cs.append(np.concatenate([np.full((len(c), 1), b, np.float32), c.astype(np.float32)], 1))
            out = model_ref.forward(sd, cfg, np.concatenate(vs), np.concatenate(cs), np.concatenate(ns)).numpy()
            assert np.abs(out - z[key]).max() <= 2e-5


IO_CASES = ["step16", "step18_unaligned", "step12_identity", "empty", "all_nan"]


@pytest.mark.parametrize("name", IO_CASES)
def test_cloud_unpack_matches_reference_golden(name):
    """msg bytes -> xyz -> transformCloud -> NaN filter, against the clouds the unmodified
    reference produced in the build container (oracle/ref_harness_io.py)."""
    with np.load(cases.GOLDEN + "/cloud_io.npz") as z:
        raw, trafo, want = z[name + "__msg"], z[name + "__trafo"], z[name + "__cloud"]
    got = cpu_ops.unpack_cloud(raw, raw.shape[1], trafo.reshape(4, 4) if trafo.size else None, True, 3)
    assert got.shape == want.shape and got.tobytes() == want.tobytes()


def test_pack_segmented_matches_reference_golden():
    """np2ros_pub_2's 16-byte wire rows."""
    with np.load(cases.GOLDEN + "/cloud_io.npz") as z:
        pts, lab, wire = z["pack__points"], z["pack__labels"], z["pack__wire16"]
    assert cpu_ops.pack_segmented(pts, lab).tobytes() == wire.tobytes()


def test_eval_counts_match_reference_golden():
    """get_GndSeg + remove_outliers + the four sums, against the reference functions' output."""
    with np.load(cases.GOLDEN + "/cloud_io.npz") as z:
        sem, pred, want, metrics = z["eval__sem"], z["eval__pred"], z["eval__counts"], z["eval__metrics"]
    got = cpu_ops.ground_counts(pred.astype(np.float64), sem)
    assert np.array_equal(got, want)
    assert np.array_equal(np.array([got[0] / got[1], got[0] / got[2], got[0] / got[3]]), metrics)


@pytest.mark.parametrize("name,mp", [("real", 100), ("hdl64", 100), ("hdl64_mp5", 5)])
def test_rasterisation_matches_reference_golden(name, mp):
    """lidar_to_img / lidar_to_heightmap (one-slot overrun into the next cell included) against the
    maps the reference's numba functions produced."""
    cfg = synth.make_cfg("kitti")
    pts = _raster_points(name)
    with np.load(cases.GOLDEN + "/cloud_io.npz") as z:
        img, hm = z["raster__%s__img" % name], z["raster__%s__height" % name]
    gr = np.asarray(cfg.grid_range)
    assert np.array_equal(cpu_ops.lidar_to_img(pts, gr, cfg.voxel_size[0], 1), img.astype(np.float64))
    assert np.abs(cpu_ops.lidar_to_heightmap(pts, gr, cfg.voxel_size[0], mp) - hm).max() <= 1e-12


def _raster_points(name):
    if name == "real":
        with np.load(cases.GOLDEN + "/real_000000.npz") as z:
            return z["points"].astype(np.float32)[:, :3].copy()
    pts = synth.hdl64_scan(3)[:, :3].copy()
    pts[:, 2] += np.float32(1.733)
    return pts


def test_multi_layer_pillar_feature_net_matches_reference_heights():
    """cfg.vfe_filters with more than one entry (pointpillars.py:57-65, :96-107): the oracle against the unmodified
    reference's outputs (oracle/ref_harness_vfe.py wrote tests/golden/model_heights_vfe.npz)."""
    from oracle.ref_harness_vfe import VFE_CASES, case_key
    with np.load(cases.GOLDEN + "/model_heights_vfe.npz") as z:
        for cfgname, name, filters, use_norm in VFE_CASES:
            cfg = synth.make_cfg(cfgname, batch_size=1, vfe_filters=filters, use_norm=use_norm)
            sd = synth.make_state_dict(cfg, 0, synth.load_bn_stats(cfgname))
            _, raw, _ = cases.load_case(name)
            p = raw[:, :cfg.input_features].copy()
            cpu_ops.shift_cloud(p, cfg.lidar_height)
            v, c, n = cpu_ops.points_to_voxel(p, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
            c4 = np.concatenate([np.zeros((len(c), 1), np.float32), c.astype(np.float32)], 1)
            out = model_ref.forward(sd, cfg, v, c4, n.astype(np.float32)).numpy()
            assert np.abs(out - z[case_key(cfgname, name, filters, use_norm)]).max() <= 2e-5
