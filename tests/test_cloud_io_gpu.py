"""GPU parity for the steps either side of the hot path, through the C ABI:
message bytes -> xyz -> float64 rigid transform -> NaN filter upstream; wire-row packing downstream.
Bit-exact against the reference's goldens and against the C oracle on seeded inputs."""
import numpy as np
import pytest
import torch

from oracle import cpu_ops
from gndnet_b200 import synth
from tests import cases, helpers
from gndnet_b200 import ros_io

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["step16", "step18_unaligned", "step12_identity", "empty", "all_nan"])
def test_unpack_matches_reference_golden(name, cuda):
    with np.load(cases.GOLDEN + "/cloud_io.npz") as z:
        raw, trafo, want = z[name + "__msg"], z[name + "__trafo"], z[name + "__cloud"]
    t = trafo.reshape(4, 4) if trafo.size else None
    pts, cnt = ros_io.unpack_cloud(raw, raw.shape[1], t, True, 3)
    m = int(cnt.item())
    assert m == len(want)
    assert pts[:m].cpu().numpy().tobytes() == want.tobytes()


@pytest.mark.parametrize("n,step,feat,drop", [(120000, 32, 4, True), (1000003, 16, 3, True), (50000, 22, 3, False),
                                              (1, 12, 3, True), (1025, 20, 4, True)])
def test_unpack_matches_oracle(n, step, feat, drop, cuda):
    rng = np.random.default_rng(n + step)
    raw = rng.integers(0, 256, size=(n, step), dtype=np.uint8)
    xyz = (rng.random((n, 3)) * 200 - 100).astype(np.float32)
    xyz[rng.random(n) < 0.05, 1] = np.nan
    xyz[rng.random(n) < 0.01, 2] = -np.inf
    raw[:, :12] = xyz.view(np.uint8).reshape(n, 12)
    ts = synth.SimpleNamespace(transform=synth.SimpleNamespace(
        translation=synth.SimpleNamespace(x=0.3, y=-2.5, z=1.733),
        rotation=synth.SimpleNamespace(x=0.0616284, y=0.0616284, z=0.7044160, w=0.7044160)))
    trafo = ros_io.transformStampedToTransformationMatrix(ts)
    want = cpu_ops.unpack_cloud(raw, step, trafo, drop, feat)
    pts, cnt = ros_io.unpack_cloud(torch.from_numpy(raw).cuda(), step, trafo, drop, feat)
    m = int(cnt.item())
    assert m == len(want)
    got = pts[:m].cpu().numpy()
    # NaN payload bits are not part of the contract when rows are kept (drop=False): compare values
    assert np.array_equal(got, want, equal_nan=True)
    if drop:
        assert got.tobytes() == want.tobytes()


def test_pack_segmented_and_obstacle_rows(cuda):
    with np.load(cases.GOLDEN + "/cloud_io.npz") as z:
        pts, lab, wire = z["pack__points"], z["pack__labels"], z["pack__wire16"]
    p = torch.from_numpy(pts).cuda()
    for l in (torch.from_numpy(lab).cuda(), torch.from_numpy(lab.astype(np.int8)).cuda()):
        rows = ros_io.pack_segmented(p, l)
        assert rows.cpu().numpy().tobytes() == wire.tobytes()
    obs, cnt = ros_io.obstacle_rows(p, torch.from_numpy(lab.astype(np.int8)).cuda())
    m = int(cnt.item())
    assert obs[:m].cpu().numpy().tobytes() == np.ascontiguousarray(pts[lab == 1]).tobytes()   # array_to_pointcloud2 rows


def test_message_to_labels_chain(cuda):
    """ros_node.py:242-288 end to end on the device: message bytes -> cloud (ragged count stays on the GPU) ->
    infer -> labels -> wire rows, against the same chain on the oracle."""
    cfg = synth.make_cfg("kitti", batch_size=1)
    model, sd = helpers.load_model(cfg, "kitti", "fp32", cuda)
    scan = synth.hdl64_scan(321)[::3].copy()
    n = len(scan)
    rng = np.random.default_rng(5)
    raw = rng.integers(0, 256, size=(n, 16), dtype=np.uint8)
    xyz = scan[:, :3].copy()
    xyz[rng.random(n) < 0.02, 0] = np.nan
    raw[:, :12] = xyz.view(np.uint8).reshape(n, 12)
    c, s = np.cos(0.1), np.sin(0.1)
    trafo = np.array([[c, -s, 0, 0.5], [s, c, 0, -0.25], [0, 0, 1, 0.0], [0, 0, 0, 1]], dtype=np.float64)
    want_cloud = cpu_ops.unpack_cloud(raw, 16, trafo, True, 4)
    pts, cnt = ros_io.unpack_cloud(raw, 16, trafo, True, 4)
    with torch.no_grad():
        height, lab = model.infer(pts.unsqueeze(0), threshold=0.16, counts=cnt, labels="int8")
    m = int(cnt.item())
    assert m == len(want_cloud)
    # oracle chain on the kept rows
    ref_pts = want_cloud.copy()
    cpu_ops.shift_cloud(ref_pts, cfg.lidar_height)
    exp = cpu_ops.segment_cloud(ref_pts.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0],
                                height[0].cpu().numpy().T, 0.16)
    got = lab[0, :m].cpu().numpy().astype(np.float64)
    assert np.array_equal(got, exp)
    rows = ros_io.pack_segmented(pts[:m].contiguous(), lab[0, :m].contiguous())
    assert rows.cpu().numpy().tobytes() == cpu_ops.pack_segmented(want_cloud, exp).tobytes()


def test_eval_metrics(cuda):
    """evaluation reductions -- integer counts bit-exact against the reference golden and the
    oracle (1 M points), masked MSE within 1e-12 relative (float64 sums, different association)."""
    from gndnet_b200 import evaluate
    with np.load(cases.GOLDEN + "/cloud_io.npz") as z:
        sem, pred, want, metrics = z["eval__sem"], z["eval__pred"], z["eval__counts"], z["eval__metrics"]
    for p in (torch.from_numpy(pred).cuda(), torch.from_numpy(pred.astype(np.float64)).cuda()):
        got = evaluate.ground_counts(p, torch.from_numpy(sem.view(np.int32)).cuda())
        assert np.array_equal(got.cpu().numpy(), want)
    assert np.array_equal(np.array(evaluate.ground_metrics(torch.from_numpy(pred).cuda(), torch.from_numpy(sem.view(np.int32)).cuda())), metrics)
    rng = np.random.default_rng(9)
    n = 1_000_003
    sem2 = rng.choice(np.array([0, 1, 9, 40, 44, 48, 49, 60, 71, 72, 99], dtype=np.uint32), size=n)
    pred2 = rng.integers(-1, 2, size=n).astype(np.int8)
    got = evaluate.ground_counts(torch.from_numpy(pred2).cuda(), torch.from_numpy(sem2.view(np.int32)).cuda())
    assert np.array_equal(got.cpu().numpy(), cpu_ops.ground_counts(pred2, sem2))
    assert np.array_equal(evaluate.ground_counts(torch.zeros(0, dtype=torch.int8, device="cuda"),
                                                 torch.zeros(0, dtype=torch.int32, device="cuda")).cpu().numpy(), np.zeros(4, np.int64))
    # masked MSE on the kitti and (non-square transpose check) a 50 x 40 map
    for ny, nx in ((100, 100), (40, 50)):
        pg = rng.standard_normal((ny, nx)).astype(np.float32)
        tg = rng.standard_normal((nx, ny))
        mk = (rng.random((nx, ny)) < 0.4).astype(np.float64)
        want_mse = cpu_ops.masked_mse(pg, tg, mk)
        got_mse = evaluate.masked_mse(torch.from_numpy(pg).cuda(), torch.from_numpy(tg).cuda(), torch.from_numpy(mk).cuda())
        assert abs(got_mse - want_mse) <= 1e-12 * abs(want_mse)


@pytest.mark.parametrize("name,mp", [("real", 100), ("hdl64", 100), ("hdl64_mp5", 5)])
def test_rasterisation(name, mp, cuda):
    """GT rasterisation on the device against the reference goldens and the oracle; numpy in ->
    numpy out and CUDA in -> CUDA out; inputs untouched."""
    from gndnet_b200 import evaluate
    from tests.test_oracle import _raster_points
    cfg = synth.make_cfg("kitti")
    gr = np.asarray(cfg.grid_range)
    pts = _raster_points(name)
    keep = pts.copy()
    with np.load(cases.GOLDEN + "/cloud_io.npz") as z:
        img, hm = z["raster__%s__img" % name].astype(np.float64), z["raster__%s__height" % name]
    assert np.array_equal(evaluate.lidar_to_img(pts, gr, cfg.voxel_size[0], 1), img)
    got = evaluate.lidar_to_heightmap(pts, gr, cfg.voxel_size[0], mp)
    assert got.shape == hm.shape and np.abs(got - hm).max() <= 1e-12
    assert np.array_equal(pts, keep)
    g2 = evaluate.lidar_to_heightmap(torch.from_numpy(pts).cuda(), gr, cfg.voxel_size[0], mp)
    assert g2.is_cuda and np.array_equal(g2.cpu().numpy(), got)


def test_rasterisation_random_and_edges(cuda):
    from gndnet_b200 import evaluate
    rng = np.random.default_rng(12)
    for gr, vs, mp, n in (([0, -5, 10, 5], 0.2, 100, 300000), ([-50, -50, 50, 50], 1, 3, 50000), ([-10, -20, 30, 20], 0.5, 1, 1),
                          ([-50, -50, 50, 50], 1, 100, 0)):
        gr = np.asarray(gr)
        pts = (rng.random((n, 4)) * np.array([gr[2] - gr[0] + 4, gr[3] - gr[1] + 4, 16, 1]) + np.array([gr[0] - 2, gr[1] - 2, -4, 0])).astype(np.float32)
        if n > 1000:      # keep the very last cell below max_points + 1 points (undefined in the reference beyond that)
            nx, ny = int((gr[2] - gr[0]) / vs), int((gr[3] - gr[1]) / vs)
            last = (np.floor((pts[:, 0] - np.float32(gr[0])).astype(np.float64) / vs) == nx - 1) & \
                   (np.floor((pts[:, 1] - np.float32(gr[1])).astype(np.float64) / vs) == ny - 1)
            pts = pts[~last]
        assert np.array_equal(evaluate.lidar_to_img(pts, gr, vs, 1), cpu_ops.lidar_to_img(pts, gr, vs, 1))
        a, b = evaluate.lidar_to_heightmap(pts, gr, vs, mp), cpu_ops.lidar_to_heightmap(pts, gr, vs, mp)
        assert a.shape == b.shape and np.abs(a - b).max() <= 1e-12
