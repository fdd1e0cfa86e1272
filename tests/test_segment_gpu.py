"""GPU parity: per-point labelling against the CPU oracle (bit-exact) and the reference's golden digests."""
import numpy as np
import pytest
import torch

from oracle import cpu_ops
from gndnet_b200 import synth
from tests import cases

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", [n for n in cases.CASES if n not in ("empty",)])
def test_segment_cloud_bit_exact(name, cuda):
    from gndnet_b200.utils import segment_cloud, segment_cloud_noground
    cfg, raw, gold = cases.load_case(name)
    pts = raw.copy()
    cpu_ops.shift_cloud(pts, cfg.lidar_height)
    pred = cases.seg_test_map(cfg)
    grid = np.asarray(cfg.grid_range)
    for thr in (0.08, 0.16, 0.0, 0.2):
        a, b = pts.copy(), pts.copy()
        got = segment_cloud(a, grid, cfg.voxel_size[0], elevation_map=pred.T, threshold=thr)
        exp = cpu_ops.segment_cloud(b, grid, cfg.voxel_size[0], elevation_map=pred.T, threshold=thr)
        assert got.dtype == np.float64 and got.tobytes() == exp.tobytes()
        assert a.tobytes() == b.tobytes(), "the in-place xy shift (utils.py:204) must match too"
        assert synth.digest(got.astype(np.int8)) == gold["segment"][str(thr)]["digest"]
    a, ao = pts[:, :3].copy(), pts[:, :3].copy()
    b, bo = pts[:, :3].copy(), pts[:, :3].copy()
    lg, og = segment_cloud_noground(a, ao, grid, cfg.voxel_size[0], pred.T, 0.16)
    le, oe = cpu_ops.segment_cloud_noground(b, bo, grid, cfg.voxel_size[0], pred.T, 0.16)
    assert lg.tobytes() == le.tobytes() and og.shape == oe.shape and og.tobytes() == oe.tobytes()
    assert synth.digest(og) == gold["segment"]["noground_0.16"]["digest"]


def test_segment_cuda_tensors_and_boundaries(cuda):
    """CUDA tensors in/out; cell index 0 is out of range; f32-subtract / f64-divide cell arithmetic."""
    from gndnet_b200.utils import segment_cloud
    cfg = synth.make_cfg("camera")
    xs = np.array([1.4, 0.2, 0.1999999, 0.4, 9.999999, 10.0, 5.0, -0.0, np.nextafter(np.float32(10), np.float32(0))], np.float32)
    pts = np.stack([xs, xs - 5.0, np.linspace(-1, 1, len(xs)).astype(np.float32)], 1).astype(np.float32)
    pred = cases.seg_test_map(cfg)
    exp = cpu_ops.segment_cloud(pts.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], pred.T, 0.08)
    got = segment_cloud(torch.from_numpy(pts.copy()).cuda(), np.asarray(cfg.grid_range), cfg.voxel_size[0],
                        torch.from_numpy(pred).cuda().T, 0.08)
    assert got.is_cuda and got.dtype == torch.float64
    assert np.array_equal(got.cpu().numpy(), exp)
