"""Shared input set of the parity tests: the same named cases oracle/ref_harness.py ran through the
unmodified reference when it wrote tests/golden/ (inputs are regenerated here, bit-reproducibly)."""
import json
import os

import numpy as np

from gndnet_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

with open(os.path.join(GOLDEN, "voxel_cases.json")) as f:
    VOXEL_GOLDEN = json.load(f)
with open(os.path.join(GOLDEN, "model_cases.json")) as f:
    MODEL_GOLDEN = json.load(f)


def _real0():
    with np.load(os.path.join(GOLDEN, "real_000000.npz")) as z:
        return z["points"].astype(np.float32)


_K = synth.make_cfg("kitti")
_C = synth.make_cfg("camera")

# name -> (cfg name, generator of the raw sensor-frame cloud, cfg overrides)
CASES = {
    "hdl64_s0": ("kitti", lambda: synth.hdl64_scan(0), {}),
    "hdl64_s1": ("kitti", lambda: synth.hdl64_scan(1), {}),
    "uniform_s0": ("kitti", lambda: synth.uniform_scan(0), {}),
    "real_000000": ("kitti", _real0, {}),
    "boundary_kitti": ("kitti", lambda: synth.boundary_scan(_K, 0), {}),
    "single_cell": ("kitti", lambda: synth.single_cell_scan(_K, 0), {}),
    "overflow_64": ("kitti", lambda: synth.hdl64_scan(2), dict(max_voxels=64)),
    "overflow_1000": ("kitti", lambda: synth.uniform_scan(3), dict(max_voxels=1000)),
    "overflow_5000": ("kitti", lambda: synth.uniform_scan(4), dict(max_voxels=5000)),
    "maxpts_5": ("kitti", lambda: synth.hdl64_scan(5), dict(max_points_voxel=5)),
    "small_kitti": ("kitti", lambda: synth.hdl64_scan(6)[::60].copy(), {}),
    "empty": ("kitti", lambda: np.zeros((0, 4), dtype=np.float32), {}),
    "all_outside": ("kitti", lambda: synth.uniform_scan(7, n=1000, lo=(200, 200, 50), hi=(300, 300, 60)), {}),
    "one_point": ("kitti", lambda: np.array([[1.5, 2.5, -1.0, 0.5]], dtype=np.float32), {}),
    "boundary_camera": ("camera", lambda: synth.boundary_scan(_C, 1), {}),
    "small_camera": ("camera", lambda: synth.camera_scan(8, n=30000), {}),
    "camera_1m": ("camera", lambda: synth.camera_scan(0), {}),
}

SMALL = ["boundary_kitti", "small_kitti", "empty", "all_outside", "one_point", "boundary_camera", "small_camera",
         "single_cell", "maxpts_5", "overflow_64"]


def load_case(name):
    """-> (cfg, raw points [N, F] float32 in the sensor frame, golden record)."""
    cfgname, gen, ov = CASES[name]
    cfg = synth.make_cfg(cfgname, **ov)
    raw = np.ascontiguousarray(gen()[:, :cfg.input_features], dtype=np.float32)
    return cfg, raw, VOXEL_GOLDEN[name]


def seg_test_map(cfg):
    """The deterministic height map ref_harness.py used for the segmentation goldens."""
    from oracle import model_ref
    nz, ny, nx = model_ref.grid_shape(cfg)
    yy, xx = np.meshgrid(np.arange(ny, dtype=np.float32), np.arange(nx, dtype=np.float32), indexing="ij")
    return (0.02 * (xx - nx / 2) - 0.015 * (yy - ny / 2) + 0.3 * ((xx * 7 + yy * 13) % 5 - 2) / 5).astype(np.float32)
