"""Deterministic synthetic workload: the scans (SURVEY.md section 8d recipes) and the seeded synthetic
checkpoint that bench.py, tools/, the tests and the oracle harness all feed to BOTH sides.  Pure numpy input
generation -- no compute of the hot path lives here.

Everything here is bit-reproducible across machines: only numpy's PCG64
``Generator.random`` (integer -> double conversion) and correctly-rounded
IEEE ops (+ - * / sqrt) are used -- no libm sin/cos/log, whose last bit can
change with the CPU's SIMD dispatch.  That matters because the golden
fixtures in tests/golden/ store digests of outputs computed on these inputs
in the build container, and the GPU box regenerates the inputs itself.

Recipes follow SURVEY.md section 8(d).
"""
import hashlib
import os
from types import SimpleNamespace

import numpy as np

# ---------------------------------------------------------------------------
# configs (values restated from the reference YAMLs, cited)
# ---------------------------------------------------------------------------
_CFG = {
    # config/config_kittiSem.yaml:3-25
    "kitti": dict(shift_cloud=True, lidar_height=1.733, num_points=100000,
                  grid_range=[-50, -50, 50, 50], pc_range=[-47, -50, -4, 53, 50, 4],
                  voxel_size=[1, 1, 8], max_points_voxel=100, max_voxels=10000,
                  input_features=4, batch_size=1, use_norm=False, vfe_filters=[64],
                  with_distance=False),
    # config/config_camera.yaml:3-25
    "camera": dict(shift_cloud=True, lidar_height=1.733, num_points=10000,
                   grid_range=[0, -5, 10, 5], pc_range=[0, -5, -4, 10, 5, 4],
                   voxel_size=[.2, .2, 8], max_points_voxel=100, max_voxels=2500,
                   input_features=3, batch_size=1, use_norm=False, vfe_filters=[64],
                   with_distance=False),
}


def make_cfg(name="kitti", **overrides):
    """Attribute-bag config like the reference's ConfigClass (predict_ground.py:65-74)."""
    d = dict(_CFG[name])
    d.update(overrides)
    return SimpleNamespace(**d)


# ---------------------------------------------------------------------------
# deterministic math helpers
# ---------------------------------------------------------------------------
def _sincos(theta):
    """sin, cos of float64 array using only + and *: Taylor on theta/32, then 5 doublings."""
    t = np.asarray(theta, dtype=np.float64) * (1.0 / 32.0)
    t2 = t * t
    # sin(t) = t (1 - t2/6 (1 - t2/20 (1 - t2/42 (1 - t2/72 (1 - t2/110)))))
    s = t * (1.0 - t2 / 6.0 * (1.0 - t2 / 20.0 * (1.0 - t2 / 42.0 * (1.0 - t2 / 72.0 * (1.0 - t2 / 110.0)))))
    c = 1.0 - t2 / 2.0 * (1.0 - t2 / 12.0 * (1.0 - t2 / 30.0 * (1.0 - t2 / 56.0 * (1.0 - t2 / 90.0 * (1.0 - t2 / 132.0)))))
    for _ in range(5):
        s, c = 2.0 * s * c, c * c - s * s
    return s, c


def _gauss(rng, n):
    """Unit-variance bell-shaped noise from 4 uniforms (Irwin-Hall), exact ops only."""
    u = rng.random((4, n))
    return (u[0] + u[1] + u[2] + u[3] - 2.0) * np.sqrt(3.0)


_PI = 3.141592653589793


# ---------------------------------------------------------------------------
# scans
# ---------------------------------------------------------------------------
def hdl64_scan(seed, n_az=1875, n_el=64, lidar_height=1.733):
    """Synthetic HDL-64 scan, [n_el*n_az, 4] float32 (x,y,z,intensity), sensor frame
    (ground near z=-lidar_height), ring-major order.  The caller applies the +lidar_height
    z-shift like predict_ground.py:112."""
    rng = np.random.default_rng(seed)
    el = (-24.8 + (2.0 + 24.8) * np.arange(n_el, dtype=np.float64) / (n_el - 1)) * (_PI / 180.0)
    phase = rng.random() * 2.0 * _PI
    az = phase + 2.0 * _PI * np.arange(n_az, dtype=np.float64) / n_az - _PI
    se, ce = _sincos(el)
    sa, ca = _sincos(az)
    dx = ce[:, None] * ca[None, :]
    dy = ce[:, None] * sa[None, :]
    dz = np.broadcast_to(se[:, None], dx.shape)
    sx = (rng.random() * 2.0 - 1.0) * 0.035            # ground tilt, about +-2 deg
    sy = (rng.random() * 2.0 - 1.0) * 0.035
    denom = dz - sx * dx - sy * dy
    hit_gnd = denom < -1e-6
    t = np.where(hit_gnd, -lidar_height / np.where(hit_gnd, denom, -1.0), np.inf)
    has_obs = rng.random(n_az) < 0.35
    r_obs = 4.0 + 66.0 * rng.random(n_az)
    t_obs = np.where(has_obs[None, :], r_obs[None, :] / ce[:, None], np.inf)
    t = np.minimum(t, t_obs)
    clutter = 20.0 + 100.0 * rng.random(t.shape)
    t = np.where(t > 120.0, clutter, t)
    t = t * (1.0 + 0.002 * _gauss(rng, t.size).reshape(t.shape))
    pts = np.empty((n_el, n_az, 4), dtype=np.float32)
    pts[..., 0] = (t * dx).astype(np.float32)
    pts[..., 1] = (t * dy).astype(np.float32)
    pts[..., 2] = (t * dz).astype(np.float32)
    pts[..., 3] = rng.random(t.shape).astype(np.float32)
    return pts.reshape(-1, 4)


def uniform_scan(seed, n=120000, cols=4, lo=(-60.0, -60.0, -3.0), hi=(60.0, 60.0, 1.0)):
    """Uniform-random cloud: worst case for pillar count (M ~ every cell)."""
    rng = np.random.default_rng(seed)
    u = rng.random((n, cols))
    out = np.empty((n, cols), dtype=np.float32)
    for j in range(3):
        out[:, j] = (lo[j] + (hi[j] - lo[j]) * u[:, j]).astype(np.float32)
    for j in range(3, cols):
        out[:, j] = u[:, j].astype(np.float32)
    return out


def camera_scan(seed, n=1_000_000):
    """Dense depth-camera cloud for config_camera (F=3): uniform in the 10x10 m box plus a
    20 % out-of-range margin; every one of the 2500 cells overflows max_points."""
    return uniform_scan(seed, n=n, cols=3, lo=(-1.0, -6.0, -3.0), hi=(11.0, 6.0, 1.0))


def boundary_scan(cfg, seed=0, reps=4):
    """Points exactly on cell / range boundaries and their float32 neighbours, duplicates,
    +-inf, -0.0, plus a sprinkle of interior points (no NaN)."""
    rng = np.random.default_rng(seed)
    lo = np.asarray(cfg.pc_range[:3], dtype=np.float32)
    hi = np.asarray(cfg.pc_range[3:], dtype=np.float32)
    vs = np.asarray(cfg.voxel_size, dtype=np.float32)
    cols = cfg.input_features
    pts = []
    for j in range(3):
        n_cells = int(np.round((hi[j] - lo[j]) / vs[j]))
        edges = (lo[j] + vs[j] * np.arange(n_cells + 1, dtype=np.float32)).astype(np.float32)
        cand = np.concatenate([edges, np.nextafter(edges, np.float32(-np.inf)),
                               np.nextafter(edges, np.float32(np.inf))])
        for v in cand:
            p = (lo + (hi - lo) * rng.random(3).astype(np.float32)).astype(np.float32)
            p[j] = v
            pts.append(p)
    pts = np.asarray(pts, dtype=np.float32)
    special = np.array([[np.inf, 0, 0], [-np.inf, 0, 0], [0, np.inf, 0], [0, 0, -np.inf],
                        [-0.0, -0.0, -0.0], [0.0, 0.0, 0.0], [1e30, 1e30, 0], [-1e30, 0, 0]],
                       dtype=np.float32)
    pts = np.concatenate([pts, special] + [pts[:: max(1, len(pts) // 50)]] * reps, axis=0)
    perm = np.argsort(rng.random(len(pts)), kind="stable")
    pts = pts[perm]
    out = np.zeros((len(pts), cols), dtype=np.float32)
    out[:, :3] = pts
    if cols > 3:
        out[:, 3:] = rng.random((len(pts), cols - 3)).astype(np.float32)
    return out


def single_cell_scan(cfg, seed=0, n=5000):
    """All points inside one cell (max_points truncation with a long tail)."""
    rng = np.random.default_rng(seed)
    lo = np.asarray(cfg.pc_range[:3], dtype=np.float64)
    vs = np.asarray(cfg.voxel_size, dtype=np.float64)
    base = lo + vs * np.array([7, 3, 0])
    out = np.zeros((n, cfg.input_features), dtype=np.float32)
    out[:, :3] = (base + vs * (0.05 + 0.9 * rng.random((n, 3)))).astype(np.float32)
    if cfg.input_features > 3:
        out[:, 3] = rng.random(n).astype(np.float32)
    return out


def digest(*arrays):
    """sha256 over the raw bytes (dtype + shape included) of the given arrays."""
    h = hashlib.sha256()
    for a in arrays:
        a = np.ascontiguousarray(a)
        h.update(str(a.dtype).encode())
        h.update(str(a.shape).encode())
        h.update(a.tobytes())
    return h.hexdigest()


# ---------------------------------------------------------------------------
# synthetic checkpoint (the shipped trained_models/checkpoint.pth.tar is absent
# from the build container: .MISSING_LARGE_BLOBS)
# ---------------------------------------------------------------------------
CONV_BLOCKS = [  # (state_dict prefix under encoder_decoder., Cin, Cout)  modules/segnet.py:118-131
    ("down1.conv1", 64, 128), ("down1.conv2", 128, 128),
    ("down2.conv1", 128, 256), ("down2.conv2", 256, 256),
    ("up2.conv1", 256, 256), ("up2.conv2", 256, 128),
    ("up1.conv1", 128, 128), ("up1.conv2", 128, 64),
]

_GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _u(rng, shape, b):
    return ((rng.random(shape) * 2.0 - 1.0) * b).astype(np.float32)


def make_state_dict(cfg, seed=0, bn_stats=None):
    """Seeded state_dict with exactly the reference's keys/shapes.

    Conv/linear weights: uniform, He-scaled.  BatchNorm affine: weight U(.75,1.25),
    bias 0.1*g.  BatchNorm running stats: ``bn_stats`` (a dict produced by
    calibrate(), normally loaded from tests/golden/ckpt_<name>_bn.npz) so that signal
    propagates like in a trained net; if absent, un-calibrated placeholders.
    Returns {key: np.ndarray}.
    """
    rng = np.random.default_rng(1000 + seed)
    sd = {}
    nin = cfg.input_features + 5 + (1 if cfg.with_distance else 0)
    filters = [nin] + list(cfg.vfe_filters)                # pointpillars.py:96-107: non-last layers have out // 2 units
    for i in range(len(filters) - 1):
        last = i == len(filters) - 2
        cin, units = filters[i], (filters[i + 1] if last else filters[i + 1] // 2)
        pre = "voxel_feature_extractor.pfn_layers.%d." % i
        sd[pre + "linear.weight"] = _u(rng, (units, cin), 0.4 if i == 0 else float(np.sqrt(6.0 / cin)))
        if cfg.use_norm:
            sd[pre + "norm.weight"] = (0.75 + 0.5 * rng.random(units)).astype(np.float32)
            sd[pre + "norm.bias"] = (0.1 * _gauss(rng, units)).astype(np.float32)
            sd[pre + "norm.running_mean"] = (2.0 * _gauss(rng, units)).astype(np.float32)
            sd[pre + "norm.running_var"] = (20.0 + 20.0 * rng.random(units)).astype(np.float32)
            sd[pre + "norm.num_batches_tracked"] = np.array(1000, dtype=np.int64)
        else:
            sd[pre + "linear.bias"] = _u(rng, (units,), 0.3)
    for name, cin, cout in CONV_BLOCKS:
        p = "encoder_decoder." + name + ".cbr_unit."
        b = float(np.sqrt(6.0 / (9 * cin)))
        sd[p + "0.weight"] = _u(rng, (cout, cin, 3, 3), b)
        sd[p + "0.bias"] = _u(rng, (cout,), 0.1)
        sd[p + "1.weight"] = (0.75 + 0.5 * rng.random(cout)).astype(np.float32)
        sd[p + "1.bias"] = (0.1 * _gauss(rng, cout)).astype(np.float32)
        if bn_stats is not None and (name + ".mean") in bn_stats:
            sd[p + "1.running_mean"] = np.asarray(bn_stats[name + ".mean"], dtype=np.float32)
            sd[p + "1.running_var"] = np.asarray(bn_stats[name + ".var"], dtype=np.float32)
        else:
            sd[p + "1.running_mean"] = (0.1 * _gauss(rng, cout)).astype(np.float32)
            sd[p + "1.running_var"] = (0.75 + 0.5 * rng.random(cout)).astype(np.float32)
        sd[p + "1.num_batches_tracked"] = np.array(1000, dtype=np.int64)
    scale = float(bn_stats["regressor.scale"]) if (bn_stats is not None and "regressor.scale" in bn_stats) else 1.0
    sd["encoder_decoder.regressor.weight"] = (_u(rng, (1, 64, 3, 3), float(np.sqrt(6.0 / 576))) * np.float32(scale))
    sd["encoder_decoder.regressor.bias"] = np.array([-0.05], dtype=np.float32)
    return sd


def bn_stats_path(name):
    return os.path.join(_GOLDEN, "ckpt_%s_bn.npz" % name)


def load_bn_stats(name):
    p = bn_stats_path(name)
    if not os.path.exists(p):
        return None
    with np.load(p) as z:
        return {k: z[k] for k in z.files}


def checkpoint(cfg_name="kitti", seed=0, cfg=None):
    """The parity checkpoint: real file if $GNDNET_CKPT points at one, else the seeded synthetic
    state_dict with the committed calibration.  Returns ({key: ndarray}, description)."""
    path = os.environ.get("GNDNET_CKPT", "")
    if path and os.path.exists(path):
        import torch
        ck = torch.load(path, map_location="cpu", weights_only=False)
        sd = ck["state_dict"] if "state_dict" in ck else ck
        return {k: v.numpy() for k, v in sd.items()}, "file:" + path
    cfg = cfg or make_cfg(cfg_name)
    return make_state_dict(cfg, seed, load_bn_stats(cfg_name)), "synthetic(seed=%d,%s)" % (seed, cfg_name)
