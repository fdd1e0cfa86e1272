"""Helpers shared by the GPU parity tests."""
import numpy as np
import torch

from oracle import cpu_ops, model_ref, synth


def oracle_pillars(cfg, raws):
    """Reference call sequence on the CPU oracle: shift -> points_to_voxel -> batch concat
    (predict_ground.py:112-118, training.py:138-147).  Returns (voxels, coors4, num_points, shifted clouds)."""
    vs, cs, ns, shifted = [], [], [], []
    for b, raw in enumerate(raws):
        p = np.ascontiguousarray(raw[:, :cfg.input_features], dtype=np.float32).copy()
        cpu_ops.shift_cloud(p, cfg.lidar_height)
        v, c, n = cpu_ops.points_to_voxel(p, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
        vs.append(v)
        cs.append(np.concatenate([np.full((len(c), 1), b, dtype=np.float32), c.astype(np.float32)], axis=1))
        ns.append(n.astype(np.float32))
        shifted.append(p)
    return np.concatenate(vs), np.concatenate(cs), np.concatenate(ns), shifted


def load_model(cfg, cfgname, precision, device):
    from gndnet_b200.model import GroundEstimatorNet
    sd = synth.make_state_dict(cfg, 0, synth.load_bn_stats(cfgname))
    model = GroundEstimatorNet(cfg, precision=precision)
    missing = model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}, strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    return model.to(device).eval(), sd


def _dilate(mask, r):
    """Binary dilation by a (2r+1)^2 square, [B,H,W] bool tensor."""
    if r <= 0:
        return mask
    m = torch.nn.functional.max_pool2d(mask.float().unsqueeze(1), 2 * r + 1, 1, r)
    return m.squeeze(1) > 0


def unstable_mask(st, rel=1e-4, abs_=1e-5):
    """Output pixels whose value depends on a max-pool argmax that is a near-tie in the oracle.

    MaxPool2d keeps ONE position per window (segnet.py:54,60) and MaxUnpool2d puts the value back there
    (segnet.py:84,89): when the two largest values of a window agree to rounding error, any implementation
    with a different summation order (the reference's own CUDA path included) may pick the other one, and
    the output changes by O(activation) in the receptive field of that window.  Such pixels are excluded
    from the height comparison (and counted)."""
    def near_tie(x):                           # x [B,C,H,W] -> [B,H/2,W/2] any channel near-tied
        B, C, H, W = x.shape
        w = x[:, :, :H // 2 * 2, :W // 2 * 2].reshape(B, C, H // 2, 2, W // 2, 2).permute(0, 1, 2, 4, 3, 5).reshape(B, C, H // 2, W // 2, 4)
        top = torch.topk(w, 2, dim=-1).values
        gap = top[..., 0] - top[..., 1]
        tie = (gap <= abs_ + rel * top[..., 0].abs()) & (top[..., 0] > 0)
        # an exact tie (gap == 0) between different positions is deterministic (first max) -> only rounding-level ties count
        tie &= gap > 0
        return tie.any(dim=1)
    t1 = near_tie(st["d1b"])                   # [B,H1,W1]
    t2 = near_tie(st["d2b"])                   # [B,H2,W2]
    B, H, W = st["out"].shape[0], st["out"].shape[2], st["out"].shape[3]
    up = lambda m, h, w: torch.nn.functional.pad(m.repeat_interleave(2, 1).repeat_interleave(2, 2),
                                                 (0, w - 2 * m.shape[2], 0, h - 2 * m.shape[1]))
    H1, W1 = st["d2b"].shape[2], st["d2b"].shape[3]
    m2 = _dilate(up(t2, H1, W1), 2)            # through up2.conv1, up2.conv2 at half resolution
    m = up(m2, H, W) | up(t1, H, W)
    return _dilate(m, 3)                       # through up1.conv1, up1.conv2, regressor


def compare_heights(out, ref, st, tol):
    """-> (max |d| on stable pixels, fraction masked, max |d| overall)."""
    out = torch.as_tensor(out).reshape(st["out"].shape[0], st["out"].shape[2], st["out"].shape[3]).cpu()
    ref = torch.as_tensor(ref).reshape(out.shape)
    d = (out - ref).abs()
    mask = unstable_mask(st)
    stable = d[~mask]
    return float(stable.max()) if stable.numel() else 0.0, float(mask.float().mean()), float(d.max())
