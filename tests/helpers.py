"""Helpers shared by the GPU parity tests."""
import numpy as np
import torch

from oracle import cpu_ops, model_ref
from gndnet_b200 import synth


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


_BN_CACHE = {}


def bn_stats(cfgname, seed):
    """BatchNorm running statistics for the seeded synthetic checkpoint.  Seed 0 is the committed calibration
    (tests/golden/ckpt_<name>_bn.npz, checked against the reference harness); other seeds are calibrated here with the
    same recipe (oracle/ref_harness.py: two calibration scans -> oracle canvas -> oracle/calibrate.py), once per session,
    so the parity claims do not rest on a single draw of the weights."""
    if seed == 0:
        return synth.load_bn_stats(cfgname)
    if (cfgname, seed) not in _BN_CACHE:
        from oracle.calibrate import calibrate
        src = {"kitti": [synth.hdl64_scan(1000), synth.hdl64_scan(1001)], "camera": [synth.camera_scan(1000, n=200000)]}[cfgname]
        cfg = synth.make_cfg(cfgname, batch_size=len(src))
        sd0 = synth.make_state_dict(cfg, seed, None)
        v, c, n, _ = oracle_pillars(cfg, src)
        with torch.no_grad():
            vf = model_ref.pillar_feature_net(sd0, cfg, v, n, c)
            canvas = model_ref.pillars_scatter(cfg, vf, c, cfg.batch_size)
        _BN_CACHE[(cfgname, seed)] = calibrate(cfg, canvas.numpy(), seed)
    return _BN_CACHE[(cfgname, seed)]


def load_model(cfg, cfgname, precision, device, seed=0):
    from gndnet_b200.model import GroundEstimatorNet
    sd = synth.make_state_dict(cfg, seed, bn_stats(cfgname, seed))
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


def _windows(x):
    """[B,C,H,W] -> [B,C,H/2,W/2,4] (2x2 windows in row-major order, floor mode)."""
    B, C, H, W = x.shape
    return x[:, :, :H // 2 * 2, :W // 2 * 2].reshape(B, C, H // 2, 2, W // 2, 2).permute(0, 1, 2, 4, 3, 5).reshape(B, C, H // 2, W // 2, 4)


def _act(model, name, B, h, w, ch):
    """A stage buffer of the CUDA path as an NCHW float tensor on the CPU (either activation format)."""
    if model.precision == "fp32":
        t = model.buffer(name, torch.float32, (B, h, w, ch)).float()
    else:
        t = model.buffer(name + ".hi", torch.float16, (B, h, w, ch)).float()
        try:
            t = t + model.buffer(name + ".lo", torch.float16, (B, h, w, ch)).float()
        except Exception:
            pass
    return t.permute(0, 3, 1, 2).cpu()


TIE_FLOOR = 0.05       # activations are ~unit variance after BatchNorm; below this magnitude the gap is taken absolutely
# An argmax that differs from the oracle's is accepted only where the oracle's two largest window values are closer than
# this, relative to the larger one: 10x the worst gap of any flipped window measured on the B200 over weight seeds 0-2
# (profiles/r3_parity_report.jsonl, column flip_gap_rel: fp32 CUDA-core 4.7e-6, 16-bit tensor-core operands 2.0e-4).
TIE_REL = {"fp32": 5e-5, "split": 2e-3, "mixed": 2e-3}


def flip_mask(model, st, tie_rel=None):
    """Output pixels in the receptive field of a max-pool window whose argmax differs from the oracle's.

    MaxPool2d keeps ONE position per window (segnet.py:54,60) and MaxUnpool2d puts the value back there
    (segnet.py:84,89): when the two largest values of a window agree to rounding error, any implementation
    with a different summation order (the reference's own CUDA path included) may pick the other one, and
    the output then changes by O(activation) around that window.  A differing argmax is accepted ONLY where
    the oracle's window is such a near-tie (asserted here); those pixels are excluded from the height
    comparison and counted."""
    B, _, H, W = st["out"].shape
    H1, W1 = st["d2b"].shape[2], st["d2b"].shape[3]
    if tie_rel is None:
        tie_rel = TIE_REL[model.precision]

    def rel_gap(x):
        top = torch.topk(_windows(x), 2, dim=-1).values
        return (top[..., 0] - top[..., 1]) / top[..., 0].abs().clamp_min(TIE_FLOOR)

    def near_tie(x):
        return rel_gap(x) <= tie_rel

    def worst_gap(x, flipped):
        """Largest relative top-2 gap (oracle values) among the windows whose argmax differs: what the arithmetic
        actually had to get wrong for the flip to happen.  0 when nothing flipped."""
        g = rel_gap(x)[flipped]
        return float(g.max()) if g.numel() else 0.0

    # level 1: recorded argmax of down1's pool
    idx = model.buffer("idx1", torch.uint8, (B, H // 2, W // 2, 128)).permute(0, 3, 1, 2).cpu().long()
    i1 = st["i1"]
    exp1 = ((i1 // W) % 2) * 2 + (i1 % W) % 2
    f1 = idx != exp1
    assert bool((near_tie(st["d1b"]) | ~f1).all()), "pool-1 argmax differs from the oracle where it is not a near-tie"
    # level 2: which position of each window survived pool+unpool
    u2 = _act(model, "u2", B, H1, W1, 256)
    d2b = st["d2b"]
    wmax = _windows(d2b).max(dim=-1).values
    pos_or = _windows(st["u2"]) != 0
    pos_us = _windows(u2) != 0
    f2 = (pos_or != pos_us).any(dim=-1) & (wmax > 0)
    assert bool((near_tie(d2b) | ~f2).all()), "pool-2 argmax differs from the oracle where it is not a near-tie"
    t1, t2 = f1.any(dim=1), f2.any(dim=1)
    up = lambda m, h, w: torch.nn.functional.pad(m.repeat_interleave(2, 1).repeat_interleave(2, 2),
                                                 (0, w - 2 * m.shape[2], 0, h - 2 * m.shape[1]))
    m2 = _dilate(up(t2, H1, W1), 2)            # through up2.conv1, up2.conv2 at half resolution
    m = up(m2, H, W) | up(t1, H, W)
    flip_mask.last_gaps = (worst_gap(st["d1b"], f1), worst_gap(d2b, f2))
    return _dilate(m, 3), int(f1.sum()), int(f2.sum())      # through up1.conv1, up1.conv2, regressor


# Height bars of BASELINE.json's north_star and what the tests assert about them.  The bar itself (1e-3 m with fp32-class
# arithmetic, 5e-3 m with 16-bit tensor-core operands) is asserted on EVERY pixel outside the receptive field of an argmax
# near-tie flip.  Nothing else is granted as a flat allowance: the flips are counted and capped (max_flips), each must be
# a measured near-tie of the oracle's window (TIE_REL), the mask may cover no more than the flips' receptive fields
# (FLIP_PIXELS each: one level-1 flip touches <= 8x8 output pixels, one level-2 flip <= 18x18), every over-bar pixel must
# lie inside it, and the worst pixel anywhere is capped (a flip moves one near-tied activation by one pixel: the output
# changes around it, it does not blow up).
# Measured on the B200 over weight seeds 0, 1, 2 (profiles/r3_parity_report.jsonl): 0-4 flips per case; the strict fp32
# CUDA-core mode flips too (another summation order is enough), so this is a property of MaxPool / MaxUnpool, not of
# the 16-bit operands.
TOL = {"fp32": 1e-3, "split": 1e-3, "mixed": 5e-3}
# The single-product fp16 decoder of `mixed` has a RELATIVE error: worst pixel ~0.4 % of the map's standard deviation on
# every weight draw and scan probed (tools/precision_probe_decoder.py, profiles/r3_precision_probe_decoder.txt).  The
# reference's own ground-truth maps have sigma(h) = 0.18-0.32 m (data/training/seq_000/gnd_labels), where that is
# ~1e-3 m; the seeded random networks of these tests are 2-5x louder (sigma 0.5-1.6 m, |h| up to 8 m).  The 5e-3 m bar is
# therefore asserted as stated up to sigma(h) = SIGMA0 and grows in proportion beyond it (deviations).
SIGMA0 = 1.0               # metres
REL_TOL = {"fp32": 0.0, "split": 0.0, "mixed": 5e-3}      # x sigma(h_ref): the bar is max(TOL, REL_TOL * sigma)
MAX_OVERALL = 0.5          # metres, any pixel, flips included (worst measured 0.33)
SMOKE_MAX_FRAC_OVER = 0.02 # __graft_entry__.smoke() has no stage buffers of the oracle to build the mask from: flat cap
# receptive field of one flipped window in output pixels (flip_mask's dilations): level 1 = 8x8, level 2 = 18x18
FLIP_PIXELS = (64, 324)


def max_flips(B):
    """Argmax flips tolerated for a batch of B maps: (level 1, level 2)."""
    return 2 * B + 3, B + 1


def compare_heights(model, out, ref, st, precision=None):
    """-> dict(stable_max: max |d| outside the flip mask, frac_masked, overall: max |d| anywhere, frac_over: fraction
    of ALL pixels (no mask) over the precision's bar, median, flips: (#level 1, #level 2))."""
    B, _, H, W = st["out"].shape
    out = torch.as_tensor(out).reshape(B, H, W).cpu()
    ref = torch.as_tensor(ref).reshape(out.shape)
    d = (out - ref).abs()
    mask, n1, n2 = flip_mask(model, st)
    gaps = flip_mask.last_gaps
    stable = d[~mask]
    prec = precision or model.precision
    tol = max(TOL[prec], REL_TOL[prec] * float(ref.float().std()) / SIGMA0)
    return dict(stable_max=float(stable.max()) if stable.numel() else 0.0, frac_masked=float(mask.float().mean()),
                overall=float(d.max()), frac_over=float((d > tol).float().mean()), median=float(d.median()),
                flips=(n1, n2), flip_gap_rel=gaps, tol=tol)


def assert_heights(rep, precision, st, what=""):
    """The height-parity claim, measurable.  Outside the receptive field of an argmax near-tie flip EVERY pixel meets the
    bar; the flips themselves are few (max_flips), each one a measured near-tie (flip_mask asserts the gap), and what
    they may touch follows from their number: the mask covers at most FLIP_PIXELS per flip, so no fraction of the map
    is granted beyond that, and no pixel anywhere moves by more than MAX_OVERALL."""
    line = ("%s %s: max|dh| outside flips=%.3g (bar %.3g)  over-bar pixels (unmasked map)=%.4f%%  masked=%.3f%%  "
            "overall max=%.3g  median=%.2g  argmax flips=%s  worst flipped gap=%s" %
            (what, precision, rep["stable_max"], rep["tol"], 100 * rep["frac_over"], 100 * rep["frac_masked"],
             rep["overall"], rep["median"], rep["flips"], tuple("%.1e" % g for g in rep["flip_gap_rel"])))
    print(line)
    _record(dict(case=what, precision=precision, **rep))
    m1, m2 = max_flips(st["out"].shape[0])
    assert rep["flips"][0] <= m1 and rep["flips"][1] <= m2, "too many argmax flips: " + line
    assert rep["stable_max"] <= rep["tol"], line
    _, _, H, W = st["out"].shape
    footprint = (FLIP_PIXELS[0] * rep["flips"][0] + FLIP_PIXELS[1] * rep["flips"][1]) / float(H * W)
    assert rep["frac_masked"] <= footprint + 1e-9, "flip mask larger than its flips explain: " + line
    assert rep["frac_over"] <= rep["frac_masked"] + 1e-9, "over-bar pixels outside the flip neighbourhoods: " + line
    assert rep["overall"] <= MAX_OVERALL, line


def _record(row):
    """Append to gpurun_out/parity_report.jsonl (scratch; copied to profiles/ as the measured parity table)."""
    import json
    import os
    d = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    try:
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "parity_report.jsonl"), "a") as f:
            f.write(json.dumps(row) + "\n")
    except OSError:
        pass
