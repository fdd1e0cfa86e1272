"""Emulate split-fp16 (hi+lo, 3 products) down path + fp16 up path. See DESIGN.md 'precision'."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.nn.functional as F
from oracle import cpu_ops, model_ref as mr
from gndnet_b200 import synth
from tools.precision_probe import fold

def split(t, scale=1.0):
    hi = (t*scale).half().float()
    lo = ((t*scale) - hi).half().float()
    return hi/scale, lo/scale

def conv_mode(x, w, b, mode):
    if mode == "fp32":
        return F.conv2d(x, w, b, padding=1)
    if mode == "fp16":
        return F.conv2d(x.half().float(), w.half().float(), b, padding=1)
    if mode == "bf16":
        return F.conv2d(x.bfloat16().float(), w.bfloat16().float(), b, padding=1)
    if mode == "split":   # fp16 hi/lo, weights pre-scaled by 2^k so lo stays normal
        k = 2.0 ** np.floor(np.log2(8.0 / float(w.abs().max())))
        xh, xl = split(x); wh, wl = split(w, k)
        return F.conv2d(xh, wh, None, padding=1) + (F.conv2d(xh, wl, None, padding=1) + F.conv2d(xl, wh, None, padding=1)) + b.view(1,-1,1,1)
    if mode == "bsplit":  # bf16 hi/lo
        xh = x.bfloat16().float(); xl = (x-xh).bfloat16().float()
        wh = w.bfloat16().float(); wl = (w-wh).bfloat16().float()
        return F.conv2d(xh, wh, None, padding=1) + (F.conv2d(xh, wl, None, padding=1) + F.conv2d(xl, wh, None, padding=1)) + b.view(1,-1,1,1)

def store(y, mode):
    if mode in ("fp16",): return y.half().float()
    if mode == "bf16": return y.bfloat16().float()
    if mode == "split":
        h = y.half().float(); return h + (y-h).half().float()
    if mode == "bsplit":
        h = y.bfloat16().float(); return h + (y-h).bfloat16().float()
    return y

def run(sd, canvas, down, up):
    def cbr(name, x, mode, nxt):
        w, b = fold(sd, name)
        return store(F.relu(conv_mode(x, w, b, mode)), nxt)
    x = store(canvas, down)
    x = cbr("down1.conv1", x, down, down); x = cbr("down1.conv2", x, down, down); s1 = x.shape
    x, i1 = F.max_pool2d(x, 2, 2, return_indices=True)
    x = cbr("down2.conv1", x, down, down); x = cbr("down2.conv2", x, down, down); s2 = x.shape
    x, i2 = F.max_pool2d(x, 2, 2, return_indices=True)
    x = store(F.max_unpool2d(x, i2, 2, 2, output_size=s2), up)
    x = cbr("up2.conv1", x, up, up); x = cbr("up2.conv2", x, up, up)
    x = F.max_unpool2d(x, i1, 2, 2, output_size=s1)
    x = cbr("up1.conv1", x, up, up); x = cbr("up1.conv2", x, up, "fp32")
    return F.conv2d(x, torch.from_numpy(sd["encoder_decoder.regressor.weight"]), torch.from_numpy(sd["encoder_decoder.regressor.bias"]), padding=1), i1, i2

for cfgname, gen in (("kitti", lambda s: synth.hdl64_scan(s)), ("kitti", lambda s: synth.uniform_scan(s)), ("camera", lambda s: synth.camera_scan(s, n=100000))):
    cfg = synth.make_cfg(cfgname)
    sd, _ = synth.checkpoint(cfgname)
    for seed in (0, 1):
        raw = gen(seed)[:, :cfg.input_features].copy()
        cpu_ops.shift_cloud(raw, cfg.lidar_height)
        v, c, n = cpu_ops.points_to_voxel(raw, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
        c4 = np.concatenate([np.zeros((len(c),1),np.float32), c.astype(np.float32)],1)
        with torch.no_grad():
            ref, st = mr.forward(sd, cfg, v, c4, n.astype(np.float32), return_stages=True)
            canvas = st["canvas"]
            for down, up in (("split","fp16"), ("split","split"), ("split","bf16")):
                out, i1, i2 = run(sd, canvas, down, up)
                out = out.squeeze()
                d = (out - ref).abs()
                f1 = int((i1 != st["i1"]).sum()); f2 = int((i2 != st["i2"]).sum())
                lr = cpu_ops.segment_cloud(raw.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], ref.numpy().T, 0.08)
                lq = cpu_ops.segment_cloud(raw.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], out.numpy().T, 0.08)
                print("%-6s s%d down=%-6s up=%-5s max|dh|=%.2e mean=%.2e p99.9=%.2e frac>5e-3=%.4f%% flips=%d/%d label mismatch=%.4f%%" % (
                    cfgname, seed, down, up, d.max(), d.mean(), float(torch.quantile(d.flatten(), 0.999)), 100*float((d>5e-3).float().mean()), f1, f2, 100*(lr!=lq).mean()))
