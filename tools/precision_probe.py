"""Emulate 16-bit-operand / fp32-accumulate convs on CPU to choose the tensor-core operand format.
Prints max|dh| and label flip rate vs the fp32 oracle for fp16 / bf16 operands."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.nn.functional as F
from oracle import cpu_ops, model_ref as mr
from gndnet_b200 import synth

def fold(sd, name):
    p = "encoder_decoder." + name + ".cbr_unit."
    w, b = torch.from_numpy(sd[p+"0.weight"]), torch.from_numpy(sd[p+"0.bias"])
    g, beta = torch.from_numpy(sd[p+"1.weight"]), torch.from_numpy(sd[p+"1.bias"])
    mu, var = torch.from_numpy(sd[p+"1.running_mean"]), torch.from_numpy(sd[p+"1.running_var"])
    s = g / torch.sqrt(var + 1e-5)
    return w * s.view(-1,1,1,1), (b - mu) * s + beta

def run(sd, canvas, q, last_fp32=False):
    def cbr(name, x, keep=False):
        w, b = fold(sd, name)
        y = F.relu(F.conv2d(q(x), q(w), b, padding=1))
        return y if keep else q(y)
    x = q(canvas)
    x = cbr("down1.conv1", x); x = cbr("down1.conv2", x); s1 = x.shape
    x, i1 = F.max_pool2d(x, 2, 2, return_indices=True)
    x = cbr("down2.conv1", x); x = cbr("down2.conv2", x); s2 = x.shape
    x, i2 = F.max_pool2d(x, 2, 2, return_indices=True)
    x = F.max_unpool2d(x, i2, 2, 2, output_size=s2)
    x = cbr("up2.conv1", x); x = cbr("up2.conv2", x)
    x = F.max_unpool2d(x, i1, 2, 2, output_size=s1)
    x = cbr("up1.conv1", x); x = cbr("up1.conv2", x, keep=last_fp32)
    return F.conv2d(x, torch.from_numpy(sd["encoder_decoder.regressor.weight"]), torch.from_numpy(sd["encoder_decoder.regressor.bias"]), padding=1)

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
            for nm, q, lf in (("fp32", lambda t: t, False), ("fp16", lambda t: t.half().float(), False), ("fp16+fp32last", lambda t: t.half().float(), True), ("bf16", lambda t: t.bfloat16().float(), False)):
                out = run(sd, canvas, q, lf).squeeze()
                d = (out - ref).abs()
                lr = cpu_ops.segment_cloud(raw.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], ref.numpy().T, 0.08)
                lq = cpu_ops.segment_cloud(raw.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], out.numpy().T, 0.08)
                print("%-7s seed %d %-14s max|dh|=%.2e mean=%.2e  label mismatch=%.4f%%" % (cfgname, seed, nm, d.max(), d.mean(), 100*(lr!=lq).mean()))
