"""Round-3 probe of the DECODER arithmetic of the `mixed` mode on several weight draws (CPU emulation, fp32 oracle).

Why: the GPU parity run on weight seeds 1 and 2 (tests/test_model_gpu.py::test_fused_infer_other_weight_draws) found a
draw -- camera config, seed 1 -- where the shipped mixed mode (split-fp16 encoder, single-product fp16 decoder) misses
the 5e-3 m bar OUTSIDE every argmax-flip neighbourhood (7.1e-3 m), while seed 0 sat at 3.2-3.7e-3.  This asks how many
of the four decoder convolutions have to run as split-fp16 (hi + lo operands, three products) for the bar to hold with
margin on every draw.  The encoder is exact here and both argmaxes are the oracle's, so no flip enters the numbers.

    python tools/precision_probe_decoder.py > profiles/r3_precision_probe_decoder.txt
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.nn.functional as F

from gndnet_b200 import synth
from oracle import model_ref as mr
from tests import cases, helpers
from tools.precision_probe import fold

torch.set_num_threads(8)
f16 = lambda t: t.half().float()
DEC = ["up2.conv1", "up2.conv2", "up1.conv1", "up1.conv2"]


def conv(x, w, b, split):
    c = lambda a, ww: F.conv2d(a, ww, None, padding=1)
    xh, wh = f16(x), f16(w)
    y = c(xh, wh)
    if split:
        xl, wl = f16(x - xh), f16(w - wh)
        y = y + (c(xh, wl) + c(xl, wh))
    return y + b.view(1, -1, 1, 1)


def store(x, two_planes):
    """What the producing layer's epilogue leaves in HBM: fp16 hi (+ fp16 lo when the consumer is a split conv)."""
    h = f16(x)
    return h + f16(x - h) if two_planes else h


def decoder(sd, st, n_split):
    """The last `n_split` decoder convolutions in split-fp16, the others single-product fp16."""
    split = [i >= 4 - n_split for i in range(4)]
    x = store(st["u2"], split[0])
    for i, name in enumerate(DEC):
        w, b = fold(sd, name)
        y = F.relu(conv(x, w, b, split[i]))
        if i == 1:
            y = F.max_unpool2d(store(y, split[2]), st["i1"], 2, 2, output_size=st["d1b"].shape)
            x = y
        elif i < 3:
            x = store(y, split[i + 1])
        else:
            x = y                                   # never stored: the regressor is fused into the last conv (fp32)
    return F.conv2d(x, torch.from_numpy(sd["encoder_decoder.regressor.weight"]),
                    torch.from_numpy(sd["encoder_decoder.regressor.bias"]), padding=1)


def main():
    print("decoder convs in split-fp16 (last k of 4) | cfg scan weight-seed | max|dh| m   mean|dh|   frac > 5e-3")
    worst = {}
    for cfgname, scans in (("kitti", ["hdl64_s0", "real_000000", "uniform_s0"]), ("camera", ["small_camera", "camera_1m"])):
        for seed in (0, 1, 2):
            cfg = synth.make_cfg(cfgname, batch_size=1)
            sd = synth.make_state_dict(cfg, seed, helpers.bn_stats(cfgname, seed))
            for scan in scans:
                raw = cases.load_case(scan)[1]
                v, c, n, _ = helpers.oracle_pillars(cfg, [raw])
                with torch.no_grad():
                    ref, st = mr.forward(sd, cfg, v, c, n, return_stages=True)
                    for k in range(5):
                        out = decoder(sd, st, k).reshape(ref.shape)
                        d = (out - ref).abs()
                        worst[k] = max(worst.get(k, 0.0), float(d.max()))
                        print("k=%d | %-6s %-12s seed %d | %.2e   %.2e   %.4f%%   (sigma(h) %.2f m, max|h| %.2f m)" %
                              (k, cfgname, scan, seed, float(d.max()), float(d.mean()), 100 * float((d > 5e-3).float().mean()),
                               float(ref.std()), float(ref.abs().max())))
    print()
    for k, w in worst.items():
        print("last %d decoder convs split: worst max|dh| over all draws and scans %.2e m -> %s (bar 5e-3 m)" %
              (k, w, "meets the bar" if w <= 5e-3 else "MISSES the bar"))


if __name__ == "__main__":
    main()
