"""Short driver for ncu captures: a few fused steps of the hot path at a given batch (no timing claims)."""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gndnet_b200 import synth  # noqa: E402
from gndnet_b200.model import GroundEstimatorNet  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--precision", default="mixed")
ap.add_argument("--cfg", default="kitti")
ap.add_argument("--pfn-path", type=int, default=None, help="1: CUDA-core pillar encoder, 2: mma.sync one")
ap.add_argument("--l2-order", type=int, default=None, help="1: segnet kernels walk the scans in alternating directions")
ap.add_argument("--vox-path", type=int, default=0, help="1: order-exact voxelizer kernels inside infer()")
args = ap.parse_args()
cfg = synth.make_cfg(args.cfg)
sd, _ = synth.checkpoint(args.cfg)
model = GroundEstimatorNet(cfg, precision=args.precision)
model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
model = model.cuda().eval()
if args.pfn_path is not None:
    model.set_option("pfn_path", args.pfn_path)
if args.l2_order is not None:
    model.set_option("l2_order", args.l2_order)
if args.vox_path:
    model.set_option("vox_path", args.vox_path)
base = [synth.hdl64_scan(s) if args.cfg == "kitti" else synth.camera_scan(s) for s in range(4)]
pts = torch.from_numpy(np.stack([base[i % 4] for i in range(args.batch)])).cuda()
with torch.no_grad():
    for _ in range(args.steps):
        h, l = model.infer(pts)
torch.cuda.synchronize()
print("ok", float(h.mean()), int((l == 1).sum()))
