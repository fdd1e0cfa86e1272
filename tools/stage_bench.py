"""Per-stage device times of the fused path as infer() runs it (the event ring of gnd_conv_timing_begin, inside a long
back-to-back run) at one batch size; --pfn-path / --vox-path / --conv-geometry select kernel variants for A/B runs."""
import argparse, json, os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gndnet_b200 import synth  # noqa: E402
from gndnet_b200.model import GroundEstimatorNet  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--steps", type=int, default=100)
ap.add_argument("--soak", type=float, default=0.4)
ap.add_argument("--precision", default="mixed")
ap.add_argument("--cfg", default="kitti")
ap.add_argument("--pfn-path", type=int, default=None)
ap.add_argument("--vox-path", type=int, default=None)
ap.add_argument("--conv-geometry", type=int, default=None)
ap.add_argument("--l2-order", type=int, default=None)
ap.add_argument("--tag", default="")
args = ap.parse_args()
cfg = synth.make_cfg(args.cfg)
sd, _ = synth.checkpoint(args.cfg)
model = GroundEstimatorNet(cfg, precision=args.precision)
model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
model = model.cuda().eval()
for name in ("pfn_path", "vox_path", "conv_geometry", "l2_order"):
    v = getattr(args, name)
    if v is not None:
        model.set_option(name, v)
B = args.batch
gen = synth.hdl64_scan if args.cfg == "kitti" else synth.camera_scan
base = [gen(s) for s in range(min(B, 16) * 2)]
pts = [torch.from_numpy(np.stack([base[(j * min(B, 16) + s) % len(base)] for s in range(B)])).cuda() for j in range(2)]
h = torch.empty((B, model.ny, model.nx), dtype=torch.float32, device="cuda")
l = torch.empty((B, pts[0].shape[1]), dtype=torch.int8, device="cuda")
with torch.no_grad():
    t0 = time.time()
    i = 0
    while time.time() - t0 < args.soak:
        for _ in range(10):
            model.infer(pts[i & 1], threshold=0.08, out=(h, l)); i += 1
        torch.cuda.synchronize()
    model.conv_timing_begin(args.steps)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        model.infer(pts[i & 1], threshold=0.08, out=(h, l))
    e1.record()
    torch.cuda.synchronize()
    conv, seg, stg = model.conv_timing_end(stages=True)
sm = stg.mean(axis=0)
out = dict(tag=args.tag, batch=B, step_ms=e0.elapsed_time(e1) / args.steps, voxelize=float(sm[0]), pfn_scatter=float(sm[1]),
           segnet=float(sm[2]), segment=float(sm[3]), convs=float(conv.sum(axis=1).mean()),
           conv_layers=[round(float(x), 4) for x in conv.mean(axis=0)], launches=model.launches)
out["scans_per_s"] = B / out["step_ms"] * 1e3
print(json.dumps(out))
