"""Per-stage device times of the fused path (CUDA events; gnd_segnet_profile for the convs) at one batch size."""
import argparse, json, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gndnet_b200 import synth  # noqa: E402
from gndnet_b200.model import GroundEstimatorNet  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--reps", type=int, default=20)
ap.add_argument("--precision", default="mixed")
args = ap.parse_args()
cfg = synth.make_cfg("kitti")
sd, _ = synth.checkpoint("kitti")
model = GroundEstimatorNet(cfg, precision=args.precision)
model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
model = model.cuda().eval()
B = args.batch
pts = [torch.from_numpy(np.stack([synth.hdl64_scan(1000 * j + s) for s in range(B)])).cuda() for j in range(2)]
h = torch.empty((B, model.ny, model.nx), dtype=torch.float32, device="cuda")
l = torch.empty((B, pts[0].shape[1]), dtype=torch.int8, device="cuda")
with torch.no_grad():
    for i in range(10):
        model.infer(pts[i & 1], threshold=0.08, out=(h, l))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(100):
        model.infer(pts[i & 1], threshold=0.08, out=(h, l))
    e1.record()
    torch.cuda.synchronize()
    st = model.stage_times(pts[1], h, l, reps=args.reps)
st["step_ms"] = e0.elapsed_time(e1) / 100
st["scans_per_s"] = B / st["step_ms"] * 1e3
print(json.dumps(st))
