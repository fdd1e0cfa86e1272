"""Batch-64 step: 19 direct launches per step against one CUDA-graph replay per step (same kernels), after a soak."""
import os, sys, time, json
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gndnet_b200 import synth
from gndnet_b200.model import GroundEstimatorNet
from gndnet_b200.pipeline import GraphedInfer
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
cfg = synth.make_cfg("kitti"); sd, _ = synth.checkpoint("kitti")
model = GroundEstimatorNet(cfg, precision="mixed")
model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
model = model.cuda().eval()
base = [synth.hdl64_scan(s) for s in range(16)]
pts = torch.from_numpy(np.stack([base[s % 16] for s in range(B)])).cuda()
h = torch.empty((B, model.ny, model.nx), device="cuda"); l = torch.empty((B, pts.shape[1]), dtype=torch.int8, device="cuda")
def run(fn, steps=200):
    t0 = time.time()
    while time.time() - t0 < 0.5:
        for _ in range(10): fn()
        torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / steps
with torch.no_grad():
    d = run(lambda: model.infer(pts, threshold=0.08, out=(h, l)))
    g = GraphedInfer(model, B, pts.shape[1]); g.points.copy_(pts)
    r = run(lambda: g())
    d2 = run(lambda: model.infer(pts, threshold=0.08, out=(h, l)))
print(json.dumps(dict(batch=B, direct_ms=d, graph_ms=r, direct_again_ms=d2)))
