"""Batch sweep (BASELINE.json configs[4]) and the dense camera frame (configs[3]): scans/s and stage split per batch.
   python tools/sweep.py [--cfg kitti|camera] [--batches 1,2,4,...]   -> one JSON line per batch size."""
import argparse, json, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gndnet_b200 import synth  # noqa: E402
from gndnet_b200.model import GroundEstimatorNet  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--cfg", default="kitti")
ap.add_argument("--batches", default="1,2,4,8,16,32,64,128,256,512")
ap.add_argument("--precision", default="mixed")
ap.add_argument("--max-voxels", type=int, default=0, help="override cfg.max_voxels (overflow handling at scale)")
args = ap.parse_args()
ov = dict(max_voxels=args.max_voxels) if args.max_voxels else {}
cfg = synth.make_cfg(args.cfg, **ov)
sd, _ = synth.checkpoint(args.cfg)
gen = synth.hdl64_scan if args.cfg == "kitti" else synth.camera_scan
pool = [gen(7000 + s)[:, :cfg.input_features].copy() for s in range(8)]
for B in [int(x) for x in args.batches.split(",")]:
    model = GroundEstimatorNet(cfg, precision=args.precision)
    model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
    model = model.cuda().eval()
    pts = [torch.from_numpy(np.stack([pool[(j + s) % 8] for s in range(B)])).cuda() for j in range(2)]
    N = pts[0].shape[1]
    h = torch.empty((B, model.ny, model.nx), dtype=torch.float32, device="cuda")
    l = torch.empty((B, N), dtype=torch.int8, device="cuda")
    steps = max(10, min(400, 4096 // B))
    with torch.no_grad():
        for i in range(5):
            model.infer(pts[i & 1], threshold=0.08, out=(h, l))
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            model.infer(pts[i & 1], threshold=0.08, out=(h, l))
        e1.record()
        torch.cuda.synchronize()
        st = model.stage_times(pts[1], h, l, reps=5)
    ms = e0.elapsed_time(e1) / steps
    M = model.pillars(B)["voxel_count"].float().mean().item()
    print(json.dumps({"cfg": args.cfg, "batch": B, "points": N, "pillars_per_scan": M, "ms_per_step": round(ms, 4),
                      "scans_per_s": round(B / ms * 1e3, 1), "conv_tflops_alg": round((17.694 if args.cfg == "kitti" else 4.424) * B / st["segnet_convs"], 1),
                      "stage_ms": {k: round(v, 4) for k, v in st.items() if isinstance(v, float)}}), flush=True)
    model.release()
    del model, pts, h, l
    torch.cuda.empty_cache()
