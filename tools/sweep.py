"""Batch sweep (BASELINE.json configs[4]) and the dense camera frame (configs[3]): scans/s and stage split per batch.
   python tools/sweep.py [--cfg kitti|camera] [--batches 1,2,4,...] [--max-voxels M]      one GPU
   python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P tools/sweep.py ...
One JSON line per batch size (rank 0).  With several ranks every GPU runs `batch` scans per step (weak scaling, scans
sharded by rank, no collective on the data path); the step time is the MAX over ranks and scans/s the whole-job figure.
Stage times are those of infer() itself (event ring of gnd_conv_timing_begin), rank 0's."""
import argparse, json, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gndnet_b200 import synth  # noqa: E402
from gndnet_b200.model import GroundEstimatorNet  # noqa: E402
from gndnet_b200.pipeline import max_over_ranks  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--cfg", default="kitti")
ap.add_argument("--batches", default="1,2,4,8,16,32,64,128,256,512")
ap.add_argument("--precision", default="mixed")
ap.add_argument("--max-voxels", type=int, default=0, help="override cfg.max_voxels (overflow handling at scale)")
args = ap.parse_args()
world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
ov = dict(max_voxels=args.max_voxels) if args.max_voxels else {}
cfg = synth.make_cfg(args.cfg, **ov)
sd, _ = synth.checkpoint(args.cfg)
gen = synth.hdl64_scan if args.cfg == "kitti" else synth.camera_scan
pool = [gen(7000 + 8 * rank + s)[:, :cfg.input_features].copy() for s in range(8)]
gflop = 17.694 if args.cfg == "kitti" else 4.424
for B in [int(x) for x in args.batches.split(",")]:
    model = GroundEstimatorNet(cfg, precision=args.precision)
    model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
    model = model.to(dev).eval()
    pts = [torch.from_numpy(np.stack([pool[(j + s) % 8] for s in range(B)])).to(dev) for j in range(2)]
    N = pts[0].shape[1]
    h = torch.empty((B, model.ny, model.nx), dtype=torch.float32, device=dev)
    l = torch.empty((B, N), dtype=torch.int8, device=dev)
    steps = max(10, min(400, 4096 // B))
    with torch.no_grad():
        for i in range(5):
            model.infer(pts[i & 1], threshold=0.08, out=(h, l))
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        model.conv_timing_begin(steps)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            model.infer(pts[i & 1], threshold=0.08, out=(h, l))
        e1.record()
        torch.cuda.synchronize()
        conv, seg, stg = model.conv_timing_end(stages=True)
    ms = max_over_ranks(e0.elapsed_time(e1) / steps, dev)
    sm = stg.mean(axis=0)
    M = model.pillars(B)["voxel_count"].float().mean().item()
    if rank == 0:
        print(json.dumps({"cfg": args.cfg, "n_gpus": world, "batch_per_gpu": B, "points": N, "pillars_per_scan": M,
                          "ms_per_step": round(ms, 4), "scans_per_s": round(world * B / ms * 1e3, 1),
                          "conv_tflops_alg_per_gpu": round(gflop * B / float(conv.sum(axis=1).mean()), 1),
                          "stage_ms": dict(voxelize=round(float(sm[0]), 4), pfn_scatter=round(float(sm[1]), 4),
                                           segnet=round(float(sm[2]), 4), segnet_convs=round(float(conv.sum(axis=1).mean()), 4),
                                           segment=round(float(sm[3]), 4))}), flush=True)
    model.release()
    del model, pts, h, l
    torch.cuda.empty_cache()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
