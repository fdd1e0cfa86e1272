"""Wall time of the three drop-in callables exactly as predict_ground.py:110-121, :165-168 uses them (one 120 k-point
scan, batch 1), numpy in/out like the reference and CUDA tensors in/out; the C oracle port of the same calls beside it."""
import json, os, sys, time
import numpy as np
import torch
import torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gndnet_b200 import synth  # noqa: E402
from gndnet_b200.model import GroundEstimatorNet  # noqa: E402
from gndnet_b200.point_cloud_ops import points_to_voxel  # noqa: E402
from gndnet_b200.utils import segment_cloud  # noqa: E402


def wall(fn, reps=30, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ts.append(1e3 * (time.perf_counter() - t0))
    return round(float(np.median(ts)), 4)


cfg = synth.make_cfg("kitti")
sd, _ = synth.checkpoint("kitti")
model = GroundEstimatorNet(cfg, precision="mixed")
model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
model = model.cuda().eval()
pts = synth.hdl64_scan(77)
pts[:, 2] += np.float32(cfg.lidar_height)
pts_d = torch.from_numpy(pts).cuda()
out = {}
out["points_to_voxel numpy->numpy ms"] = wall(lambda: points_to_voxel(pts, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels))
out["points_to_voxel cuda->cuda ms"] = wall(lambda: points_to_voxel(pts_d, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels))
v, c, n = points_to_voxel(pts_d, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
vf, cf, nf = v.float(), F.pad(c, (1, 0), "constant", 0).float(), n.float()      # predict_ground.py:115-118
with torch.no_grad():
    out["model(voxels, coors, num_points) cuda ms"] = wall(lambda: model(vf, cf, nf))
    pred = model(vf, cf, nf)
pred_np = pred.cpu().numpy()
out["segment_cloud numpy->numpy ms"] = wall(lambda: segment_cloud(pts.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], pred_np.T, 0.08))
out["segment_cloud cuda->cuda ms"] = wall(lambda: segment_cloud(pts_d.clone(), np.asarray(cfg.grid_range), cfg.voxel_size[0], pred.T, 0.08))
with torch.no_grad():
    raw = torch.from_numpy(synth.hdl64_scan(77)).cuda()
    out["fused model.infer(points) cuda ms"] = wall(lambda: model.infer(raw, threshold=0.08))
# the oracle port of the same three calls on this host (test infrastructure, for scale only)
from oracle import cpu_ops, model_ref  # noqa: E402
t0 = time.perf_counter(); vv, cc, nn = cpu_ops.points_to_voxel(pts, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
out["cpu port points_to_voxel ms"] = round(1e3 * (time.perf_counter() - t0), 2)
c4 = np.concatenate([np.zeros((len(cc), 1), np.float32), cc.astype(np.float32)], 1)
t0 = time.perf_counter(); h = model_ref.forward(sd, cfg, vv, c4, nn.astype(np.float32)).numpy()
out["cpu port model forward ms (%d threads)" % torch.get_num_threads()] = round(1e3 * (time.perf_counter() - t0), 2)
t0 = time.perf_counter(); cpu_ops.segment_cloud(pts.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], h.T, 0.08)
out["cpu port segment_cloud ms"] = round(1e3 * (time.perf_counter() - t0), 2)
print(json.dumps(out, indent=1))
