"""Drop-in for the hot-path functions of the reference's ``utils/utils.py``.

  segment_cloud(points, grid_size, voxel_size, elevation_map, threshold=0.2) -> float64 [N] in {1, 0, -1}
      utils/utils.py:198-220
  segment_cloud_noground(points, points_obstacle, grid_size, voxel_size, elevation_map, threshold=0.2)
      -> (labels, points_obstacle[:n_obs])                                   utils/utils.py:222-248
  shift_cloud_func(cloud, height)                                            utils/utils.py:137-139

numpy in -> numpy out, CUDA tensors in -> CUDA tensors out.  Like the reference, ``points[:, :2]`` is
overwritten with the grid-shifted coordinates (:204) -- callers pass ``points.copy()``
(predict_ground.py:168, ros_node.py:274).
"""
import ctypes

import numpy as np
import torch

from . import _lib


def _to_cuda(a, name):
    if isinstance(a, np.ndarray):
        if a.dtype != np.float32:
            raise TypeError("%s must be float32" % name)
        return torch.from_numpy(a).cuda(), True      # from_numpy keeps strides (pred.T is F-ordered)
    if isinstance(a, torch.Tensor) and a.is_cuda and a.dtype == torch.float32:
        return a, False
    raise TypeError("%s must be a float32 numpy array or CUDA tensor" % name)


def _segment(points, grid_size, voxel_size, elevation_map, threshold, want_i8=False):
    if points.ndim != 2 or points.shape[1] < 3:
        raise ValueError("points must be [N, >=3]")
    if isinstance(points, np.ndarray) and not points.flags.c_contiguous:
        raise ValueError("points must be C-contiguous")
    pts, pts_np = _to_cuda(points, "points")
    if not pts.is_contiguous():
        raise ValueError("points must be contiguous")
    emap, _ = _to_cuda(elevation_map, "elevation_map")
    if emap.dim() != 2:
        raise ValueError("elevation_map must be 2-D")
    L = _lib.lib()
    n, f = pts.shape
    sp = _lib.SegmentParams()
    sp.grid_lo[0], sp.grid_lo[1] = float(grid_size[0]), float(grid_size[1])
    sp.voxel_size = float(voxel_size)
    sp.threshold = float(threshold)
    sp.map_nx, sp.map_ny = int(emap.shape[0]), int(emap.shape[1])
    sp.map_sx, sp.map_sy = int(emap.stride(0)), int(emap.stride(1))
    sp.map_batch_stride = 0
    sp.num_features = int(f)
    sp.apply_z_shift, sp.z_shift, sp.mutate_points = 0, 0.0, 1
    with torch.cuda.device(pts.device):
        labels = torch.empty((n,), dtype=torch.float64, device=pts.device)
        lab8 = torch.empty((n,), dtype=torch.int8, device=pts.device) if want_i8 else None
        _lib.check(L.gnd_segment_cloud(ctypes.byref(sp), _lib.ptr(pts), None, 1, n, _lib.ptr(emap),
                                       _lib.ptr(labels), _lib.ptr(lab8), _lib.stream_ptr()), "gnd_segment_cloud")
    if pts_np:
        points[:, :2] = pts[:, :2].cpu().numpy()      # the reference's in-place update (:204)
    return labels, lab8, pts_np


def segment_cloud(points, grid_size, voxel_size, elevation_map, threshold=0.2):
    labels, _, is_np = _segment(points, grid_size, voxel_size, elevation_map, threshold)
    return labels.cpu().numpy() if is_np else labels


def segment_cloud_noground(points, points_obstacle, grid_size, voxel_size, elevation_map, threshold=0.2):
    labels, lab8, is_np = _segment(points, grid_size, voxel_size, elevation_map, threshold, want_i8=True)
    obs, obs_np = _to_cuda(points_obstacle, "points_obstacle")
    if obs.dim() != 2 or obs.shape[0] != labels.shape[0] or not obs.is_contiguous():
        raise ValueError("points_obstacle must be a contiguous [N, k] array with the same N as points")
    L = _lib.lib()
    n, k = obs.shape
    need = ctypes.c_size_t()
    with torch.cuda.device(obs.device):
        _lib.check(L.gnd_compact_workspace_bytes(n, ctypes.byref(need)), "gnd_compact_workspace_bytes")
        ws = torch.empty(max(need.value, 1), dtype=torch.uint8, device=obs.device)
        out = torch.empty_like(obs)
        cnt = torch.zeros((1,), dtype=torch.int32, device=obs.device)
        _lib.check(L.gnd_compact_obstacles(_lib.ptr(lab8), _lib.ptr(obs), n, k, _lib.ptr(ws), need.value,
                                           _lib.ptr(out), _lib.ptr(cnt), _lib.stream_ptr()), "gnd_compact_obstacles")
        n_obs = int(cnt.item())
        obs[:n_obs] = out[:n_obs]                      # :242-243 compaction lands in points_obstacle itself
    if obs_np:
        points_obstacle[:n_obs] = out[:n_obs].cpu().numpy()
        res_obs = points_obstacle[:n_obs]
    else:
        res_obs = obs[:n_obs]
    return (labels.cpu().numpy() if is_np else labels), res_obs


def shift_cloud_func(cloud, height):
    """utils/utils.py:137-139 / predict_ground.py:104-107: cloud += [0, 0, height, 0], in place."""
    L = _lib.lib()
    if isinstance(cloud, np.ndarray):
        t = torch.from_numpy(np.ascontiguousarray(cloud, dtype=np.float32)).cuda()
    else:
        t = cloud
    if not (t.is_cuda and t.dtype == torch.float32 and t.dim() == 2 and t.is_contiguous()):
        raise TypeError("cloud must be a contiguous float32 [N, F] array")
    with torch.cuda.device(t.device):
        _lib.check(L.gnd_shift_cloud(_lib.ptr(t), t.shape[0], t.shape[1], float(np.float32(height)), _lib.stream_ptr()),
                   "gnd_shift_cloud")
    if isinstance(cloud, np.ndarray):
        cloud[...] = t.cpu().numpy()
    return cloud
