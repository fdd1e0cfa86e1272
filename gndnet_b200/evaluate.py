"""Device-side evaluation reductions of the reference's evaluate_SemanticKITTI.py.

  ground_metrics(pred_labels, sem_label, GndClasses) -> (iou, prec, recall)
        get_GndSeg (:94-100) + remove_outliers (:102-111) + the ratios of :198-204, from four exact integer counts
  masked_mse(pred_gnd, target_gnd, gnd_mask) -> mse                                   :224-226
"""
import ctypes

import numpy as np
import torch

from . import _lib

GND_CLASSES = (40, 44, 48, 49, 60, 72)        # evaluate_SemanticKITTI.py:190


def ground_counts(pred_labels, sem_label, GndClasses=GND_CLASSES):
    """int64 CUDA tensor [4]: intersection, union, predicted ground, GT ground (see include/gndnet_b200.h)."""
    if not (pred_labels.is_cuda and pred_labels.dtype in (torch.int8, torch.float64) and pred_labels.is_contiguous()):
        raise TypeError("pred_labels must be a contiguous int8 / float64 CUDA tensor")
    if not (sem_label.is_cuda and sem_label.dtype in (torch.int32, torch.uint32) and sem_label.is_contiguous()
            and sem_label.numel() == pred_labels.numel()):
        raise TypeError("sem_label must be a contiguous 32-bit CUDA tensor with one entry per point")
    cls = (ctypes.c_uint32 * len(GndClasses))(*[int(c) for c in GndClasses])
    L = _lib.lib()
    with torch.cuda.device(pred_labels.device):
        out = torch.empty((4,), dtype=torch.int64, device=pred_labels.device)
        l8 = pred_labels if pred_labels.dtype == torch.int8 else None
        l64 = pred_labels if pred_labels.dtype == torch.float64 else None
        _lib.check(L.gnd_eval_ground_counts(_lib.ptr(l8), _lib.ptr(l64), _lib.ptr(sem_label), pred_labels.numel(), cls,
                                            len(GndClasses), _lib.ptr(out), _lib.stream_ptr()), "gnd_eval_ground_counts")
    return out


def ground_metrics(pred_labels, sem_label, GndClasses=GND_CLASSES):
    """(iou, prec, recall) as the reference computes them (numpy float64 division of integer sums, :200-204)."""
    c = ground_counts(pred_labels, sem_label, GndClasses).cpu().numpy()
    with np.errstate(divide="ignore", invalid="ignore"):
        return (np.float64(c[0]) / np.float64(c[1]), np.float64(c[0]) / np.float64(c[2]), np.float64(c[0]) / np.float64(c[3]))


def masked_mse(pred_gnd, target_gnd, gnd_mask):
    """pred_gnd [ny, nx] float32 CUDA (model output); target_gnd, gnd_mask [nx, ny] float64 CUDA -> python float."""
    if not (pred_gnd.is_cuda and pred_gnd.dtype == torch.float32 and pred_gnd.dim() == 2 and pred_gnd.is_contiguous()):
        raise TypeError("pred_gnd must be a contiguous float32 CUDA tensor [ny, nx]")
    ny, nx = pred_gnd.shape
    for t in (target_gnd, gnd_mask):
        if not (t.is_cuda and t.dtype == torch.float64 and tuple(t.shape) == (nx, ny) and t.is_contiguous()):
            raise TypeError("target_gnd / gnd_mask must be contiguous float64 CUDA tensors [nx, ny]")
    L = _lib.lib()
    with torch.cuda.device(pred_gnd.device):
        sums = torch.empty((2,), dtype=torch.float64, device=pred_gnd.device)
        _lib.check(L.gnd_eval_masked_sqerr(_lib.ptr(pred_gnd), _lib.ptr(target_gnd), _lib.ptr(gnd_mask), nx, ny,
                                           _lib.ptr(sums), _lib.stream_ptr()), "gnd_eval_masked_sqerr")
    s = sums.cpu().numpy()
    return float(s[0] / s[1])


# ---------------------------------------------------------------------------------------- ground-truth rasterisation
def _raster_args(points, grid_size):
    is_np = isinstance(points, np.ndarray)
    if is_np:
        if points.dtype != np.float32 or points.ndim != 2 or points.shape[1] < 3:
            raise TypeError("points must be float32 [N, >=3]")
        t = torch.from_numpy(np.ascontiguousarray(points)).cuda()
    else:
        if not (points.is_cuda and points.dtype == torch.float32 and points.dim() == 2 and points.shape[1] >= 3):
            raise TypeError("points must be a float32 [N, >=3] numpy array or CUDA tensor")
        t = points.contiguous()
    gs = (ctypes.c_double * 4)(*[float(x) for x in np.asarray(grid_size).reshape(4)])
    return t, gs, is_np


def _map_shape(grid_size, voxel_size):
    g = np.asarray(grid_size, dtype=np.float64)
    shp = (g[2:] - g[:2]) / float(voxel_size)
    return int(shp[0]), int(shp[1])


def lidar_to_img(points, grid_size, voxel_size, fill):
    """utils/utils.py:250-270: [nx, ny] float64 occupancy mask (``fill`` where a point with z + 1.732 < 10 falls).
    Unlike the numba original it does not overwrite ``points[:, :2]`` (the reference's callers pass a copy)."""
    t, gs, is_np = _raster_args(points, grid_size)
    nx, ny = _map_shape(grid_size, voxel_size)
    with torch.cuda.device(t.device):
        img = torch.empty((nx, ny), dtype=torch.float64, device=t.device)
        _lib.check(_lib.lib().gnd_lidar_to_img(_lib.ptr(t), t.shape[0], t.shape[1], gs, float(voxel_size), float(fill),
                                               _lib.ptr(img), _lib.stream_ptr()), "gnd_lidar_to_img")
    return img.cpu().numpy() if is_np else img


def lidar_to_heightmap(points, grid_size, voxel_size, max_points):
    """utils/utils.py:273-295: [nx, ny] float64 mean height of the first points of each cell in input order, with the
    reference's divisor (count + 1) and its one-slot overrun into the next cell reproduced (csrc/raster.cu)."""
    t, gs, is_np = _raster_args(points, grid_size)
    nx, ny = _map_shape(grid_size, voxel_size)
    L = _lib.lib()
    need = ctypes.c_size_t()
    with torch.cuda.device(t.device):
        _lib.check(L.gnd_raster_workspace_bytes(t.shape[0], gs, float(voxel_size), int(max_points), ctypes.byref(need)),
                   "gnd_raster_workspace_bytes")
        ws = torch.empty(need.value + 256, dtype=torch.uint8, device=t.device)
        off = (-ws.data_ptr()) % 256
        out = torch.empty((nx, ny), dtype=torch.float64, device=t.device)
        _lib.check(L.gnd_lidar_to_heightmap(_lib.ptr(t), t.shape[0], t.shape[1], gs, float(voxel_size), int(max_points),
                                            ctypes.c_void_p(ws.data_ptr() + off), need.value, _lib.ptr(out), _lib.stream_ptr()),
                   "gnd_lidar_to_heightmap")
    return out.cpu().numpy() if is_np else out


def get_target_gnd(cloud, sem_label, grid_range, voxel_size, GndClasses=GND_CLASSES, max_points=100):
    """evaluate_SemanticKITTI.py:118-127: (gnd_heightmap, gnd_mask) of the points labelled as ground; CUDA tensors."""
    if cloud.shape[0] != sem_label.shape[0]:
        raise Exception("Points and label MisMatch")
    sem = sem_label.to(torch.int64) & 0xFFFFFFFF
    index = torch.isin(sem, torch.tensor([int(c) for c in GndClasses], device=sem.device))
    gnd = cloud[index].contiguous()
    return lidar_to_heightmap(gnd, grid_range, voxel_size, max_points), lidar_to_img(gnd, grid_range, voxel_size, 1)
