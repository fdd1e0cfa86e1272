"""Drop-in for the reference's ``utils/point_cloud_ops.py`` (points_to_voxel), running on the B200.

Same signature, same outputs (utils/point_cloud_ops.py:107-167): ``voxels [M, max_points, ndim]``
(points.dtype), ``coordinates [M, 3]`` int32 (zyx when ``reverse_index``), ``num_points_per_voxel [M]``
int32 -- bit-identical to the numba loop, including ``max_points`` truncation and the ``max_voxels``
``break``.  numpy in -> numpy out (like the reference); CUDA tensor in -> CUDA tensors out.
Does not mutate ``points``.  The kernels compute in float32; the reference computes in ``points.dtype``, so
float64 arrays (predict_ground.py:129-131 makes one for short .npy files) are cast with a warning and the
pillars come back as float64 ("cast-and-warn or reject").
"""
import ctypes
import warnings

import numpy as np
import torch

from . import _lib


def _voxel_params(voxel_size, coors_range, max_points, max_voxels, reverse_index, num_features,
                  z_shift=None):
    p = _lib.VoxelParams()
    vs = np.asarray(voxel_size, dtype=np.float32).reshape(3)       # :138-139 cast to points.dtype
    cr = np.asarray(coors_range, dtype=np.float32).reshape(6)      # :140-141
    for j in range(3):
        p.voxel_size[j] = float(vs[j])
    for j in range(6):
        p.coors_range[j] = float(cr[j])
    p.max_points, p.max_voxels = int(max_points), int(max_voxels)
    p.reverse_index = 1 if reverse_index else 0
    p.num_features = int(num_features)
    p.apply_z_shift = 0 if z_shift is None else 1
    p.z_shift = 0.0 if z_shift is None else float(np.float32(z_shift))
    return p


def voxelize_batch(points, voxel_size, coors_range, max_points, max_voxels, reverse_index=True,
                   counts=None, z_shift=None, want_voxels=False, want_slot=False):
    """Batched core: ``points`` [B, N, F] float32 CUDA tensor.  Returns a dict of CUDA tensors
    (coors [B,max_voxels,3], num_points [B,max_voxels], voxel_count [B], point_offsets
    [B,max_voxels+1], point_list [B,N], optionally voxels [B,max_voxels,P,F] and point_slot [B,N]).
    No host synchronisation."""
    if not (isinstance(points, torch.Tensor) and points.is_cuda and points.dtype == torch.float32
            and points.dim() == 3 and points.is_contiguous()):
        raise TypeError("voxelize_batch needs a contiguous float32 CUDA tensor [B, N, F]")
    L = _lib.lib()
    B, N, F = points.shape
    dev = points.device
    p = _voxel_params(voxel_size, coors_range, max_points, max_voxels, reverse_index, F, z_shift)
    need = ctypes.c_size_t()
    with torch.cuda.device(dev):
        _lib.check(L.gnd_voxel_workspace_bytes(ctypes.byref(p), B, N, ctypes.byref(need)), "gnd_voxel_workspace_bytes")
        ws = torch.empty(need.value, dtype=torch.uint8, device=dev)
        out = dict(
            coors=torch.empty((B, max_voxels, 3), dtype=torch.int32, device=dev),
            num_points=torch.empty((B, max_voxels), dtype=torch.int32, device=dev),
            voxel_count=torch.empty((B,), dtype=torch.int32, device=dev),
            point_offsets=torch.empty((B, max_voxels + 1), dtype=torch.int32, device=dev),
            point_list=torch.empty((B, max(N, 1)), dtype=torch.int32, device=dev),
        )
        if want_voxels:   # :149-150 zero-filled max-size buffer, sliced by the caller
            out["voxels"] = torch.zeros((B, max_voxels, max_points, F), dtype=torch.float32, device=dev)
        if want_slot:
            out["point_slot"] = torch.empty((B, max(N, 1)), dtype=torch.int32, device=dev)
        _lib.check(L.gnd_points_to_voxel(
            ctypes.byref(p), _lib.ptr(points), _lib.ptr(counts), B, N, _lib.ptr(ws), need.value,
            _lib.ptr(out["coors"]), _lib.ptr(out["num_points"]), _lib.ptr(out["voxel_count"]),
            _lib.ptr(out["point_offsets"]), _lib.ptr(out["point_list"]), _lib.ptr(out.get("voxels")),
            _lib.ptr(out.get("point_slot")), _lib.stream_ptr()), "gnd_points_to_voxel")
    return out


def points_to_voxel(points, voxel_size, coors_range, max_points=35, reverse_index=True, max_voxels=20000):
    """utils/point_cloud_ops.py:107-167, same arguments and return values."""
    is_np = isinstance(points, np.ndarray)
    if is_np:
        if points.ndim != 2 or points.shape[1] < 3:
            raise ValueError("points must be [N, >=3]")
        src_dtype = points.dtype
        if src_dtype == np.float64:
            # The reference computes in points.dtype (:138-141), and predict_ground.py:129-131 itself builds a float64
            # array when a .npy file has too few columns.  The kernels are float32: cast, say so, and hand the pillars
            # back in the caller's dtype.  A point within one float32 ulp of a cell edge may land in the neighbouring
            # cell compared with a float64 run.
            warnings.warn("gndnet_b200.points_to_voxel: float64 points are voxelized in float32 (the reference would "
                          "compute in float64); pass float32 to silence this", stacklevel=2)
        elif src_dtype != np.float32:
            raise TypeError("points_to_voxel: float32 (or float64, cast with a warning) points only, got %s" % src_dtype)
        t = torch.from_numpy(np.ascontiguousarray(points, dtype=np.float32)).cuda()
    elif isinstance(points, torch.Tensor):
        if not points.is_cuda:
            raise TypeError("points_to_voxel: torch input must be a CUDA tensor (there is no CPU path)")
        if points.dtype != torch.float32 or points.dim() != 2 or points.shape[1] < 3:
            raise TypeError("points must be a float32 [N, >=3] tensor")
        t = points.contiguous()
    else:
        raise TypeError("points must be a numpy array or a CUDA tensor")
    out = voxelize_batch(t.unsqueeze(0), voxel_size, coors_range, int(max_points), int(max_voxels),
                         bool(reverse_index), want_voxels=True)
    m = int(out["voxel_count"][0].item())           # the reference returns sliced arrays: one sync
    voxels = out["voxels"][0, :m]
    coors = out["coors"][0, :m]
    num = out["num_points"][0, :m]
    if is_np:
        return voxels.cpu().numpy().astype(src_dtype, copy=False), coors.cpu().numpy(), num.cpu().numpy()
    return voxels, coors, num
