"""Device-side versions of the steps either side of the hot path in the reference's streaming node
(ros_node.py:242-288).  No ROS types here: the message is its ``data`` bytes and
``point_step``; headers / field lists stay with the caller.

  cloud_msg_to_numpy(msg, camera_height, shift_cloud=False)["xyz"]      utils/utils.py:148-195
  transformStampedToTransformationMatrix / transformation_matrix        utils/transform.py:99-111, :63-93
  transformCloud(points, trafo)                                          utils/transform.py:148-165
  cloud[~np.isnan(cloud).any(axis=1)]                                    ros_node.py:263
        -> unpack_cloud(data, point_step, trafo=None, drop_nan=True)     one call, on the GPU
  np2ros_pub_2's 16-byte rows                                            utils/ros_utils.py:171-191
        -> pack_segmented(points, labels)
  array_to_pointcloud2(cloud_obs, header)'s 12-byte rows                 utils/ros_utils.py:210-248
        -> obstacle_rows(points_obstacle, labels)   (gnd_compact_obstacles, k = 3)
"""
import ctypes

import numpy as np
import torch

from . import _lib


def quaternion_rotation_matrix(Q):
    """utils/transform.py:63-93: Q = (qx, qy, qz, qw) -> 3x3 float64, same expression order."""
    q0, q1, q2, q3 = float(Q[3]), float(Q[0]), float(Q[1]), float(Q[2])
    return np.array([[2 * (q0 * q0 + q1 * q1) - 1, 2 * (q1 * q2 - q0 * q3), 2 * (q1 * q3 + q0 * q2)],
                     [2 * (q1 * q2 + q0 * q3), 2 * (q0 * q0 + q2 * q2) - 1, 2 * (q2 * q3 - q0 * q1)],
                     [2 * (q1 * q3 - q0 * q2), 2 * (q2 * q3 + q0 * q1), 2 * (q0 * q0 + q3 * q3) - 1]])


def transformation_matrix(T, R):
    """utils/transform.py:99-105."""
    return np.array([[R[0][0], R[0][1], R[0][2], T[0]],
                     [R[1][0], R[1][1], R[1][2], T[1]],
                     [R[2][0], R[2][1], R[2][2], T[2]],
                     [0, 0, 0, 1]], dtype=np.float64)


def transformStampedToTransformationMatrix(ts):
    """utils/transform.py:107-111 (duck-typed geometry_msgs/TransformStamped)."""
    t, r = ts.transform.translation, ts.transform.rotation
    return transformation_matrix(np.array([t.x, t.y, t.z]), quaternion_rotation_matrix(np.array([r.x, r.y, r.z, r.w])))


def _bytes_to_cuda(data, device):
    if isinstance(data, torch.Tensor):
        if not (data.is_cuda and data.dtype == torch.uint8 and data.is_contiguous()):
            raise TypeError("msg data tensor must be a contiguous uint8 CUDA tensor")
        return data.reshape(-1)
    a = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else np.ascontiguousarray(data).view(np.uint8).reshape(-1)
    return torch.from_numpy(a.copy() if not a.flags.writeable else a).to(device)


def unpack_cloud(data, point_step, trafo=None, drop_nan=True, out_features=3, device=None, out=None):
    """msg.data -> (points [n, out_features] float32 CUDA tensor, count int32 CUDA tensor [1]).

    The first ``count`` rows of ``points`` are the cloud ros_node.py:242-263 would hand to the model: xyz of every
    row, moved by the float64 4x4 ``trafo`` if given, rows with a NaN dropped, order kept.  ``count`` stays on the
    device (pass it as ``counts`` to GroundEstimatorNet.infer); ``int(count)`` synchronises.
    """
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    buf = _bytes_to_cuda(data, device)
    point_step = int(point_step)
    if point_step < 12 or buf.numel() % point_step:
        raise ValueError("data length %d is not a multiple of point_step %d (>= 12)" % (buf.numel(), point_step))
    n = buf.numel() // point_step
    L = _lib.lib()
    p = _lib.UnpackParams()
    p.point_step, p.drop_nan, p.out_features = point_step, int(bool(drop_nan)), int(out_features)
    p.has_transform = 0 if trafo is None else 1
    if trafo is not None:
        t = np.asarray(trafo, dtype=np.float64)
        if t.shape != (4, 4):
            raise ValueError("trafo must be a 4x4 matrix")
        for i, v in enumerate(t.reshape(16)):
            p.transform[i] = float(v)
    need = ctypes.c_size_t()
    with torch.cuda.device(buf.device):
        _lib.check(L.gnd_unpack_workspace_bytes(n, ctypes.byref(need)), "gnd_unpack_workspace_bytes")
        ws = torch.empty(max(need.value, 1), dtype=torch.uint8, device=buf.device)
        pts = out if out is not None else torch.empty((n, int(out_features)), dtype=torch.float32, device=buf.device)
        if pts.shape != (n, int(out_features)) or not pts.is_contiguous() or pts.dtype != torch.float32:
            raise ValueError("out must be a contiguous float32 [n, out_features] CUDA tensor")
        cnt = torch.empty((1,), dtype=torch.int32, device=buf.device)
        _lib.check(L.gnd_unpack_cloud(ctypes.byref(p), _lib.ptr(buf), n, _lib.ptr(ws), need.value, _lib.ptr(pts),
                                      _lib.ptr(cnt), _lib.stream_ptr()), "gnd_unpack_cloud")
    return pts, cnt


def pack_segmented(points, labels):
    """[N, >=3] float32 CUDA points + labels (int8 or float64 CUDA) -> [N, 4] float32 rows (x, y, z, label): the
    ``data`` of the PointCloud2 np2ros_pub_2 publishes (point_step 16; ``rows.view(torch.uint8)`` are the bytes)."""
    if not (points.is_cuda and points.dtype == torch.float32 and points.dim() == 2 and points.is_contiguous()):
        raise TypeError("points must be a contiguous float32 CUDA tensor [N, F]")
    if labels.dtype not in (torch.int8, torch.float64) or labels.shape != (points.shape[0],) or not labels.is_cuda:
        raise TypeError("labels must be an int8 or float64 CUDA tensor [N]")
    L = _lib.lib()
    with torch.cuda.device(points.device):
        out = torch.empty((points.shape[0], 4), dtype=torch.float32, device=points.device)
        l8 = labels if labels.dtype == torch.int8 else None
        l64 = labels if labels.dtype == torch.float64 else None
        _lib.check(L.gnd_pack_segmented(_lib.ptr(points), points.shape[0], points.shape[1], _lib.ptr(l8), _lib.ptr(l64),
                                        _lib.ptr(out), _lib.stream_ptr()), "gnd_pack_segmented")
    return out


def obstacle_rows(points_obstacle, labels_i8):
    """The 12-byte rows of array_to_pointcloud2(cloud_obs): points with label 1, input order kept.
    Returns (rows [N, k] float32 CUDA tensor, count int32 CUDA tensor [1]); the first ``count`` rows are valid."""
    if not (points_obstacle.is_cuda and points_obstacle.dtype == torch.float32 and points_obstacle.is_contiguous()):
        raise TypeError("points_obstacle must be a contiguous float32 CUDA tensor [N, k]")
    L = _lib.lib()
    n, k = points_obstacle.shape
    need = ctypes.c_size_t()
    with torch.cuda.device(points_obstacle.device):
        _lib.check(L.gnd_compact_workspace_bytes(n, ctypes.byref(need)), "gnd_compact_workspace_bytes")
        ws = torch.empty(max(need.value, 1), dtype=torch.uint8, device=points_obstacle.device)
        out = torch.empty_like(points_obstacle)
        cnt = torch.zeros((1,), dtype=torch.int32, device=points_obstacle.device)
        _lib.check(L.gnd_compact_obstacles(_lib.ptr(labels_i8), _lib.ptr(points_obstacle), n, k, _lib.ptr(ws), need.value,
                                           _lib.ptr(out), _lib.ptr(cnt), _lib.stream_ptr()), "gnd_compact_obstacles")
    return out, cnt
