"""ctypes wrappers around oracle/libgnd_oracle.so (TEST INFRASTRUCTURE ONLY).

The functions keep the reference's Python signatures so the parity tests read
like calls into the reference:

  points_to_voxel        utils/point_cloud_ops.py:107-167
  shift_cloud            predict_ground.py:104-107 (_shift_cloud)
  segment_cloud          utils/utils.py:198-220
  segment_cloud_noground utils/utils.py:222-248
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libgnd_oracle.so")
_lib = None


def build(force=False):
    """Compile the C restatement (gcc); called by __graft_entry__.build()."""
    src = os.path.join(_HERE, "gnd_oracle.c")
    if (not force and os.path.exists(_LIB_PATH)
            and os.path.getmtime(_LIB_PATH) >= os.path.getmtime(src)):
        return _LIB_PATH
    subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB_PATH)
        c_fp = ctypes.POINTER(ctypes.c_float)
        c_ip = ctypes.POINTER(ctypes.c_int32)
        c_lp = ctypes.POINTER(ctypes.c_int64)
        c_dp = ctypes.POINTER(ctypes.c_double)
        L.gnd_oracle_shift_cloud.argtypes = [c_fp, ctypes.c_int64, ctypes.c_int, ctypes.c_float]
        L.gnd_oracle_shift_cloud.restype = None
        L.gnd_oracle_grid_size.argtypes = [c_fp, c_fp, c_ip]
        L.gnd_oracle_grid_size.restype = None
        L.gnd_oracle_points_to_voxel.argtypes = [
            c_fp, ctypes.c_int64, ctypes.c_int, c_fp, c_fp,
            ctypes.c_int32, ctypes.c_int32, ctypes.c_int,
            c_fp, c_ip, c_ip, c_ip, c_lp]
        L.gnd_oracle_points_to_voxel.restype = ctypes.c_int32
        L.gnd_oracle_segment_cloud.argtypes = [
            c_fp, ctypes.c_int64, ctypes.c_int, c_dp, ctypes.c_double,
            c_fp, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64,
            ctypes.c_double, c_dp, c_fp, ctypes.c_int]
        L.gnd_oracle_segment_cloud.restype = ctypes.c_int64
        L.gnd_oracle_unpack_cloud.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32, c_dp, ctypes.c_int,
                                              ctypes.c_int, c_fp]
        L.gnd_oracle_unpack_cloud.restype = ctypes.c_int64
        L.gnd_oracle_pack_segmented.argtypes = [c_fp, ctypes.c_int64, ctypes.c_int, c_dp, c_fp]
        L.gnd_oracle_pack_segmented.restype = None
        _lib = L
    return _lib


def _fp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _ip(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))


def shift_cloud(cloud, height):
    """In place, like the reference; returns the same array."""
    assert cloud.dtype == np.float32 and cloud.flags.c_contiguous and cloud.ndim == 2
    lib().gnd_oracle_shift_cloud(_fp(cloud), cloud.shape[0], cloud.shape[1], np.float32(height))
    return cloud


def grid_size(voxel_size, coors_range):
    vs = np.ascontiguousarray(voxel_size, dtype=np.float32)
    cr = np.ascontiguousarray(coors_range, dtype=np.float32)
    g = np.zeros(3, dtype=np.int32)
    lib().gnd_oracle_grid_size(_fp(vs), _fp(cr), _ip(g))
    return g


def points_to_voxel(points, voxel_size, coors_range, max_points=35, reverse_index=True,
                    max_voxels=20000, return_point_slot=False):
    """Same signature / outputs as the reference.  float32 points only."""
    points = np.ascontiguousarray(points)
    if points.dtype != np.float32:
        raise TypeError("oracle points_to_voxel restates the float32 path only")
    vs = np.ascontiguousarray(voxel_size, dtype=np.float32)
    cr = np.ascontiguousarray(coors_range, dtype=np.float32)
    g = grid_size(vs, cr)
    n, cols = points.shape
    voxels = np.zeros((max_voxels, max_points, cols), dtype=np.float32)
    coors = np.zeros((max_voxels, 3), dtype=np.int32)
    npv = np.zeros((max_voxels,), dtype=np.int32)
    c2v = -np.ones(int(g[0]) * int(g[1]) * int(g[2]), dtype=np.int32)
    slot = np.empty((n,), dtype=np.int64) if return_point_slot else None
    m = lib().gnd_oracle_points_to_voxel(
        _fp(points), n, cols, _fp(vs), _fp(cr), int(max_points), int(max_voxels),
        1 if reverse_index else 0, _fp(voxels), _ip(coors), _ip(npv), _ip(c2v),
        slot.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)) if slot is not None else None)
    out = (voxels[:m], coors[:m], npv[:m])
    if return_point_slot:
        return out + (slot,)
    return out


def _segment(points, points_obstacle, grid_size_, voxel_size, elevation_map, threshold):
    assert points.dtype == np.float32 and points.flags.c_contiguous and points.ndim == 2
    emap = elevation_map
    assert emap.dtype == np.float32 and emap.ndim == 2
    gl = np.array([float(grid_size_[0]), float(grid_size_[1])], dtype=np.float64)
    labels = np.zeros(points.shape[0], dtype=np.float64)
    isz = emap.itemsize
    if points_obstacle is not None:
        assert points_obstacle.dtype == np.float32 and points_obstacle.flags.c_contiguous
        po, oc = _fp(points_obstacle), points_obstacle.shape[1]
    else:
        po, oc = None, 0
    # keep the (possibly transposed) view alive; pass its base pointer + strides
    cnt = lib().gnd_oracle_segment_cloud(
        _fp(points), points.shape[0], points.shape[1],
        gl.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), float(voxel_size),
        ctypes.cast(emap.ctypes.data, ctypes.POINTER(ctypes.c_float)),
        emap.shape[0], emap.shape[1], emap.strides[0] // isz, emap.strides[1] // isz,
        float(threshold), labels.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), po, oc)
    return labels, cnt


def segment_cloud(points, grid_size, voxel_size, elevation_map, threshold=0.2):
    return _segment(points, None, grid_size, voxel_size, elevation_map, threshold)[0]


def segment_cloud_noground(points, points_obstacle, grid_size, voxel_size, elevation_map,
                           threshold=0.2):
    labels, cnt = _segment(points, points_obstacle, grid_size, voxel_size, elevation_map, threshold)
    return labels, points_obstacle[:cnt]


def unpack_cloud(data, point_step, trafo=None, drop_nan=True, out_cols=3):
    """ros_node.py:242-263 up to the NaN filter: msg.data bytes -> xyz rows (utils/utils.py:179-180), optional
    transformCloud with a 4x4 float64 matrix (utils/transform.py:148-165), NaN rows dropped in order."""
    buf = np.ascontiguousarray(np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data.view(np.uint8).reshape(-1))
    n = buf.size // point_step
    out = np.zeros((n, out_cols), dtype=np.float32)
    t = None if trafo is None else np.ascontiguousarray(np.asarray(trafo, dtype=np.float64).reshape(16))
    tp = t.ctypes.data_as(ctypes.POINTER(ctypes.c_double)) if t is not None else None
    m = lib().gnd_oracle_unpack_cloud(buf.ctypes.data, n, point_step, tp, 1 if drop_nan else 0, out_cols, _fp(out))
    return out[:m]


def pack_segmented(points, labels):
    """utils/ros_utils.py:171-191 np2ros_pub_2: the 16-byte (x, y, z, intensity=label) wire rows, as bytes."""
    points = np.ascontiguousarray(points, dtype=np.float32)
    lab = np.ascontiguousarray(labels, dtype=np.float64)
    out = np.empty((points.shape[0], 4), dtype=np.float32)
    lib().gnd_oracle_pack_segmented(_fp(points), points.shape[0], points.shape[1],
                                    lab.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), _fp(out))
    return out


def ground_counts(pred_labels, sem_label, gnd_classes=(40, 44, 48, 49, 60, 72)):
    """evaluate_SemanticKITTI.py:94-111, :198-204 restated with numpy (integer work): intersection, union,
    predicted-ground and GT-ground counts over the points remove_outliers keeps."""
    pred = np.asarray(pred_labels)
    sem = np.asarray(sem_label)
    gt = np.ones(sem.shape)                       # get_GndSeg :95-99
    gt[np.isin(sem, list(gnd_classes))] = 0
    gt[np.isin(sem, [0, 1])] = -1
    keep = (pred >= 0) & (gt >= 0)                # remove_outliers :103-110
    p, g = 1 - pred[keep], 1 - gt[keep]
    inter, union = np.logical_and(g, p), np.logical_or(g, p)
    return np.array([inter.sum(), union.sum(), p.sum(), g.sum()], dtype=np.int64)


def masked_mse(pred_gnd, target_gnd, gnd_mask):
    """evaluate_SemanticKITTI.py:224-226."""
    mse = (np.square(target_gnd - pred_gnd.T) * gnd_mask).sum()
    return mse / gnd_mask.sum()


def _map_shape(grid_size, voxel_size):
    gs = np.asarray(grid_size, dtype=np.float64)
    shp = (gs[2:] - gs[:2]) / float(voxel_size)           # utils/utils.py:261-262, :284-285 (int() truncation)
    return int(shp[0]), int(shp[1])


def lidar_to_img(points, grid_size, voxel_size, fill):
    """utils/utils.py:250-270 (does not modify ``points``; the reference is handed a copy)."""
    L = lib()
    pts = np.ascontiguousarray(points, dtype=np.float32)
    nx, ny = _map_shape(grid_size, voxel_size)
    gs = np.ascontiguousarray(np.asarray(grid_size, dtype=np.float64))
    img = np.empty((nx, ny), dtype=np.float64)
    L.gnd_oracle_lidar_to_img.argtypes = [ctypes.POINTER(ctypes.c_float), ctypes.c_int64, ctypes.c_int, ctypes.POINTER(ctypes.c_double),
                                          ctypes.c_double, ctypes.c_double, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
    L.gnd_oracle_lidar_to_img.restype = None
    L.gnd_oracle_lidar_to_img(_fp(pts), pts.shape[0], pts.shape[1], gs.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
                              float(voxel_size), float(fill), nx, ny, img.ctypes.data_as(ctypes.POINTER(ctypes.c_double)))
    return img


def lidar_to_heightmap(points, grid_size, voxel_size, max_points):
    """utils/utils.py:273-295, including the one-past-the-end slot write (see gnd_oracle.c)."""
    L = lib()
    pts = np.ascontiguousarray(points, dtype=np.float32)
    nx, ny = _map_shape(grid_size, voxel_size)
    gs = np.ascontiguousarray(np.asarray(grid_size, dtype=np.float64))
    slots = np.empty(nx * ny * int(max_points) + 1, dtype=np.float64)
    num = np.empty(nx * ny, dtype=np.int32)
    out = np.empty((nx, ny), dtype=np.float64)
    dp = ctypes.POINTER(ctypes.c_double)
    L.gnd_oracle_lidar_to_heightmap.argtypes = [ctypes.POINTER(ctypes.c_float), ctypes.c_int64, ctypes.c_int, dp, ctypes.c_double,
                                                ctypes.c_int, ctypes.c_int, ctypes.c_int, dp, ctypes.POINTER(ctypes.c_int32), dp]
    L.gnd_oracle_lidar_to_heightmap.restype = ctypes.c_int
    L.gnd_oracle_lidar_to_heightmap(_fp(pts), pts.shape[0], pts.shape[1], gs.ctypes.data_as(dp), float(voxel_size), int(max_points),
                                    nx, ny, slots.ctypes.data_as(dp), num.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), out.ctypes.data_as(dp))
    return out
