"""ctypes binding of libgndnet_b200.so (include/gndnet_b200.h).

The library is the product: there is no Python / CPU fallback.  If it is missing or a call fails,
the error is raised here, loudly.  ctypes releases the GIL around every foreign call, which is what
the reference's threading model needs (inference runs on a daemon thread, ros_node.py:51-60).
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgndnet_b200.so")

GND_OK = 0
PRECISIONS = {"fp32": 0, "mixed": 1, "split": 2}

c_f, c_d, c_i32, c_i64, c_vp, c_sz = (ctypes.c_float, ctypes.c_double, ctypes.c_int32, ctypes.c_int64,
                                      ctypes.c_void_p, ctypes.c_size_t)


class GndError(RuntimeError):
    def __init__(self, code, where, msg):
        super().__init__("%s failed (%d): %s" % (where, code, msg))
        self.code = code


class VoxelParams(ctypes.Structure):
    _fields_ = [("voxel_size", c_f * 3), ("coors_range", c_f * 6), ("max_points", c_i32),
                ("max_voxels", c_i32), ("reverse_index", c_i32), ("num_features", c_i32),
                ("apply_z_shift", c_i32), ("z_shift", c_f)]


class SegmentParams(ctypes.Structure):
    _fields_ = [("grid_lo", c_d * 2), ("voxel_size", c_d), ("threshold", c_d), ("map_nx", c_i32),
                ("map_ny", c_i32), ("map_sx", c_i64), ("map_sy", c_i64), ("map_batch_stride", c_i64),
                ("num_features", c_i32), ("apply_z_shift", c_i32), ("z_shift", c_f), ("mutate_points", c_i32)]


class Config(ctypes.Structure):
    _fields_ = [("pc_range", c_f * 6), ("voxel_size", c_f * 3), ("pc_range_f64", c_d * 6),
                ("voxel_size_f64", c_d * 3), ("grid_range", c_d * 4), ("lidar_height", c_f),
                ("max_points_voxel", c_i32), ("max_voxels", c_i32), ("input_features", c_i32),
                ("use_norm", c_i32), ("with_distance", c_i32), ("precision", c_i32), ("bn_batch_stats", c_i32),
                ("n_vfe", c_i32), ("vfe_filters", c_i32 * 4)]


class UnpackParams(ctypes.Structure):
    _fields_ = [("point_step", c_i32), ("has_transform", c_i32), ("transform", c_d * 16), ("drop_nan", c_i32),
                ("out_features", c_i32)]


class Weight(ctypes.Structure):
    _fields_ = [("name", ctypes.c_char_p), ("h_data", ctypes.POINTER(c_f)), ("numel", c_i64)]


# name -> (restype, argtypes); every name here must be exported (tests/test_abi.py checks against the header)
_SIGS = {
    "gnd_last_error_string": (ctypes.c_char_p, []),
    "gnd_version": (ctypes.c_int, []),
    "gnd_voxel_grid_size": (ctypes.c_int, [ctypes.POINTER(VoxelParams), ctypes.POINTER(c_i32)]),
    "gnd_voxel_workspace_bytes": (ctypes.c_int, [ctypes.POINTER(VoxelParams), c_i32, c_i64, ctypes.POINTER(c_sz)]),
    "gnd_points_to_voxel": (ctypes.c_int, [ctypes.POINTER(VoxelParams), c_vp, c_vp, c_i32, c_i64, c_vp, c_sz,
                                           c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "gnd_shift_cloud": (ctypes.c_int, [c_vp, c_i64, c_i32, c_f, c_vp]),
    "gnd_segment_cloud": (ctypes.c_int, [ctypes.POINTER(SegmentParams), c_vp, c_vp, c_i32, c_i64, c_vp, c_vp, c_vp, c_vp]),
    "gnd_compact_workspace_bytes": (ctypes.c_int, [c_i64, ctypes.POINTER(c_sz)]),
    "gnd_compact_obstacles": (ctypes.c_int, [c_vp, c_vp, c_i64, c_i32, c_vp, c_sz, c_vp, c_vp, c_vp]),
    "gnd_compact_obstacles_static": (ctypes.c_int, [c_vp, c_vp, c_i64, c_i32, c_vp, c_vp, c_vp]),
    "gnd_create": (ctypes.c_int, [ctypes.POINTER(Config), c_i32, ctypes.POINTER(c_vp)]),
    "gnd_destroy": (None, [c_vp]),
    "gnd_load_weights": (ctypes.c_int, [c_vp, ctypes.POINTER(Weight), c_i32]),
    "gnd_workspace_bytes": (ctypes.c_int, [c_vp, c_i32, c_i64, ctypes.POINTER(c_sz)]),
    "gnd_bind_workspace": (ctypes.c_int, [c_vp, c_vp, c_sz, c_i32, c_i64]),
    "gnd_grid_shape": (ctypes.c_int, [c_vp, ctypes.POINTER(c_i32)]),
    "gnd_voxelize": (ctypes.c_int, [c_vp, c_vp, c_vp, c_i32, c_i64, c_vp]),
    "gnd_pfn_scatter": (ctypes.c_int, [c_vp, c_vp, c_i32, c_i64, c_vp]),
    "gnd_pfn_scatter_dense": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i32, c_vp]),
    "gnd_segnet": (ctypes.c_int, [c_vp, c_i32, c_vp, c_vp]),
    "gnd_segnet_profile": (ctypes.c_int, [c_vp, c_i32, c_vp, ctypes.POINTER(c_f), ctypes.POINTER(c_f), c_vp]),
    "gnd_segment": (ctypes.c_int, [c_vp, c_vp, c_vp, c_i32, c_i64, c_vp, c_d, c_vp, c_vp, c_vp]),
    "gnd_infer": (ctypes.c_int, [c_vp, c_vp, c_vp, c_i32, c_i64, c_d, c_vp, c_vp, c_vp, c_vp]),
    "gnd_unpack_workspace_bytes": (ctypes.c_int, [c_i64, ctypes.POINTER(c_sz)]),
    "gnd_unpack_cloud": (ctypes.c_int, [ctypes.POINTER(UnpackParams), c_vp, c_i64, c_vp, c_sz, c_vp, c_vp, c_vp]),
    "gnd_pack_segmented": (ctypes.c_int, [c_vp, c_i64, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "gnd_eval_ground_counts": (ctypes.c_int, [c_vp, c_vp, c_vp, c_i64, ctypes.POINTER(ctypes.c_uint32), c_i32, c_vp, c_vp]),
    "gnd_eval_masked_sqerr": (ctypes.c_int, [c_vp, c_vp, c_vp, c_i32, c_i32, c_vp, c_vp]),
    "gnd_raster_workspace_bytes": (ctypes.c_int, [c_i64, ctypes.POINTER(c_d), c_d, c_i32, ctypes.POINTER(c_sz)]),
    "gnd_lidar_to_img": (ctypes.c_int, [c_vp, c_i64, c_i32, ctypes.POINTER(c_d), c_d, c_d, c_vp, c_vp]),
    "gnd_lidar_to_heightmap": (ctypes.c_int, [c_vp, c_i64, c_i32, ctypes.POINTER(c_d), c_d, c_i32, c_vp, c_sz, c_vp, c_vp]),
    "gnd_get_buffer": (ctypes.c_int, [c_vp, ctypes.c_char_p, ctypes.POINTER(c_vp), ctypes.POINTER(c_sz)]),
    "gnd_last_launch_count": (ctypes.c_int, [c_vp]),
    "gnd_set_option": (ctypes.c_int, [c_vp, ctypes.c_char_p, c_i32]),
    "gnd_conv_timing_begin": (ctypes.c_int, [c_vp, c_i32]),
    "gnd_conv_timing_end": (ctypes.c_int, [c_vp, ctypes.POINTER(c_f), ctypes.POINTER(c_f), ctypes.POINTER(c_f), ctypes.POINTER(c_i32)]),
}

_lib = None


def lib():
    """Load the shared library (once).  Raises if it has not been built: there is no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "gndnet_b200: %s is missing -- build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or gndnet_b200/csrc/build.sh.  There is no CPU fallback." % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)          # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc, where):
    if rc != GND_OK:
        msg = lib().gnd_last_error_string()
        raise GndError(rc, where, msg.decode() if msg else "?")


def ptr(t):
    """Device (or host) address of a torch tensor / None."""
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
