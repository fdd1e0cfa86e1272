"""gndnet_b200 -- B200-native (sm_100a) implementation of the GndNet inference hot path.

voxelize -> pillar-encode -> scatter -> segnet -> per-point labels, as hand-written CUDA kernels in
``libgndnet_b200.so`` (C ABI: include/gndnet_b200.h) behind the reference's own Python surface:

    from gndnet_b200.point_cloud_ops import points_to_voxel      # utils/point_cloud_ops.py
    from gndnet_b200.model import GroundEstimatorNet              # model.py
    from gndnet_b200.utils import segment_cloud, segment_cloud_noground, shift_cloud_func   # utils/utils.py

There is no CPU fallback: importing works anywhere, calling needs the built library and a B200.
"""
from . import _lib  # noqa: F401
from .model import GroundEstimatorNet  # noqa: F401
from .point_cloud_ops import points_to_voxel, voxelize_batch  # noqa: F401
from .utils import segment_cloud, segment_cloud_noground, shift_cloud_func  # noqa: F401

__all__ = ["GroundEstimatorNet", "points_to_voxel", "voxelize_batch", "segment_cloud",
           "segment_cloud_noground", "shift_cloud_func"]
