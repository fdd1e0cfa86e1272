"""CPU: the C-ABI library loads without a GPU and exports every symbol include/gndnet_b200.h declares;
host-only entry points (sizes, argument validation) behave.  No compute call is made here."""
import ctypes
import os
import re

import pytest

from gndnet_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "gndnet_b200.h")).read()
    return sorted(set(re.findall(r"GND_API\s+[\w\s\*]+?\b(gnd_\w+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    names = _declared()
    assert len(names) >= 20
    L = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(L, n), "libgndnet_b200.so does not export %s" % n
    assert sorted(_lib._SIGS) == names, "ctypes table and header disagree"
    assert _lib.lib().gnd_version() >= 100


def _params(**kw):
    p = _lib.VoxelParams()
    for j, v in enumerate(kw.get("vs", [1, 1, 8])):
        p.voxel_size[j] = v
    for j, v in enumerate(kw.get("cr", [-47, -50, -4, 53, 50, 4])):
        p.coors_range[j] = v
    p.max_points, p.max_voxels = kw.get("max_points", 100), kw.get("max_voxels", 10000)
    p.reverse_index, p.num_features = 1, kw.get("f", 4)
    return p


def test_grid_size_is_float32_round_half_even():
    L = _lib.lib()
    g = (ctypes.c_int32 * 3)()
    assert L.gnd_voxel_grid_size(ctypes.byref(_params()), g) == 0 and list(g) == [100, 100, 1]
    p = _params(vs=[.2, .2, 8], cr=[0, -5, -4, 10, 5, 4])
    assert L.gnd_voxel_grid_size(ctypes.byref(p), g) == 0 and list(g) == [50, 50, 1]


def test_workspace_and_validation_without_gpu():
    L = _lib.lib()
    n = ctypes.c_size_t()
    assert L.gnd_voxel_workspace_bytes(ctypes.byref(_params()), 64, 120000, ctypes.byref(n)) == 0
    assert 40e6 < n.value < 200e6
    assert L.gnd_voxel_workspace_bytes(ctypes.byref(_params()), 1, 0, ctypes.byref(n)) == 0
    # unsupported envelope is reported, not silently mishandled
    rc = L.gnd_voxel_workspace_bytes(ctypes.byref(_params(vs=[.1, .1, 8])), 1, 1000, ctypes.byref(n))
    assert rc == -4 and b"cells" in L.gnd_last_error_string()
    rc = L.gnd_voxel_workspace_bytes(ctypes.byref(_params(f=2)), 1, 1000, ctypes.byref(n))
    assert rc == -1
    rc = L.gnd_voxel_workspace_bytes(ctypes.byref(_params()), 0, 1000, ctypes.byref(n))
    assert rc == -1
    assert L.gnd_compact_workspace_bytes(120000, ctypes.byref(n)) == 0 and n.value >= 118 * 4


def test_calls_fail_loudly_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("box has a GPU")
    L = _lib.lib()
    h = ctypes.c_void_p()
    rc = L.gnd_create(ctypes.byref(_lib.Config()), 0, ctypes.byref(h))
    assert rc != 0 and not h.value
    from gndnet_b200.point_cloud_ops import points_to_voxel
    import numpy as np
    with pytest.raises(Exception):
        points_to_voxel(np.zeros((10, 4), np.float32), [1, 1, 8], [-47, -50, -4, 53, 50, 4], 100, True, 10000)
