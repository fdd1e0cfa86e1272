"""GPU: gnd_compact_obstacles_static -- the obstacle compaction for callers that cannot allocate device memory.

Some users of the library only ever receive device buffers: a callback living inside another component's memory
pool, a plug-in in a process whose GPU memory was handed out completely at start-up, a real-time loop where
cudaMalloc is forbidden because it stalls the device.  They cannot give gnd_compact_obstacles its workspace, so this
variant keeps the block counters in a fixed 16 KB array inside the library (one per GPU) and takes inputs and outputs
only.  The array is shared by every call and never cleared, so what has to hold is: same bytes as the workspace
version, call after call, whatever the previous call left in the array.

This file is named to be collected last: with `pytest -x` a failure here cannot hide any other test.  It does not
depend on anything the other tests did (it passes alone in a fresh process and after the whole suite).
"""
import ctypes

import numpy as np
import pytest
import torch

from gndnet_b200 import _lib

pytestmark = pytest.mark.gpu

MAX_POINTS = 4194304      # GND_COMPACT_STATIC_MAX_POINTS in include/gndnet_b200.h


def _case(n, k, seed, frac_obstacle):
    g = np.random.default_rng(seed)
    labels = g.choice(np.array([-1, 0, 1], np.int8), size=n,
                      p=[0.1, 0.9 - frac_obstacle, frac_obstacle]).astype(np.int8)
    rows = g.standard_normal((n, k)).astype(np.float32)
    return torch.from_numpy(labels).cuda(), torch.from_numpy(rows).cuda()


def _with_workspace(L, labels, rows):
    """The existing entry point: the caller brings the scratch."""
    n, k = rows.shape
    need = ctypes.c_size_t()
    _lib.check(L.gnd_compact_workspace_bytes(n, ctypes.byref(need)), "gnd_compact_workspace_bytes")
    ws = torch.empty(max(need.value, 256), dtype=torch.uint8, device=rows.device)
    out = torch.full_like(rows, -7.0)
    cnt = torch.full((1,), -1, dtype=torch.int32, device=rows.device)
    _lib.check(L.gnd_compact_obstacles(_lib.ptr(labels), _lib.ptr(rows), n, k, _lib.ptr(ws), need.value,
                                       _lib.ptr(out), _lib.ptr(cnt), _lib.stream_ptr()), "gnd_compact_obstacles")
    return out, cnt


def _without_workspace(L, labels, rows):
    """The entry point under test: inputs and outputs only."""
    n, k = rows.shape
    out = torch.full_like(rows, -7.0)
    cnt = torch.full((1,), -1, dtype=torch.int32, device=rows.device)
    _lib.check(L.gnd_compact_obstacles_static(_lib.ptr(labels), _lib.ptr(rows), n, k, _lib.ptr(out), _lib.ptr(cnt),
                                              _lib.stream_ptr()), "gnd_compact_obstacles_static")
    return out, cnt


def test_static_scratch_compaction_matches_workspace_version(cuda):
    L = _lib.lib()
    # back-to-back calls of different sizes: a big one first, so that the smaller ones after it find the array full
    # of that call's counters, then a big one again.  Compaction is index work: the bar is bit equality.
    for n, k, seed, frac in [(300000, 3, 0, 0.4), (120000, 3, 1, 0.05), (1023, 4, 2, 0.9), (1025, 1, 3, 0.5),
                             (300000, 3, 4, 0.0), (120000, 3, 1, 0.05)]:
        labels, rows = _case(n, k, seed, frac)
        exp_out, exp_cnt = _with_workspace(L, labels, rows)
        got_out, got_cnt = _without_workspace(L, labels, rows)
        torch.cuda.synchronize()
        m = int(exp_cnt.item())
        assert m == int((labels == 1).sum().item())
        assert int(got_cnt.item()) == m
        assert got_out.cpu().numpy().tobytes() == exp_out.cpu().numpy().tobytes()   # untouched tail (-7) included
        # and against the definition: the rows labelled 1, in input order
        assert torch.equal(got_out[:m], rows[labels == 1])


def test_static_scratch_limits_and_empty_cloud(cuda):
    L = _lib.lib()
    cnt = torch.full((1,), -1, dtype=torch.int32, device=cuda)
    one = torch.zeros(4, dtype=torch.float32, device=cuda)
    lab = torch.ones(4, dtype=torch.int8, device=cuda)
    # an empty cloud is answered with a zero count
    assert L.gnd_compact_obstacles_static(None, None, 0, 3, None, _lib.ptr(cnt), _lib.stream_ptr()) == 0
    torch.cuda.synchronize()
    assert int(cnt.item()) == 0
    # one point past the limit is refused before anything is launched, with a message that names the limit
    rc = L.gnd_compact_obstacles_static(_lib.ptr(lab), _lib.ptr(one), MAX_POINTS + 1, 1, _lib.ptr(one), _lib.ptr(cnt),
                                        _lib.stream_ptr())
    assert rc == -4 and str(MAX_POINTS).encode() in L.gnd_last_error_string()
    with pytest.raises(_lib.GndError):
        _lib.check(rc, "gnd_compact_obstacles_static")
    # exactly the limit runs (labels only: one column of payload)
    labels, rows = _case(MAX_POINTS, 1, 5, 0.3)
    exp_out, exp_cnt = _with_workspace(L, labels, rows)
    got_out, got_cnt = _without_workspace(L, labels, rows)
    torch.cuda.synchronize()
    assert int(got_cnt.item()) == int(exp_cnt.item()) and torch.equal(got_out, exp_out)
