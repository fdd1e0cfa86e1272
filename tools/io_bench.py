"""Device times and achieved GB/s of the SURVEY 8f kernels (message unpack + transform + NaN filter, wire packing,
obstacle compaction, evaluation counts) -- algorithmic bytes / CUDA-event time, inputs rotated through > L2."""
import json, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gndnet_b200 import ros_io, evaluate  # noqa: E402


def timeit(fn, sets, reps=40):
    for i in range(5):
        fn(sets[i % len(sets)])
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps):
        fn(sets[i % len(sets)])
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


out = {}
rng = np.random.default_rng(0)
c, s = np.cos(0.3), np.sin(0.3)
trafo = np.array([[c, -s, 0, 1.0], [s, c, 0, 2.0], [0, 0, 1, 0.5], [0, 0, 0, 1]])
ONLY = int(os.environ.get("IO_N", "0"))       # restrict to one size (clean per-kernel ncu numbers)
for n, step in ((120000, 16), (120000, 32), (1000000, 16), (8000000, 16)):
    if ONLY and n != ONLY:
        continue
    nset = max(2, int(300e6 // (n * step)) + 1) if n * step < 300e6 else 2
    sets = []
    for k in range(min(nset, 24)):
        raw = rng.integers(0, 256, size=(n, step), dtype=np.uint8)
        xyz = (rng.random((n, 3)) * 100 - 50).astype(np.float32)
        xyz[rng.random(n) < 0.02, 0] = np.nan
        raw[:, :12] = xyz.view(np.uint8).reshape(n, 12)
        sets.append(torch.from_numpy(raw).cuda())
    buf = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    ms = timeit(lambda d: ros_io.unpack_cloud(d, step, trafo, True, 4, out=buf), sets)
    alg = n * (2 * 12 + 16)            # xyz read twice (count + write passes), 16-byte rows written
    out["unpack n=%d step=%d" % (n, step)] = {"ms": round(ms, 4), "alg_mb": alg / 1e6, "gbs": round(alg / 1e6 / ms, 1)}
for n in (120000, 1000000, 8000000):
    if ONLY and n != ONLY:
        continue
    k = 24 if n <= 1000000 else 3
    pts = [torch.from_numpy((rng.random((n, 4)) * 100 - 50).astype(np.float32)).cuda() for _ in range(k)]
    lab = torch.from_numpy(rng.integers(-1, 2, size=n).astype(np.int8)).cuda()
    sem = torch.from_numpy(rng.choice(np.array([0, 1, 40, 44, 50, 72, 99], dtype=np.int32), size=n)).cuda()
    ms = timeit(lambda p: ros_io.pack_segmented(p, lab), pts)
    out["pack_segmented n=%d" % n] = {"ms": round(ms, 4), "gbs": round(n * (12 + 1 + 16) / 1e6 / ms, 1)}
    ms = timeit(lambda p: ros_io.obstacle_rows(p, lab), pts)
    out["obstacle_rows n=%d" % n] = {"ms": round(ms, 4), "gbs": round(n * (2 * 1 + 16 / 3 + 16 / 3) / 1e6 / ms, 1)}
    ms = timeit(lambda p: evaluate.ground_counts(lab, sem), pts)
    out["eval_counts n=%d" % n] = {"ms": round(ms, 4), "gbs": round(n * 5 / 1e6 / ms, 1)}
print(json.dumps(out, indent=1))
