"""Does running two independent contexts on two streams (each half the batch) beat one context on one stream?
Fills conv wave-quantisation tails / launch bubbles of one with the front-end kernels of the other."""
import json, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gndnet_b200 import synth  # noqa: E402
from gndnet_b200.model import GroundEstimatorNet  # noqa: E402

cfg = synth.make_cfg("kitti")
sd, _ = synth.checkpoint("kitti")


def mk():
    m = GroundEstimatorNet(cfg, precision="mixed")
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
    return m.cuda().eval()


scans = np.stack([synth.hdl64_scan(100 + s) for s in range(128)])
res = {}
for nctx, B in ((1, 64), (2, 32), (2, 64), (4, 16), (3, 64)):
    models = [mk() for _ in range(nctx)]
    streams = [torch.cuda.Stream() for _ in range(nctx)]
    pts = [[torch.from_numpy(scans[(j * 64 + k * B) % 128:][:B] if (j * 64 + k * B) % 128 + B <= 128 else scans[:B]).cuda() for j in range(2)] for k in range(nctx)]
    outs = [(torch.empty((B, 100, 100), device="cuda"), torch.empty((B, 120000), dtype=torch.int8, device="cuda")) for _ in range(nctx)]
    steps = 100

    def run(n):
        for i in range(n):
            for k in range(nctx):
                with torch.cuda.stream(streams[k]):
                    models[k].infer(pts[k][i & 1], threshold=0.08, out=outs[k])
    with torch.no_grad():
        run(5)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in streams:
            s.wait_event(e0)
        run(steps)
        for s in streams:
            torch.cuda.current_stream().wait_stream(s)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    res["%d ctx x batch %d" % (nctx, B)] = round(nctx * B * steps / ms * 1e3, 0)
    for m in models:
        m.release()
    del models, pts, outs
    torch.cuda.empty_cache()
print(json.dumps(res, indent=1))
