#!/usr/bin/env python
"""bench.py -- scans/s end to end (120k-point HDL-64 scans) for the GndNet inference hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision fp32|mixed|split]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

A "step" is one pass of the whole hot path (z-shift -> voxelize -> pillar-encode -> scatter -> segnet ->
per-point labels) over one batch of `--batch` synthetic 120 000-point HDL-64 scans per GPU (BASELINE.json
configs[2], "streaming throughput, batch 64", sharded by scan with no collective: weak scaling).
  value   scans/s with the batch already resident in HBM (CUDA events, max over ranks)
  e2e     scans/s through the public API with HOST buffers: pinned H2D of the step's scans and D2H of its
          height maps + labels inside the timed region
  roofline    the segnet convolution kernel (the only dense contraction): algorithmic FLOPs / measured time
  cpu_baseline    the CPU oracle port of the same path on this box's host cores (bounded sample)
--impl reference times that CPU path as its own arm (rank 0 only).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_POINTS = 120000
GFLOP_PER_SCAN = 17.706          # SURVEY.md 8a layer table, kittiSem grid
CONV_GFLOP = [1.4746, 2.9491, 1.4746, 2.9491, 2.9491, 1.4746, 2.9491, 1.4746]
METRIC = "scans/s end-to-end (120k-pt scan)"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


def conv_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of the conv kernel, per launch, from the committed ncu --set full
    capture of this workload (profiles/conv_traffic.json, written by tools/make_profile_summary.py)."""
    p = os.path.join(ROOT, "profiles", "conv_traffic.json")
    if not os.path.exists(p):
        return None, None
    with open(p) as f:
        d = json.load(f)
    return d["bytes_per_launch"], d["source"]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def make_scans(seeds):
    from gndnet_b200 import synth
    return np.stack([synth.hdl64_scan(int(s)) for s in seeds]).astype(np.float32)


def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1; the CPU arm is meant to use every host core it can."""
    import torch
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        n = os.cpu_count() or 1
    if torch.get_num_threads() < n:
        torch.set_num_threads(n)
    return torch.get_num_threads()


def cpu_path_scan(sd, cfg, raw):
    """The reference's CPU path on the oracle port: numba loop -> C restatement, torch CPU forward, segment."""
    import torch
    from oracle import cpu_ops, model_ref
    t0 = time.perf_counter()
    pts = raw.copy()
    cpu_ops.shift_cloud(pts, cfg.lidar_height)
    v, c, n = cpu_ops.points_to_voxel(pts, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
    t1 = time.perf_counter()
    c4 = np.concatenate([np.zeros((len(c), 1), np.float32), c.astype(np.float32)], 1)
    with torch.no_grad():
        h = model_ref.forward(sd, cfg, v, c4, n.astype(np.float32)).numpy()
    t2 = time.perf_counter()
    cpu_ops.segment_cloud(pts.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], h.T, 0.08)
    t3 = time.perf_counter()
    return t3 - t0, (t1 - t0, t2 - t1, t3 - t2)


def cpu_baseline(seconds, warm=2, pool=16):
    """The CPU path on a bounded sample: distinct synthetic scans (a pool of `pool`, cycled) until `seconds`
    of CPU work have been timed."""
    import torch
    from gndnet_b200 import synth
    use_all_host_threads()
    cfg = synth.make_cfg("kitti")
    sd, ck = synth.checkpoint("kitti")
    scans = [synth.hdl64_scan(9000 + i) for i in range(pool)]
    for s in scans[:warm]:
        cpu_path_scan(sd, cfg, s)
    times, parts = [], np.zeros(3)
    while sum(times) < seconds:
        t, p = cpu_path_scan(sd, cfg, scans[len(times) % pool])
        times.append(t)
        parts += np.array(p)
    n_scans = len(times)
    total = float(np.sum(times))
    return {"value": n_scans / total, "unit": "scans/s", "cores": int(torch.get_num_threads()), "kind": "port",
            "sample": "%d synthetic HDL-64 scans of %d points, batch 1, %.1f s of CPU work (voxelize %.0f%% / torch-CPU "
                      "forward %.0f%% / segment %.0f%%); checkpoint %s" % (n_scans, N_POINTS, total, 100 * parts[0] / total,
                                                                           100 * parts[1] / total, 100 * parts[2] / total, ck)}, total


def torch_cuda_baseline(seconds=3.0):
    """The reference's GPU deployment shape (ros_node.py:61-66, 96-126) as a side baseline: CPU voxelization ->
    pageable H2D of the dense [M,100,4] voxels -> eager PyTorch CUDA forward (cuDNN, torch's default TF32 policy)
    -> D2H of the map -> CPU segment_cloud, one scan at a time; plus the forward alone on resident tensors."""
    import torch
    from oracle import cpu_ops, model_ref
    from gndnet_b200 import synth
    cfg = synth.make_cfg("kitti")
    sd, ck = synth.checkpoint("kitti")
    dev = torch.device("cuda", torch.cuda.current_device())
    with model_ref.on_device(dev):
        sd_dev = {k: model_ref._t(v) for k, v in sd.items()}
        scans = [synth.hdl64_scan(9100 + i) for i in range(8)]

        def one(raw, fwd_only=None):
            pts = raw.copy()
            cpu_ops.shift_cloud(pts, cfg.lidar_height)
            v, c, n = cpu_ops.points_to_voxel(pts, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
            c4 = np.concatenate([np.zeros((len(c), 1), np.float32), c.astype(np.float32)], 1)
            with torch.no_grad():
                vt, ct, nt = (torch.from_numpy(x).to(dev) for x in (v, c4, n.astype(np.float32)))
                if fwd_only is not None:
                    return vt, ct, nt
                h = model_ref.forward(sd_dev, cfg, vt, ct, nt).cpu().numpy()
            cpu_ops.segment_cloud(pts.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], h.T, 0.08)

        for s_ in scans[:3]:
            one(s_)
        torch.cuda.synchronize()
        t0, k = time.perf_counter(), 0
        while time.perf_counter() - t0 < seconds:
            one(scans[k % 8]); k += 1
        torch.cuda.synchronize()
        e2e = k / (time.perf_counter() - t0)
        vt, ct, nt = one(scans[0], fwd_only=True)
        with torch.no_grad():
            for _ in range(5):
                model_ref.forward(sd_dev, cfg, vt, ct, nt)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(50):
                model_ref.forward(sd_dev, cfg, vt, ct, nt)
            b.record()
            torch.cuda.synchronize()
        fwd_ms = a.elapsed_time(b) / 50
    return {"value": e2e, "unit": "scans/s", "kind": "port",
            "what": "ros_node.py shape, batch 1: C voxelizer (numba port) on the host -> pageable H2D of dense voxels -> eager "
                    "torch CUDA forward (cuDNN, allow_tf32=%s) -> D2H map -> C segment_cloud; %d scans in %.1f s" %
                    (torch.backends.cudnn.allow_tf32, k, seconds),
            "forward_only_ms_batch1": fwd_ms, "checkpoint": ck}


def run_reference(args, rank):
    if rank != 0:
        return
    cfgd = {"workload": "GndNet kittiSem inference, synthetic 120k-point HDL-64 scans, streaming batch %d per GPU" % args.batch,
            "points_per_scan": N_POINTS, "batch_per_gpu": args.batch,
            "sample_per_step": "%d scans of the batch (bounded sample), one at a time like predict_ground.py" % args.ref_scans_per_step}
    from gndnet_b200 import synth
    use_all_host_threads()
    cfg = synth.make_cfg("kitti")
    sd, ck = synth.checkpoint("kitti")
    per = args.ref_scans_per_step
    pool = [synth.hdl64_scan(9000 + i) for i in range(16)]
    k = 0
    for _ in range(args.warmup):
        cpu_path_scan(sd, cfg, pool[k % 16]); k += 1
    t0 = time.perf_counter()
    for _ in range(args.steps * per):
        cpu_path_scan(sd, cfg, pool[k % 16]); k += 1
    total = time.perf_counter() - t0
    import torch
    v = args.steps * per / total
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "scans/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfgd,
            "cpu_baseline": {"value": v, "unit": "scans/s", "cores": int(torch.get_num_threads()), "kind": "port",
                             "sample": "%d steps x %d scans (oracle port of the reference CPU path: C voxelize + torch-CPU "
                                       "forward + C segment); checkpoint %s" % (args.steps, args.ref_scans_per_step, ck)},
            "e2e": {"value": v, "unit": "scans/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("GNDNET_PRECISION", "mixed"))
    ap.add_argument("--batch", type=int, default=64, help="scans per step per GPU")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU work timed for cpu_baseline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ref-scans-per-step", type=int, default=8, help="--impl reference: scans timed per step (bounded sample of the batch)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        args.steps = args.steps or 20
        if args.steps * args.ref_scans_per_step > 4000:      # keep the arm within a few minutes at ~25 scans/s
            args.ref_scans_per_step = max(1, 4000 // args.steps)
        args.warmup = 1 if args.warmup is None else max(1, args.warmup)
        run_reference(args, rank)
        return
    args.steps = args.steps or 200            # ~0.5 s timed region: long enough for the power-capped clocks to settle
    args.warmup = 5 if args.warmup is None else max(3, args.warmup)

    import torch
    import torch.distributed as dist
    from gndnet_b200 import synth
    from gndnet_b200.model import GroundEstimatorNet
    assert torch.cuda.is_available(), "bench.py needs a CUDA device: there is no CPU fallback"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    cfg = synth.make_cfg("kitti")
    sd, ck = synth.checkpoint("kitti")
    model = GroundEstimatorNet(cfg, precision=args.precision)
    model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
    model = model.to(dev).eval()
    B = args.batch
    # two distinct batches per rank, rotated every step: 2 x 123 MB of inputs > 126 MB L2
    # the global batch of world * B scans is sharded by scan: rank r owns [lo, hi) (gndnet_b200.pipeline.shard_range)
    from gndnet_b200.pipeline import shard_range, max_over_ranks
    lo, hi = shard_range(world * B, rank, world)
    seeds = 100000 + np.concatenate([np.arange(lo, hi), world * B + np.arange(lo, hi)])
    host = make_scans(seeds).reshape(2, B, N_POINTS, 4)
    pinned = [torch.from_numpy(host[i]).pin_memory() for i in range(2)]
    resident = [p.to(dev) for p in pinned]
    height = torch.empty((B, model.ny, model.nx), dtype=torch.float32, device=dev)
    labels = torch.empty((B, N_POINTS), dtype=torch.int8, device=dev)
    from gndnet_b200.pipeline import ScanStream
    pipe = ScanStream(model, B, N_POINTS, depth=2, labels="int8", threshold=0.08)

    def step_resident(i):
        model.infer(resident[i & 1], threshold=0.08, out=(height, labels))

    def step_e2e(i):
        # host buffers in, host buffers out: pinned H2D of this step's scans and D2H of its height maps + labels,
        # double-buffered over three streams (gndnet_b200.pipeline.ScanStream)
        pipe.submit(pinned[i & 1])

    def timed(fn, steps, warmup, sample_clocks=False, after=None):
        with torch.no_grad():
            for i in range(warmup):
                fn(i)
            if after:
                after()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            sampler = ClockSampler(local) if sample_clocks else None
            if sampler:
                sampler.start()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(steps):
                fn(i)
            if after:
                after()
            e1.record()
            torch.cuda.synchronize()
            clocks = sampler.stop() if sampler else None
            ms = e0.elapsed_time(e1)
            if world > 1:
                ms = max_over_ranks(ms, device=dev)
                dist.barrier()
        return ms, clocks

    ms, clocks = timed(step_resident, args.steps, args.warmup, sample_clocks=(rank == 0))
    launches_per_step = model.launches
    ms_e2e, _ = timed(step_e2e, args.steps, args.warmup, after=pipe.drain)
    value = world * B * args.steps / (ms / 1e3)
    e2e = world * B * args.steps / (ms_e2e / 1e3)

    # batch-1 latency (BASELINE.json configs[1]) and the conv kernel's live timing, rank 0
    extra = {}
    if rank == 0:
        one = resident[0][:1].contiguous()
        h1 = torch.empty((1, model.ny, model.nx), dtype=torch.float32, device=dev)
        l1 = torch.empty((1, N_POINTS), dtype=torch.int8, device=dev)
        lat = []
        with torch.no_grad():
            for i in range(1030):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                model.infer(one, threshold=0.08, out=(h1, l1))
                b.record()
                torch.cuda.synchronize()
                if i >= 30:
                    lat.append(a.elapsed_time(b))
        extra["latency_ms_p50_batch1"] = float(np.median(lat))
        extra["latency_ms_p95_batch1"] = float(np.percentile(lat, 95))
        # the same scan through a CUDA-graph replay of the 17 launches (gndnet_b200.pipeline.GraphedInfer), resident
        # input, and with the 1.92 MB pinned H2D + D2H of the map and labels inside the timed region
        from gndnet_b200.pipeline import GraphedInfer
        gi = GraphedInfer(model, 1, N_POINTS, threshold=0.08, labels="int8")
        gi.points.copy_(one)
        hp = pinned[0][:1]
        hh = torch.empty((1, model.ny, model.nx), dtype=torch.float32).pin_memory()
        hl = torch.empty((1, N_POINTS), dtype=torch.int8).pin_memory()
        lat_g, lat_h = [], []
        with torch.no_grad():
            for i in range(1030):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                gi()
                b.record()
                torch.cuda.synchronize()
                if i >= 30:
                    lat_g.append(a.elapsed_time(b))
            for i in range(230):
                t0 = time.perf_counter()
                gi(hp)
                hh.copy_(gi.height, non_blocking=True)
                hl.copy_(gi.labels, non_blocking=True)
                torch.cuda.synchronize()
                if i >= 30:
                    lat_h.append(1e3 * (time.perf_counter() - t0))
        extra["latency_ms_p50_batch1_graph"] = float(np.median(lat_g))
        extra["latency_ms_p95_batch1_graph"] = float(np.percentile(lat_g, 95))
        extra["latency_ms_p50_batch1_host_to_host"] = float(np.median(lat_h))
        extra["latency_ms_p95_batch1_host_to_host"] = float(np.percentile(lat_h, 95))
        # restore the batch binding and time the stages of one step with events
        with torch.no_grad():
            model.infer(resident[0], threshold=0.08, out=(height, labels))
            torch.cuda.synchronize()
            st = model.stage_times(resident[1], height, labels, reps=3)
        extra["stage_ms"] = st
        pk = peaks()
        conv_ms = st["segnet_convs"]
        achieved = GFLOP_PER_SCAN * (1 - 0.0115 / GFLOP_PER_SCAN) * B / conv_ms     # TFLOP/s (GFLOP / ms)
        peak = pk["tf_sust"]
        issued = sum(g * (3.0 if (args.precision == "split" or (args.precision == "mixed" and i < 4)) else 1.0)
                     for i, g in enumerate(CONV_GFLOP)) * B / conv_ms if args.precision != "fp32" else achieved
        traffic, traffic_src = conv_traffic()
        extra["roofline"] = {"bound": "tensor", "kernel": st["conv_kernel"], "achieved": achieved, "peak": peak,
                             "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic, "traffic_unit": "bytes per launch (dram read + write)",
                             "traffic_source": traffic_src,
                             "peak_source": pk["src"] + " sustained bf16 (kernel timed inside a long step)",
                             "launch_ms_avg": conv_ms / 8.0,
                             "issued_tflops": issued, "issued_frac": issued / peak,
                             "issued_note": "MMA work actually issued: the four encoder layers run three fp16 products per "
                                            "algorithmic product (split-fp16, DESIGN.md 3)"}
        # the HBM-bound stages: algorithmic bytes (SURVEY.md 8d / DESIGN.md 5) over the measured stage time
        pil = model.pillars(B)
        M = pil["voxel_count"].long()
        S = int(torch.gather(pil["point_offsets"].long(), 1, M.view(-1, 1)).sum().item())
        Msum = int(M.sum().item())
        e_canvas = 4 if args.precision == "fp32" else 4          # fp32 plane, or fp16 hi + lo planes
        alg = {"voxelize": B * N_POINTS * 16 + 16 * Msum + 4 * S,
               "pfn_scatter": 16 * S + 16 * Msum + B * 64 * model.ny * model.nx * e_canvas,
               "segment": B * (N_POINTS * 12 + model.ny * model.nx * 4 + N_POINTS)}
        extra["hbm_stages"] = {k: {"algorithmic_mb": alg[k] / 1e6, "ms": st[k], "achieved_gbs": alg[k] / 1e6 / st[k],
                                   "frac_of_measured_hbm": alg[k] / 1e6 / st[k] / pk["hbm"],
                                   "frac_of_nominal_8tbs": alg[k] / 1e6 / st[k] / 8000.0} for k in alg}
        extra["hbm_stages"]["pillars_per_scan"] = Msum / B
        extra["hbm_stages"]["kept_points_per_scan"] = S / B
        extra["hbm_stages"]["peak_gbs"] = pk["hbm"]
        if not args.no_cpu_baseline and world == 1:
            extra["cpu_baseline"], _ = cpu_baseline(args.cpu_seconds)
            try:
                extra["torch_cuda_baseline"] = torch_cuda_baseline()
            except Exception as ex:        # a side baseline must never take the bench line down
                extra["torch_cuda_baseline"] = {"unavailable": repr(ex)[:200]}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "scans/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": {"fp32": "f32", "mixed": "f16x3/f16 (f32 accumulate)", "split": "f16x3 (f32 accumulate)"}[args.precision],
                "data": "synthetic",
                "config": {"workload": "GndNet kittiSem inference, synthetic 120k-point HDL-64 scans, streaming batch %d per GPU" % B,
                           "points_per_scan": N_POINTS, "batch_per_gpu": B, "global_batch": B * world, "precision": args.precision,
                           "checkpoint": ck, "l2": "inputs larger than L2: two distinct %d MB batches rotated every step" % (B * N_POINTS * 16 // 1000000),
                           "parallelism": "scans sharded by rank, no collective"},
                "e2e": {"value": e2e, "unit": "scans/s", "h2d_bytes_per_step": int(B * N_POINTS * 16),
                        "d2h_bytes_per_step": int(B * (model.ny * model.nx * 4 + N_POINTS)), "ms_per_step": ms_e2e / args.steps},
                "gpu_launches": int(launches_per_step * args.steps), "gpu_launches_per_step": int(launches_per_step),
                "clocks": clocks}
        line.update(extra)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
