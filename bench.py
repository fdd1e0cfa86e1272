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
          height maps + labels inside the timed region; with several ranks the global batch is split in
          proportion to each rank's measured host link rate (the links of a multi-GPU host are not equal)
  roofline    the segnet convolution kernel (the only dense contraction): algorithmic FLOPs / the launch durations
              measured INSIDE a long back-to-back run (sustained clocks), against the sustained bf16 peak
  cpu_baseline    the reference's own CPU code (numba voxelizer + nn.Module forward + numba labelling) from the
                  git-ignored baseline/_ref copy when it imports, else the oracle port, on this box's host cores
Both timed regions are preceded by an un-timed soak of the same step (>= 0.5 s, `soak_s`), so that a short
--steps run is measured at the sustained, power-capped clocks of a long one.
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
SCAN_BYTES = N_POINTS * 16


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d["bf16_tflops_sustained"], src="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback (B200_PROFILING.md)")


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
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def stop(self, t_from=None, t_to=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.1)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, sm_all, mx, reasons, pw = [], [], None, set(), []
        for t, r in self.rows:
            try:
                v = float(r[0])
                mx = float(r[1])
            except Exception:
                continue
            sm_all.append(v)
            if t_from is not None and not (t_from <= t <= t_to + 0.03):
                continue
            sm.append(v)
            try:
                pw.append(float(r[2]))
            except Exception:
                pass
            for name, x in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if x.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            sm = sm_all[-3:]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw) if pw else None,
                "window": "samples taken during the soak + timed region of `value`"}


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


# ---------------------------------------------------------------------------------------------- the CPU path
class CpuPath:
    """One scan at a time like predict_ground.py:110-121, :165-168.  kind "reference": the unmodified reference from
    baseline/_ref (numba points_to_voxel, GroundEstimatorNet nn.Module in eval mode, numba segment_cloud);
    kind "port": the oracle restatement (C voxelizer / labelling, functional torch forward).  With `device` the
    nn.Module lives on that CUDA device (the reference's own GPU path); that needs kind "reference"."""

    def __init__(self, device=None):
        import torch
        from gndnet_b200 import synth
        self.torch = torch
        self.cfg = synth.make_cfg("kitti")
        self.sd, self.ck = synth.checkpoint("kitti")
        self.kind, self.why_port = "port", None
        self.device = device
        try:
            from baseline import refimpl
            self.ops, refmodel, self.utils = refimpl.load()
            m = refmodel.GroundEstimatorNet(self.cfg)
            m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in self.sd.items()}, strict=True)
            self.model = (m.to(device) if device is not None else m.cpu()).eval()
            self.kind = "reference"
        except Exception as ex:          # numba or the copy missing on this box: say so, fall back to the port
            if device is not None:
                raise
            self.why_port = repr(ex)[:160]
            from oracle import cpu_ops, model_ref
            self.cpu_ops, self.model_ref = cpu_ops, model_ref

    def describe(self):
        if self.kind == "reference":
            return ("unmodified reference (baseline/_ref copy): numba points_to_voxel + GroundEstimatorNet.eval() on the CPU "
                    "+ numba segment_cloud")
        return "oracle port (C voxelize + torch-CPU functional forward + C segment); reference not importable: %s" % self.why_port

    def voxelize(self, raw):
        cfg = self.cfg
        pts = raw.copy()
        if self.kind == "reference":
            self.utils.shift_cloud_func(pts, np.float32(cfg.lidar_height))                 # utils/utils.py:137-139
            v, c, n = self.ops.points_to_voxel(pts, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
        else:
            self.cpu_ops.shift_cloud(pts, cfg.lidar_height)
            v, c, n = self.cpu_ops.points_to_voxel(pts, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
        return pts, v, c, n

    def scan(self, raw):
        """-> (seconds, (voxelize, forward, segment) seconds, height [ny, nx], labels float64 [N])."""
        torch, cfg = self.torch, self.cfg
        t0 = time.perf_counter()
        pts, v, c, n = self.voxelize(raw)
        t1 = time.perf_counter()
        with torch.no_grad():
            if self.kind == "reference":
                tv = torch.from_numpy(v).float()
                tc = torch.nn.functional.pad(torch.from_numpy(c), (1, 0), "constant", 0).float()      # predict_ground.py:115-118
                tn = torch.from_numpy(n).float()
                h = self.model(tv, tc, tn).numpy()
            else:
                c4 = np.concatenate([np.zeros((len(c), 1), np.float32), c.astype(np.float32)], 1)
                h = self.model_ref.forward(self.sd, cfg, v, c4, n.astype(np.float32)).numpy()
        t2 = time.perf_counter()
        if self.kind == "reference":
            lab = self.utils.segment_cloud(pts.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0],
                                           elevation_map=h.T, threshold=0.08)
        else:
            lab = self.cpu_ops.segment_cloud(pts.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], h.T, 0.08)
        t3 = time.perf_counter()
        return t3 - t0, (t1 - t0, t2 - t1, t3 - t2), h, lab


def cpu_baseline(seconds, gpu_model=None, device=None, warm=2, pool=16):
    """The CPU path on a bounded sample: distinct synthetic scans (a pool of `pool`, cycled) until `seconds` of CPU
    work have been timed.  The same scans then go through the CUDA path (gpu_model) and the two results are compared:
    the parity figures of the bench line come from this leg."""
    import torch
    from gndnet_b200 import synth
    use_all_host_threads()
    path = CpuPath()
    scans = [synth.hdl64_scan(9000 + i) for i in range(pool)]
    for s in scans[:warm]:
        path.scan(s)                                  # numba JIT + torch warm-up, untimed
    times, parts, outs = [], np.zeros(3), {}
    while sum(times) < seconds:
        k = len(times) % pool
        t, p, h, lab = path.scan(scans[k])
        times.append(t)
        parts += np.array(p)
        outs[k] = (h, lab)
    n_scans = len(times)
    total = float(np.sum(times))
    res = {"value": n_scans / total, "unit": "scans/s", "cores": int(torch.get_num_threads()), "kind": path.kind,
           "what": path.describe(),
           "sample": "%d synthetic HDL-64 scans of %d points, batch 1, %.1f s of CPU work (voxelize %.0f%% / forward %.0f%% / "
                     "segment %.0f%%); checkpoint %s" % (n_scans, N_POINTS, total, 100 * parts[0] / total,
                                                         100 * parts[1] / total, 100 * parts[2] / total, path.ck)}
    parity = None
    if gpu_model is not None and outs:
        ks = sorted(outs)
        pts = torch.from_numpy(np.stack([scans[k] for k in ks])).to(device)
        with torch.no_grad():
            hg, lg = gpu_model.infer(pts, threshold=0.08, labels="int8")
        hg, lg = hg.cpu().numpy(), lg.cpu().numpy().astype(np.float64)
        hc = np.stack([outs[k][0] for k in ks])
        lc = np.stack([outs[k][1] for k in ks])
        d = np.abs(hg - hc)
        tol = 1e-3 if gpu_model.precision in ("fp32", "split") else 5e-3
        parity = {"against": "cpu_baseline (%s)" % path.kind, "scans": len(ks), "height_tolerance_m": tol,
                  "max_abs_dh_m": float(d.max()), "median_abs_dh_m": float(np.median(d)),
                  "frac_pixels_over_tolerance": float((d > tol).mean()),
                  "labels_equal_frac": float((lg == lc).mean()),
                  "note": "no mask: pixels over the tolerance sit around max-pool argmax near-ties (DESIGN.md section 2)"}
    return res, parity


def other_configs(dev, precision, B=64, steps=10, warm=4):
    """BASELINE.json configs[3] (dense 1 M-point camera frames) and configs[4] (kittiSem with max_voxels = 1000, the
    voxelizer's overflow break) at batch `B` on this GPU, inputs resident, the same measurement as tools/sweep.py in a
    few steps: scans/s and the stage split of gnd_infer.  Parity of both shapes is the GPU suite's business
    (tests/test_model_gpu.py: camera_1m and overflow_1000 / overflow_5000 through forward() and infer())."""
    import torch
    from gndnet_b200 import synth
    from gndnet_b200.model import GroundEstimatorNet
    out = {}
    for key, cfgname, ov, gen, npool in (("configs[3] camera 1M-point frames", "camera", {}, synth.camera_scan, 2),
                                         ("configs[4] kittiSem max_voxels=1000 (overflow)", "kitti", dict(max_voxels=1000), synth.hdl64_scan, 8)):
        try:
            cfg = synth.make_cfg(cfgname, **ov)
            sd, _ = synth.checkpoint(cfgname)
            pool = [gen(7000 + s_)[:, :cfg.input_features].copy() for s_ in range(npool)]
            model = GroundEstimatorNet(cfg, precision=precision)
            model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
            model = model.to(dev).eval()
            dpool = [torch.from_numpy(p_).to(dev) for p_ in pool]
            pts = [torch.stack([dpool[(j + s_) % npool] for s_ in range(B)]) for j in range(2)]
            N = pts[0].shape[1]
            h = torch.empty((B, model.ny, model.nx), dtype=torch.float32, device=dev)
            l = torch.empty((B, N), dtype=torch.int8, device=dev)
            with torch.no_grad():
                for i in range(warm):
                    model.infer(pts[i & 1], threshold=0.08, out=(h, l))
                torch.cuda.synchronize()
                model.conv_timing_begin(steps)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for i in range(steps):
                    model.infer(pts[i & 1], threshold=0.08, out=(h, l))
                e1.record()
                torch.cuda.synchronize()
                conv, seg, stg = model.conv_timing_end(stages=True)
            ms = e0.elapsed_time(e1) / steps
            sm = stg.mean(axis=0)
            out[key] = {"batch": B, "points_per_scan": int(N), "steps": steps, "ms_per_step": ms, "scans_per_s": B / ms * 1e3,
                        "pillars_per_scan": float(model.pillars(B)["voxel_count"].float().mean().item()),
                        "stage_ms": {"voxelize": float(sm[0]), "pfn_scatter": float(sm[1]), "segnet": float(sm[2]),
                                     "segnet_convs": float(conv.sum(axis=1).mean()), "segment": float(sm[3])},
                        "voxelize_algorithmic_gbs": B * N * 12.0 / 1e6 / float(sm[0]),
                        "l2": "two distinct batches rotated every step (%d MB each)" % (B * N * cfg.input_features * 4 // 1000000)}
            model.release()
            del model, pts, dpool, h, l
            torch.cuda.empty_cache()
        except Exception as ex:               # a side measurement must never take the bench line down
            out[key] = {"unavailable": repr(ex)[:200]}
    return out


def torch_cuda_baseline(device, gpu_model, resident, seconds=2.0):
    """The reference's GPU path as side baselines (north_star: "the reference's PyTorch CUDA forward on the same GPUs").
    (1) its deployment shape, batch 1 (ros_node.py:61-66, 96-126): CPU voxelization -> pageable H2D of the dense
    [M,100,4] voxels -> eager forward (cuDNN) -> D2H of the map -> CPU segment_cloud;  (2) the bar SURVEY.md 2.2 sets for
    the segnet: segnetGndEst.forward (modules/segnet.py:133-142) on [64,64,100,100] and the whole model() at batch 64
    (batch convention of training.py:138-147), cuDNN with allow_tf32 True and False, CUDA events."""
    import torch
    from gndnet_b200 import synth
    from gndnet_b200.point_cloud_ops import points_to_voxel
    path = CpuPath(device=device)             # raises if the reference cannot be imported: reported as unavailable
    cfg = path.cfg
    out = {"kind": path.kind, "what": "reference nn.Module (baseline/_ref copy), eager torch %s / cuDNN %s" %
           (torch.__version__, torch.backends.cudnn.version()), "checkpoint": path.ck}
    scans = [synth.hdl64_scan(9100 + i) for i in range(8)]

    def one(raw):
        pts, v, c, n = path.voxelize(raw)
        with torch.no_grad():
            tv = torch.from_numpy(v).float().to(device)
            tc = torch.nn.functional.pad(torch.from_numpy(c), (1, 0), "constant", 0).float().to(device)
            tn = torch.from_numpy(n).float().to(device)
            h = path.model(tv, tc, tn).cpu().numpy()
        path.utils.segment_cloud(pts.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], elevation_map=h.T, threshold=0.08)
        return tv, tc, tn

    for s_ in scans[:3]:
        tv, tc, tn = one(s_)
    torch.cuda.synchronize()
    t0, k = time.perf_counter(), 0
    while time.perf_counter() - t0 < seconds:
        one(scans[k % 8]); k += 1
    torch.cuda.synchronize()
    out["deployment_shape_batch1_scans_s"] = k / (time.perf_counter() - t0)

    def ev_ms(fn, reps, warm=2):
        with torch.no_grad():
            for _ in range(warm):
                fn()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(reps):
                fn()
            b.record()
            torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    out["forward_only_ms_batch1"] = ev_ms(lambda: path.model(tv, tc, tn), 50, 5)
    # Does the reference agree with ITSELF across devices?  The same module and weights on the CPU and on this GPU
    # (ros_node.py:207-216,227 deploys it with .cuda()), same scans, strict fp32 (allow_tf32 off) and torch's default
    # (cuDNN may use TF32): max-pool argmax near-ties flip under any change of summation order,
    # and this is how far the reference's own two paths are apart because of it.
    try:
        cpu_path = CpuPath()
        if cpu_path.kind == "reference":
            par_scans = [synth.hdl64_scan(9000 + i) for i in range(8)]
            vox = [path.voxelize(s_) for s_ in par_scans]
            with torch.no_grad():
                h_cpu = [cpu_path.model(torch.from_numpy(v).float(),
                                        torch.nn.functional.pad(torch.from_numpy(c), (1, 0), "constant", 0).float(),
                                        torch.from_numpy(n).float()).numpy() for _, v, c, n in vox]
            agree = {}
            for tf32 in (False, True):
                torch.backends.cudnn.allow_tf32 = tf32
                with torch.no_grad():
                    h_gpu = [path.model(torch.from_numpy(v).float().to(device),
                                        torch.nn.functional.pad(torch.from_numpy(c), (1, 0), "constant", 0).float().to(device),
                                        torch.from_numpy(n).float().to(device)).cpu().numpy() for _, v, c, n in vox]
                d = np.abs(np.stack(h_gpu) - np.stack(h_cpu))
                same = tot = 0
                for (pts, _, _, _), hc, hg in zip(vox, h_cpu, h_gpu):
                    a = path.utils.segment_cloud(pts.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], elevation_map=hc.T, threshold=0.08)
                    b = path.utils.segment_cloud(pts.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], elevation_map=hg.T, threshold=0.08)
                    same += int((a == b).sum()); tot += len(a)
                agree["allow_tf32=%s" % tf32] = {
                    "max_abs_dh_m": float(d.max()), "median_abs_dh_m": float(np.median(d)),
                    "frac_pixels_over_1e-3_m": float((d > 1e-3).mean()), "frac_pixels_over_5e-3_m": float((d > 5e-3).mean()),
                    "scans_with_a_pixel_over_5e-3_m": int(sum(float(x.max()) > 5e-3 for x in d)),
                    "labels_equal_frac": same / max(tot, 1)}
            torch.backends.cudnn.allow_tf32 = True
            out["reference_cuda_vs_reference_cpu"] = dict(agree, scans=len(par_scans), what=(
                "the unmodified reference module on this GPU against the same module on the CPU, batch 1, no mask"))
    except Exception as ex:
        torch.backends.cudnn.allow_tf32 = True
        out["reference_cuda_vs_reference_cpu"] = {"unavailable": repr(ex)[:200]}
    # batch 64: voxels / coors / num_points of the bench's own resident batch, made by the drop-in voxelizer (untimed)
    B = resident.shape[0]
    vs, cs, ns = [], [], []
    for b in range(B):
        pts = resident[b].clone()
        pts[:, 2] += np.float32(cfg.lidar_height)
        v, c, n = points_to_voxel(pts, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
        vs.append(v)
        cs.append(torch.nn.functional.pad(c, (1, 0), "constant", b).float())               # training.py:141-143
        ns.append(n.float())
    V, C, Nn = torch.cat(vs), torch.cat(cs), torch.cat(ns)
    del vs, cs, ns
    canvas = torch.randn(B, 64, 100, 100, device=device)
    old_bs = cfg.batch_size
    cfg.batch_size = B
    rows = {}
    try:
        for tf32 in (True, False):
            torch.backends.cudnn.allow_tf32 = tf32
            torch.backends.cuda.matmul.allow_tf32 = tf32
            seg_ms = ev_ms(lambda: path.model.encoder_decoder(canvas), 5)
            full_ms = ev_ms(lambda: path.model(V, C, Nn), 3, 1)
            rows["allow_tf32=%s" % tf32] = {
                "segnet_ms_batch%d" % B: seg_ms, "segnet_tflops": GFLOP_PER_SCAN * B / seg_ms,
                "model_forward_ms_batch%d" % B: full_ms, "model_forward_scans_s": B / (full_ms / 1e3)}
    finally:
        cfg.batch_size = old_bs
        torch.backends.cudnn.allow_tf32 = True
        torch.backends.cuda.matmul.allow_tf32 = False
    out["batch%d" % B] = rows
    # the same two quantities here (segnet stage alone; pillar-encode + segnet on the same dense tensors)
    hbuf = torch.empty((B, gpu_model.ny, gpu_model.nx), dtype=torch.float32, device=device)
    lbuf = torch.empty((B, N_POINTS), dtype=torch.int8, device=device)
    st = gpu_model.stage_times(resident, hbuf, lbuf, reps=3)
    gcfg_bs = gpu_model.cfg.batch_size
    gpu_model.cfg.batch_size = B
    try:
        ours_full = ev_ms(lambda: gpu_model(V, C, Nn), 5)
    finally:
        gpu_model.cfg.batch_size = gcfg_bs
    out["ours_same_shapes"] = {"segnet_ms_batch%d" % B: st["segnet"], "segnet_tflops": GFLOP_PER_SCAN * B / st["segnet"],
                               "model_forward_ms_batch%d" % B: ours_full, "precision": gpu_model.precision}
    best = min(r["segnet_ms_batch%d" % B] for r in rows.values())
    out["segnet_speedup_vs_best_cudnn"] = best / st["segnet"]
    out["model_forward_speedup_vs_best_cudnn"] = min(r["model_forward_ms_batch%d" % B] for r in rows.values()) / ours_full
    return out


def run_reference(args, rank):
    if rank != 0:
        return
    import torch
    use_all_host_threads()
    from gndnet_b200 import synth
    path = CpuPath()
    per = args.ref_scans_per_step
    cfgd = {"workload": "GndNet kittiSem inference, synthetic 120k-point HDL-64 scans, streaming batch %d per GPU" % args.batch,
            "points_per_scan": N_POINTS, "batch_per_gpu": args.batch,
            "sample_per_step": "%d scans of the batch (bounded sample), one at a time like predict_ground.py" % per}
    pool = [synth.hdl64_scan(9000 + i) for i in range(16)]
    k = 0
    for _ in range(max(args.warmup, 2)):              # includes the numba JIT of the two kernels
        path.scan(pool[k % 16]); k += 1
    t0 = time.perf_counter()
    for _ in range(args.steps * per):
        path.scan(pool[k % 16]); k += 1
    total = time.perf_counter() - t0
    v = args.steps * per / total
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "scans/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfgd,
            "cpu_baseline": {"value": v, "unit": "scans/s", "cores": int(torch.get_num_threads()), "kind": path.kind,
                             "what": path.describe(),
                             "sample": "%d steps x %d scans; checkpoint %s" % (args.steps, per, path.ck)},
            "e2e": {"value": v, "unit": "scans/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------- host link
def h2d_probe(pinned, dev_buf, seconds=0.25, d2h=None):
    """Pinned host -> device copy rate of THIS rank's link, GB/s (all ranks call it at the same time, so the figure is
    the rate under the same contention the e2e run sees).  ``d2h`` = (device tensor, pinned host tensor): the step's
    result read-back runs next to every copy on a second stream, i.e. the run's own traffic mix."""
    import torch
    s, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    nbytes = pinned.numel() * pinned.element_size()
    with torch.cuda.stream(s):
        dev_buf.copy_(pinned, non_blocking=True)
        s.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0, reps = time.perf_counter(), 0
        a.record()
        while time.perf_counter() - t0 < seconds:
            dev_buf.copy_(pinned, non_blocking=True)
            if d2h is not None:
                with torch.cuda.stream(s2):
                    d2h[1].copy_(d2h[0], non_blocking=True)
            reps += 1
            if reps % 4 == 0:
                s.synchronize()
        b.record()
        s.synchronize()
        s2.synchronize()
    return reps * nbytes / 1e6 / a.elapsed_time(b)


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
    ap.add_argument("--no-extras", action="store_true", help="only value / e2e (profiling runs)")
    ap.add_argument("--soak", type=float, default=0.6, help="seconds of un-timed soak before each timed region")
    ap.add_argument("--no-graph", action="store_true", help="resident steps and ScanStream as direct launches instead of CUDA-graph replays")
    ap.add_argument("--split", type=int, default=1, help="e2e: sub-batches per submitted batch (default 1: measured best at N = 1 and 2 -- 26.4k / 24.7k / 21.4k / 15.5k scans/s for 1 / 2 / 4 / 8 on one GPU: the extra enqueues cost more than the shorter ramp saves)")
    ap.add_argument("--equal-shards", action="store_true", help="e2e: equal shards instead of link-rate-proportional ones")
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
    args.steps = args.steps or 200
    args.warmup = 5 if args.warmup is None else max(3, args.warmup)

    import torch
    import torch.distributed as dist
    from gndnet_b200 import synth
    from gndnet_b200.model import GroundEstimatorNet
    from gndnet_b200.pipeline import ScanStream, GraphedInfer, shard_range, max_over_ranks, rate_shards, gather_rates
    assert torch.cuda.is_available(), "bench.py needs a CUDA device: there is no CPU fallback"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    cfg = synth.make_cfg("kitti")
    sd, ck = synth.checkpoint("kitti")

    def build_model(precision):
        m = GroundEstimatorNet(cfg, precision=precision)
        m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
        return m.to(dev).eval()

    model = build_model(args.precision)
    B = args.batch
    Bmax = 2 * B if world > 1 else B          # link-rate-proportional e2e shards may reach twice the nominal batch
    # two distinct batches per rank, rotated every step: 2 x 123 MB of inputs > 126 MB L2
    # the global batch of world * B scans is sharded by scan: rank r owns [lo, hi) (gndnet_b200.pipeline.shard_range)
    lo, hi = shard_range(world * B, rank, world)
    seeds = 100000 + np.concatenate([np.arange(lo, hi), world * B + np.arange(lo, hi)])
    host = make_scans(seeds)                                   # [2B, N, 4] distinct scans
    idx = [np.arange(Bmax) % (2 * B), (np.arange(Bmax) + B) % (2 * B)]
    pinned = [torch.from_numpy(host[i]).pin_memory() for i in idx]          # [Bmax, N, 4] each; rows [:B] are the nominal batches
    resident = [p[:B].to(dev) for p in pinned]
    height = torch.empty((B, model.ny, model.nx), dtype=torch.float32, device=dev)
    labels = torch.empty((B, N_POINTS), dtype=torch.int8, device=dev)
    model.reserve(Bmax, N_POINTS)

    # ---- host link: every rank measures its pinned H2D rate at the same time; e2e shards follow the rates
    probe_buf = torch.empty((B, N_POINTS, 4), dtype=torch.float32, device=dev)
    if world > 1:
        dist.barrier()
    my_rate = h2d_probe(pinned[0][:B], probe_buf)
    rates_h2d_only = gather_rates(my_rate, device=dev)
    # ... and next to the run's own device -> host traffic (labels + height maps of the same batch): the ceiling e2e is
    # quoted against, and the rates the shards follow
    d2h_src = torch.empty(B * (N_POINTS + model.ny * model.nx * 4), dtype=torch.uint8, device=dev)
    d2h_dst = torch.empty(d2h_src.numel(), dtype=torch.uint8).pin_memory()
    if world > 1:
        dist.barrier()
    my_rate = h2d_probe(pinned[0][:B], probe_buf, d2h=(d2h_src, d2h_dst))
    del probe_buf, d2h_src, d2h_dst
    rates = gather_rates(my_rate, device=dev)
    shards = [B] * world if (world == 1 or args.equal_shards) else rate_shards(rates, world * B, max(8, B // 2), Bmax)
    my_b = shards[rank]
    pipe = ScanStream(model, Bmax, N_POINTS, depth=2, labels="int8", threshold=0.08, split=max(1, args.split), use_graph=not args.no_graph)

    def step_direct(i):
        model.infer(resident[i & 1], threshold=0.08, out=(height, labels))

    # the resident step through the public API's graph front end (gndnet_b200.pipeline.GraphedInfer): the 19 launches of a
    # 64-scan step captured once per resident batch, one replay per step (ScanStream does the same for host-fed batches)
    gres = None
    if not args.no_graph:
        with torch.no_grad():
            gres = [GraphedInfer(model, B, N_POINTS, threshold=0.08, labels="int8") for _ in range(2)]
            for k in range(2):
                gres[k].points.copy_(resident[k])
        torch.cuda.synchronize()

    def step_resident(i):
        if gres is not None:
            gres[i & 1]()
        else:
            step_direct(i)

    def step_e2e(i):
        # host buffers in, host buffers out: pinned H2D of this step's scans and D2H of its height maps + labels,
        # double-buffered over three streams (gndnet_b200.pipeline.ScanStream)
        pipe.submit(pinned[i & 1][:my_b])

    def soak(fn, seconds, after=None):
        """Un-timed back-to-back steps for >= `seconds` (the device never idles for long; the host syncs every few
        steps only to look at the clock)."""
        t0, n = time.perf_counter(), 0
        while True:
            for _ in range(8):
                fn(n); n += 1
            if after:
                after()
            torch.cuda.synchronize()
            if time.perf_counter() - t0 >= seconds:
                break
        return time.perf_counter() - t0, n

    def timed(fn, steps, warmup, sample_clocks=False, after=None):
        with torch.no_grad():
            for i in range(warmup):
                fn(i)
            if after:
                after()
            torch.cuda.synchronize()
            sampler = ClockSampler(local) if sample_clocks else None
            if sampler:
                sampler.start()
            t_from = time.perf_counter()
            soak_s, _ = soak(fn, args.soak, after) if args.soak > 0 else (0.0, 0)
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(steps):
                fn(i)
            if after:
                after()
            e1.record()
            torch.cuda.synchronize()
            t_to = time.perf_counter()
            clocks = sampler.stop(t_from, t_to) if sampler else None
            ms_own = e0.elapsed_time(e1)
            ms = ms_own
            if world > 1:
                ms = max_over_ranks(ms, device=dev)
                dist.barrier()
        return ms, ms_own, clocks, soak_s

    ms, _, clocks, soak_v = timed(step_resident, args.steps, args.warmup, sample_clocks=(rank == 0))
    launches_per_step = model.launches
    # ---- e2e shards, second pass: the 0.25 s copy probe over-states what some links sustain next to the D2H traffic and
    # the other ranks' real copy pattern (8-GPU box: 36 GB/s probed, 29.7 GB/s in the run), so the shards are re-balanced
    # twice from each rank's OWN measured e2e rate over a few un-timed steps (no barrier inside: a rank's time is its link's)
    shard_history = [list(shards)]
    if world > 1 and not args.equal_shards:
        for _ in range(2):
            with torch.no_grad():
                for i in range(2):
                    step_e2e(i)
                pipe.drain()
                torch.cuda.synchronize()
                dist.barrier()
                c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                c0.record()
                for i in range(8):
                    step_e2e(i)
                pipe.drain()
                c1.record()
                torch.cuda.synchronize()
            shards = rate_shards(gather_rates(my_b * 8 / c0.elapsed_time(c1), device=dev), world * B, max(8, B // 2), Bmax)
            my_b = shards[rank]
            shard_history.append(list(shards))
    ms_e2e, ms_e2e_own, _, soak_e = timed(step_e2e, args.steps, args.warmup, after=pipe.drain)
    value = world * B * args.steps / (ms / 1e3)
    e2e = sum(shards) * args.steps / (ms_e2e / 1e3)
    own_gbs = my_b * SCAN_BYTES * args.steps / 1e6 / ms_e2e_own
    run_gbs = [own_gbs]
    if world > 1:
        t = torch.tensor([own_gbs], dtype=torch.float64, device=dev)
        allr = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allr, t)
        run_gbs = [float(x.item()) for x in allr]
    link_ceiling = sum(rates) * 1e9 / SCAN_BYTES

    # batch-1 latency (BASELINE.json configs[1]), the conv kernel's live timing, side baselines: rank 0
    extra = {}
    if rank == 0 and not args.no_extras:
        pk = peaks()
        with torch.no_grad():
            # ---- roofline: the 8 conv launches timed INSIDE a long back-to-back run (sustained clocks)
            soak(step_direct, max(args.soak, 0.5))
            n_t = 96
            model.conv_timing_begin(n_t)
            for i in range(n_t):
                step_direct(i)
            conv_ms_all, seg_ms_all, stage_ms_all = model.conv_timing_end(stages=True)
        conv_layers = conv_ms_all.mean(axis=0)
        conv_ms = float(conv_layers.sum())
        alg_gflop = (GFLOP_PER_SCAN - 0.0115) * B
        achieved = alg_gflop / conv_ms                                  # TFLOP/s (GFLOP / ms)
        issued_gflop = sum(g * (3.0 if (args.precision == "split" or (args.precision == "mixed" and i < 4)) else 1.0)
                           for i, g in enumerate(CONV_GFLOP)) * B if args.precision != "fp32" else alg_gflop
        traffic, traffic_src = conv_traffic()
        extra["roofline"] = {
            "bound": "tensor", "kernel": "conv3x3_f32_kernel" if args.precision == "fp32" else "conv3x3_halo2_kernel",
            "achieved": achieved, "peak": pk["tf_sust"], "unit": "TFLOP/s", "frac": achieved / pk["tf_sust"],
            "traffic": traffic, "traffic_unit": "bytes per launch (dram read + write)", "traffic_source": traffic_src,
            "peak_source": pk["src"] + ": SUSTAINED bf16 -- the launches are timed inside a long back-to-back run "
                           "(>= 0.5 s soak, then %d consecutive steps with CUDA events around every conv launch)" % n_t,
            "launch_ms_avg": conv_ms / 8.0, "launches_timed": int(8 * len(conv_ms_all)),
            "algorithmic_gflop_per_launch_avg": alg_gflop / 8.0,
            "conv_layers_ms": [float(x) for x in conv_layers],
            "conv_layers_tflops": [float(CONV_GFLOP[i] * B / conv_layers[i]) for i in range(8)],
            "segnet_ms": float(seg_ms_all.mean()), "segnet_minus_convs_ms": float(seg_ms_all.mean() - conv_ms),
            "frac_of_burst_peak": achieved / pk["tf_burst"],
            "issued_tflops": issued_gflop / conv_ms, "issued_frac": issued_gflop / conv_ms / pk["tf_sust"],
            "issued_note": "MMA work actually issued: the four encoder layers run three fp16 products per algorithmic "
                           "product (split-fp16, DESIGN.md 3)"}
        one = resident[0][:1].contiguous()
        h1 = torch.empty((1, model.ny, model.nx), dtype=torch.float32, device=dev)
        l1 = torch.empty((1, N_POINTS), dtype=torch.int8, device=dev)

        def lat_events(fn, n=1000, skip=30):
            out = []
            with torch.no_grad():
                for i in range(n + skip):
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                    fn()
                    b.record()
                    torch.cuda.synchronize()
                    if i >= skip:
                        out.append(a.elapsed_time(b))
            return float(np.median(out)), float(np.percentile(out, 95))

        extra["latency_ms_p50_batch1"], extra["latency_ms_p95_batch1"] = lat_events(
            lambda: model.infer(one, threshold=0.08, out=(h1, l1)))
        # the same scan through a CUDA-graph replay of the launches (gndnet_b200.pipeline.GraphedInfer), resident
        # input; with the 1.92 MB pinned H2D + D2H of the map and labels inside the timed region; and with clouds of a
        # different size every call (ros_node.py drops NaN rows per message) through ONE graph with a device-side count
        gi = GraphedInfer(model, 1, N_POINTS, threshold=0.08, labels="int8")
        gi.points.copy_(one)
        extra["latency_ms_p50_batch1_graph"], extra["latency_ms_p95_batch1_graph"] = lat_events(lambda: gi())
        hp = pinned[0][:1]
        hh = torch.empty((1, model.ny, model.nx), dtype=torch.float32).pin_memory()
        hl = torch.empty((1, N_POINTS), dtype=torch.int8).pin_memory()
        lat_h = []
        with torch.no_grad():
            for i in range(230):
                t0 = time.perf_counter()
                gi(hp)
                hh.copy_(gi.height, non_blocking=True)
                hl.copy_(gi.labels, non_blocking=True)
                torch.cuda.synchronize()
                if i >= 30:
                    lat_h.append(1e3 * (time.perf_counter() - t0))
        extra["latency_ms_p50_batch1_host_to_host"] = float(np.median(lat_h))
        extra["latency_ms_p95_batch1_host_to_host"] = float(np.percentile(lat_h, 95))
        gv = GraphedInfer(model, 1, N_POINTS, threshold=0.08, labels="int8", variable_n=True)
        sizes = [N_POINTS - 97 * (k % 50) - 1 for k in range(64)]
        state = {"k": 0}

        def var_call():
            n = sizes[state["k"] % 64]
            state["k"] += 1
            gv(one[:, :n])
        extra["latency_ms_p50_batch1_graph_variable_n"], extra["latency_ms_p95_batch1_graph_variable_n"] = lat_events(var_call, 500)
        extra["latency_note"] = ("graph_variable_n: every call has a different point count (<= 120 000); same captured graph, "
                                 "device-side copy of the rows + count inside the timed region, no re-bind")
        # restore the batch binding and time the stages of one step with events
        with torch.no_grad():
            model.infer(resident[0], threshold=0.08, out=(height, labels))
            torch.cuda.synchronize()
            # the same resident step as 19 direct launches per step (what the graph replay saves at batch 64)
            soak(step_direct, 0.5)
            ga, gb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ga.record()
            for i in range(args.steps):
                step_direct(i)
            gb.record()
            torch.cuda.synchronize()
            ms_d = ga.elapsed_time(gb) / args.steps
            extra["value_direct_launches"] = {"value": B / ms_d * 1e3, "unit": "scans/s", "ms_per_step": ms_d,
                                              "note": "inputs resident; model.infer() called per step (19 kernel launches) instead of "
                                                      "one CUDA-graph replay" if gres is not None else "same as value (--no-graph)"}
            soak(step_direct, 0.3)
            st = model.stage_times(resident[1], height, labels, reps=5)
        extra["stage_ms_stage_calls"] = st
        extra["stage_ms_stage_calls_note"] = ("the four stage entry points of the C ABI (gnd_voxelize = the ORDER-exact voxelizer that "
                                              "also fills point_list), 5 steps with a host sync between steps, after a 0.3 s soak")
        # the stages as infer() ran them inside the long back-to-back run above (set-exact voxelizer on this workload)
        sm = stage_ms_all[stage_ms_all[:, 0] >= 0].mean(axis=0)
        st = dict(st, voxelize=float(sm[0]), pfn_scatter=float(sm[1]), segnet=float(sm[2]), segment=float(sm[3]))
        extra["stage_ms"] = st
        extra["stage_ms_note"] = ("voxelize / pfn_scatter / segnet / segment: CUDA events between the stages of gnd_infer, averaged over "
                                  "%d consecutive steps of a long run (sustained clocks)" % len(stage_ms_all))
        # the HBM-bound stages: algorithmic bytes (SURVEY.md 8d / DESIGN.md 5) over the measured stage time
        pil = model.pillars(B)
        M = pil["voxel_count"].long()
        S = int(torch.gather(pil["point_offsets"].long(), 1, M.view(-1, 1)).sum().item())
        Msum = int(M.sum().item())
        alg = {"voxelize": B * N_POINTS * 16 + 16 * Msum + 4 * S,
               "pfn_scatter": 16 * S + 16 * Msum + B * 64 * model.ny * model.nx * 4,       # fp32 plane, or fp16 hi + lo planes
               "segment": B * (N_POINTS * 12 + model.ny * model.nx * 4 + N_POINTS)}
        extra["hbm_stages"] = {k: {"algorithmic_mb": alg[k] / 1e6, "ms": st[k], "achieved_gbs": alg[k] / 1e6 / st[k],
                                   "frac_of_measured_hbm": alg[k] / 1e6 / st[k] / pk["hbm"],
                                   "frac_of_nominal_8tbs": alg[k] / 1e6 / st[k] / 8000.0} for k in alg}
        extra["hbm_stages"]["pillars_per_scan"] = Msum / B
        extra["hbm_stages"]["kept_points_per_scan"] = S / B
        extra["hbm_stages"]["peak_gbs"] = pk["hbm"]
        if world == 1:
            # ---- the mode that meets the 1e-3 m bar (all eight convs split-fp16), same workload, same run
            if args.precision != "split":
                try:
                    m2 = build_model("split")
                    m2.reserve(B, N_POINTS)
                    with torch.no_grad():
                        soak(lambda i: m2.infer(resident[i & 1], threshold=0.08, out=(height, labels)), 0.5)
                        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        a.record()
                        for i in range(args.steps):
                            m2.infer(resident[i & 1], threshold=0.08, out=(height, labels))
                        b.record()
                        torch.cuda.synchronize()
                    ms2 = a.elapsed_time(b)
                    extra["precision_split"] = {"value": B * args.steps / (ms2 / 1e3), "unit": "scans/s", "ms_per_step": ms2 / args.steps,
                                                "dtype": "f16x3 (f32 accumulate) in all eight convs", "height_bar_m": 1e-3,
                                                "note": "inputs resident, same batches, same steps, after a 0.5 s soak"}
                    m2.release()
                    del m2
                except Exception as ex:
                    extra["precision_split"] = {"unavailable": repr(ex)[:200]}
            if world == 1 and B == 64:
                extra["configs"] = other_configs(dev, args.precision)
            if not args.no_cpu_baseline:
                extra["cpu_baseline"], extra["parity"] = cpu_baseline(args.cpu_seconds, gpu_model=model, device=dev)
                try:
                    extra["torch_cuda_baseline"] = torch_cuda_baseline(dev, model, resident[0])
                except Exception as ex:        # a side baseline must never take the bench line down
                    extra["torch_cuda_baseline"] = {"unavailable": repr(ex)[:300]}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "scans/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": {"fp32": "f32", "mixed": "f16x3/f16 (f32 accumulate)", "split": "f16x3 (f32 accumulate)"}[args.precision],
                "data": "synthetic", "soak_s": soak_v,
                "config": {"workload": "GndNet kittiSem inference, synthetic 120k-point HDL-64 scans, streaming batch %d per GPU" % B,
                           "points_per_scan": N_POINTS, "batch_per_gpu": B, "global_batch": B * world, "precision": args.precision,
                           "checkpoint": ck, "l2": "inputs larger than L2: two distinct %d MB batches rotated every step" % (B * N_POINTS * 16 // 1000000),
                           "parallelism": "scans sharded by rank, no collective",
                           "launch": ("one CUDA-graph replay per step (19 kernels each; GraphedInfer / ScanStream)" if gres is not None
                                      else "19 direct kernel launches per step")},
                "e2e": {"value": e2e, "unit": "scans/s", "h2d_bytes_per_step": int(sum(shards) * SCAN_BYTES // world),
                        "d2h_bytes_per_step": int(sum(shards) * (model.ny * model.nx * 4 + N_POINTS) // world),
                        "bytes_note": "per rank, average over ranks", "ms_per_step": ms_e2e / args.steps, "soak_s": soak_e,
                        "shards_scans_per_step": shards,
                        "shard_rule": "equal" if (world == 1 or args.equal_shards) else
                                      ("global batch of %d scans split in proportion to each rank's measured pinned H2D rate, then "
                                       "re-balanced twice from each rank's own e2e rate over 8 un-timed steps" % (world * B)),
                        "shard_history": shard_history,
                        "h2d_gbs_probe_per_rank": [round(r, 2) for r in rates],
                        "h2d_gbs_in_run_per_rank": [round(r, 2) for r in run_gbs],
                        "h2d_gbs_probe_h2d_only_per_rank": [round(r, 2) for r in rates_h2d_only],
                        "link_ceiling_scans_s": link_ceiling,
                        "frac_of_link_ceiling": e2e / link_ceiling,
                        "link_note": "ceiling = sum over ranks of the pinned host->device rate measured in this run with all "
                                     "ranks copying at once NEXT TO the step's device->host read-back (h2d_gbs_probe_per_rank; "
                                     "the h2d-only probe is printed beside it), / %d bytes per scan; no kernel can feed "
                                     "host-resident scans faster" % SCAN_BYTES},
                "gpu_launches": int(launches_per_step * args.steps), "gpu_launches_per_step": int(launches_per_step),
                "clocks": clocks}
        line.update(extra)
        # The driver's record keeps the sub-dicts config / roofline / cpu_baseline / e2e / clocks whole and every other
        # extra only by key name: mirror the numbers the metric and BASELINE's configs ask for into those sub-dicts (the
        # top-level keys stay where earlier rounds' readers expect them).
        lat = {k: v for k, v in extra.items() if k.startswith("latency_ms_")}
        if lat:
            line["e2e"]["latency"] = lat
        if isinstance(line.get("roofline"), dict):
            for k in ("hbm_stages", "stage_ms", "precision_split", "value_direct_launches", "configs"):
                if k in extra:
                    line["roofline"][k] = extra[k]
        if isinstance(line.get("cpu_baseline"), dict):
            for k in ("parity", "torch_cuda_baseline"):
                if k in extra:
                    line["cpu_baseline"][k] = extra[k]
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
