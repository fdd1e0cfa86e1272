"""BUILD-CONTAINER ONLY: run the unmodified reference next to the oracle.

    python -m oracle.ref_harness [--write]

1. imports the unmodified reference as the package ``gnd_net`` (verbatim copy under the git-ignored
   baseline/_ref/, stub modules for the ROS / plotting imports it cannot satisfy:
   SURVEY.md section 8c; baseline/refimpl.py),
2. checks every stage of the restatement (oracle/gnd_oracle.c, oracle/model_ref.py) against the
   reference's own numba / torch code on the section 8d input set -- bit-exact for voxels, coors,
   num_points and labels, float-equal for the model stages,
3. with --write, (re)generates the fixtures in tests/golden/: BatchNorm calibration of the
   synthetic checkpoint, digests for the full-size cases, full arrays for the small ones.

Nothing in tests/ -m gpu, smoke() or bench.py reads /root/reference; they read the fixtures.
"""
import argparse
import json
import os
import sys
import types

import numpy as np

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def import_reference():
    """The unmodified reference as the package ``gnd_net`` (baseline/refimpl.py: verbatim copy under the git-ignored
    baseline/_ref/, refreshed from /root/reference here)."""
    sys.dont_write_bytecode = True
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    from baseline import refimpl
    refimpl.install(force=True)
    return refimpl.load()


def ref_shift(points, h):
    """predict_ground.py:104-107 (numba body is a numpy broadcast add)."""
    v = points[:, :4] if points.shape[1] >= 4 else points
    add = np.zeros(v.shape[1], dtype=np.float32)
    add[2] = h
    v += add
    return points


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--write", action="store_true")
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--verify", action="store_true",
                    help="regenerate every fixture in memory and require equality with the committed tests/golden/ files")
    args = ap.parse_args()

    import torch
    sys.path.insert(0, ROOT)
    from oracle import cpu_ops, model_ref
    from gndnet_b200 import synth

    ops, refmodel, refutils = import_reference()
    torch.set_num_threads(8)
    report = {}

    # ------------------------------------------------------------------ inputs
    K = synth.make_cfg("kitti")
    C = synth.make_cfg("camera")
    real0 = np.load(os.path.join(REF, "data/training/seq_000/reduced_velo/000000.npy")).astype(np.float32)
    cases = []     # (name, cfg_name, cfg, raw points (sensor frame, pre-shift), max_voxels override)

    def add(name, cfgname, pts, **ov):
        cases.append((name, cfgname, pts, ov))

    add("hdl64_s0", "kitti", synth.hdl64_scan(0))
    add("hdl64_s1", "kitti", synth.hdl64_scan(1))
    add("uniform_s0", "kitti", synth.uniform_scan(0))
    add("real_000000", "kitti", real0)
    for i in range(1, 5):
        add("real_%06d" % i, "kitti",
            np.load(os.path.join(REF, "data/training/seq_000/reduced_velo/%06d.npy" % i)).astype(np.float32))
    add("boundary_kitti", "kitti", synth.boundary_scan(K, 0))
    add("single_cell", "kitti", synth.single_cell_scan(K, 0))
    add("overflow_64", "kitti", synth.hdl64_scan(2), max_voxels=64)
    add("overflow_1000", "kitti", synth.uniform_scan(3), max_voxels=1000)
    add("overflow_5000", "kitti", synth.uniform_scan(4), max_voxels=5000)
    add("maxpts_5", "kitti", synth.hdl64_scan(5), max_points_voxel=5)
    add("small_kitti", "kitti", synth.hdl64_scan(6)[::60].copy())
    add("empty", "kitti", np.zeros((0, 4), dtype=np.float32))
    add("all_outside", "kitti", synth.uniform_scan(7, n=1000, lo=(200, 200, 50), hi=(300, 300, 60)))
    add("one_point", "kitti", np.array([[1.5, 2.5, -1.0, 0.5]], dtype=np.float32))
    add("boundary_camera", "camera", synth.boundary_scan(C, 1))
    add("small_camera", "camera", synth.camera_scan(8, n=30000))
    if not args.quick:
        add("camera_1m", "camera", synth.camera_scan(0))

    # ------------------------------------------------------------------ voxelizer + segment parity
    vox_golden = {}
    small_arrays = {}
    for name, cfgname, raw, ov in cases:
        cfg = synth.make_cfg(cfgname, **ov)
        F_ = cfg.input_features
        pts_ref = raw[:, :F_].copy()
        pts_orc = raw[:, :F_].copy()
        if len(pts_ref):
            ref_shift(pts_ref, np.float32(cfg.lidar_height))
        cpu_ops.shift_cloud(pts_orc, cfg.lidar_height)
        assert pts_ref.tobytes() == pts_orc.tobytes(), name + ": shift mismatch"
        rv, rc, rn = ops.points_to_voxel(pts_ref, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel,
                                         True, cfg.max_voxels)
        ov_, oc, on, slot = cpu_ops.points_to_voxel(pts_orc, cfg.voxel_size, cfg.pc_range,
                                                    cfg.max_points_voxel, True, cfg.max_voxels,
                                                    return_point_slot=True)
        assert rv.shape == ov_.shape, (name, rv.shape, ov_.shape)
        assert rv.tobytes() == ov_.tobytes(), name + ": voxels differ"
        assert rc.tobytes() == oc.tobytes() and rc.dtype == oc.dtype, name + ": coors differ"
        assert rn.tobytes() == on.tobytes() and rn.dtype == on.dtype, name + ": num_points differ"
        # non-reversed index order (unused by callers, signature flag kept)
        if len(pts_ref) and name in ("small_kitti", "boundary_kitti"):
            a = ops.points_to_voxel(pts_ref, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, False, cfg.max_voxels)
            b = cpu_ops.points_to_voxel(pts_orc, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, False, cfg.max_voxels)
            assert all(x.tobytes() == y.tobytes() for x, y in zip(a, b)), name + ": non-reverse differs"
        # reconstruct voxels from point_slot (what the GPU tests do for the fused path)
        rec = np.zeros((len(on) * cfg.max_points_voxel, F_), dtype=np.float32)
        kept = slot >= 0
        rec[slot[kept]] = pts_orc[kept]
        assert rec.tobytes() == ov_.tobytes(), name + ": point_slot reconstruction differs"

        # segmentation against a deterministic height map (independent of the model)
        nz, ny, nx = model_ref.grid_shape(cfg)
        yy, xx = np.meshgrid(np.arange(ny, dtype=np.float32), np.arange(nx, dtype=np.float32), indexing="ij")
        pred = (0.02 * (xx - nx / 2) - 0.015 * (yy - ny / 2) + 0.3 * ((xx * 7 + yy * 13) % 5 - 2) / 5).astype(np.float32)
        seg_digest = {}
        for thr in (0.08, 0.16, 0.0, 0.2):
            if len(pts_ref) == 0:
                continue
            lr = refutils.segment_cloud(pts_ref.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0],
                                        elevation_map=pred.T, threshold=thr)
            lo_ = cpu_ops.segment_cloud(pts_orc.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0],
                                        elevation_map=pred.T, threshold=thr)
            assert lr.dtype == lo_.dtype == np.float64 and lr.tobytes() == lo_.tobytes(), \
                "%s: segment labels differ at thr=%s" % (name, thr)
            seg_digest[str(thr)] = dict(digest=synth.digest(lo_.astype(np.int8)),
                                        counts=[int((lo_ == v).sum()) for v in (-1, 0, 1)])
        if len(pts_ref):
            a_pts, a_obs = pts_ref[:, :3].copy(), pts_ref[:, :3].copy()
            b_pts, b_obs = pts_orc[:, :3].copy(), pts_orc[:, :3].copy()
            lr, obr = refutils.segment_cloud_noground(a_pts, a_obs, np.asarray(cfg.grid_range), cfg.voxel_size[0],
                                                      pred.T, 0.16)
            lo_, obo = cpu_ops.segment_cloud_noground(b_pts, b_obs, np.asarray(cfg.grid_range), cfg.voxel_size[0],
                                                      pred.T, 0.16)
            assert lr.tobytes() == lo_.tobytes() and obr.tobytes() == obo.tobytes(), name + ": noground differs"
            assert a_pts.tobytes() == b_pts.tobytes(), name + ": mutated points differ"
            seg_digest["noground_0.16"] = dict(n_obs=int(len(obo)), digest=synth.digest(obo))

        vox_golden[name] = dict(
            cfg=cfgname, overrides=ov, n=int(len(raw)), input_digest=synth.digest(raw),
            M=int(len(on)), kept=int(kept.sum()), full=int((on == cfg.max_points_voxel).sum()),
            voxels_digest=synth.digest(ov_), coors_digest=synth.digest(oc), num_points_digest=synth.digest(on),
            slot_digest=synth.digest(slot), segment=seg_digest)
        print("[voxel+segment ok] %-16s n=%7d M=%5d kept=%7d full=%4d" %
              (name, len(raw), len(on), kept.sum(), (on == cfg.max_points_voxel).sum()))
        if name.startswith("small_") or name.startswith("boundary_") or name in ("one_point",):
            small_arrays[name] = dict(points=raw, coors=oc, num_points=on, slot=slot)

    # ------------------------------------------------------------------ checkpoint calibration
    def voxel_inputs(cfg, raws):
        vs, cs, ns = [], [], []
        for b, raw in enumerate(raws):
            p = raw[:, :cfg.input_features].copy()
            cpu_ops.shift_cloud(p, cfg.lidar_height)
            v, c, n = cpu_ops.points_to_voxel(p, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
            vs.append(v)
            cs.append(np.concatenate([np.full((len(c), 1), b, dtype=np.float32), c.astype(np.float32)], axis=1))
            ns.append(n.astype(np.float32))
        return np.concatenate(vs), np.concatenate(cs), np.concatenate(ns)

    calib_src = {"kitti": [synth.hdl64_scan(1000), synth.hdl64_scan(1001)],
                 "camera": [synth.camera_scan(1000, n=200000)]}
    for cfgname in ("kitti", "camera"):
        cfg = synth.make_cfg(cfgname, batch_size=len(calib_src[cfgname]))
        path = synth.bn_stats_path(cfgname)
        if args.write or not os.path.exists(path):
            sd0 = synth.make_state_dict(cfg, 0, None)
            v, c, n = voxel_inputs(cfg, calib_src[cfgname])
            with torch.no_grad():
                vf = model_ref.pillar_feature_net(sd0, cfg, v, n, c)
                canvas = model_ref.pillars_scatter(cfg, vf, c, cfg.batch_size)
            from oracle.calibrate import calibrate
            stats = calibrate(cfg, canvas.numpy(), 0)
            if args.write:
                np.savez(path, **stats)
                print("[wrote]", path)

    # ------------------------------------------------------------------ model parity (reference modules, eval)
    model_golden = {}
    model_arrays = {}
    model_cases = [("kitti", ["hdl64_s0"], {}), ("kitti", ["hdl64_s1", "uniform_s0", "real_000000"], {}),
                   ("camera", ["small_camera"], {}), ("kitti", ["small_kitti"], dict(use_norm=True)),
                   # BASELINE configs[4]: max_voxels overflow (the loop's break, point_cloud_ops.py:44-45) through the model
                   ("kitti", ["overflow_1000"], dict(max_voxels=1000))]
    if not args.quick:   # BASELINE configs[3]: the dense 1 M-point camera frame, 50 -> 25 -> 12 pooling
        model_cases.append(("camera", ["camera_1m"], {}))
    raw_by_name = {n: r for n, _, r, _ in cases}
    for cfgname, names, ov in model_cases:
        cfg = synth.make_cfg(cfgname, batch_size=len(names), **ov)
        sd = synth.make_state_dict(cfg, 0, synth.load_bn_stats(cfgname))
        ref = refmodel.GroundEstimatorNet(cfg).cpu()
        ref.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}, strict=True)
        ref.eval()
        v, c, n = voxel_inputs(cfg, [raw_by_name[x] for x in names])
        tv, tc, tn = torch.from_numpy(v), torch.from_numpy(c), torch.from_numpy(n)
        with torch.no_grad():
            r_vf = ref.voxel_feature_extractor(tv, tn, tc)
            r_canvas = ref.middle_feature_extractor(r_vf, tc, cfg.batch_size)
            r_out = ref(tv, tc, tn)
        o_out, st = model_ref.forward(sd, cfg, v, c, n, return_stages=True)
        d_vf = float((r_vf - st["pillar_features"]).abs().max())
        d_cv = float((r_canvas - st["canvas"]).abs().max())
        d_out = float((r_out - o_out).abs().max())
        assert d_vf == 0.0 and d_cv == 0.0 and d_out <= 1e-6, (names, d_vf, d_cv, d_out)
        key = cfgname + ":" + "+".join(names) + ("+norm" if "use_norm" in ov else "")
        o = r_out.numpy().reshape(o_out.shape)          # the REFERENCE's heights are the golden (oracle == them to 1e-6)
        model_golden[key] = dict(cfg=cfgname, scans=names, overrides=ov, M=int(len(n)),
                                 height_mean=float(o.mean()), height_std=float(o.std()),
                                 height_min=float(o.min()), height_max=float(o.max()),
                                 stages={k: dict(mean=float(t.float().mean()), std=float(t.float().std()),
                                                 absmax=float(t.float().abs().max()))
                                         for k, t in st.items() if k not in ("i1", "i2")})
        model_arrays[key.replace(":", "__").replace("+", "_")] = o.astype(np.float32)
        # per-point labels of the whole reference call sequence (predict_ground.py:110-121, :165-168) on these scans
        lab_digest = []
        for b, x in enumerate(names):
            p_ = raw_by_name[x][:, :cfg.input_features].copy()
            ref_shift(p_, np.float32(cfg.lidar_height))
            hb = o.reshape(len(names), *o.shape[-2:])[b]
            lr = refutils.segment_cloud(p_.copy(), np.asarray(cfg.grid_range), cfg.voxel_size[0], elevation_map=hb.T, threshold=0.08)
            lab_digest.append(dict(digest=synth.digest(lr.astype(np.int8)), counts=[int((lr == q).sum()) for q in (-1, 0, 1)]))
        model_golden[key]["labels_thr0.08"] = lab_digest
        print("[model ok] %-40s M=%6d |d_pfn|=%g |d_canvas|=%g |d_out|=%g  height std=%.3f range=[%.3f,%.3f]" %
              (key, len(n), d_vf, d_cv, d_out, o.std(), o.min(), o.max()))
        for k, t in st.items():
            if k not in ("i1", "i2"):
                print("      %-16s absmax=%9.3f std=%8.3f" % (k, float(t.float().abs().max()), float(t.float().std())))

    if args.verify:
        with open(os.path.join(GOLDEN, "voxel_cases.json")) as f:
            assert json.load(f) == json.loads(json.dumps(vox_golden)), "voxel_cases.json is stale"
        with open(os.path.join(GOLDEN, "model_cases.json")) as f:
            assert json.load(f) == json.loads(json.dumps(model_golden)), "model_cases.json is stale"
        flat = {name + "." + k: a for name, d in small_arrays.items() for k, a in d.items()}
        for fname, arrays in (("small_cases.npz", flat), ("model_heights.npz", model_arrays), ("real_000000.npz", dict(points=real0))):
            with np.load(os.path.join(GOLDEN, fname)) as z:
                assert sorted(z.files) == sorted(arrays), fname + ": key set differs"
                for k in z.files:
                    assert z[k].dtype == np.asarray(arrays[k]).dtype and z[k].tobytes() == np.asarray(arrays[k]).tobytes(), \
                        "%s[%s] is stale" % (fname, k)
        for cfgname in ("kitti", "camera"):
            cfg = synth.make_cfg(cfgname, batch_size=len(calib_src[cfgname]))
            sd0 = synth.make_state_dict(cfg, 0, None)
            v, c, n = voxel_inputs(cfg, calib_src[cfgname])
            with torch.no_grad():
                vf = model_ref.pillar_feature_net(sd0, cfg, v, n, c)
                canvas = model_ref.pillars_scatter(cfg, vf, c, cfg.batch_size)
            from oracle.calibrate import calibrate
            stats = calibrate(cfg, canvas.numpy(), 0)
            with np.load(synth.bn_stats_path(cfgname)) as z:
                for k in z.files:
                    assert np.asarray(stats[k]).astype(z[k].dtype).tobytes() == z[k].tobytes(), "ckpt_%s_bn.npz[%s] is stale" % (cfgname, k)
        print("[verify] the committed tests/golden/ fixtures are reproduced array for array")
    if args.write:
        os.makedirs(GOLDEN, exist_ok=True)
        with open(os.path.join(GOLDEN, "voxel_cases.json"), "w") as f:
            json.dump(vox_golden, f, indent=1, sort_keys=True)
        with open(os.path.join(GOLDEN, "model_cases.json"), "w") as f:
            json.dump(model_golden, f, indent=1, sort_keys=True)
        flat = {}
        for name, d in small_arrays.items():
            for k, a in d.items():
                flat[name + "." + k] = a
        np.savez_compressed(os.path.join(GOLDEN, "small_cases.npz"), **flat)
        np.savez_compressed(os.path.join(GOLDEN, "model_heights.npz"), **model_arrays)
        np.savez_compressed(os.path.join(GOLDEN, "real_000000.npz"), points=real0)
        print("[wrote] fixtures in", GOLDEN)
    print("reference parity of the oracle: OK")


if __name__ == "__main__":
    main()
