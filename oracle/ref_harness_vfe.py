"""BUILD-CONTAINER ONLY: pins the multi-layer PillarFeatureNet of the oracle (modules/pointpillars.py:57-65, :96-107:
non-last PFNLayers have out // 2 units and hand concat(x, max over the pillar's slots) to the next layer, padded slots
included) against the unmodified reference model built with vfe_filters != [64], and writes
tests/golden/model_heights_vfe.npz.

    python -m oracle.ref_harness_vfe [--write]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

# (config name, case, vfe_filters, use_norm)
VFE_CASES = [("kitti", "small_kitti", [32, 64], False), ("kitti", "small_kitti", [64, 64], True),
             ("camera", "small_camera", [16, 32, 64], False)]


def case_key(cfgname, name, filters, use_norm):
    return "%s__%s__vfe_%s%s" % (cfgname, name, "_".join(str(f) for f in filters), "_norm" if use_norm else "")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--write", action="store_true")
    args = ap.parse_args()
    sys.path.insert(0, ROOT)
    import torch
    from oracle import cpu_ops, model_ref
    from gndnet_b200 import synth
    from oracle.ref_harness import import_reference
    from tests import cases
    _, refmodel, _ = import_reference()
    torch.set_num_threads(8)
    out = {}
    for cfgname, name, filters, use_norm in VFE_CASES:
        cfg = synth.make_cfg(cfgname, batch_size=1, vfe_filters=filters, use_norm=use_norm)
        sd = synth.make_state_dict(cfg, 0, synth.load_bn_stats(cfgname))
        ref = refmodel.GroundEstimatorNet(cfg).cpu()
        ref.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}, strict=True)
        ref.eval()
        raw = cases.load_case(name)[1]
        p = raw[:, :cfg.input_features].copy()
        cpu_ops.shift_cloud(p, cfg.lidar_height)
        v, c, n = cpu_ops.points_to_voxel(p, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
        c4 = np.concatenate([np.zeros((len(c), 1), np.float32), c.astype(np.float32)], 1)
        with torch.no_grad():
            r_out = ref(torch.from_numpy(v), torch.from_numpy(c4), torch.from_numpy(n.astype(np.float32)))
            r_pfn = ref.voxel_feature_extractor(torch.from_numpy(v), torch.from_numpy(n.astype(np.float32)), torch.from_numpy(c4))
        o_out = model_ref.forward(sd, cfg, v, c4, n.astype(np.float32))
        o_pfn = model_ref.pillar_feature_net(sd, cfg, v, n.astype(np.float32), c4)
        d_pfn = float((r_pfn.reshape(o_pfn.shape) - o_pfn).abs().max())
        d_out = float((r_out.reshape(o_out.shape) - o_out).abs().max())
        key = case_key(cfgname, name, filters, use_norm)
        print("%-44s |d_pfn| = %.3g  |d_out| = %.3g  (pillars %d, |pfn| max %.3g)" % (key, d_pfn, d_out, len(n), float(o_pfn.abs().max())))
        assert d_pfn <= 1e-6 and d_out <= 1e-5, (d_pfn, d_out)
        out[key] = r_out.numpy().astype(np.float32).reshape(o_out.shape)
    if args.write:
        np.savez_compressed(os.path.join(GOLDEN, "model_heights_vfe.npz"), **out)
        print("wrote tests/golden/model_heights_vfe.npz")


if __name__ == "__main__":
    main()
