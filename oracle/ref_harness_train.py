"""BUILD-CONTAINER ONLY: pins the train-mode-BatchNorm variant of the oracle against the
unmodified reference model left in training mode -- predict_ground.py's actual behaviour, it never calls .eval() --
and writes tests/golden/model_heights_train.npz.

    python -m oracle.ref_harness_train [--write]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


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
    for cfgname, names in (("kitti", ["hdl64_s0"]), ("kitti", ["hdl64_s1", "uniform_s0", "real_000000"]),
                           ("camera", ["small_camera"])):
        cfg = synth.make_cfg(cfgname, batch_size=len(names))
        sd = synth.make_state_dict(cfg, 0, synth.load_bn_stats(cfgname))
        ref = refmodel.GroundEstimatorNet(cfg).cpu()
        ref.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}, strict=True)
        ref.train()                                   # what predict_ground.py leaves the module in
        vs, cs, ns = [], [], []
        for b, nm in enumerate(names):
            raw = cases.load_case(nm)[1]
            p = raw[:, :cfg.input_features].copy()
            cpu_ops.shift_cloud(p, cfg.lidar_height)
            v, c, n = cpu_ops.points_to_voxel(p, cfg.voxel_size, cfg.pc_range, cfg.max_points_voxel, True, cfg.max_voxels)
            vs.append(v); ns.append(n.astype(np.float32))
            cs.append(np.concatenate([np.full((len(c), 1), b, np.float32), c.astype(np.float32)], 1))
        v, c, n = np.concatenate(vs), np.concatenate(cs), np.concatenate(ns)
        with torch.no_grad():
            r_out = ref(torch.from_numpy(v), torch.from_numpy(c), torch.from_numpy(n))
        o_out = model_ref.forward(sd, cfg, v, c, n, bn_batch_stats=True)
        o_eval = model_ref.forward(sd, cfg, v, c, n)
        d = float((r_out - o_out).abs().max())
        print("%-8s %-40s |reference(train) - oracle(batch stats)| = %.3g   (train vs eval output differ by %.3g)" %
              (cfgname, "+".join(names), d, float((o_out - o_eval).abs().max())))
        assert d <= 1e-5, d
        key = cfgname + "__" + "_".join(names)
        out[key] = o_out.numpy().astype(np.float32)
        # the training-mode side effect: running buffers of the eight BatchNorm2d after this ONE forward (momentum 0.1,
        # unbiased batch variance), straight from the reference module
        rsd = ref.state_dict()
        bn = ["down1.conv1", "down1.conv2", "down2.conv1", "down2.conv2", "up2.conv1", "up2.conv2", "up1.conv1", "up1.conv2"]
        rm = np.zeros((8, 256), np.float32); rv = np.zeros((8, 256), np.float32)
        for l, nm in enumerate(bn):
            m = rsd["encoder_decoder.%s.cbr_unit.1.running_mean" % nm].numpy()
            v2 = rsd["encoder_decoder.%s.cbr_unit.1.running_var" % nm].numpy()
            rm[l, :len(m)] = m; rv[l, :len(v2)] = v2
            k0 = "encoder_decoder.%s.cbr_unit.1.num_batches_tracked" % nm
            assert int(rsd[k0]) == int(np.asarray(sd.get(k0, 0))) + 1
        out[key + "__running_mean"] = rm
        out[key + "__running_var"] = rv
    if args.write:
        np.savez_compressed(os.path.join(GOLDEN, "model_heights_train.npz"), **out)
        print("wrote tests/golden/model_heights_train.npz")


if __name__ == "__main__":
    main()
