"""TEST INFRASTRUCTURE (build container): BatchNorm calibration of the seeded synthetic checkpoint.

Lives under oracle/ because it runs the torch restatement (oracle/model_ref.py); the product package
gndnet_b200/ never imports oracle/.  oracle/ref_harness.py --write calls it once and commits the result as
tests/golden/ckpt_<name>_bn.npz, which gndnet_b200.synth.make_state_dict() then loads everywhere."""
import numpy as np

from gndnet_b200.synth import _gauss, make_state_dict


def calibrate(cfg, canvases, seed=0):
    """Pick BatchNorm running stats so every block sees roughly unit-variance inputs, the way
    a trained network would: run the torch restatement block by block on calibration canvases
    [B,64,H,W] and record (perturbed) per-channel batch statistics.  Also picks the regressor
    scale that gives sigma(height) ~ 0.3 m.  Run once in the build container; the result is
    committed as tests/golden/ckpt_<name>_bn.npz so the checkpoint is reproducible anywhere."""
    import torch
    import torch.nn.functional as F
    from oracle import model_ref as mr
    rng = np.random.default_rng(77 + seed)
    stats = {}
    sd = make_state_dict(cfg, seed, None)
    x = torch.as_tensor(canvases, dtype=torch.float32)

    def block(name, x):
        p = "encoder_decoder." + name + ".cbr_unit."
        y = F.conv2d(x, mr._t(sd[p + "0.weight"]), mr._t(sd[p + "0.bias"]), padding=1)
        m = y.mean(dim=(0, 2, 3)).numpy().astype(np.float64)
        v = y.var(dim=(0, 2, 3), unbiased=False).numpy().astype(np.float64)
        c = len(m)
        mean = (m + 0.2 * np.sqrt(v) * _gauss(rng, c)).astype(np.float32)
        var = (v * (0.75 + 0.5 * rng.random(c)) + 1e-6).astype(np.float32)
        stats[name + ".mean"], stats[name + ".var"] = mean, var
        sd[p + "1.running_mean"], sd[p + "1.running_var"] = mean, var
        return mr._cbr(sd, "encoder_decoder." + name, x)

    with torch.no_grad():
        x = block("down1.conv1", x)
        x = block("down1.conv2", x)
        s1 = x.shape
        x, i1 = F.max_pool2d(x, 2, 2, return_indices=True)
        x = block("down2.conv1", x)
        x = block("down2.conv2", x)
        s2 = x.shape
        x, i2 = F.max_pool2d(x, 2, 2, return_indices=True)
        x = F.max_unpool2d(x, i2, 2, 2, output_size=s2)
        x = block("up2.conv1", x)
        x = block("up2.conv2", x)
        x = F.max_unpool2d(x, i1, 2, 2, output_size=s1)
        x = block("up1.conv1", x)
        x = block("up1.conv2", x)
        y = F.conv2d(x, mr._t(sd["encoder_decoder.regressor.weight"]), None, padding=1)
        stats["regressor.scale"] = np.float32(0.3 / float(y.std()))
    return stats
