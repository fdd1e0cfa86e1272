"""torch-fp32 CPU restatement of the GndNet model forward (TEST INFRASTRUCTURE ONLY).

Functional (no nn.Module), driven by a ``state_dict`` with exactly the
reference's key names, so the same weights feed the oracle and the CUDA path.
All BatchNorm is evaluated in eval mode (running statistics), which is how the
deployed node runs the model (ros_node.py:227); predict_ground.py's missing
``.eval()`` is a documented reference quirk.

  pillar_feature_net   modules/pointpillars.py:115-146 + PFNLayer :52-65
  pillars_scatter      modules/pointpillars.py:168-199
  segnet               modules/segnet.py:133-142 (+ :11-44, :49-61, :81-92)
  forward              model.py:36-42
"""
import numpy as np
import torch
import torch.nn.functional as F

BN2D_EPS = 1e-5   # nn.BatchNorm2d default, modules/segnet.py:35-38
BN1D_EPS = 1e-3   # modules/pointpillars.py:42


_DEV = torch.device("cpu")     # the oracle runs on the CPU; bench.py's torch-CUDA side baseline sets "cuda"


class on_device:
    """with on_device("cuda"): the same eager torch ops on the GPU (cuDNN convolutions) -- the reference's
    PyTorch CUDA forward (ros_node.py:61-66, 211) as a timed side baseline, never a checker."""

    def __init__(self, dev):
        self.dev = torch.device(dev)

    def __enter__(self):
        global _DEV
        self.prev, _DEV = _DEV, self.dev

    def __exit__(self, *exc):
        global _DEV
        _DEV = self.prev


def _t(x):
    if isinstance(x, torch.Tensor):
        return x.detach().to(_DEV, torch.float32)
    return torch.from_numpy(np.ascontiguousarray(x)).to(_DEV, torch.float32)


def grid_shape(cfg):
    """model.py:26-28: (nz, ny, nx) from pc_range / voxel_size in float64 + round."""
    g = (np.asarray(cfg.pc_range[3:], dtype=np.float64) - np.asarray(cfg.pc_range[:3], dtype=np.float64)) \
        / np.asarray(cfg.voxel_size, dtype=np.float64)
    g = np.round(g).astype(np.int64)
    return int(g[2]), int(g[1]), int(g[0])


def pillar_feature_net(sd, cfg, voxels, num_points, coors):
    """voxels [M,P,F], num_points [M], coors [M,4]=(b,z,y,x) float -> [M,64]."""
    features = _t(voxels)
    num_voxels = _t(num_points)
    coors = _t(coors)
    vx, vy = cfg.voxel_size[0], cfg.voxel_size[1]
    x_offset = vx / 2 + cfg.pc_range[0]          # pointpillars.py:110-113
    y_offset = vy / 2 + cfg.pc_range[1]
    points_mean = features[:, :, :3].sum(dim=1, keepdim=True) / num_voxels.view(-1, 1, 1)   # :120
    f_cluster = features[:, :, :3] - points_mean                                             # :121
    f_center = torch.zeros_like(features[:, :, :2])
    f_center[:, :, 0] = features[:, :, 0] - (coors[:, 3].unsqueeze(1) * vx + x_offset)       # :125
    f_center[:, :, 1] = features[:, :, 1] - (coors[:, 2].unsqueeze(1) * vy + y_offset)       # :126
    ls = [features, f_cluster, f_center]
    if getattr(cfg, "with_distance", False):
        ls.append(torch.norm(features[:, :, :3], 2, 2, keepdim=True))                        # :130-132
    features = torch.cat(ls, dim=-1)
    P = features.shape[1]
    mask = (num_voxels.int().unsqueeze(1) > torch.arange(P, dtype=torch.int, device=_DEV).view(1, -1))    # :202-223
    features = features * mask.unsqueeze(-1).to(features.dtype)                              # :137-140
    n_layers = len(getattr(cfg, "vfe_filters", [64]))                                        # :96-107
    for i in range(n_layers):
        pre = "voxel_feature_extractor.pfn_layers.%d." % i
        w = _t(sd[pre + "linear.weight"])
        if getattr(cfg, "use_norm", False):
            x = F.linear(features, w)                                                        # bias=False
            x = F.batch_norm(x.permute(0, 2, 1).contiguous(), _t(sd[pre + "norm.running_mean"]),
                             _t(sd[pre + "norm.running_var"]), _t(sd[pre + "norm.weight"]),
                             _t(sd[pre + "norm.bias"]), False, 0.0, BN1D_EPS).permute(0, 2, 1).contiguous()
        else:
            x = F.linear(features, w, _t(sd[pre + "linear.bias"]))                           # :54
        x = F.relu(x)
        x_max = torch.max(x, dim=1, keepdim=True)[0]                                         # :58
        if i == n_layers - 1:
            return x_max.reshape(x_max.shape[0], -1)  # reference .squeeze() (M>=2) ; sane for M in {0,1}
        features = torch.cat([x, x_max.repeat(1, P, 1)], dim=2)                              # :60-65 (padded slots included)


def pillars_scatter(cfg, voxel_features, coors, batch_size):
    """-> [B,64,ny,nx] (pointpillars.py:168-199)."""
    nz, ny, nx = grid_shape(cfg)
    coors = _t(coors)
    C = voxel_features.shape[1]
    out = []
    for b in range(batch_size):
        canvas = torch.zeros(C, nx * ny, dtype=voxel_features.dtype, device=voxel_features.device)
        m = coors[:, 0] == b
        tc = coors[m, :]
        idx = (tc[:, 2] * nx + tc[:, 3]).long()
        canvas[:, idx] = voxel_features[m, :].t()
        out.append(canvas)
    return torch.stack(out, 0).view(batch_size, C, ny, nx)


def _cbr(sd, prefix, x, batch_stats=False):
    """conv2DBatchNormRelu (segnet.py:11-44): conv3x3 s1 p1 bias -> BN2d -> ReLU.  BN2d in eval mode (running
    statistics) unless batch_stats: then the module's train-mode arithmetic (per-channel mean / biased variance of
    this very batch), which is what predict_ground.py runs because it never calls .eval()."""
    x = F.conv2d(x, _t(sd[prefix + ".cbr_unit.0.weight"]), _t(sd[prefix + ".cbr_unit.0.bias"]),
                 stride=1, padding=1)
    if batch_stats:
        x = F.batch_norm(x, None, None, _t(sd[prefix + ".cbr_unit.1.weight"]), _t(sd[prefix + ".cbr_unit.1.bias"]),
                         True, 0.0, BN2D_EPS)
        return F.relu(x)
    x = F.batch_norm(x, _t(sd[prefix + ".cbr_unit.1.running_mean"]), _t(sd[prefix + ".cbr_unit.1.running_var"]),
                     _t(sd[prefix + ".cbr_unit.1.weight"]), _t(sd[prefix + ".cbr_unit.1.bias"]),
                     False, 0.0, BN2D_EPS)
    return F.relu(x)


def segnet(sd, x, return_stages=False, batch_stats=False):
    """segnetGndEst.forward (segnet.py:133-142).  x [B,64,H,W] -> [B,1,H,W]."""
    e = "encoder_decoder."
    st = {}
    d1a = _cbr(sd, e + "down1.conv1", x, batch_stats)
    d1b = _cbr(sd, e + "down1.conv2", d1a, batch_stats)
    shape1 = d1b.shape
    p1, i1 = F.max_pool2d(d1b, 2, 2, return_indices=True)          # :54-60
    d2a = _cbr(sd, e + "down2.conv1", p1, batch_stats)
    d2b = _cbr(sd, e + "down2.conv2", d2a, batch_stats)
    shape2 = d2b.shape
    p2, i2 = F.max_pool2d(d2b, 2, 2, return_indices=True)
    u2 = F.max_unpool2d(p2, i2, 2, 2, output_size=shape2)          # :84-89
    u2a = _cbr(sd, e + "up2.conv1", u2, batch_stats)
    u2b = _cbr(sd, e + "up2.conv2", u2a, batch_stats)
    u1 = F.max_unpool2d(u2b, i1, 2, 2, output_size=shape1)
    u1a = _cbr(sd, e + "up1.conv1", u1, batch_stats)
    u1b = _cbr(sd, e + "up1.conv2", u1a, batch_stats)
    out = F.conv2d(u1b, _t(sd[e + "regressor.weight"]), _t(sd[e + "regressor.bias"]), padding=1)   # :131,:140
    if return_stages:
        st.update(d1a=d1a, d1b=d1b, p1=p1, i1=i1, d2a=d2a, d2b=d2b, p2=p2, i2=i2, u2=u2,
                  u2a=u2a, u2b=u2b, u1=u1, u1a=u1a, u1b=u1b, out=out)
        return out, st
    return out


def forward(sd, cfg, voxels, coors, num_points, batch_size=None, return_stages=False, bn_batch_stats=False):
    """GroundEstimatorNet.forward (model.py:36-42) in eval mode."""
    B = int(cfg.batch_size if batch_size is None else batch_size)
    with torch.no_grad():
        vf = pillar_feature_net(sd, cfg, voxels, num_points, coors)
        canvas = pillars_scatter(cfg, vf, coors, B)
        if return_stages:
            out, st = segnet(sd, canvas, True, bn_batch_stats)
            st.update(pillar_features=vf, canvas=canvas)
            return torch.squeeze(out), st
        return torch.squeeze(segnet(sd, canvas, False, bn_batch_stats))
