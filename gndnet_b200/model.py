"""Drop-in for the reference's ``model.py`` (GroundEstimatorNet) plus the fused device pipeline.

``GroundEstimatorNet(cfg)`` keeps the reference's surface (model.py:13-42):
  * an ``nn.Module`` whose ``state_dict()`` has exactly the reference's keys and shapes, so
    ``model.load_state_dict(checkpoint['state_dict'])`` (predict_ground.py:196, ros_node.py:223) works;
  * sub-module attribute names ``voxel_feature_extractor`` / ``middle_feature_extractor`` /
    ``encoder_decoder`` (model.py:19,31,34);
  * ``forward(voxels [M,P,F], coors [M,4] (b,z,y,x), num_points [M]) -> torch.squeeze([B,1,ny,nx])``,
    with ``cfg.batch_size`` read at every call (model.py:36-42).
The torch modules below are parameter containers only: the arithmetic runs in libgndnet_b200.so.
BatchNorm is evaluated with running statistics (eval mode, as the deployed node does,
ros_node.py:227); the kernels fold it into the conv / linear weights at load time.

``GroundEstimatorNet.infer(points)`` is the additive fused entry: raw scans in,
height maps and per-point labels out, no dense ``voxels`` tensor and no host round trip.
"""
import ctypes
import threading
import warnings

import numpy as np
import torch
import torch.nn as nn

from . import _lib


class _Empty(nn.Module):
    """torchplus.nn.Empty (torchplus/nn/modules/common.py:7-16): parameter-free placeholder."""

    def forward(self, *args):
        return args[0] if len(args) == 1 else args


class PFNLayer(nn.Module):
    """Parameter layout of modules/pointpillars.py:19-50."""

    def __init__(self, in_channels, out_channels, use_norm=True, last_layer=True):
        super().__init__()
        if not last_layer:
            out_channels = out_channels // 2            # :37-39: the other half is the pillar maximum, concatenated
        self.units = out_channels
        if use_norm:
            self.linear = nn.Linear(in_channels, out_channels, bias=False)
            self.norm = nn.BatchNorm1d(out_channels, eps=1e-3, momentum=0.01)
        else:
            self.linear = nn.Linear(in_channels, out_channels, bias=True)
            self.norm = _Empty()


class PillarFeatureNet(nn.Module):
    """Parameter layout of modules/pointpillars.py:68-113."""

    def __init__(self, num_input_features=4, use_norm=True, num_filters=(64,), with_distance=False,
                 voxel_size=(0.2, 0.2, 4), pc_range=(0, -40, -3, 70.4, 40, 1)):
        super().__init__()
        num_filters = [int(f) for f in num_filters]
        if not (1 <= len(num_filters) <= 4) or num_filters[-1] != 64 or any(f < 2 or f > 128 or f % 2 for f in num_filters[:-1]):
            raise ValueError("vfe_filters: 1..4 layers, the last one 64 (PointPillarsScatter and the segnet take 64 "
                             "channels), the others even and <= 128; got %r" % (num_filters,))
        n_in = num_input_features + 5 + (1 if with_distance else 0)
        filters = [n_in] + num_filters                  # pointpillars.py:96-107
        self.pfn_layers = nn.ModuleList([PFNLayer(filters[i], filters[i + 1], use_norm, last_layer=(i == len(filters) - 2))
                                         for i in range(len(filters) - 1)])


class PointPillarsScatter(nn.Module):
    """modules/pointpillars.py:149-166: no parameters; the scatter is fused into the PFN kernel."""

    def __init__(self, output_shape, num_input_features=64):
        super().__init__()
        self.output_shape = output_shape
        self.ny, self.nx = output_shape[2], output_shape[3]
        self.nchannels = num_input_features


class _ConvBNReLU(nn.Module):
    """conv2DBatchNormRelu (modules/segnet.py:11-44)."""

    def __init__(self, cin, cout):
        super().__init__()
        self.cbr_unit = nn.Sequential(nn.Conv2d(cin, cout, 3, 1, 1, bias=True), nn.BatchNorm2d(cout),
                                      nn.ReLU(inplace=True))


class _Block2(nn.Module):
    def __init__(self, c1_in, c1_out, c2_out):
        super().__init__()
        self.conv1 = _ConvBNReLU(c1_in, c1_out)
        self.conv2 = _ConvBNReLU(c1_out, c2_out)


class SegnetGndEst(nn.Module):
    """Parameter layout of segnetGndEst (modules/segnet.py:118-131)."""

    def __init__(self, in_channels=64):
        super().__init__()
        self.down1 = _Block2(in_channels, 128, 128)     # segnetDown2(64, 128)
        self.down2 = _Block2(128, 256, 256)             # segnetDown2(128, 256)
        self.up2 = _Block2(256, 256, 128)               # segnetUp2(256, 128)
        self.up1 = _Block2(128, 128, 64)                # segnetUp2(128, 64)
        self.regressor = nn.Conv2d(64, 1, kernel_size=3, padding=1)


class GroundEstimatorNet(nn.Module):
    def __init__(self, cfg, precision=None, bn_mode="running", update_running_stats=True):
        """``bn_mode``: "running" (default) evaluates every BatchNorm2d with its running statistics, i.e. the module
        after ``.eval()`` as the deployed node runs it (ros_node.py:227); "batch" reproduces the module left in
        TRAINING mode -- statistics of the batch in flight -- which is what predict_ground.py computes because it
        never calls ``.eval()``.  "batch" needs ``precision="fp32"``.  Like the reference in training
        mode (modules/segnet.py:35-38: nn.BatchNorm2d, momentum 0.1), every forward / infer in "batch" mode then blends
        the batch mean and unbiased variance into ``running_mean`` / ``running_var`` and counts ``num_batches_tracked``,
        so ``state_dict()`` after k forwards equals the reference's; ``update_running_stats=False`` leaves the buffers
        alone."""
        super().__init__()
        self.cfg = cfg
        self.precision = precision or getattr(cfg, "precision", "mixed")
        if self.precision not in _lib.PRECISIONS:
            raise ValueError("precision must be one of %s" % sorted(_lib.PRECISIONS))
        if bn_mode not in ("running", "batch"):
            raise ValueError('bn_mode must be "running" or "batch"')
        if bn_mode == "batch" and self.precision != "fp32":
            raise ValueError('bn_mode="batch" needs precision="fp32"')
        self.bn_mode = bn_mode
        self.update_running_stats = bool(update_running_stats)
        self.voxel_feature_extractor = PillarFeatureNet(
            num_input_features=cfg.input_features, use_norm=cfg.use_norm, num_filters=cfg.vfe_filters,
            with_distance=cfg.with_distance, voxel_size=cfg.voxel_size, pc_range=cfg.pc_range)
        grid = (np.asarray(cfg.pc_range[3:]) - np.asarray(cfg.pc_range[:3])) / np.asarray(cfg.voxel_size)
        grid = np.round(grid).astype(np.int64)                       # model.py:26-28
        dense_shape = [1] + grid[::-1].tolist() + [cfg.vfe_filters[-1]]
        self.middle_feature_extractor = PointPillarsScatter(output_shape=dense_shape,
                                                            num_input_features=cfg.vfe_filters[-1])
        self.encoder_decoder = SegnetGndEst(in_channels=64)
        self.nz, self.ny, self.nx = int(grid[2]), int(grid[1]), int(grid[0])
        self._ctx = None
        self._ctx_device = None
        self._weights_version = None
        self._weights_epoch = 0
        self._ws = None
        self._bound = None
        self._warned_train = False
        self.launches = 0
        self._last_n = 0
        self._slots = None
        # a gnd_ctx owns mutable device scratch: one stage call in flight per ctx (include/gndnet_b200.h, threading
        # contract).  The reference runs inference on a daemon thread next to its callback thread (ros_node.py:51-60);
        # calls from several threads on ONE model are serialised here.  ctypes releases the GIL inside the C calls.
        self._lock = threading.RLock()

    # ------------------------------------------------------------------ context management
    def _config_struct(self):
        cfg = self.cfg
        c = _lib.Config()
        for j in range(6):
            c.pc_range[j] = float(np.float32(cfg.pc_range[j]))
            c.pc_range_f64[j] = float(cfg.pc_range[j])
        for j in range(3):
            c.voxel_size[j] = float(np.float32(cfg.voxel_size[j]))
            c.voxel_size_f64[j] = float(cfg.voxel_size[j])
        gr = getattr(cfg, "grid_range", [cfg.pc_range[0], cfg.pc_range[1], cfg.pc_range[3], cfg.pc_range[4]])
        for j in range(4):
            c.grid_range[j] = float(gr[j])
        c.lidar_height = float(np.float32(getattr(cfg, "lidar_height", 0.0)))
        c.max_points_voxel = int(getattr(cfg, "max_points_voxel", 100))
        c.max_voxels = int(getattr(cfg, "max_voxels", self.nx * self.ny * self.nz))
        c.input_features = int(cfg.input_features)
        c.use_norm = 1 if cfg.use_norm else 0
        c.with_distance = 1 if cfg.with_distance else 0
        c.precision = _lib.PRECISIONS[self.precision]
        c.bn_batch_stats = 1 if self.bn_mode == "batch" else 0
        vf = [int(f) for f in cfg.vfe_filters]
        c.n_vfe = len(vf)
        for j, f in enumerate(vf):
            c.vfe_filters[j] = f
        return c

    def _device(self):
        dev = next(self.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("gndnet_b200.GroundEstimatorNet runs on CUDA only -- call .cuda() first "
                               "(there is no CPU path)")
        return dev

    def _params_version(self):
        # identity AND in-place version of every parameter / buffer, looked up afresh on every call through the owning
        # modules' dicts (the module tree is fixed, the tensor objects are not): catches load_state_dict (in place),
        # load_state_dict(assign=True), re-assigned Parameters and dtype / device moves.  ~60 dict lookups, a few
        # microseconds -- it sits on the batch-1 latency path (state_dict() here cost ~75 us per call).
        # NOT caught: writes that go through ``p.data`` (p.data.copy_(w), p.data.add_(g) -- the idiom of the reference's
        # torchplus/train/optim.py:110).  ``.data`` is a detached alias with a version counter of its own, so neither
        # id(p) nor p._version moves.  After such a write call invalidate_weights(); the epoch below then changes.
        if self._slots is None:
            # batch-statistics mode never reads the BatchNorm2d running buffers (and updates them after every call)
            skip = (lambda m, d: self.bn_mode == "batch" and isinstance(m, nn.BatchNorm2d) and d is m._buffers)
            self._slots = [(d, n) for m in self.modules() for d in (m._parameters, m._buffers) for n in d if not skip(m, d)]
        return (self._weights_epoch,) + tuple((id(t), t._version) for t in (d[n] for d, n in self._slots) if t is not None)

    def invalidate_weights(self):
        """Tell the model that parameters or buffers were written behind autograd's back (``p.data.copy_(...)``,
        ``p.data.add_(...)``, a raw pointer write): such writes do not bump ``Tensor._version``, so the automatic check
        cannot see them and the device copy of the folded weights would stay as it was.  The next forward() / infer()
        folds and uploads them again; ScanStream drops its captured graphs and a GraphedInfer refuses to replay."""
        self._weights_epoch += 1

    def _ensure_ctx(self, defer_version_check=False):
        """``defer_version_check``: the latency path (infer) compares the parameter versions AFTER its launches are
        enqueued -- the ~20 us walk then overlaps the device work instead of delaying the first kernel -- and re-runs
        in the rare case they changed (_weights_stale)."""
        dev = self._device()
        L = _lib.lib()
        if self._ctx is None or self._ctx_device != dev:
            self.release()
            handle = ctypes.c_void_p()
            c = self._config_struct()
            _lib.check(L.gnd_create(ctypes.byref(c), dev.index if dev.index is not None else torch.cuda.current_device(),
                                    ctypes.byref(handle)), "gnd_create")
            self._ctx, self._ctx_device = handle, dev
            self._weights_version = None
            self._bound = None
        if defer_version_check and self._weights_version is not None:
            return self._ctx
        ver = self._params_version()
        if ver != self._weights_version:
            self._upload_weights()
            self._weights_version = ver
        return self._ctx

    def _weights_stale(self):
        """True (after re-uploading) if a parameter changed since the last upload."""
        ver = self._params_version()
        if ver == self._weights_version:
            return False
        self._upload_weights()
        self._weights_version = ver
        return True

    def _upload_weights(self):
        """gnd_load_weights: the C side of load_state_dict (BatchNorm folding + repack)."""
        L = _lib.lib()
        sd = {k: v.detach().to("cpu", torch.float32).contiguous() for k, v in self.state_dict().items()
              if not k.endswith("num_batches_tracked")}
        arr = (_lib.Weight * len(sd))()
        keep = []
        for i, (k, v) in enumerate(sd.items()):
            name = k.encode()
            keep.append((name, v))
            arr[i].name = name
            arr[i].h_data = ctypes.cast(v.data_ptr(), ctypes.POINTER(ctypes.c_float))
            arr[i].numel = v.numel()
        with torch.cuda.device(self._ctx_device):
            torch.cuda.synchronize()
            _lib.check(L.gnd_load_weights(self._ctx, arr, len(sd)), "gnd_load_weights")

    def _ensure_workspace(self, batch, n_points):
        """The workspace is bound to a CAPACITY (batch, n_points): smaller batches and shorter clouds reuse it as it is
        (no synchronise, no reallocation, tensor maps and captured graphs stay valid); a larger request re-binds to the
        element-wise maximum.  ``reserve()`` binds ahead of time."""
        L = _lib.lib()
        if self._bound is not None and self._bound[0] >= batch and self._bound[1] >= n_points:
            return
        if self._bound is not None:
            batch, n_points = max(batch, self._bound[0]), max(n_points, self._bound[1])
        need = ctypes.c_size_t()
        _lib.check(L.gnd_workspace_bytes(self._ctx, batch, n_points, ctypes.byref(need)), "gnd_workspace_bytes")
        with torch.cuda.device(self._ctx_device):
            torch.cuda.synchronize()
            self._ws = None
            ws = torch.empty(need.value + 1024, dtype=torch.uint8, device=self._ctx_device)
            off = (-ws.data_ptr()) % 1024
            self._ws = ws
            _lib.check(L.gnd_bind_workspace(self._ctx, ctypes.c_void_p(ws.data_ptr() + off), need.value, batch, n_points),
                       "gnd_bind_workspace")
        self._bound = (batch, n_points)

    def reserve(self, batch, n_points):
        """Bind the workspace for up to ``batch`` scans of up to ``n_points`` points each, so that clouds of varying size
        (ros_node.py:263 drops NaN rows per message; predict_ground.py loads files of different lengths) never re-bind."""
        with self._lock:
            self._ensure_ctx()
            self._ensure_workspace(int(batch), int(n_points))

    def set_option(self, name, value):
        """gnd_set_option: "conv_path" (0 CTA-pair halo kernel, 1 single-CTA tap kernel), "conv_geometry" (0 auto, 1
        throughput, 2 small-batch), "pfn_path" (0 auto, 1 CUDA-core, 2 mma.sync), "vox_path" (0 auto, 1 order-exact, 2
        set-exact): include/gndnet_b200.h."""
        with self._lock:
            self._ensure_ctx()
            _lib.check(_lib.lib().gnd_set_option(self._ctx, name.encode(), int(value)), "gnd_set_option")

    def release(self):
        if self._ctx is not None:
            _lib.lib().gnd_destroy(self._ctx)
        self._ctx, self._ws, self._bound = None, None, None

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass

    def _check_mode(self):
        if self.training and self.bn_mode != "batch" and not self._warned_train:
            warnings.warn("gndnet_b200 evaluates BatchNorm with running statistics (eval mode); the module is in "
                          "training mode -- call .eval() (predict_ground.py omits it, ros_node.py:227 does not), or "
                          'construct with bn_mode="batch", precision="fp32" to reproduce the training-mode arithmetic')
            self._warned_train = True

    def _bn2d_modules(self):
        ed = self.encoder_decoder
        return [blk.cbr_unit[1] for blk in (ed.down1.conv1, ed.down1.conv2, ed.down2.conv1, ed.down2.conv2,
                                            ed.up2.conv1, ed.up2.conv2, ed.up1.conv1, ed.up1.conv2)]

    def _blend_running_stats(self):
        """nn.BatchNorm2d's training-mode side effect (momentum update with the batch mean / unbiased variance the
        kernels left in "bn_stats"), enqueued on the current stream behind the forward."""
        if self.bn_mode != "batch" or not self.update_running_stats:
            return
        stats = self.buffer("bn_stats", torch.float32, (8, 256, 2))
        with torch.no_grad():
            for l, bn in enumerate(self._bn2d_modules()):
                C, mom = bn.num_features, (0.1 if bn.momentum is None else bn.momentum)
                bn.running_mean.mul_(1.0 - mom).add_(stats[l, :C, 0], alpha=mom)
                bn.running_var.mul_(1.0 - mom).add_(stats[l, :C, 1], alpha=mom)
                bn.num_batches_tracked.add_(1)

    def buffer(self, name, dtype, shape=None):
        """Test / profiling hook: a workspace buffer as a tensor view (gnd_get_buffer)."""
        L = _lib.lib()
        p, n = ctypes.c_void_p(), ctypes.c_size_t()
        _lib.check(L.gnd_get_buffer(self._ctx, name.encode(), ctypes.byref(p), ctypes.byref(n)), "gnd_get_buffer")
        off = p.value - self._ws.data_ptr()
        t = self._ws[off:off + n.value].view(dtype)
        return t.view(shape) if shape is not None else t

    def last_launch_count(self):
        return int(_lib.lib().gnd_last_launch_count(self._ctx)) if self._ctx is not None else 0

    # ------------------------------------------------------------------ reference API
    def forward(self, voxels, coors, num_points):
        """model.py:36-42 on the drop-in tensors."""
        with self._lock:
            return self._forward_locked(voxels, coors, num_points)

    def _forward_locked(self, voxels, coors, num_points):
        self._check_mode()
        ctx = self._ensure_ctx()
        L = _lib.lib()
        B = int(self.cfg.batch_size)
        dev = self._ctx_device
        voxels = voxels.to(dev, torch.float32).contiguous()
        coors = coors.to(dev, torch.float32).contiguous()
        num_points = num_points.to(dev, torch.float32).contiguous()
        if voxels.dim() != 3 or voxels.shape[2] != self.cfg.input_features:
            raise ValueError("voxels must be [M, P, %d]" % self.cfg.input_features)
        if voxels.shape[1] != int(getattr(self.cfg, "max_points_voxel", voxels.shape[1])):
            raise ValueError("voxels.shape[1] must equal cfg.max_points_voxel")
        if coors.shape != (voxels.shape[0], 4) or num_points.shape != (voxels.shape[0],):
            raise ValueError("coors must be [M, 4] (b, z, y, x) and num_points [M]")
        self._ensure_workspace(B, self._bound[1] if self._bound else 0)
        with torch.cuda.device(dev):
            height = torch.empty((B, 1, self.ny, self.nx), dtype=torch.float32, device=dev)
            st = _lib.stream_ptr()
            _lib.check(L.gnd_pfn_scatter_dense(ctx, _lib.ptr(voxels), _lib.ptr(coors), _lib.ptr(num_points),
                                               voxels.shape[0], B, st), "gnd_pfn_scatter_dense")
            n1 = self.last_launch_count()
            _lib.check(L.gnd_segnet(ctx, B, _lib.ptr(height), st), "gnd_segnet")
            self.launches = n1 + self.last_launch_count()
            self._blend_running_stats()
        return torch.squeeze(height)

    # ------------------------------------------------------------------ fused path
    def infer(self, points, threshold=0.08, counts=None, labels="int8", out=None):
        """Raw scans -> (height [B, ny, nx] float32, labels [B, N]).

        ``points``: float32 CUDA tensor [B, N, F] (or [N, F]) in the sensor frame; the +lidar_height z-shift
        (predict_ground.py:112), voxelization, pillar encoding, scatter, segnet and the point labelling
        against ``cfg.grid_range`` (predict_ground.py:168) all run on the device, enqueued back to back.
        ``labels``: "int8", "float64" (the reference's dtype) or None.  ``out``: optional (height, labels)
        tensors to write into.  ``counts``: optional int32 CUDA tensor [B], the number of valid rows of each scan
        (rows beyond it are ignored and labelled -1); values are clamped to [0, N] on the device.
        """
        with self._lock:
            return self._infer_locked(points, threshold, counts, labels, out)

    def _infer_locked(self, points, threshold, counts, labels, out):
        self._check_mode()
        ctx = self._ensure_ctx(defer_version_check=True)
        L = _lib.lib()
        single = points.dim() == 2
        if single:
            points = points.unsqueeze(0)
        if not (points.is_cuda and points.dtype == torch.float32 and points.is_contiguous()
                and points.dim() == 3 and points.shape[2] == self.cfg.input_features):
            raise TypeError("points must be a contiguous float32 CUDA tensor [B, N, %d]" % self.cfg.input_features)
        if points.device != self._ctx_device:
            raise ValueError("points are on %s but the model is on %s" % (points.device, self._ctx_device))
        B, N, _ = points.shape
        if counts is not None:
            if not (torch.is_tensor(counts) and counts.dtype == torch.int32 and counts.is_contiguous()
                    and counts.device == self._ctx_device and tuple(counts.shape) == (B,)):
                raise TypeError("counts must be a contiguous int32 tensor of shape [%d] on %s" % (B, self._ctx_device))
        if out is not None:
            oh, ol = out
            if not (oh.is_cuda and oh.device == self._ctx_device and oh.dtype == torch.float32 and oh.is_contiguous()
                    and oh.numel() == B * self.ny * self.nx):
                raise TypeError("out[0] must be a contiguous float32 tensor of %d x %d x %d on %s" % (B, self.ny, self.nx, self._ctx_device))
            if ol is not None and not (ol.device == self._ctx_device and ol.dtype in (torch.int8, torch.float64)
                                       and ol.is_contiguous() and ol.numel() == B * N):
                raise TypeError("out[1] must be a contiguous int8 / float64 tensor of %d x %d on %s" % (B, N, self._ctx_device))
        self._ensure_workspace(B, N)
        self._last_n = N
        dev = self._ctx_device
        with torch.cuda.device(dev):
            if out is not None:
                height, lab = out
            else:
                height = torch.empty((B, self.ny, self.nx), dtype=torch.float32, device=dev)
                lab = None
                if labels == "int8":
                    lab = torch.empty((B, N), dtype=torch.int8, device=dev)
                elif labels == "float64":
                    lab = torch.empty((B, N), dtype=torch.float64, device=dev)
            l8 = lab if (lab is not None and lab.dtype == torch.int8) else None
            l64 = lab if (lab is not None and lab.dtype == torch.float64) else None
            for _ in range(2):
                _lib.check(L.gnd_infer(ctx, _lib.ptr(points), _lib.ptr(counts), B, N, float(threshold), _lib.ptr(height),
                                       _lib.ptr(l8), _lib.ptr(l64), _lib.stream_ptr()), "gnd_infer")
                if not self._weights_stale():         # checked behind the launches; a changed parameter re-runs once
                    break
            self.launches = self.last_launch_count()
            self._blend_running_stats()
        if single:
            return height[0], (lab[0] if lab is not None else None)
        return height, lab

    def stage_times(self, points, height, labels, threshold=0.08, reps=3):
        """Device milliseconds of each stage of one infer() over `points` (CUDA events around the stage entry
        points of the C ABI; the convolution launches are timed individually by gnd_segnet_profile)."""
        L = _lib.lib()
        ctx = self._ensure_ctx()
        B, N, _ = points.shape
        self._ensure_workspace(B, N)
        self._last_n = N
        st = _lib.stream_ptr()
        l8 = labels if labels.dtype == torch.int8 else None
        l64 = labels if labels.dtype == torch.float64 else None
        acc = dict(voxelize=0.0, pfn_scatter=0.0, segnet=0.0, segnet_convs=0.0, segment=0.0)
        conv = np.zeros(8)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        for _ in range(reps):
            ev[0].record()
            _lib.check(L.gnd_voxelize(ctx, _lib.ptr(points), None, B, N, st), "gnd_voxelize")
            ev[1].record()
            _lib.check(L.gnd_pfn_scatter(ctx, _lib.ptr(points), B, N, st), "gnd_pfn_scatter")
            ev[2].record()
            cms = (ctypes.c_float * 8)()
            tot = ctypes.c_float()
            _lib.check(L.gnd_segnet_profile(ctx, B, _lib.ptr(height), cms, ctypes.byref(tot), st), "gnd_segnet_profile")
            ev[3].record()
            _lib.check(L.gnd_segment(ctx, _lib.ptr(points), None, B, N, _lib.ptr(height), float(threshold),
                                     _lib.ptr(l8), _lib.ptr(l64), st), "gnd_segment")
            end = torch.cuda.Event(enable_timing=True)
            end.record()
            torch.cuda.synchronize()
            acc["voxelize"] += ev[0].elapsed_time(ev[1])
            acc["pfn_scatter"] += ev[1].elapsed_time(ev[2])
            acc["segnet"] += tot.value
            acc["segnet_convs"] += float(sum(cms))
            acc["segment"] += ev[3].elapsed_time(end)
            conv += np.array(list(cms))
        out = {k: v / reps for k, v in acc.items()}
        out["conv_layers_ms"] = [float(x) / reps for x in conv]
        out["conv_kernel"] = "conv3x3_f32_kernel" if self.precision == "fp32" else "conv3x3_halo2_kernel"
        return out

    def conv_timing_begin(self, slots):
        """Arm gnd_conv_timing_begin: the next ``slots`` infer() / forward() calls record CUDA events around their eight
        convolution launches, asynchronously (bench.py times the kernel inside a long back-to-back run this way)."""
        with self._lock:
            self._ensure_ctx()
            _lib.check(_lib.lib().gnd_conv_timing_begin(self._ctx, int(slots)), "gnd_conv_timing_begin")
            self._tim_slots = int(slots)

    def conv_timing_end(self, stages=False):
        """-> (conv_ms [recorded, 8], segnet_ms [recorded]) float32 arrays; waits for the recorded events.  With
        ``stages`` also stage_ms [recorded, 4] (voxelize, pillar-encode + scatter, segnet, labels as infer() ran them)."""
        with self._lock:
            n = self._tim_slots
            conv = (ctypes.c_float * (8 * n))()
            seg = (ctypes.c_float * n)()
            stg = (ctypes.c_float * (4 * n))()
            rec = ctypes.c_int32()
            _lib.check(_lib.lib().gnd_conv_timing_end(self._ctx, conv, seg, stg, ctypes.byref(rec)), "gnd_conv_timing_end")
            r = rec.value
            out = (np.array(conv[:8 * r], dtype=np.float32).reshape(r, 8), np.array(seg[:r], dtype=np.float32))
            if stages:
                out = out + (np.array(stg[:4 * r], dtype=np.float32).reshape(r, 4),)
            return out

    def pillars(self, batch):
        """Voxelizer outputs of the last infer() as tensors (views of the workspace): coors, num_points, voxel_count,
        point_offsets, and the kept points grouped by CELL (cell_offsets [B, cells + 1], cell_points [B, N, 4]).
        ``point_list`` (per-pillar point indices in slot order) and ``point_offsets[b, :M]`` (their per-pillar CSR
        offsets) are filled only when the order-exact voxelizer ran (``set_option("vox_path", 1)``, or the stage /
        drop-in APIs): the fused path keeps sets, not slot order, and writes just ``point_offsets[b, M]`` -- the number
        of kept points of scan b, M = voxel_count[b].  Entries below M are then stale (or uninitialised) and must not be
        read; per-pillar sizes are ``num_points``, per-cell groups are ``cell_offsets`` / ``cell_points``."""
        mv = int(getattr(self.cfg, "max_voxels", self.nx * self.ny * self.nz))
        Bb = self._bound[0]
        N = self._last_n                     # row stride of the per-point buffers = the n_points of that call
        return dict(
            coors=self.buffer("coors", torch.int32, (Bb, mv, 3))[:batch],
            num_points=self.buffer("num_points", torch.int32, (Bb, mv))[:batch],
            voxel_count=self.buffer("voxel_count", torch.int32, (Bb,))[:batch],
            point_offsets=self.buffer("point_offsets", torch.int32, (Bb, mv + 1))[:batch],
            point_list=self.buffer("point_list", torch.int32)[:batch * max(N, 1)].view(batch, max(N, 1)),
            cell_offsets=self.buffer("cell_offsets", torch.int32, (Bb, self.nx * self.ny * self.nz + 1))[:batch],
            cell_points=self.buffer("cell_points", torch.float32)[:batch * max(N, 1) * 4].view(batch, max(N, 1), 4),
        )
