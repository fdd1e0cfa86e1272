"""Streaming front end for host-resident scans: pinned H2D copy, the fused device path and the D2H read of
the results run on three CUDA streams, so consecutive batches overlap (copy-in of batch i+1 and copy-out
of batch i-1 under the compute of batch i).  This is the deployment shape of ros_node.py's loop
(:96-126: take the latest cloud -> voxelize -> H2D -> model -> D2H of the map), batched.

No collective and no cross-GPU traffic: one ScanStream per GPU, scans sharded by the caller.
"""
import torch


class ScanStream:
    """``split`` > 1 cuts every submitted batch into that many sub-batches, each with its own H2D copy -> infer -> D2H
    chain: compute starts after the first sub-batch has landed instead of after the whole batch, which shortens the
    pipeline's ramp-up and tail when the host link, not the GPU, is the bottleneck (several GPUs fed by one host)."""

    def __init__(self, model, batch, n_points, depth=2, labels="int8", threshold=0.08, split=1, use_graph=True):
        """``use_graph``: the 19 launches of a (slot, row range) are captured into a CUDA graph the second time that
        range is submitted and replayed from then on (measured at batch 64: 2.25 ms per step against 2.30 ms for the
        direct enqueues); a re-bound workspace or changed weights drop the captured graphs."""
        self.model, self.depth, self.threshold = model, depth, threshold
        self.use_graph = bool(use_graph)
        self._graphs = {}
        self._graph_state = None
        self.split = max(1, min(int(split), batch))
        dev = model._device()
        self.dev = dev
        F = model.cfg.input_features
        ldt = {"int8": torch.int8, "float64": torch.float64}[labels]
        with torch.cuda.device(dev):
            self.s_in, self.s_cmp, self.s_out = (torch.cuda.Stream(dev) for _ in range(3))
            self.slots = []
            for _ in range(depth):
                self.slots.append(dict(
                    points=torch.empty((batch, n_points, F), dtype=torch.float32, device=dev),
                    height=torch.empty((batch, model.ny, model.nx), dtype=torch.float32, device=dev),
                    labels=torch.empty((batch, n_points), dtype=ldt, device=dev),
                    h_height=torch.empty((batch, model.ny, model.nx), dtype=torch.float32).pin_memory(),
                    h_labels=torch.empty((batch, n_points), dtype=ldt).pin_memory(),
                    ev_in=torch.cuda.Event(), ev_done=torch.cuda.Event(), ev_out=torch.cuda.Event(), used=False))
        self.n = 0
        model.reserve(batch, n_points)
        self.h2d_bytes = batch * n_points * F * 4
        self.d2h_bytes = batch * (model.ny * model.nx * 4 + n_points * (1 if ldt == torch.int8 else 8))

    def submit(self, host_points):
        """Enqueue one batch of pinned host scans [b, N, F] (b <= the stream's batch); returns the slot holding its
        results (rows [:b] of the slot's tensors)."""
        k = self.n % self.depth
        s = self.slots[k]
        b = host_points.shape[0]
        cur = torch.cuda.current_stream(self.dev)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.s_in.wait_event(ev)                       # ordered after the caller's earlier work
        if s["used"]:
            self.s_in.wait_event(s["ev_done"])         # the slot's previous scans have been consumed
            self.s_cmp.wait_event(s["ev_out"])         # the slot's previous results have been read out
        bounds = [(i * b) // self.split for i in range(self.split + 1)]
        for lo, hi in zip(bounds[:-1], bounds[1:]):
            if hi == lo:
                continue
            with torch.cuda.stream(self.s_in):
                s["points"][lo:hi].copy_(host_points[lo:hi], non_blocking=True)
                e_in = torch.cuda.Event()
                e_in.record()
            self.s_cmp.wait_event(e_in)
            self._infer(k, s, lo, hi)
            with torch.cuda.stream(self.s_cmp):
                s["ev_done"].record()                   # after the last sub-batch: the slot's scans have been consumed
                e_done = torch.cuda.Event()
                e_done.record()
            self.s_out.wait_event(e_done)
            with torch.cuda.stream(self.s_out):
                s["h_height"][lo:hi].copy_(s["height"][lo:hi], non_blocking=True)
                s["h_labels"][lo:hi].copy_(s["labels"][lo:hi], non_blocking=True)
                s["ev_out"].record()
        s["used"] = True
        s["rows"] = b
        self.n += 1
        return k

    def _infer(self, k, s, lo, hi):
        """The fused path over rows [lo, hi) of slot k on the compute stream: direct launches the first time a range is
        seen (binds the workspace, uploads weights), captured into a CUDA graph the second time, replayed afterwards."""
        m = self.model

        def direct():
            with torch.cuda.stream(self.s_cmp):
                m.infer(s["points"][lo:hi], threshold=self.threshold, out=(s["height"][lo:hi], s["labels"][lo:hi]))
        if not self.use_graph:
            return direct()
        state = (m._bound, m._weights_version)
        if state != self._graph_state or m._params_version() != m._weights_version:
            self._graphs.clear()                       # re-bound workspace / new weights: captured launches are stale
            self._graph_state = None
        key = (k, lo, hi)
        g = self._graphs.get(key)
        if g is None:
            direct()
            self._graphs[key] = "seen"
            self._graph_state = (m._bound, m._weights_version)
            return
        if g == "seen":
            try:
                torch.cuda.synchronize(self.dev)       # one stall per (slot, range): capture needs a quiet device
                g = torch.cuda.CUDAGraph()
                with torch.no_grad(), torch.cuda.graph(g, stream=self.s_cmp, capture_error_mode="thread_local"):
                    m.infer(s["points"][lo:hi], threshold=self.threshold, out=(s["height"][lo:hi], s["labels"][lo:hi]))
                self._graphs[key] = g
            except Exception:                          # capture refused (e.g. another capture in progress): stay direct
                self._graphs[key] = "direct"
                return direct()
        if g == "direct":
            return direct()
        with torch.cuda.stream(self.s_cmp):
            self._graphs[key].replay()

    def drain(self):
        """Make the caller's current stream wait for everything submitted so far."""
        cur = torch.cuda.current_stream(self.dev)
        for s in self.slots:
            if s["used"]:
                cur.wait_event(s["ev_out"])

    def result(self, slot):
        """Host tensors (height [B,ny,nx], labels [B,N]) of a submitted batch; blocks until they are there."""
        s = self.slots[slot]
        s["ev_out"].synchronize()
        b = s.get("rows", s["h_height"].shape[0])
        return s["h_height"][:b], s["h_labels"][:b]


class GraphedInfer:
    """The fused path (gnd_infer: 19 kernels) captured once into a CUDA graph and replayed per scan batch.

    For the batch-1 latency case (BASELINE.json configs[1]; ros_node.py's one-cloud-at-a-time loop) the launch
    overhead of 17 back-to-back enqueues is a third of the device time; a replay is one launch.  The graph is
    bound to fixed buffers: write the scan(s) into ``self.points`` (or pass them to ``__call__``, which copies
    on the current stream), results appear in ``self.height`` / ``self.labels``.
    """

    def __init__(self, model, batch, n_points, threshold=0.08, labels="int8", variable_n=False):
        """``variable_n``: the graph also takes a device-side per-scan row count, so clouds of ANY size <= n_points
        replay the same graph (``__call__`` takes [b, n, F] with n <= n_points and writes the counts)."""
        dev = model._device()
        self.model, self.dev = model, dev
        self.n_points = int(n_points)
        F = model.cfg.input_features
        ldt = {"int8": torch.int8, "float64": torch.float64}[labels]
        with torch.cuda.device(dev), torch.no_grad():
            self.points = torch.zeros((batch, n_points, F), dtype=torch.float32, device=dev)
            self.height = torch.empty((batch, model.ny, model.nx), dtype=torch.float32, device=dev)
            self.labels = torch.empty((batch, n_points), dtype=ldt, device=dev)
            self.counts = torch.full((batch,), n_points, dtype=torch.int32, device=dev) if variable_n else None
            side = torch.cuda.Stream(dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):             # binds the workspace / uploads weights outside the capture
                model.infer(self.points, threshold=threshold, counts=self.counts, out=(self.height, self.labels))
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                model.infer(self.points, threshold=threshold, counts=self.counts, out=(self.height, self.labels))
            self.launches = model.launches
            self._bound = model._bound
            self._wver = model._weights_version

    def __call__(self, points=None):
        if self.model._bound != self._bound:
            raise RuntimeError("the model's workspace was re-bound after the graph was captured; build a new GraphedInfer")
        if points is not None:
            if self.counts is not None:
                if points.dim() == 2:
                    points = points.unsqueeze(0)
                b, n = points.shape[0], points.shape[1]
                if b != self.points.shape[0] or n > self.n_points:
                    raise ValueError("expected [%d, <=%d, F] scans" % (self.points.shape[0], self.n_points))
                self.points[:, :n].copy_(points, non_blocking=True)
                self.counts.fill_(n)
            else:
                self.points.copy_(points.reshape(self.points.shape), non_blocking=True)
        self.graph.replay()
        if self.model._params_version() != self._wver:      # behind the replay: off the latency path
            raise RuntimeError("the model's weights changed after the graph was captured; build a new GraphedInfer")
        return self.height, self.labels


# ---------------------------------------------------------------------------------------- multi-GPU plumbing
# The path shards by scan with no exchange step: one process per GPU, rank r owns the scans
# shard_range(n, r, world) and a full weight replica.  The only collectives are the barrier and the MAX of the
# timed milliseconds in bench.py, and the optional gather of per-rank result counts below.

def shard_range(n_items, rank, world):
    """Contiguous balanced split: rank r gets items [lo, hi); sizes differ by at most one, earlier ranks larger."""
    if world < 1 or not (0 <= rank < world) or n_items < 0:
        raise ValueError("bad shard request: n=%d rank=%d world=%d" % (n_items, rank, world))
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def max_over_ranks(value, device=None):
    """MAX of a per-rank scalar (the timed milliseconds) over the default process group; identity without one."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_shard_sizes(n_local, device=None):
    """Every rank's item count (e.g. scans processed), as a list indexed by rank."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [int(n_local)]
    t = torch.tensor([int(n_local)], dtype=torch.int64, device=device if device is not None else "cpu")
    out = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [int(x.item()) for x in out]


def gather_rates(own_rate, device=None):
    """Every rank's measured rate (host link GB/s, or scans per millisecond end to end), as a list indexed by rank."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [float(own_rate)]
    t = torch.tensor([float(own_rate)], dtype=torch.float64, device=device if device is not None else "cpu")
    out = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [float(x.item()) for x in out]


def rate_shards(rates, total, lo, hi):
    """Split `total` scans over the ranks in proportion to their measured rates (largest remainder), each shard within
    [lo, hi]: the host links of a multi-GPU box are not equal (measured: 20.5 GB/s to four GPUs, 36 GB/s to the other
    four), and with equal shards the slowest link sets the step time of a host-fed run.  Deterministic: every rank
    computes the same list from the same gathered rates."""
    import numpy as np
    r = np.asarray(rates, dtype=np.float64)
    if len(r) < 1 or not np.all(np.isfinite(r)) or np.any(r <= 0) or lo * len(r) > total or hi * len(r) < total:
        raise ValueError("bad shard request: rates=%r total=%d bounds=[%d, %d]" % (list(rates), total, lo, hi))
    # water-filling: ranks whose proportional share falls outside [lo, hi] are pinned to the bound, the rest share what
    # is left in proportion to their rates, until no share violates a bound (at most len(r) rounds)
    want = np.zeros(len(r))
    free = np.ones(len(r), dtype=bool)
    left = float(total)
    while free.any():
        share = left * r / r[free].sum()
        under, over = free & (share < lo), free & (share > hi)
        if not under.any() and not over.any():
            want[free] = share[free]
            break
        # pin the side that is further out of bounds first; the other side is re-examined with the new remainder
        pin_lo = (lo - share[under]).sum() >= (share[over] - hi).sum()
        pin, bound = (under, lo) if pin_lo else (over, hi)
        want[pin] = bound
        free &= ~pin
        left -= bound * int(pin.sum())
    base = np.clip(np.floor(want + 1e-9).astype(int), lo, hi)
    rem = total - int(base.sum())
    order = np.argsort(-(want - base), kind="stable")
    # the rounding remainder goes out one scan at a time, largest fraction first, over the ranks that are not at a
    # bound; the request fails only when no rank can take (or give) another scan
    while rem != 0:
        step = 1 if rem > 0 else -1
        moved = False
        for k in (order if step > 0 else order[::-1]):
            if rem != 0 and lo <= base[k] + step <= hi:
                base[k] += step
                rem -= step
                moved = True
        if not moved:
            raise ValueError("cannot place %d scans within [%d, %d] per rank" % (total, lo, hi))
    return [int(x) for x in base]

