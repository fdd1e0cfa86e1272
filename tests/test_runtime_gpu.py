"""GPU tests of the host-side contract of SURVEY.md 8b: the reference's threading model, concurrent contexts / streams,
clouds of varying size in one bound workspace, weight re-upload, argument validation."""
import threading

import numpy as np
import pytest
import torch

from gndnet_b200 import synth
from oracle import cpu_ops
from tests import helpers

pytestmark = pytest.mark.gpu


def _scan(seed, n=30000):
    return synth.hdl64_scan(seed)[:: 120000 // n].copy()


def test_daemon_thread_inference_with_concurrent_segmentation(cuda):
    """ros_node.py's threading model: a daemon thread loops cloud -> voxelize -> model -> latest map
    (ros_node.py:51-60, :96-126) while the callback thread takes the latest map, calls segment_cloud_noground and reads
    results back with .cpu() (:270-274).  Thread A runs model.infer on its own stream; thread B runs the ctx-free
    segmentation on another stream; both must reproduce the single-threaded bytes on every iteration."""
    from gndnet_b200.utils import segment_cloud_noground
    cfg = synth.make_cfg("kitti", batch_size=1)
    model, sd = helpers.load_model(cfg, "kitti", "mixed", cuda)
    scans = [torch.from_numpy(_scan(70 + k)).cuda() for k in range(4)]
    with torch.no_grad():
        expect = []
        for s in scans:
            h, l = model.infer(s, threshold=0.16, labels="int8")
            expect.append((h.clone(), l.clone()))
    torch.cuda.synchronize()
    # what thread B computes: labels + compacted obstacles against a FIXED published map (the "latest" one)
    pub_map = expect[0][0].clone()
    shifted = []
    for s in scans:
        p = s.cpu().numpy().copy()
        cpu_ops.shift_cloud(p, cfg.lidar_height)
        shifted.append(p)
    grid = np.asarray(cfg.grid_range)
    ref_b = []
    for p in shifted:
        lab, obs = cpu_ops.segment_cloud_noground(p[:, :3].copy(), p[:, :3].copy(), grid, cfg.voxel_size[0],
                                                  pub_map.cpu().numpy().T, 0.16)
        ref_b.append((lab, obs))
    errors = []
    iters = 40

    def inference_thread():
        try:
            st = torch.cuda.Stream()
            with torch.cuda.stream(st), torch.no_grad():
                for i in range(iters):
                    k = i % len(scans)
                    h, l = model.infer(scans[k], threshold=0.16, labels="int8")
                    st.synchronize()
                    if not (torch.equal(h, expect[k][0]) and torch.equal(l, expect[k][1])):
                        errors.append("inference thread: iteration %d differs" % i)
        except Exception as ex:      # noqa: BLE001
            errors.append("inference thread: %r" % (ex,))

    def callback_thread():
        try:
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                for i in range(iters):
                    k = (3 * i) % len(scans)
                    pts = torch.from_numpy(shifted[k][:, :3].copy()).cuda()
                    obs = pts.clone()
                    lab, out = segment_cloud_noground(pts, obs, grid, cfg.voxel_size[0], pub_map.T, 0.16)
                    lab_h, out_h = lab.cpu().numpy(), out.cpu().numpy()          # the .cpu() of ros_node.py:272
                    if not (np.array_equal(lab_h, ref_b[k][0]) and np.array_equal(out_h, ref_b[k][1])):
                        errors.append("callback thread: iteration %d differs" % i)
        except Exception as ex:      # noqa: BLE001
            errors.append("callback thread: %r" % (ex,))

    ta = threading.Thread(target=inference_thread, daemon=True)
    tb = threading.Thread(target=callback_thread, daemon=True)
    ta.start(); tb.start()
    ta.join(120); tb.join(120)
    assert not ta.is_alive() and not tb.is_alive(), "a thread hung"
    assert not errors, errors[:3]


def test_two_threads_share_one_model(cuda):
    """Two host threads calling infer() on the SAME model (one ctx): the wrapper serialises the calls (one stage call in
    flight per ctx), results stay byte-identical to the single-threaded ones."""
    cfg = synth.make_cfg("kitti", batch_size=1)
    model, sd = helpers.load_model(cfg, "kitti", "mixed", cuda)
    scans = [torch.from_numpy(_scan(80 + k)).cuda() for k in range(3)]
    with torch.no_grad():
        expect = [tuple(t.clone() for t in model.infer(s, threshold=0.08, labels="int8")) for s in scans]
    torch.cuda.synchronize()
    errors = []

    def worker(tid):
        try:
            with torch.no_grad():
                for i in range(30):
                    k = (i + tid) % 3
                    # same stream for both threads (the default one): calls on one ctx must also be stream-ordered
                    h, l = model.infer(scans[k], threshold=0.08, labels="int8")
                    h, l = h.clone(), l.clone()
                    torch.cuda.synchronize()
                    if not (torch.equal(h, expect[k][0]) and torch.equal(l, expect[k][1])):
                        errors.append("thread %d iteration %d differs" % (tid, i))
        except Exception as ex:      # noqa: BLE001
            errors.append("thread %d: %r" % (tid, ex))

    ts = [threading.Thread(target=worker, args=(t,), daemon=True) for t in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(120)
    assert not any(t.is_alive() for t in ts), "a thread hung"
    assert not errors, errors[:3]


def test_two_contexts_on_two_streams_run_concurrently(cuda):
    """Two models (two gnd_ctx, two workspaces) driven from two threads on two streams at batch 16, where every
    convolution launch fills all 74 CTA pairs: kernels of the two contexts interleave on the device.  DESIGN.md
    section 11 records that a co-resident variant of the conv kernel deadlocked in tcgen05.alloc; the shipped kernel
    owns its SM, so this must finish, and with the single-context bytes."""
    cfg = synth.make_cfg("kitti", batch_size=1)
    models = [helpers.load_model(cfg, "kitti", "mixed", cuda)[0] for _ in range(2)]
    batch = torch.from_numpy(np.stack([_scan(90 + k) for k in range(16)])).cuda()
    with torch.no_grad():
        expect = tuple(t.clone() for t in models[0].infer(batch, threshold=0.08, labels="int8"))
    torch.cuda.synchronize()
    errors = []
    barrier = threading.Barrier(2)

    def worker(m):
        try:
            st = torch.cuda.Stream()
            barrier.wait(30)
            with torch.cuda.stream(st), torch.no_grad():
                outs = [tuple(t.clone() for t in models[m].infer(batch, threshold=0.08, labels="int8")) for _ in range(25)]
                st.synchronize()
            for i, (h, l) in enumerate(outs):
                if not (torch.equal(h, expect[0]) and torch.equal(l, expect[1])):
                    errors.append("context %d iteration %d differs" % (m, i))
        except Exception as ex:      # noqa: BLE001
            errors.append("context %d: %r" % (m, ex))

    ts = [threading.Thread(target=worker, args=(m,), daemon=True) for m in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(180)
    assert not any(t.is_alive() for t in ts), "two contexts on two streams hung"
    assert not errors, errors[:3]


def test_variable_cloud_sizes_reuse_one_binding(cuda):
    """ros_node.py filters NaN rows per message (:263) and predict_ground.py loads files of different lengths: N changes
    from call to call.  The workspace is bound to a capacity; shorter clouds must neither re-bind nor change results."""
    cfg = synth.make_cfg("kitti", batch_size=1)
    model, sd = helpers.load_model(cfg, "kitti", "mixed", cuda)
    full = synth.hdl64_scan(33)
    sizes = [120000, 30000, 119999, 4, 77777, 120000, 1]
    with torch.no_grad():
        alone = {}
        for n in sorted(set(sizes)):
            fresh, _ = helpers.load_model(cfg, "kitti", "mixed", cuda)       # bound exactly to n
            h, l = fresh.infer(torch.from_numpy(full[:n].copy()).cuda(), threshold=0.08, labels="int8")
            alone[n] = (h.clone(), l.clone(), int(fresh.pillars(1)["voxel_count"][0]))
            fresh.release()
        model.reserve(1, 120000)
        ws_ptr = model._ws.data_ptr()
        for n in sizes:
            h, l = model.infer(torch.from_numpy(full[:n].copy()).cuda(), threshold=0.08, labels="int8")
            assert model._ws.data_ptr() == ws_ptr and model._bound == (1, 120000), "the workspace was re-bound"
            assert torch.equal(h, alone[n][0]) and torch.equal(l, alone[n][1]), "n=%d differs from a workspace bound to n" % n
            assert int(model.pillars(1)["voxel_count"][0]) == alone[n][2]
        # the same through ONE captured graph: capacity-sized input buffer + a device-side count
        from gndnet_b200.pipeline import GraphedInfer
        gi = GraphedInfer(model, 1, 120000, threshold=0.08, labels="int8", variable_n=True)
        for n in sizes:
            h, l = gi(torch.from_numpy(full[:n].copy()).cuda())
            assert torch.equal(h[0], alone[n][0]), "graph replay, n=%d" % n
            assert torch.equal(l[0, :n], alone[n][1]) and bool((l[0, n:] == -1).all())


def test_replaced_parameters_are_re_uploaded(cuda):
    """load_state_dict(assign=True) and re-assigned Parameters replace the tensor OBJECTS: the folded device weights
    must follow (a cached tensor list went stale)."""
    cfg = synth.make_cfg("kitti", batch_size=1)
    model, sd = helpers.load_model(cfg, "kitti", "mixed", cuda)
    pts = torch.from_numpy(_scan(12)).cuda()
    with torch.no_grad():
        h0, _ = model.infer(pts, labels=None)
        h0 = h0.clone()
        sd2 = {k: torch.from_numpy(np.asarray(v)).clone() for k, v in sd.items()}
        sd2["encoder_decoder.regressor.bias"] = sd2["encoder_decoder.regressor.bias"] + 0.25
        model.load_state_dict({k: v.cuda() for k, v in sd2.items()}, assign=True)
        h1, _ = model.infer(pts, labels=None)
        assert torch.allclose(h1, h0 + 0.25, atol=1e-6), "assign=True state_dict was not re-uploaded"
        model.encoder_decoder.regressor.bias = torch.nn.Parameter(model.encoder_decoder.regressor.bias.detach() + 0.5)
        h2, _ = model.infer(pts, labels=None)
        assert torch.allclose(h2, h0 + 0.75, atol=1e-6), "a re-assigned Parameter was not re-uploaded"
        model.encoder_decoder.regressor.bias.sub_(0.75)                     # in place: the _version path
        h3, _ = model.infer(pts, labels=None)
        assert torch.allclose(h3, h0, atol=1e-6)
        # writes through .data (torchplus/train/optim.py:110 in the reference) move neither id() nor _version: the
        # documented limitation, and invalidate_weights() is the way out
        bias = model.encoder_decoder.regressor.bias
        v = bias._version
        bias.data.add_(0.125)
        assert bias._version == v
        model.invalidate_weights()
        h4, _ = model.infer(pts, labels=None)
        assert torch.allclose(h4, h0 + 0.125, atol=1e-6), "invalidate_weights() did not re-upload"
        bias.data.copy_(bias.data - 0.125)
        model.invalidate_weights()
        h5, _ = model.infer(pts, labels=None)
        assert torch.allclose(h5, h0, atol=1e-6)


def test_infer_argument_validation(cuda):
    cfg = synth.make_cfg("kitti", batch_size=1)
    model, sd = helpers.load_model(cfg, "kitti", "mixed", cuda)
    pts = torch.from_numpy(np.stack([_scan(1), _scan(2)])).cuda()
    n = pts.shape[1]
    with torch.no_grad():
        for bad in (torch.tensor([n, n], dtype=torch.int64).cuda(), torch.tensor([n, n], dtype=torch.int32),
                    torch.tensor([n], dtype=torch.int32).cuda(), [n, n]):
            with pytest.raises(TypeError):
                model.infer(pts, counts=bad)
        with pytest.raises(TypeError):
            model.infer(pts.double())
        with pytest.raises(TypeError):
            model.infer(pts, out=(torch.empty(1, model.ny, model.nx, device="cuda"), None))
        # out-of-range counts are clamped on the device: negative = empty scan, too large = the whole row block
        h_ref, l_ref = model.infer(pts, counts=torch.tensor([0, n], dtype=torch.int32).cuda(), labels="int8")
        h_ref, l_ref = h_ref.clone(), l_ref.clone()
        h, l = model.infer(pts, counts=torch.tensor([-5, n + 1000], dtype=torch.int32).cuda(), labels="int8")
        assert torch.equal(h, h_ref) and torch.equal(l, l_ref)
        assert bool((l[0] == -1).all())


def test_scan_stream_host_buffers_match_direct_infer_with_and_without_graphs(cuda):
    """gndnet_b200.pipeline.ScanStream (pinned host scans in, host results out, three streams): every submitted batch
    gives byte for byte what infer() gives on the same scans -- with direct launches, and with the CUDA-graph replay
    that takes over the third time a (slot, row range) is submitted; shorter batches and changed weights in between."""
    from gndnet_b200.pipeline import ScanStream
    cfg = synth.make_cfg("kitti", batch_size=1)
    model, sd = helpers.load_model(cfg, "kitti", "mixed", cuda)
    N = 20000
    scans = [synth.hdl64_scan(40 + i)[::6][:N].copy() for i in range(6)]
    batches = [torch.from_numpy(np.stack(scans[i:i + 3])).pin_memory() for i in (0, 3)]
    expect = []
    with torch.no_grad():
        for hb in batches:
            h, l = model.infer(hb.cuda(), threshold=0.08, labels="int8")
            expect.append((h.cpu().clone(), l.cpu().clone()))
    for use_graph in (False, True):
        pipe = ScanStream(model, 3, N, depth=2, use_graph=use_graph)
        with torch.no_grad():
            for it in range(8):                         # each slot / range is seen four times: direct, capture, replay, replay
                k = pipe.submit(batches[it & 1])
                h, l = pipe.result(k)
                assert torch.equal(h, expect[it & 1][0]) and torch.equal(l, expect[it & 1][1]), (use_graph, it)
            k = pipe.submit(batches[1][:2])             # a shorter batch through the same slots
            h, l = pipe.result(k)
            assert torch.equal(h, expect[1][0][:2]) and torch.equal(l, expect[1][1][:2])
        if use_graph:
            assert any(not isinstance(g, str) for g in pipe._graphs.values()), "no graph was captured"
            # new weights: the captured graphs are dropped, results follow the new weights
            with torch.no_grad():
                model.encoder_decoder.regressor.bias.add_(0.25)
                k = pipe.submit(batches[0])
                h, l = pipe.result(k)
                assert torch.allclose(h, expect[0][0] + 0.25, atol=1e-6)
                model.encoder_decoder.regressor.bias.sub_(0.25)
    model.release()
