// abi.cu -- workspace carving and the stage / pipeline entry points of the C ABI.
//
// gnd_infer restates the reference's call sequence (predict_ground.py:110-121 InferGround, :165-168):
//   _shift_cloud -> points_to_voxel -> model(voxels, coors, num_points) -> segment_cloud(points, grid_range,
//   voxel_size[0], pred.T, threshold)
// as four enqueues on one stream with no host round trip and no dense [M,P,F] voxel tensor in between.
#include "model.cuh"
#include <string.h>

using namespace gnd;

namespace {

struct NamedBuf { std::string name; void *ptr; size_t bytes; };

void set_act(Act &a, int H, int W, int C) { a.H = H; a.W = W; a.C = C; }

// Lays out every stage buffer in the caller's workspace.  With ws == nullptr only sizes are computed.
size_t carve_all(gnd_ctx *ctx, int B, int64_t N, void *ws, std::vector<NamedBuf> *names) {
    Carver c(ws);
    const gnd_config &cfg = ctx->cfg;
    // voxelizer scratch + outputs
    c.off += vox_carve(ctx->vplan, B, N, ws, ws ? &ctx->vws : nullptr, true);
    const size_t mv = (size_t)cfg.max_voxels;
    int32_t *coors = c.take<int32_t>((size_t)B * mv * 3);
    int32_t *num = c.take<int32_t>((size_t)B * mv);
    int32_t *cnt = c.take<int32_t>((size_t)B);
    int32_t *off = c.take<int32_t>((size_t)B * (mv + 1));
    int32_t *list = c.take<int32_t>((size_t)B * (size_t)(N < 1 ? 1 : N));
    if (ws) { ctx->coors = coors; ctx->num_points = num; ctx->voxel_count = cnt; ctx->offsets = off; ctx->list = list; }

    const int prec = cfg.precision;
    const int H = ctx->ny, W = ctx->nx, H1 = ctx->H1, W1 = ctx->W1;
    struct Spec { Act *a; const char *name; int H, W, C; bool enc; bool stored; };
    Spec specs[] = {
        {&ctx->canvas, "canvas", H, W, 64, true, true},  {&ctx->d1a, "d1a", H, W, 128, true, true},
        {&ctx->d1b, "d1b", H, W, 128, true, true},       {&ctx->p1, "p1", H1, W1, 128, true, true},
        {&ctx->d2a, "d2a", H1, W1, 256, true, true},     {&ctx->d2b, "d2b", H1, W1, 256, true, true},
        {&ctx->u2, "u2", H1, W1, 256, false, true},      {&ctx->u2a, "u2a", H1, W1, 256, false, true},
        {&ctx->u2b, "u2b", H1, W1, 128, false, true},    {&ctx->u1, "u1", H, W, 128, false, true},
        {&ctx->u1a, "u1a", H, W, 128, false, true},      {&ctx->u1b, "u1b", H, W, 64, false, prec == GND_PREC_FP32},
    };
    for (Spec &s : specs) {
        Act a;
        set_act(a, s.H, s.W, s.C);
        const size_t n = a.elems(B);
        if (s.stored) {
            if (prec == GND_PREC_FP32) {
                a.f32 = c.take<float>(n);
            } else {
                a.hi = c.take<__half>(n);
                // a lo plane exists where the consumer is a split-fp16 conv or an argmax decision
                if (prec == GND_PREC_SPLIT || s.enc) a.lo = c.take<__half>(n);
            }
        }
        if (ws) *s.a = a;
        if (ws && names) {
            if (a.f32) names->push_back({std::string(s.name), a.f32, n * 4});
            if (a.hi) names->push_back({std::string(s.name) + ".hi", a.hi, n * 2});
            if (a.lo) names->push_back({std::string(s.name) + ".lo", a.lo, n * 2});
        }
    }
    uint8_t *idx1 = c.take<uint8_t>((size_t)B * H1 * W1 * 128);
    float *regpart = (prec == GND_PREC_FP32) ? nullptr : c.take<float>((size_t)B * H * W * 9);
    int32_t *pfn_ctr = c.take<int32_t>(4);
    int32_t *pfn_order = c.take<int32_t>((size_t)6 * ctx->nx * ctx->ny * ctx->nz);
    double *bn_partial = cfg.bn_batch_stats ? c.take<double>((size_t)BN_BLOCKS * 256 * 2) : nullptr;
    float *bn_ab = cfg.bn_batch_stats ? c.take<float>(512) : nullptr;
    float *bn_stats = cfg.bn_batch_stats ? c.take<float>(8 * 512) : nullptr;
    if (ws) {
        ctx->idx1 = idx1; ctx->regpart = regpart; ctx->pfn_ctr = pfn_ctr; ctx->pfn_order = pfn_order; ctx->bn_partial = bn_partial; ctx->bn_ab = bn_ab; ctx->bn_stats = bn_stats;
        if (names) {
            names->push_back({"idx1", idx1, (size_t)B * H1 * W1 * 128});
            if (regpart) names->push_back({"regpart", regpart, (size_t)B * H * W * 9 * 4});
            if (bn_stats) names->push_back({"bn_stats", bn_stats, (size_t)8 * 512 * 4});
            names->push_back({"coors", coors, (size_t)B * mv * 3 * 4});
            names->push_back({"num_points", num, (size_t)B * mv * 4});
            names->push_back({"voxel_count", cnt, (size_t)B * 4});
            names->push_back({"point_offsets", off, (size_t)B * (mv + 1) * 4});
            names->push_back({"point_list", list, (size_t)B * (size_t)(N < 1 ? 1 : N) * 4});
            names->push_back({"cell_voxel", ctx->vws.vid, (size_t)B * ctx->vplan.cells * 4});
            names->push_back({"cell_offsets", ctx->vws.offc, (size_t)B * (ctx->vplan.cells + 1) * 4});
            if (ctx->vws.sorted) names->push_back({"cell_points", ctx->vws.sorted, (size_t)B * (size_t)(N < 1 ? 1 : N) * 16});
        }
    }
    return c.bytes();
}

thread_local std::vector<NamedBuf> t_names;   // rebuilt by gnd_get_buffer

int check_bound(gnd_ctx *ctx, int32_t batch, int64_t n_points, bool need_weights) {
    if (!ctx) return fail(GND_ERR_INVALID, "null ctx");
    if (need_weights && !ctx->weights_loaded) return fail(GND_ERR_STATE, "gnd_load_weights has not been called");
    if (!ctx->ws) return fail(GND_ERR_STATE, "gnd_bind_workspace has not been called");
    if (batch < 1 || batch > ctx->B) return fail(GND_ERR_INVALID, "batch %d exceeds the bound batch %d", batch, ctx->B);
    // the workspace is bound to a CAPACITY (batch, n_points): any smaller batch / row count runs in it unchanged
    if (n_points > ctx->N)
        return fail(GND_ERR_INVALID, "n_points %lld exceeds the bound capacity %lld", (long long)n_points, (long long)ctx->N);
    cudaError_t e = cudaSetDevice(ctx->device);
    if (e != cudaSuccess) return fail(GND_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    return GND_OK;
}

}  // namespace

extern "C" int gnd_grid_shape(const gnd_ctx *ctx, int32_t *h) {
    if (!ctx || !h) return fail(GND_ERR_INVALID, "null argument");
    h[0] = ctx->nz; h[1] = ctx->ny; h[2] = ctx->nx;
    return GND_OK;
}

extern "C" int gnd_workspace_bytes(const gnd_ctx *cctx, int32_t batch, int64_t n_points, size_t *h_bytes) {
    if (!cctx || !h_bytes) return fail(GND_ERR_INVALID, "null argument");
    gnd_ctx tmp = *cctx;     // sizes only; never touches device memory
    tmp.tmaps = nullptr;
    tmp.h2maps = nullptr;
    int rc = vox_make_plan(&tmp.vparams, batch, n_points, &tmp.vplan);
    if (rc != GND_OK) return rc;
    *h_bytes = carve_all(&tmp, batch, n_points, nullptr, nullptr) + 1024;
    return GND_OK;
}

extern "C" int gnd_bind_workspace(gnd_ctx *ctx, void *workspace, size_t bytes, int32_t batch, int64_t n_points) {
    if (!ctx || !workspace) return fail(GND_ERR_INVALID, "null argument");
    if (((uintptr_t)workspace & 1023) != 0) return fail(GND_ERR_INVALID, "workspace must be 1024-byte aligned");
    size_t need;
    int rc = gnd_workspace_bytes(ctx, batch, n_points, &need);
    if (rc != GND_OK) return rc;
    if (bytes < need) return fail(GND_ERR_WORKSPACE, "workspace: need %zu bytes, got %zu", need, bytes);
    GND_CUDA_CHECK(cudaSetDevice(ctx->device));
    rc = vox_make_plan(&ctx->vparams, batch, n_points, &ctx->vplan);
    if (rc != GND_OK) return rc;
    ctx->ws = workspace; ctx->ws_bytes = bytes; ctx->B = batch; ctx->N = n_points;
    carve_all(ctx, batch, n_points, workspace, nullptr);
    GND_CUDA_CHECK(cudaMemset(ctx->pfn_ctr, 0, 4 * sizeof(int32_t)));
    rc = pfn_build_order(ctx);
    if (rc != GND_OK) { ctx->ws = nullptr; return rc; }
    rc = vox_lean_init(ctx->vplan, ctx->vws, batch);
    if (rc != GND_OK) { ctx->ws = nullptr; return rc; }
    if (ctx->cfg.precision != GND_PREC_FP32 && ctx->weights_loaded) {
        rc = tc_build_tensor_maps(ctx);
        if (rc != GND_OK) { ctx->ws = nullptr; return rc; }
    }
    return GND_OK;
}

extern "C" int gnd_get_buffer(gnd_ctx *ctx, const char *name, void **ptr, size_t *bytes) {
    if (!ctx || !name || !ptr || !bytes) return fail(GND_ERR_INVALID, "null argument");
    if (!ctx->ws) return fail(GND_ERR_STATE, "gnd_bind_workspace has not been called");
    t_names.clear();
    carve_all(ctx, ctx->B, ctx->N, ctx->ws, &t_names);
    for (const NamedBuf &b : t_names)
        if (b.name == name) { *ptr = b.ptr; *bytes = b.bytes; return GND_OK; }
    return fail(GND_ERR_INVALID, "no workspace buffer named %s", name);
}

extern "C" int gnd_last_launch_count(const gnd_ctx *ctx) { return ctx ? ctx->last_launches : 0; }

extern "C" int gnd_set_option(gnd_ctx *ctx, const char *name, int32_t value) {
    if (!ctx || !name) return fail(GND_ERR_INVALID, "null argument");
    if (!strcmp(name, "conv_path")) {
        if (value < 0 || value > 1) return fail(GND_ERR_INVALID, "conv_path must be 0 (CTA-pair halo kernel) or 1 (single-CTA tap kernel)");
        ctx->opt_conv_path = value;
    } else if (!strcmp(name, "pfn_path")) {
        if (value < 0 || value > 2) return fail(GND_ERR_INVALID, "pfn_path must be 0 (automatic), 1 (CUDA-core pillar encoder) or 2 (mma.sync one)");
        ctx->opt_pfn_path = value;
    } else if (!strcmp(name, "vox_path")) {
        if (value < 0 || value > 2) return fail(GND_ERR_INVALID, "vox_path must be 0 (automatic), 1 (order-exact kernels) or 2 (set-exact kernels)");
        ctx->opt_vox_path = value;
    } else if (!strcmp(name, "l2_order")) {
        if (value < 0 || value > 1) return fail(GND_ERR_INVALID, "l2_order must be 0 (every kernel walks the scans forward) or 1 (alternating)");
        ctx->opt_l2_order = value;
    } else if (!strcmp(name, "conv_geometry")) {
        if (value < 0 || value > 2) return fail(GND_ERR_INVALID, "conv_geometry must be 0 (auto), 1 (throughput) or 2 (small batch)");
        ctx->opt_conv_geom = value;
    } else {
        return fail(GND_ERR_INVALID, "unknown option %s", name);
    }
    return GND_OK;
}

// ---- asynchronous per-convolution timing (bench.py's roofline: the kernel timed INSIDE a long run) -----------------
static void timing_free(gnd_ctx *ctx) {
    for (cudaEvent_t e : ctx->tim_events) cudaEventDestroy(e);
    ctx->tim_events.clear();
    ctx->tim_staged.clear();
    ctx->tim_slots = ctx->tim_next = 0;
}

extern "C" int gnd_conv_timing_begin(gnd_ctx *ctx, int32_t slots) {
    if (!ctx || slots < 1 || slots > 4096) return fail(GND_ERR_INVALID, "slots must be 1..4096");
    GND_CUDA_CHECK(cudaSetDevice(ctx->device));
    timing_free(ctx);
    ctx->tim_events.resize((size_t)slots * TIM_STRIDE);
    for (cudaEvent_t &e : ctx->tim_events) GND_CUDA_CHECK(cudaEventCreate(&e));
    ctx->tim_staged.assign((size_t)slots, 0);
    ctx->tim_slots = slots;
    return GND_OK;
}

extern "C" int gnd_conv_timing_end(gnd_ctx *ctx, float *h_conv_ms, float *h_segnet_ms, float *h_stage_ms, int32_t *h_recorded) {
    if (!ctx || !h_conv_ms || !h_segnet_ms || !h_recorded) return fail(GND_ERR_INVALID, "null argument");
    GND_CUDA_CHECK(cudaSetDevice(ctx->device));
    const int n = ctx->tim_next;
    *h_recorded = n;
    int rc = GND_OK;
    for (int s = 0; s < n && rc == GND_OK; ++s) {
        cudaEvent_t *ev = &ctx->tim_events[(size_t)s * TIM_STRIDE];
        const bool staged = ctx->tim_staged.size() > (size_t)s && ctx->tim_staged[s];
        cudaError_t e = cudaEventSynchronize(ev[staged ? 21 : 17]);
        for (int l = 0; l < 8 && e == cudaSuccess; ++l) e = cudaEventElapsedTime(&h_conv_ms[s * 8 + l], ev[2 * l], ev[2 * l + 1]);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&h_segnet_ms[s], ev[16], ev[17]);
        if (h_stage_ms) {        // gnd_infer calls: voxelize, pillar-encode + scatter, segnet, labels; -1 for other calls
            static const int from[4] = {18, 19, 16, 20}, to[4] = {19, 16, 17, 21};
            for (int k = 0; k < 4; ++k) {
                h_stage_ms[s * 4 + k] = -1.0f;
                if (staged && e == cudaSuccess) e = cudaEventElapsedTime(&h_stage_ms[s * 4 + k], ev[from[k]], ev[to[k]]);
            }
        }
        if (e != cudaSuccess) rc = fail(GND_ERR_CUDA, "conv timing read-back: %s", cudaGetErrorString(e));
    }
    timing_free(ctx);
    return rc;
}

extern "C" int gnd_voxelize(gnd_ctx *ctx, const float *points, const int32_t *counts, int32_t batch,
                            int64_t n_points, void *stream) {
    g_launches = 0;
    int rc = check_bound(ctx, batch, n_points, false);
    if (rc != GND_OK) return rc;
    rc = vox_run(&ctx->vparams, ctx->vplan, ctx->vws, points, counts, batch, n_points, ctx->coors, ctx->num_points,
                 ctx->voxel_count, ctx->offsets, ctx->list, nullptr, nullptr, (cudaStream_t)stream);
    ctx->last_launches = g_launches;
    return rc;
}

extern "C" int gnd_pfn_scatter(gnd_ctx *ctx, const float *points, int32_t batch, int64_t n_points, void *stream) {
    g_launches = 0;
    int rc = check_bound(ctx, batch, n_points, true);
    if (rc != GND_OK) return rc;
    rc = pfn_scatter_run(ctx, points, batch, n_points, (cudaStream_t)stream);
    ctx->last_launches = g_launches;
    return rc;
}

extern "C" int gnd_pfn_scatter_dense(gnd_ctx *ctx, const float *voxels, const float *coors4, const float *num_points,
                                     int64_t n_pillars, int32_t batch, void *stream) {
    g_launches = 0;
    int rc = check_bound(ctx, batch, -1, true);
    if (rc != GND_OK) return rc;
    if (n_pillars < 0) return fail(GND_ERR_INVALID, "n_pillars < 0");
    if (n_pillars > 0 && (!voxels || !coors4 || !num_points)) return fail(GND_ERR_INVALID, "null pillar tensors");
    rc = pfn_scatter_dense_run(ctx, voxels, coors4, num_points, n_pillars, batch, (cudaStream_t)stream);
    ctx->last_launches = g_launches;
    return rc;
}

extern "C" int gnd_segnet(gnd_ctx *ctx, int32_t batch, float *height, void *stream) {
    g_launches = 0;
    int rc = check_bound(ctx, batch, -1, true);
    if (rc != GND_OK) return rc;
    if (!height) return fail(GND_ERR_INVALID, "height is null");
    rc = segnet_run(ctx, batch, height, (cudaStream_t)stream);
    ctx->last_launches = g_launches;
    return rc;
}

// Profiling aid for bench.py's roofline: one gnd_segnet with CUDA events around each of the 8 convolution
// launches.  SYNCHRONOUS.  h_conv_ms[8] = per-layer device time, *h_total_ms = the whole segnet.
extern "C" int gnd_segnet_profile(gnd_ctx *ctx, int32_t batch, float *height, float *h_conv_ms, float *h_total_ms, void *stream) {
    g_launches = 0;
    int rc = check_bound(ctx, batch, -1, true);
    if (rc != GND_OK) return rc;
    if (!height || !h_conv_ms || !h_total_ms) return fail(GND_ERR_INVALID, "null argument");
    cudaStream_t st = (cudaStream_t)stream;
    cudaEvent_t ev[18];
    for (int i = 0; i < 18; ++i) GND_CUDA_CHECK(cudaEventCreate(&ev[i]));
    GND_CUDA_CHECK(cudaEventRecord(ev[16], st));
    rc = segnet_run(ctx, batch, height, st, ev);
    if (rc == GND_OK) {
        GND_CUDA_CHECK(cudaEventRecord(ev[17], st));
        GND_CUDA_CHECK(cudaStreamSynchronize(st));
        for (int l = 0; l < 8; ++l) GND_CUDA_CHECK(cudaEventElapsedTime(&h_conv_ms[l], ev[2 * l], ev[2 * l + 1]));
        GND_CUDA_CHECK(cudaEventElapsedTime(h_total_ms, ev[16], ev[17]));
    }
    for (int i = 0; i < 18; ++i) cudaEventDestroy(ev[i]);
    ctx->last_launches = g_launches;
    return rc;
}

static int segment_with_ctx(gnd_ctx *ctx, const float *points, const int32_t *counts, int32_t batch, int64_t n_points,
                            const float *height, double threshold, int8_t *labels_i8, double *labels_f64, cudaStream_t st) {
    gnd_segment_params sp;
    memset(&sp, 0, sizeof(sp));
    sp.grid_lo[0] = ctx->cfg.grid_range[0];
    sp.grid_lo[1] = ctx->cfg.grid_range[1];
    sp.voxel_size = ctx->cfg.voxel_size_f64[0];     // predict_ground.py:168 passes cfg.voxel_size[0]
    sp.threshold = threshold;
    sp.map_nx = ctx->nx; sp.map_ny = ctx->ny;       // elevation_map = pred.T: [x, y] view of pred[y][x]
    sp.map_sx = 1; sp.map_sy = ctx->nx;
    sp.map_batch_stride = (int64_t)ctx->nx * ctx->ny;
    sp.num_features = ctx->cfg.input_features;
    sp.apply_z_shift = 1;
    sp.z_shift = ctx->cfg.lidar_height;
    sp.mutate_points = 0;
    return segment_run(&sp, const_cast<float *>(points), counts, batch, n_points, height, labels_f64, labels_i8, st);
}

extern "C" int gnd_segment(gnd_ctx *ctx, const float *points, const int32_t *counts, int32_t batch, int64_t n_points,
                           const float *height, double threshold, int8_t *labels_i8, double *labels_f64, void *stream) {
    g_launches = 0;
    if (!ctx) return fail(GND_ERR_INVALID, "null ctx");
    GND_CUDA_CHECK(cudaSetDevice(ctx->device));
    int rc = segment_with_ctx(ctx, points, counts, batch, n_points, height, threshold, labels_i8, labels_f64, (cudaStream_t)stream);
    ctx->last_launches = g_launches;
    return rc;
}

extern "C" int gnd_infer(gnd_ctx *ctx, const float *points, const int32_t *counts, int32_t batch, int64_t n_points,
                         double threshold, float *height, int8_t *labels_i8, double *labels_f64, void *stream) {
    g_launches = 0;
    int rc = check_bound(ctx, batch, n_points, true);
    if (rc != GND_OK) return rc;
    if (!height) return fail(GND_ERR_INVALID, "height is null");
    cudaStream_t st = (cudaStream_t)stream;
    cudaEvent_t *sev = nullptr;          // armed by gnd_conv_timing_begin: this call's slot also gets the stage events
    if (ctx->tim_next < ctx->tim_slots) {
        sev = &ctx->tim_events[(size_t)ctx->tim_next * TIM_STRIDE];
        cudaEventRecord(sev[18], st);
    }
    // the fused path needs per cell the SET of kept points, not the slot-ordered index lists: set-exact kernels
    // (voxelize.cu); gnd_voxelize / gnd_points_to_voxel keep producing the lists with the order-exact ones
    // ... unless the cloud is dense (the camera frames: ~400 points in every cell, all over-full) AND the batch is small:
    // measured with 1 M-point frames, order-exact against set-exact (chunked counts): 0.077 / 0.085 ms at batch 1,
    // 0.147 / 0.183 at 4, 0.264 / 0.270 at 8, 2.03 / 1.61 at 64
    const bool dense = n_points > (int64_t)32 * ctx->vplan.cells;
    if (ctx->vplan.lean_ok && ctx->vws.lfirst && ctx->opt_vox_path != 1 &&
        (ctx->opt_vox_path == 2 || !dense || (int64_t)batch * n_points >= 12000000))
        rc = vox_run_lean(&ctx->vparams, ctx->vplan, ctx->vws, points, counts, batch, n_points, ctx->coors, ctx->num_points,
                          ctx->voxel_count, ctx->offsets, st);
    else
        rc = vox_run(&ctx->vparams, ctx->vplan, ctx->vws, points, counts, batch, n_points, ctx->coors, ctx->num_points,
                     ctx->voxel_count, ctx->offsets, ctx->list, nullptr, nullptr, st);
    if (sev) cudaEventRecord(sev[19], st);
    if (rc == GND_OK) rc = pfn_scatter_run(ctx, points, batch, n_points, st);
    const int slot = ctx->tim_next;
    if (rc == GND_OK) rc = segnet_run(ctx, batch, height, st);       // takes the slot (events 0..17)
    if (sev) cudaEventRecord(sev[20], st);
    if (rc == GND_OK && (labels_i8 || labels_f64) && n_points > 0)
        rc = segment_with_ctx(ctx, points, counts, batch, n_points, height, threshold, labels_i8, labels_f64, st);
    if (sev) {
        cudaEventRecord(sev[21], st);
        if (rc == GND_OK && ctx->tim_next == slot + 1) ctx->tim_staged[slot] = 1;
    }
    ctx->last_launches = g_launches;
    return rc;
}
