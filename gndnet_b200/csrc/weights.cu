// weights.cu -- gnd_create / gnd_destroy / gnd_load_weights.
//
// gnd_load_weights is the C side of model.load_state_dict(checkpoint['state_dict'])
// (predict_ground.py:196, ros_node.py:223): it takes the reference's state_dict keys,
// folds every BatchNorm in eval mode --
//     s = gamma / sqrt(running_var + eps),  w' = w * s,  b' = (b - running_mean) * s + beta
//     eps = 1e-5 for BatchNorm2d (modules/segnet.py:35-38), 1e-3 for BatchNorm1d (modules/pointpillars.py:42)
// -- and repacks to the kernels' layouts: PFN [feature][64]; fp32 conv [tap][cin][cout]; tcgen05 conv
// [tap][cout][cin] fp16 hi/lo (K-major rows for UMMA); regressor [tap][64].
#include "model.cuh"
#include <math.h>
#include <string.h>
#include <map>

using namespace gnd;

static const struct { const char *name; int cin, cout; } kBlocks[8] = {
    {"down1.conv1", 64, 128}, {"down1.conv2", 128, 128}, {"down2.conv1", 128, 256}, {"down2.conv2", 256, 256},
    {"up2.conv1", 256, 256},  {"up2.conv2", 256, 128},   {"up1.conv1", 128, 128},   {"up1.conv2", 128, 64},
};

extern "C" int gnd_create(const gnd_config *cfg, int32_t device, gnd_ctx **out) {
    if (!cfg || !out) return fail(GND_ERR_INVALID, "null argument");
    *out = nullptr;
    if (cfg->input_features != 3 && cfg->input_features != 4)
        return fail(GND_ERR_UNSUPPORTED, "input_features must be 3 or 4, got %d", cfg->input_features);
    if (cfg->precision < GND_PREC_FP32 || cfg->precision > GND_PREC_SPLIT)
        return fail(GND_ERR_INVALID, "unknown precision %d", cfg->precision);
    if (cfg->bn_batch_stats && cfg->precision != GND_PREC_FP32)
        return fail(GND_ERR_UNSUPPORTED, "batch-statistics BatchNorm needs precision fp32 (BatchNorm cannot be folded into the tensor-core weights)");
    if (cfg->bn_batch_stats && cfg->use_norm)
        return fail(GND_ERR_UNSUPPORTED, "batch-statistics BatchNorm1d in the pillar encoder (use_norm) is not built");
    int ndev = 0;
    GND_CUDA_CHECK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(GND_ERR_INVALID, "device %d of %d", device, ndev);
    GND_CUDA_CHECK(cudaSetDevice(device));
    cudaDeviceProp prop;
    GND_CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(GND_ERR_UNSUPPORTED, "libgndnet_b200 is built for sm_100a only; device %d is sm_%d%d", device, prop.major, prop.minor);

    gnd_ctx *ctx = new gnd_ctx();
    ctx->cfg = *cfg;
    ctx->device = device;
    gnd_voxel_params &vp = ctx->vparams;
    for (int j = 0; j < 3; ++j) vp.voxel_size[j] = cfg->voxel_size[j];
    for (int j = 0; j < 6; ++j) vp.coors_range[j] = cfg->pc_range[j];
    vp.max_points = cfg->max_points_voxel;
    vp.max_voxels = cfg->max_voxels;
    vp.reverse_index = 1;
    vp.num_features = cfg->input_features;
    vp.apply_z_shift = 1;
    vp.z_shift = cfg->lidar_height;
    // model.py:26-28: grid from the Python floats, rounded
    int g[3];
    for (int j = 0; j < 3; ++j) g[j] = (int)nearbyint((cfg->pc_range_f64[3 + j] - cfg->pc_range_f64[j]) / cfg->voxel_size_f64[j]);
    int32_t gv[3];
    int rc = gnd_voxel_grid_size(&vp, gv);
    if (rc != GND_OK) { delete ctx; return rc; }
    if (g[0] != gv[0] || g[1] != gv[1] || g[2] != gv[2]) {
        delete ctx;
        return fail(GND_ERR_INVALID, "float32 voxel grid (%d,%d,%d) differs from the model's float64 grid (%d,%d,%d)",
                    gv[0], gv[1], gv[2], g[0], g[1], g[2]);
    }
    ctx->nx = g[0]; ctx->ny = g[1]; ctx->nz = g[2];
    if (ctx->nx < 4 || ctx->ny < 4) { delete ctx; return fail(GND_ERR_UNSUPPORTED, "BEV grid %dx%d too small", ctx->nx, ctx->ny); }
    ctx->H1 = ctx->ny / 2; ctx->W1 = ctx->nx / 2;
    ctx->H2 = ctx->H1 / 2; ctx->W2 = ctx->W1 / 2;
    ctx->nfeat = cfg->input_features + 5 + (cfg->with_distance ? 1 : 0);
    {   // cfg.vfe_filters (modules/pointpillars.py:96-107)
        int n = cfg->n_vfe == 0 ? 1 : cfg->n_vfe;
        int f[GND_MAX_VFE] = {64, 0, 0, 0};
        if (cfg->n_vfe != 0) for (int j = 0; j < GND_MAX_VFE; ++j) f[j] = cfg->vfe_filters[j];
        bool ok = n >= 1 && n <= GND_MAX_VFE && f[n - 1] == PFN_OUT;
        for (int j = 0; ok && j < n - 1; ++j) ok = f[j] >= 2 && f[j] <= 128 && (f[j] % 2) == 0;
        if (!ok) { delete ctx; return fail(GND_ERR_UNSUPPORTED, "vfe_filters: 1..%d layers, the last one %d, the others even and <= 128", GND_MAX_VFE, PFN_OUT); }
        if (n > 1 && cfg->bn_batch_stats) { delete ctx; return fail(GND_ERR_UNSUPPORTED, "batch-statistics mode with a multi-layer PillarFeatureNet"); }
        ctx->n_vfe = n;
        int cin = ctx->nfeat;
        for (int j = 0; j < n; ++j) {
            ctx->vfe_in[j] = cin;
            ctx->vfe_units[j] = (j == n - 1) ? f[j] : f[j] / 2;
            cin = f[j];
        }
    }
    for (int l = 0; l < 8; ++l) {
        ctx->conv[l].name = kBlocks[l].name;
        ctx->conv[l].cin = kBlocks[l].cin;
        ctx->conv[l].cout = kBlocks[l].cout;
        ctx->conv[l].split = (cfg->precision == GND_PREC_SPLIT) || (cfg->precision == GND_PREC_MIXED && l < 4);
    }
    *out = ctx;
    return GND_OK;
}

extern "C" void gnd_destroy(gnd_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    tc_free_tensor_maps(ctx);
    for (cudaEvent_t e : ctx->tim_events) cudaEventDestroy(e);
    if (ctx->weight_blob) cudaFree(ctx->weight_blob);
    delete ctx;
}

namespace {
struct Table {
    std::map<std::string, const gnd_weight *> m;
    const float *get(const std::string &key, int64_t numel, int *rc) const {
        auto it = m.find(key);
        if (it == m.end()) { *rc = fail(GND_ERR_INVALID, "state_dict is missing key %s", key.c_str()); return nullptr; }
        if (it->second->numel != numel || !it->second->h_data) {
            *rc = fail(GND_ERR_INVALID, "state_dict key %s has %lld elements, expected %lld", key.c_str(),
                       (long long)it->second->numel, (long long)numel);
            return nullptr;
        }
        return it->second->h_data;
    }
};
}  // namespace

extern "C" int gnd_load_weights(gnd_ctx *ctx, const gnd_weight *table, int32_t n_entries) {
    if (!ctx || !table || n_entries < 1) return fail(GND_ERR_INVALID, "null argument");
    GND_CUDA_CHECK(cudaSetDevice(ctx->device));
    Table t;
    for (int i = 0; i < n_entries; ++i)
        if (table[i].name) t.m[table[i].name] = &table[i];
    int rc = GND_OK;
    const bool tc = ctx->cfg.precision != GND_PREC_FP32;

    // host staging of the whole blob
    size_t cur = 0;
    auto take = [&cur](size_t bytes) { cur = align_up(cur, 256); size_t r = cur; cur += bytes; return r; };
    size_t o_vfe_w[GND_MAX_VFE] = {0, 0, 0, 0}, o_vfe_b[GND_MAX_VFE] = {0, 0, 0, 0};
    for (int l = 0; l < ctx->n_vfe; ++l) {
        o_vfe_w[l] = take((size_t)ctx->vfe_in[l] * ctx->vfe_units[l] * sizeof(float));
        o_vfe_b[l] = take((size_t)ctx->vfe_units[l] * sizeof(float));
    }
    const size_t o_pfn_w = o_vfe_w[0], o_pfn_b = o_vfe_b[0];      // one layer: [nfeat][64] + [64]
    size_t o_w[8], o_hi[8], o_lo[8], o_b[8], o_g[8], o_be[8];
    const bool batch_bn = ctx->cfg.bn_batch_stats != 0;
    for (int l = 0; l < 8; ++l) {
        size_t n = (size_t)9 * ctx->conv[l].cin * ctx->conv[l].cout;
        o_w[l] = o_hi[l] = o_lo[l] = 0;
        if (!tc) o_w[l] = take(n * sizeof(float));
        else { o_hi[l] = take(n * sizeof(__half)); o_lo[l] = take(n * sizeof(__half)); }
        o_b[l] = take(ctx->conv[l].cout * sizeof(float));
        o_g[l] = take(ctx->conv[l].cout * sizeof(float));
        o_be[l] = take(ctx->conv[l].cout * sizeof(float));
    }
    size_t o_reg_w = take(9 * 64 * sizeof(float));
    size_t o_reg_b = take(sizeof(float));
    const size_t blob_bytes = align_up(cur, 256);
    std::vector<char> host(blob_bytes, 0);

    // ---- PillarFeatureNet (modules/pointpillars.py:42-47, :96-107): every PFNLayer's Linear, BatchNorm1d (eps 1e-3)
    // folded when use_norm, stored input-major [cin][units] ----
    for (int l = 0; l < ctx->n_vfe; ++l) {
        const int cin = ctx->vfe_in[l], units = ctx->vfe_units[l];
        const std::string pre = "voxel_feature_extractor.pfn_layers." + std::to_string(l) + ".";
        const float *w = t.get(pre + "linear.weight", (int64_t)units * cin, &rc);
        if (!w) return rc;
        float *dw = (float *)(host.data() + o_vfe_w[l]), *db = (float *)(host.data() + o_vfe_b[l]);
        if (ctx->cfg.use_norm) {
            const float *g = t.get(pre + "norm.weight", units, &rc);       if (!g) return rc;
            const float *be = t.get(pre + "norm.bias", units, &rc);        if (!be) return rc;
            const float *mu = t.get(pre + "norm.running_mean", units, &rc); if (!mu) return rc;
            const float *var = t.get(pre + "norm.running_var", units, &rc); if (!var) return rc;
            for (int ch = 0; ch < units; ++ch) {
                double s = (double)g[ch] / sqrt((double)var[ch] + 1e-3);
                for (int k = 0; k < cin; ++k) dw[k * units + ch] = (float)((double)w[ch * cin + k] * s);
                db[ch] = (float)((double)be[ch] - (double)mu[ch] * s);
            }
        } else {
            const float *b = t.get(pre + "linear.bias", units, &rc);
            if (!b) return rc;
            for (int ch = 0; ch < units; ++ch) {
                for (int k = 0; k < cin; ++k) dw[k * units + ch] = w[ch * cin + k];
                db[ch] = b[ch];
            }
        }
    }
    // ---- conv blocks (modules/segnet.py:11-44) ----
    for (int l = 0; l < 8; ++l) {
        const int cin = ctx->conv[l].cin, cout = ctx->conv[l].cout;
        const std::string p = std::string("encoder_decoder.") + ctx->conv[l].name + ".cbr_unit.";
        const float *w = t.get(p + "0.weight", (int64_t)cout * cin * 9, &rc);  if (!w) return rc;
        const float *b = t.get(p + "0.bias", cout, &rc);                       if (!b) return rc;
        const float *g = t.get(p + "1.weight", cout, &rc);                     if (!g) return rc;
        const float *be = t.get(p + "1.bias", cout, &rc);                      if (!be) return rc;
        const float *mu = t.get(p + "1.running_mean", cout, &rc);              if (!mu) return rc;
        const float *var = t.get(p + "1.running_var", cout, &rc);              if (!var) return rc;
        float *db = (float *)(host.data() + o_b[l]);
        for (int co = 0; co < cout; ++co) {
            // batch-statistics mode keeps the convolution as it is (weights and bias unscaled) and the affine apart
            const double s = batch_bn ? 1.0 : (double)g[co] / sqrt((double)var[co] + 1e-5);
            db[co] = batch_bn ? b[co] : (float)(((double)b[co] - (double)mu[co]) * s + (double)be[co]);
            ((float *)(host.data() + o_g[l]))[co] = g[co];
            ((float *)(host.data() + o_be[l]))[co] = be[co];
            for (int ci = 0; ci < cin; ++ci)
                for (int tap = 0; tap < 9; ++tap) {
                    const float wf = (float)((double)w[((size_t)co * cin + ci) * 9 + tap] * s);
                    if (!tc) {
                        ((float *)(host.data() + o_w[l]))[((size_t)tap * cin + ci) * cout + co] = wf;
                    } else {
                        const __half hi = __float2half_rn(wf);
                        const __half lo = __float2half_rn(wf - __half2float(hi));
                        ((__half *)(host.data() + o_hi[l]))[((size_t)tap * cout + co) * cin + ci] = hi;
                        ((__half *)(host.data() + o_lo[l]))[((size_t)tap * cout + co) * cin + ci] = lo;
                    }
                }
        }
    }
    // ---- regressor (modules/segnet.py:131) ----
    {
        const float *w = t.get("encoder_decoder.regressor.weight", 64 * 9, &rc);  if (!w) return rc;
        const float *b = t.get("encoder_decoder.regressor.bias", 1, &rc);         if (!b) return rc;
        float *dw = (float *)(host.data() + o_reg_w);
        for (int c = 0; c < 64; ++c)
            for (int tap = 0; tap < 9; ++tap) dw[tap * 64 + c] = w[c * 9 + tap];
        *(float *)(host.data() + o_reg_b) = b[0];
    }

    if (ctx->weight_blob) { GND_CUDA_CHECK(cudaFree(ctx->weight_blob)); ctx->weight_blob = nullptr; }
    GND_CUDA_CHECK(cudaMalloc(&ctx->weight_blob, blob_bytes));
    GND_CUDA_CHECK(cudaMemcpy(ctx->weight_blob, host.data(), blob_bytes, cudaMemcpyHostToDevice));
    char *d = (char *)ctx->weight_blob;
    ctx->pfn_w = (float *)(d + o_pfn_w);
    ctx->pfn_b = (float *)(d + o_pfn_b);
    for (int l = 0; l < ctx->n_vfe; ++l) { ctx->vfe_w[l] = (float *)(d + o_vfe_w[l]); ctx->vfe_b[l] = (float *)(d + o_vfe_b[l]); }
    for (int l = 0; l < 8; ++l) {
        ctx->conv[l].w32 = tc ? nullptr : (float *)(d + o_w[l]);
        ctx->conv[l].whi = tc ? (__half *)(d + o_hi[l]) : nullptr;
        ctx->conv[l].wlo = tc ? (__half *)(d + o_lo[l]) : nullptr;
        ctx->conv[l].bias = (float *)(d + o_b[l]);
        ctx->conv[l].gamma = (float *)(d + o_g[l]);
        ctx->conv[l].beta = (float *)(d + o_be[l]);
    }
    ctx->reg_w = (float *)(d + o_reg_w);
    ctx->reg_b = (float *)(d + o_reg_b);
    ctx->weights_loaded = true;
    if (ctx->ws && tc) {   // weights moved: the weight tensor maps must be rebuilt
        rc = tc_build_tensor_maps(ctx);
        if (rc != GND_OK) return rc;
    }
    return GND_OK;
}
