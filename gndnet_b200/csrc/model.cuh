// model.cuh -- the model context (weights, workspace carving, stage buffers) shared by the
// pillar-encode, segnet and ABI translation units.
#pragma once
#include "common.cuh"
#include "voxelize.cuh"
#include <cuda.h>
#include <string>
#include <vector>

namespace gnd {

constexpr int PFN_OUT = 64;        // cfg.vfe_filters = [64] in every shipped config
constexpr int PFN_MAX_FEAT = 10;
constexpr int TIM_STRIDE = 22;     // events per slot of the timing ring (gnd_conv_timing_begin)
constexpr int BN_BLOCKS = 592;     // partial-sum blocks of the batch-statistics reduction   // F(<=4) + 3 + 2 + with_distance

// A BEV activation tensor, NHWC ([B][H][W][C]).  fp32 mode: one float plane.  tcgen05 modes: an fp16
// "hi" plane and (for split-fp16 consumers) an fp16 "lo" plane with value = hi + lo.
struct Act {
    float *f32 = nullptr;
    __half *hi = nullptr;
    __half *lo = nullptr;
    int H = 0, W = 0, C = 0;
    size_t elems(int B) const { return (size_t)B * H * W * C; }
};

struct ConvLayer {
    const char *name;       // state_dict prefix under encoder_decoder.
    int cin, cout;
    float *w32 = nullptr;   // [9][cin][cout] fp32, BatchNorm folded           (fp32 mode)
    __half *whi = nullptr;  // [9][cout][cin] fp16 hi, K-major for UMMA        (tcgen05 modes)
    __half *wlo = nullptr;  // [9][cout][cin] fp16 lo
    float *bias = nullptr;  // [cout] folded bias (raw conv bias in batch-statistics mode)
    float *gamma = nullptr, *beta = nullptr;   // [cout] BatchNorm2d affine, batch-statistics mode only
    int split = 0;          // tcgen05: 1 = three-product split-fp16, 0 = single fp16 product
};

struct TensorMaps;   // conv_tc.cu
struct Halo2Maps;    // conv_halo2.cu

}  // namespace gnd

struct gnd_ctx {
    gnd_config cfg;
    int device = 0;
    int nx = 0, ny = 0, nz = 0;
    int H1 = 0, W1 = 0, H2 = 0, W2 = 0;      // pooled sizes (floor)
    int nfeat = 0;                           // PFN input features
    gnd_voxel_params vparams;

    // weights (device)
    bool weights_loaded = false;
    float *pfn_w = nullptr;    // [nfeat][64], BatchNorm1d folded when use_norm (= vfe_w[0] when there is one PFN layer)
    float *pfn_b = nullptr;    // [64]
    // multi-layer PillarFeatureNet (cfg.vfe_filters with more than one entry; no shipped config): layer l maps vfe_in[l]
    // inputs to vfe_units[l] units (out / 2 for the non-last layers, whose output is concat(x, max over the pillar))
    int n_vfe = 1;
    int vfe_in[4] = {0, 0, 0, 0}, vfe_units[4] = {64, 0, 0, 0};
    float *vfe_w[4] = {nullptr, nullptr, nullptr, nullptr};    // [vfe_in][vfe_units], folded
    float *vfe_b[4] = {nullptr, nullptr, nullptr, nullptr};
    gnd::ConvLayer conv[8];
    float *reg_w = nullptr;    // [9][64]
    float *reg_b = nullptr;    // [1]
    void *weight_blob = nullptr;

    // bound workspace
    void *ws = nullptr;
    size_t ws_bytes = 0;
    int B = 0;
    int64_t N = 0;
    gnd::VoxPlan vplan;
    gnd::VoxWs vws;
    int32_t *coors = nullptr, *num_points = nullptr, *voxel_count = nullptr, *offsets = nullptr, *list = nullptr;
    gnd::Act canvas, d1a, d1b, p1, d2a, d2b, u2, u2a, u2b, u1, u1a, u1b;
    uint8_t *idx1 = nullptr;   // [B][H1][W1][128] argmax (0..3) of down1's pool
    float *regpart = nullptr;  // [B][H][W][9] per-tap regressor partials (tcgen05 modes)
    double *bn_partial = nullptr;   // [BN_BLOCKS][256][2] per-block channel sums (batch-statistics mode)
    float *bn_ab = nullptr;         // [2][256] per-channel scale / shift of the current layer
    float *bn_stats = nullptr;      // [8][256][2] batch mean / unbiased variance of every conv block (bn_batch_stats)
    int32_t *pfn_ctr = nullptr; // [2] work-queue head + finished-CTA count of pfn_scatter_kernel (self-resetting)
    int32_t *pfn_order = nullptr;   // [6][cells] work units of 2^s cells (s = 0..5) sorted by distance from the sensor's cell
    gnd::TensorMaps *tmaps = nullptr;
    gnd::Halo2Maps *h2maps = nullptr;

    int last_launches = 0;

    // gnd_set_option
    int opt_conv_path = 0;      // 0: CTA-pair halo kernel (conv_halo2.cu) where it fits, 1: single-CTA tap kernel (conv_tc.cu)
    int opt_conv_geom = 0;      // 0: automatic, 1: throughput geometry, 2: small-batch geometry
    int opt_l2_order = 0;       // 1: consecutive segnet kernels walk the scans in alternating directions (L2 reuse), 0: all forward (default: no gain measured)
    int opt_pfn_path = 0;       // fused pillar encoder: 0: automatic (= the CUDA-core kernel: measured faster), 1: CUDA-core kernel, 2: mma.sync (TF32 x3) kernel
    int opt_vox_path = 0;       // gnd_infer: 0: automatic (set-exact voxelization unless the cloud is dense), 1: order-exact kernels, 2: set-exact

    // gnd_conv_timing_begin / _end: library-owned ring of CUDA events around the 8 convolution launches of every
    // gnd_segnet / gnd_infer while armed (asynchronous: nothing is read back until _end)
    std::vector<cudaEvent_t> tim_events;     // [slots][TIM_STRIDE]: 0..15 conv launches, 16/17 segnet, 18..21 gnd_infer stages
    std::vector<char> tim_staged;            // slot was filled by gnd_infer (stage events valid)
    int tim_slots = 0, tim_next = 0;
};

namespace gnd {

// pfn.cu
int pfn_build_order(gnd_ctx *ctx);          // set-up time: distance-sorted work-unit lists into ctx->pfn_order
int pfn_scatter_run(gnd_ctx *ctx, const float *points, int32_t batch, int64_t n_points, cudaStream_t st);
int pfn_scatter_dense_run(gnd_ctx *ctx, const float *voxels, const float *coors4, const float *num_points,
                          int64_t n_pillars, int32_t batch, cudaStream_t st);
// segnet.cu
int segnet_run(gnd_ctx *ctx, int32_t batch, float *height, cudaStream_t st, cudaEvent_t *conv_events = nullptr);
// conv_tc.cu (tcgen05 path)
int tc_build_tensor_maps(gnd_ctx *ctx);
void tc_free_tensor_maps(gnd_ctx *ctx);
int tc_conv_run(gnd_ctx *ctx, int layer, const Act &in, const Act &out, int batch, bool fuse_regressor, cudaStream_t st);
// conv_halo2.cu (halo staged once per channel block, CTA pair, tcgen05 cta_group::2)
int halo2_build(gnd_ctx *ctx, const Act *const inputs[8]);
void halo2_free(gnd_ctx *ctx);
bool halo2_usable(const gnd_ctx *ctx, int layer);
int halo2_conv_run(gnd_ctx *ctx, int layer, const Act &out, int batch, bool fuse_regressor, cudaStream_t st);
// segment.cu
int segment_run(const gnd_segment_params *p, float *points, const int32_t *counts, int32_t batch,
                int64_t n_points, const float *elevation, double *labels_f64, int8_t *labels_i8,
                cudaStream_t stream);

}  // namespace gnd
