// conv_tc.cu -- 3x3 convolution + folded BatchNorm + ReLU as an implicit GEMM on the 5th-generation
// tensor cores (tcgen05.mma, accumulators in TMEM, operands staged by TMA).  sm_100a only.
//
// Replaces conv2DBatchNormRelu (modules/segnet.py:11-44) for the eight segnet blocks, and -- fused into the
// last block's epilogue -- the channel contraction of the regressor (segnet.py:131,140).
//
// GEMM view per layer: D[pixel, cout] = sum over (tap, cin) A[pixel shifted by tap, cin] * W[tap][cout][cin].
//   M tile  = TW x TH output pixels (<= 128 rows of one UMMA M=128 instruction)
//   N tile  = min(Cout, 256) output channels
//   K block = one filter tap x 64 input channels (64 fp16 = one 128-byte swizzle row)
// No im2col: for every (tap, channel block) the TMA engine loads the box {64 ch, TW, TH, 1 scan} of the NHWC
// activation tensor at (w0 + kw - 1, h0 + kh - 1); coordinates outside the image are zero-filled by the
// hardware, which IS the conv's zero padding.  The box lands in shared memory as TW*TH rows of 128 bytes
// with the 128-byte swizzle, i.e. exactly the K-major operand layout a UMMA shared-memory descriptor
// (SWIZZLE_128B, stride 1024 B per 8 rows) describes.  Weights are pre-packed [tap][cout][cin] so a
// {64, N, 1} box is the K-major B operand.
//
// Precision: fp16 operands, fp32 accumulation.  "split" layers carry every activation and
// weight as hi + lo fp16 pairs and issue three MMAs per K step (hi*hi + hi*lo + lo*hi), which restores
// ~22 bits of operand mantissa; the max-pool argmax decisions of the encoder need that.
//
// Warp roles (256 threads, one CTA per SM, persistent over work items):
//   warp 0 lane 0   TMA producer     (ring of `stages` smem buffers, full/empty mbarriers)
//   warp 1 lane 0   MMA issuer       (tcgen05.mma + tcgen05.commit; two TMEM accumulators, ping-pong)
//   warp 2          TMEM allocator
//   warps 4-7       epilogue         (tcgen05.ld -> bias + ReLU -> fp16 hi/lo NHWC stores, or regressor partials)
#include "model.cuh"
#include "tc_ptx.cuh"

namespace gnd {

struct TensorMaps {
    CUtensorMap a_hi[8], a_lo[8], b_hi[8], b_lo[8];
    int TW[8], TH[8];
    bool built = false;
};

struct ConvParams {
    int H, W, Cin, Cout, B;
    int TW, TH, tiles_w, tiles_h, n_tiles, NT;
    int split, cblocks, stages;
    uint32_t a_bytes, b_bytes, stage_bytes, tx_bytes;
    int total_items;
    uint32_t tmem_cols;
    const float *bias;
    __half *out_hi, *out_lo;
    const float *reg_w;     // [9][64] when the regressor is fused
    float *regpart;         // [B][H][W][9]
};

__device__ __forceinline__ void decode_item(const ConvParams &p, int item, int &b, int &h0, int &w0, int &n0) {
    const int nt = item % p.n_tiles;
    int t = item / p.n_tiles;
    const int tw = t % p.tiles_w; t /= p.tiles_w;
    const int th = t % p.tiles_h;
    b = t / p.tiles_h;
    h0 = th * p.TH; w0 = tw * p.TW; n0 = nt * p.NT;
}

constexpr int TC_THREADS = 256;
constexpr int TC_MAX_STAGES = 8;

__global__ void __launch_bounds__(TC_THREADS, 1)
conv3x3_tc_kernel(const __grid_constant__ CUtensorMap tm_a_hi, const __grid_constant__ CUtensorMap tm_a_lo,
                  const __grid_constant__ CUtensorMap tm_b_hi, const __grid_constant__ CUtensorMap tm_b_lo,
                  const ConvParams p) {
    extern __shared__ uint8_t smem_raw[];
    // pointer arithmetic on the shared array (not an integer round trip): the compiler keeps the shared address space,
    // so s_bias / s_regw reads are LDS and not generic loads
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + (size_t)p.stages * p.stage_bytes);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * TC_MAX_STAGES + 4);
    float *s_regw = reinterpret_cast<float *>(tmem_slot + 4);          // 9*64 floats (fused regressor only)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar_full = smem_u32(bars), bar_empty = smem_u32(bars + TC_MAX_STAGES);
    const uint32_t bar_tfull = smem_u32(bars + 2 * TC_MAX_STAGES), bar_tempty = smem_u32(bars + 2 * TC_MAX_STAGES + 2);

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_a_hi) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_b_hi) : "memory");
        if (p.split) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_a_lo) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_b_lo) : "memory");
        }
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(bar_tfull + 8 * a, 1); mbar_init(bar_tempty + 8 * a, 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(p.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (p.regpart && threadIdx.x >= 128)
        for (int i = threadIdx.x - 128; i < 576; i += 128) s_regw[i] = p.reg_w[i];
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer =====
            int stage = 0; uint32_t phase = 0;
            for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
                int b, h0, w0, n0;
                decode_item(p, item, b, h0, w0, n0);
                for (int tap = 0; tap < 9; ++tap) {
                    const int kh = tap / 3, kw = tap % 3;
                    for (int cb = 0; cb < p.cblocks; ++cb) {
                        mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                        const uint32_t full = bar_full + 8 * stage;
                        const uint32_t sa = smem_u32(smem + (size_t)stage * p.stage_bytes);
                        mbar_expect_tx(full, p.tx_bytes);
                        tma_load_4d(sa, &tm_a_hi, full, cb * 64, w0 + kw - 1, h0 + kh - 1, b);
                        if (p.split) {
                            tma_load_4d(sa + 16384, &tm_a_lo, full, cb * 64, w0 + kw - 1, h0 + kh - 1, b);
                            tma_load_3d(sa + 32768, &tm_b_hi, full, cb * 64, n0, tap);
                            tma_load_3d(sa + 32768 + p.b_bytes, &tm_b_lo, full, cb * 64, n0, tap);
                        } else {
                            tma_load_3d(sa + 16384, &tm_b_hi, full, cb * 64, n0, tap);
                        }
                        if (++stage == p.stages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===== MMA issuer =====
            const uint32_t idesc = (1u << 4) | ((uint32_t)(p.NT >> 3) << 17) | ((128u >> 4) << 24);   // F16 x F16 -> F32, K-major A and B
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t acc_phase = 0;
            const int kblocks = 9 * p.cblocks;
            for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
                mbar_wait(bar_tempty + 8 * acc, acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * p.NT);
                for (int kb = 0; kb < kblocks; ++kb) {
                    mbar_wait(bar_full + 8 * stage, phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + (size_t)stage * p.stage_bytes);
                    const uint32_t a_hi = sa, a_lo = sa + 16384;
                    const uint32_t b_hi = p.split ? sa + 32768 : sa + 16384, b_lo = sa + 32768 + p.b_bytes;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint64_t da = umma_desc_sw128(a_hi + 32 * k), db = umma_desc_sw128(b_hi + 32 * k);
                        tc_mma_f16(d_tmem, da, db, idesc, (kb | k) != 0);
                        if (p.split) {
                            tc_mma_f16(d_tmem, da, umma_desc_sw128(b_lo + 32 * k), idesc, 1);
                            tc_mma_f16(d_tmem, umma_desc_sw128(a_lo + 32 * k), db, idesc, 1);
                        }
                    }
                    tc_commit(bar_empty + 8 * stage);           // frees the smem slot when these MMAs retire
                    if (++stage == p.stages) { stage = 0; phase ^= 1; }
                }
                tc_commit(bar_tfull + 8 * acc);                 // accumulator complete -> epilogue
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= 4) {
        // ===== epilogue: TMEM -> registers -> bias + ReLU -> NHWC fp16 (hi, lo) or regressor partials =====
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int tr = row / p.TW, tcw = row - tr * p.TW;
        int acc = 0; uint32_t acc_phase = 0;
        for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
            int b, h0, w0, n0;
            decode_item(p, item, b, h0, w0, n0);
            const int h = h0 + tr, w = w0 + tcw;
            const bool valid = (row < p.TW * p.TH) && (h < p.H) && (w < p.W);
            const size_t pix = ((size_t)b * p.H + h) * p.W + w;
            mbar_wait(bar_tfull + 8 * acc, acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * p.NT);
            if (p.regpart) {
                float part[9];
#pragma unroll
                for (int t = 0; t < 9; ++t) part[t] = 0.0f;
                for (int j = 0; j < 2; ++j) {
                    uint32_t v[32];
                    tc_ld_32x32(taddr + 32 * j, v);
                    tc_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const float x = fmaxf(__uint_as_float(v[i]) + __ldg(p.bias + n0 + 32 * j + i), 0.0f);
#pragma unroll
                        for (int t = 0; t < 9; ++t) part[t] = fmaf(x, s_regw[t * 64 + 32 * j + i], part[t]);
                    }
                }
                if (valid) {
#pragma unroll
                    for (int t = 0; t < 9; ++t) p.regpart[pix * 9 + t] = part[t];
                }
            } else {
                for (int j = 0; j < p.NT / 32; ++j) {
                    uint32_t v[32];
                    tc_ld_32x32(taddr + 32 * j, v);
                    tc_ld_wait();
                    uint32_t hi[16], lo[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const float x0 = fmaxf(__uint_as_float(v[2 * i]) + __ldg(p.bias + n0 + 32 * j + 2 * i), 0.0f);
                        const float x1 = fmaxf(__uint_as_float(v[2 * i + 1]) + __ldg(p.bias + n0 + 32 * j + 2 * i + 1), 0.0f);
                        const __half2 hh = __floats2half2_rn(x0, x1);
                        hi[i] = *reinterpret_cast<const uint32_t *>(&hh);
                        const float2 back = __half22float2(hh);
                        lo[i] = pack_half2(x0 - back.x, x1 - back.y);
                    }
                    if (valid) {
                        const size_t o = pix * p.Cout + n0 + 32 * j;
                        uint4 *dh = reinterpret_cast<uint4 *>(p.out_hi + o);
#pragma unroll
                        for (int i = 0; i < 4; ++i) dh[i] = make_uint4(hi[4 * i], hi[4 * i + 1], hi[4 * i + 2], hi[4 * i + 3]);
                        if (p.out_lo) {
                            uint4 *dl = reinterpret_cast<uint4 *>(p.out_lo + o);
#pragma unroll
                            for (int i = 0; i < 4; ++i) dl[i] = make_uint4(lo[4 * i], lo[4 * i + 1], lo[4 * i + 2], lo[4 * i + 3]);
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_tempty + 8 * acc);
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
    }
}


// ---- host side --------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

// NHWC fp16 activation [B][H][W][C]: box {64 channels, TW, TH, 1 scan}, 128-byte swizzle, zero fill outside.
static int encode_act(CUtensorMap *m, const __half *ptr, int B, int H, int W, int C, int TW, int TH) {
    EncodeTiledFn enc = get_encode();
    if (!enc) return fail(GND_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)TW, (cuuint32_t)TH, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, (void *)ptr, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(GND_ERR_CUDA, "cuTensorMapEncodeTiled(activation %dx%dx%d box %dx%d) failed: %d", H, W, C, TW, TH, (int)r);
    return GND_OK;
}

// packed weights [9][Cout][Cin] fp16: box {64 cin, NT cout, 1 tap}
static int encode_wgt(CUtensorMap *m, const __half *ptr, int Cin, int Cout, int NT) {
    EncodeTiledFn enc = get_encode();
    if (!enc) return fail(GND_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[3] = {(cuuint64_t)Cin, (cuuint64_t)Cout, 9};
    cuuint64_t strides[2] = {(cuuint64_t)Cin * 2, (cuuint64_t)Cout * Cin * 2};
    cuuint32_t box[3] = {64, (cuuint32_t)NT, 1};
    cuuint32_t es[3] = {1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, (void *)ptr, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(GND_ERR_CUDA, "cuTensorMapEncodeTiled(weights %dx%d) failed: %d", Cout, Cin, (int)r);
    return GND_OK;
}

// TW x TH <= 128 pixels covering H x W with the fewest 128-row tiles
static void choose_tile(int H, int W, int *TW, int *TH) {
    long best = -1;
    for (int tw = 1; tw <= (W < 128 ? W : 128); ++tw) {
        int th = 128 / tw;
        if (th > H) th = H;
        if (th < 1) continue;
        long tiles = (long)((W + tw - 1) / tw) * ((H + th - 1) / th);
        if (best < 0 || tiles <= best) { best = tiles; *TW = tw; *TH = th; }   // ties: the widest box (longest contiguous runs)
    }
}

static const Act *layer_input(const gnd_ctx *ctx, int l) {
    const Act *in[8] = {&ctx->canvas, &ctx->d1a, &ctx->p1, &ctx->d2a, &ctx->u2, &ctx->u2a, &ctx->u1, &ctx->u1a};
    return in[l];
}

static inline int layer_nt(const ConvLayer &L) { return L.cout < 256 ? L.cout : 256; }

int tc_build_tensor_maps(gnd_ctx *ctx) {
    if (!ctx->tmaps) ctx->tmaps = new TensorMaps();
    TensorMaps &t = *ctx->tmaps;
    t.built = false;
    for (int l = 0; l < 8; ++l) {
        const ConvLayer &L = ctx->conv[l];
        const Act &in = *layer_input(ctx, l);
        if (!in.hi || (L.split && !in.lo)) return fail(GND_ERR_STATE, "layer %d: input activation planes missing", l);
        choose_tile(in.H, in.W, &t.TW[l], &t.TH[l]);
        int rc = encode_act(&t.a_hi[l], in.hi, ctx->B, in.H, in.W, in.C, t.TW[l], t.TH[l]);
        if (rc != GND_OK) return rc;
        rc = encode_act(&t.a_lo[l], L.split ? in.lo : in.hi, ctx->B, in.H, in.W, in.C, t.TW[l], t.TH[l]);
        if (rc != GND_OK) return rc;
        rc = encode_wgt(&t.b_hi[l], L.whi, L.cin, L.cout, layer_nt(L));
        if (rc != GND_OK) return rc;
        rc = encode_wgt(&t.b_lo[l], L.wlo, L.cin, L.cout, layer_nt(L));
        if (rc != GND_OK) return rc;
    }
    t.built = true;
    const Act *ins[8];
    for (int l = 0; l < 8; ++l) ins[l] = layer_input(ctx, l);
    GND_CUDA_CHECK(cudaFuncSetAttribute(conv3x3_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
    return halo2_build(ctx, ins);
}

void tc_free_tensor_maps(gnd_ctx *ctx) {
    delete ctx->tmaps;
    ctx->tmaps = nullptr;
    halo2_free(ctx);
}

int tc_conv_run(gnd_ctx *ctx, int l, const Act &in, const Act &out, int B, bool fuse_regressor, cudaStream_t st) {
    if (!ctx->tmaps || !ctx->tmaps->built) return fail(GND_ERR_STATE, "tensor maps not built (bind the workspace after loading weights)");
    if (halo2_usable(ctx, l)) return halo2_conv_run(ctx, l, out, B, fuse_regressor, st);
    const TensorMaps &t = *ctx->tmaps;
    const ConvLayer &L = ctx->conv[l];
    ConvParams p;
    p.H = in.H; p.W = in.W; p.Cin = L.cin; p.Cout = L.cout; p.B = B;
    p.TW = t.TW[l]; p.TH = t.TH[l];
    p.tiles_w = (in.W + p.TW - 1) / p.TW; p.tiles_h = (in.H + p.TH - 1) / p.TH;
    p.NT = layer_nt(L); p.n_tiles = L.cout / p.NT;
    p.split = L.split; p.cblocks = L.cin / 64;
    p.a_bytes = (uint32_t)p.TW * p.TH * 128; p.b_bytes = (uint32_t)p.NT * 128;
    p.stage_bytes = p.split ? 32768 + 2 * p.b_bytes : 16384 + p.b_bytes;
    p.tx_bytes = (p.a_bytes + p.b_bytes) * (p.split ? 2 : 1);
    int stages = (int)((200 * 1024) / p.stage_bytes);
    if (stages > TC_MAX_STAGES) stages = TC_MAX_STAGES;
    if (stages < 2) return fail(GND_ERR_UNSUPPORTED, "layer %d does not fit two pipeline stages", l);
    p.stages = stages;
    p.total_items = B * p.tiles_h * p.tiles_w * p.n_tiles;
    uint32_t cols = 2 * p.NT, pow2 = 32;
    while (pow2 < cols) pow2 <<= 1;
    p.tmem_cols = pow2;
    p.bias = L.bias;
    p.out_hi = fuse_regressor ? nullptr : out.hi;
    p.out_lo = fuse_regressor ? nullptr : out.lo;
    p.reg_w = ctx->reg_w;
    p.regpart = fuse_regressor ? ctx->regpart : nullptr;
    if (fuse_regressor && p.NT != 64) return fail(GND_ERR_UNSUPPORTED, "regressor fusion needs a 64-channel last block");
    if (!fuse_regressor && !out.hi) return fail(GND_ERR_STATE, "layer %d: output activation missing", l);
    const size_t smem = (size_t)stages * p.stage_bytes + 1024 /*align*/ + (2 * TC_MAX_STAGES + 4) * 8 + 16 + 576 * 4;
    int grid = p.total_items < 148 ? p.total_items : 148;
    conv3x3_tc_kernel<<<grid, TC_THREADS, smem, st>>>(t.a_hi[l], t.a_lo[l], t.b_hi[l], t.b_lo[l], p);
    GND_LAUNCH_CHECK("conv3x3_tc_kernel");
    return GND_OK;
}

}  // namespace gnd
