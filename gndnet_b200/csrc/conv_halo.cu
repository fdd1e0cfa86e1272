// conv_halo.cu -- 3x3 convolution + folded BatchNorm + ReLU as an implicit GEMM on tcgen05, with the
// input halo staged ONCE per channel block and the nine filter taps read as nine shifted views of it.
//
// Replaces conv2DBatchNormRelu (modules/segnet.py:11-44) for the segnet blocks and -- fused into the last
// block's epilogue -- the channel contraction of the regressor (segnet.py:131,140).
//
// Why: the tap-per-TMA-load kernel (conv_tc.cu) re-reads every activation nine times and every weight once
// per 128 pixels from L2; at ~6300 B/clk chip-wide (~43 B/clk/SM) that, not the tensor pipe, bounded the
// fp16 layers.  Here a work item covers a run of consecutive positions of a vertical STRIP of the image:
//   strip   = TW image columns plus one halo column on each side (padded width Wp = TW + 2)
//   position q of a strip <-> (row q / Wp, padded column q % Wp); columns 0 and Wp-1 are halo (their
//             outputs are computed and discarded: TW / Wp of the MMA rows are useful)
//   item    = MT tiles of 128 consecutive positions (MT * NT = 256 TMEM columns) x NT output channels
// For one channel block the TMA engine loads the box {kc channels, Wp columns, NR rows} that holds every
// input any tap of the item needs (out-of-image coordinates are zero-filled by the hardware: the conv's
// zero padding).  It lands as NR*Wp rows of RB bytes with the RB-byte swizzle, so the operand of tap
// (kh, kw) for tile t is simply the 128 rows starting (q0 - r_lo*Wp) + 128 t + kh*Wp + kw - 1 rows into
// the box: the UMMA descriptor's start address moves by whole rows (the swizzle is a function of the
// absolute shared-memory address -- tools/umma_shift_probe.cu checks this on the hardware for both swizzle
// widths).  Weights stream per (tap, channel block) through their own ring and are shared by the MT tiles.
// L2->SM bytes per MMA clock drop from ~110 to ~44 (fp16, NT=128) / ~29 (split).
//
// Precision: fp16 operands, fp32 accumulation in TMEM.  "split" layers carry
// activations and weights as hi + lo fp16 pairs and issue hi*hi + hi*lo + lo*hi; they use 32-channel
// blocks (64-byte rows, SWIZZLE_64B) so that two halo stages of both planes fit in shared memory.
//
// Warp roles (256 threads, one persistent CTA per SM):
//   warp 0 lane 0  halo (A) producer      warp 3 lane 0  weight (B) producer
//   warp 1 lane 0  MMA issuer (two accumulator sets of 256 TMEM columns, ping-pong)
//   warp 2         TMEM allocator         warps 4-7      epilogue (tcgen05.ld -> bias + ReLU -> NHWC fp16
//                                                        hi/lo stores, or the regressor's per-tap partials)
#include "model.cuh"
#include "tc_ptx.cuh"
#include <stdlib.h>

namespace gnd {

constexpr int HALO_THREADS = 256;
constexpr int HALO_MAX_SA = 2;
constexpr int HALO_MAX_SB = 8;
constexpr int HALO_NBARS = 2 * HALO_MAX_SA + 2 * HALO_MAX_SB + 4;
constexpr int HALO_RING_BUDGET = 222 * 1024;

struct HaloParams {
    int H, W, Cout, B;
    int Wp, TW, strips, MT, NT, n_ntiles;
    int tiles_per_strip, groups;
    int cblocks, kc;          // channel blocks of kc channels (64 or 32)
    int split, RB, ksteps;    // operand row bytes (128 / 64), K=16 steps per block (4 / 2)
    int SA, SB;
    uint32_t a_plane, a_stage, b_plane, b_stage, a_tx, b_tx;
    int total_items;
    const float *bias;
    __half *out_hi, *out_lo;
    const float *reg_w;       // [9][64] when the regressor is fused
    float *regpart;           // [B][H][W][9]
};

struct HaloLayer {
    bool usable = false;
    HaloParams p;
    size_t smem = 0;
    CUtensorMap a_hi, a_lo, b_hi, b_lo;
};

struct HaloMaps {
    HaloLayer layer[8];
};

struct HaloItem { int b, sx0, q0, r_lo, tiles, n0; };

#ifdef HALO_STATS
__device__ unsigned long long g_halo_stats[8];   // issuer: wait t_empty, wait a_full, wait b_full, total; epilogue: wait t_full, total
#define HSTAT_T0() long long _t0 = clock64()
#define HSTAT_ADD(i) do { long long _t1 = clock64(); st[i] += _t1 - _t0; } while (0)
#else
#define HSTAT_T0()
#define HSTAT_ADD(i)
#endif

__device__ __forceinline__ HaloItem halo_decode(const HaloParams &p, int item) {
    HaloItem it;
    const int nt = item % p.n_ntiles;
    int t = item / p.n_ntiles;
    const int g = t % p.groups; t /= p.groups;
    const int s = t % p.strips;
    it.b = t / p.strips;
    it.sx0 = s * p.TW;
    it.q0 = g * p.MT * 128;
    it.r_lo = it.q0 / p.Wp;
    it.tiles = min(p.MT, p.tiles_per_strip - g * p.MT);
    it.n0 = nt * p.NT;
    return it;
}

// K-major operand descriptor: rows of RB bytes, 8-row groups 8*RB bytes apart, RB-byte swizzle.
__device__ __forceinline__ uint64_t halo_desc_hi(int RB) {
    uint64_t d = 0;
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)((8 * RB) >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)(RB == 128 ? 2 : 4) << 61;      // SWIZZLE_128B / SWIZZLE_64B
    return d;
}
__device__ __forceinline__ uint64_t halo_desc(uint64_t hi, uint32_t saddr) { return hi | (uint64_t)((saddr & 0x3FFFF) >> 4); }

// All MMAs of one (tap, channel block): `tiles` accumulators x KSTEPS K=16 steps (x3 products when SPLIT).
template <int KSTEPS, bool SPLIT>
__device__ __forceinline__ void halo_issue_tap(uint32_t d0, int tiles, uint32_t NT, uint32_t a_lo, uint32_t a_tile_step,
                                               uint32_t a_plane16, uint32_t b_lo, uint32_t b_plane16, uint32_t dhi32,
                                               uint32_t idesc, uint32_t acc_on) {
    const uint64_t hi = (uint64_t)dhi32 << 32;
    for (int t = 0; t < tiles; ++t) {
        const uint32_t al = a_lo + (uint32_t)t * a_tile_step;
        const uint32_t d = d0 + (uint32_t)t * NT;
#pragma unroll
        for (int k = 0; k < KSTEPS; ++k) {
            const uint64_t da = hi | (uint64_t)(al + 2 * k), db = hi | (uint64_t)(b_lo + 2 * k);
            if (k == 0) tc_mma_f16(d, da, db, idesc, acc_on); else tc_mma_f16_on(d, da, db, idesc);
            if (SPLIT) {
                tc_mma_f16_on(d, da, hi | (uint64_t)(b_lo + b_plane16 + 2 * k), idesc);
                tc_mma_f16_on(d, hi | (uint64_t)(al + a_plane16 + 2 * k), db, idesc);
            }
        }
    }
}

__global__ void __launch_bounds__(HALO_THREADS, 1)
conv3x3_halo_kernel(const __grid_constant__ CUtensorMap tm_a_hi, const __grid_constant__ CUtensorMap tm_a_lo,
                    const __grid_constant__ CUtensorMap tm_b_hi, const __grid_constant__ CUtensorMap tm_b_lo,
                    const HaloParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t *ring_a = smem;
    uint8_t *ring_b = smem + (size_t)p.SA * p.a_stage;
    uint64_t *bars = reinterpret_cast<uint64_t *>(ring_b + (size_t)p.SB * p.b_stage);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + HALO_NBARS);
    float *s_bias = reinterpret_cast<float *>(tmem_slot + 4);           // [256]
    float *s_regw = s_bias + 256;                                        // [64][12] (fused regressor only)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar_afull = smem_u32(bars), bar_aempty = smem_u32(bars + HALO_MAX_SA);
    const uint32_t bar_bfull = smem_u32(bars + 2 * HALO_MAX_SA), bar_bempty = smem_u32(bars + 2 * HALO_MAX_SA + HALO_MAX_SB);
    const uint32_t bar_tfull = smem_u32(bars + 2 * HALO_MAX_SA + 2 * HALO_MAX_SB), bar_tempty = bar_tfull + 16;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_a_hi) : "memory");
        if (p.split) asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_a_lo) : "memory");
    }
    if (warp == 3 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_b_hi) : "memory");
        if (p.split) asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_b_lo) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < p.SA; ++s) { mbar_init(bar_afull + 8 * s, 1); mbar_init(bar_aempty + 8 * s, 1); }
        for (int s = 0; s < p.SB; ++s) { mbar_init(bar_bfull + 8 * s, 1); mbar_init(bar_bempty + 8 * s, 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(bar_tfull + 8 * a, 1); mbar_init(bar_tempty + 8 * a, 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (threadIdx.x >= 128) {
        for (int i = threadIdx.x - 128; i < p.Cout && i < 256; i += 128) s_bias[i] = p.bias[i];
        if (p.regpart)
            for (int i = threadIdx.x - 128; i < 576; i += 128) s_regw[(i & 63) * 12 + (i >> 6)] = p.reg_w[i];   // [tap][c] -> [c][tap]
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    // The CTA owns the SM (launch bounds 1, > half the shared memory) and allocates all 512 columns, so the
    // allocation starts at column 0: using the constant keeps every tcgen05 operand in uniform registers (a
    // TMEM address held in a vector register costs an ELECT/R2UR waterfall loop in front of every MMA).
    if (*tmem_slot != 0u) __trap();
    constexpr uint32_t tmem_base = 0u;

    if (warp == 0) {
        if (lane == 0) {
            // ===== halo producer: one box per (item, channel block) =====
            int sa = 0; uint32_t ph = 0;
            for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
                const HaloItem it = halo_decode(p, item);
                for (int cb = 0; cb < p.cblocks; ++cb) {
                    mbar_wait(bar_aempty + 8 * sa, ph ^ 1);
                    const uint32_t full = bar_afull + 8 * sa;
                    const uint32_t dst = smem_u32(ring_a + (size_t)sa * p.a_stage) + 1024;
                    mbar_expect_tx(full, p.a_tx);
                    tma_load_4d(dst, &tm_a_hi, full, cb * p.kc, it.sx0 - 1, it.r_lo - 1, it.b);
                    if (p.split) tma_load_4d(dst + p.a_plane, &tm_a_lo, full, cb * p.kc, it.sx0 - 1, it.r_lo - 1, it.b);
                    if (++sa == p.SA) { sa = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 3) {
        if (lane == 0) {
            // ===== weight producer: one box per (item, channel block, tap) =====
            int sb = 0; uint32_t ph = 0;
            for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
                const int n0 = (item % p.n_ntiles) * p.NT;
                for (int cb = 0; cb < p.cblocks; ++cb)
                    for (int tap = 0; tap < 9; ++tap) {
                        mbar_wait(bar_bempty + 8 * sb, ph ^ 1);
                        const uint32_t full = bar_bfull + 8 * sb;
                        const uint32_t dst = smem_u32(ring_b + (size_t)sb * p.b_stage);
                        mbar_expect_tx(full, p.b_tx);
                        tma_load_3d(dst, &tm_b_hi, full, cb * p.kc, n0, tap);
                        if (p.split) tma_load_3d(dst + p.b_plane, &tm_b_lo, full, cb * p.kc, n0, tap);
                        if (++sb == p.SB) { sb = 0; ph ^= 1; }
                    }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            // ===== MMA issuer =====
            const uint32_t idesc = (1u << 4) | ((uint32_t)(p.NT >> 3) << 17) | ((128u >> 4) << 24);   // F16 x F16 -> F32, K-major A and B
            const uint32_t dhi32 = (uint32_t)(halo_desc_hi(p.RB) >> 32);
            int sa = 0, sb = 0, acc = 0;
            uint32_t pha = 0, phb = 0, phacc = 0;
#ifdef HALO_STATS
            long long st[4] = {0, 0, 0, 0}; const long long tstart = clock64();
#endif
            for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
                const HaloItem it = halo_decode(p, item);
                const int off0 = it.q0 - it.r_lo * p.Wp - 1;          // rows from the box start; -1 falls in the slack row
                { HSTAT_T0(); mbar_wait(bar_tempty + 8 * acc, phacc ^ 1); HSTAT_ADD(0); }
                tc_fence_after();
                const uint32_t d0 = tmem_base + (uint32_t)(acc * 256);
                for (int cb = 0; cb < p.cblocks; ++cb) {
                    { HSTAT_T0(); mbar_wait(bar_afull + 8 * sa, pha); HSTAT_ADD(1); }
                    tc_fence_after();
                    const uint32_t a_base = smem_u32(ring_a + (size_t)sa * p.a_stage) + 1024;
                    for (int tap = 0; tap < 9; ++tap) {
                        const int kh = tap / 3, kw = tap - 3 * kh;
                        { HSTAT_T0(); mbar_wait(bar_bfull + 8 * sb, phb); HSTAT_ADD(2); }
                        tc_fence_after();
                        const uint32_t b_base = smem_u32(ring_b + (size_t)sb * p.b_stage);
                        const uint32_t acc_on = (cb | tap) != 0;
                        // descriptor low words ((addr >> 4) | LBO): row shifts, tile steps and K steps are plain adds
                        const uint32_t a_lo = ((a_base + (uint32_t)((off0 + kh * p.Wp + kw) * p.RB)) >> 4) | 0x10000u;
                        const uint32_t b_lo = (b_base >> 4) | 0x10000u;
                        if (p.split) halo_issue_tap<2, true>(d0, it.tiles, (uint32_t)p.NT, a_lo, 128u * 64u / 16u, p.a_plane >> 4, b_lo, p.b_plane >> 4, dhi32, idesc, acc_on);
                        else halo_issue_tap<4, false>(d0, it.tiles, (uint32_t)p.NT, a_lo, 128u * 128u / 16u, 0u, b_lo, 0u, dhi32, idesc, acc_on);
                        tc_commit(bar_bempty + 8 * sb);          // frees the weight slot when these MMAs retire
                        if (++sb == p.SB) { sb = 0; phb ^= 1; }
                    }
                    tc_commit(bar_aempty + 8 * sa);              // frees the halo slot
                    if (++sa == p.SA) { sa = 0; pha ^= 1; }
                }
                tc_commit(bar_tfull + 8 * acc);                  // accumulators complete -> epilogue
                if (++acc == 2) { acc = 0; phacc ^= 1; }
            }
#ifdef HALO_STATS
            st[3] = clock64() - tstart;
            for (int i = 0; i < 4; ++i) atomicAdd(&g_halo_stats[i], (unsigned long long)st[i]);
#endif
        }
    } else if (warp >= 4) {
        // ===== epilogue: TMEM -> registers -> bias + ReLU -> NHWC fp16 (hi, lo) or regressor partials =====
        const int qd = warp & 3;
        const int row = qd * 32 + lane;
        int acc = 0; uint32_t phacc = 0;
#ifdef HALO_STATS
        long long st[6] = {0, 0, 0, 0, 0, 0}; const long long tstart = clock64();
#endif
        for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
            const HaloItem it = halo_decode(p, item);
            { HSTAT_T0(); mbar_wait(bar_tfull + 8 * acc, phacc); HSTAT_ADD(4); }
            tc_fence_after();
            for (int t = 0; t < it.tiles; ++t) {
                const int q = it.q0 + t * 128 + row;
                const int y = q / p.Wp, col = q - y * p.Wp;
                const int x = it.sx0 - 1 + col;
                const bool valid = (col >= 1) && (col <= p.TW) && (x < p.W) && (y < p.H);
                const size_t pix = ((size_t)it.b * p.H + y) * p.W + x;
                const uint32_t taddr = tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)(acc * 256 + t * p.NT);
                if (p.regpart) {
                    float part[9];
#pragma unroll
                    for (int k = 0; k < 9; ++k) part[k] = 0.0f;
#pragma unroll 1
                    for (int j = 0; j < 2; ++j) {
                        uint32_t v[32];
                        tc_ld_32x32(taddr + 32 * j, v);
                        tc_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            const int c = 32 * j + i;
                            const float xv = fmaxf(__uint_as_float(v[i]) + s_bias[c], 0.0f);
                            const float4 w0 = *reinterpret_cast<const float4 *>(s_regw + c * 12);
                            const float4 w1 = *reinterpret_cast<const float4 *>(s_regw + c * 12 + 4);
                            const float w8 = s_regw[c * 12 + 8];
                            part[0] = fmaf(xv, w0.x, part[0]); part[1] = fmaf(xv, w0.y, part[1]);
                            part[2] = fmaf(xv, w0.z, part[2]); part[3] = fmaf(xv, w0.w, part[3]);
                            part[4] = fmaf(xv, w1.x, part[4]); part[5] = fmaf(xv, w1.y, part[5]);
                            part[6] = fmaf(xv, w1.z, part[6]); part[7] = fmaf(xv, w1.w, part[7]);
                            part[8] = fmaf(xv, w8, part[8]);
                        }
                    }
                    if (valid) {
#pragma unroll
                        for (int k = 0; k < 9; ++k) p.regpart[pix * 9 + k] = part[k];
                    }
                } else {
#pragma unroll 1
                    for (int j = 0; j < p.NT / 32; ++j) {
                        uint32_t v[32];
                        tc_ld_32x32(taddr + 32 * j, v);
                        tc_ld_wait();
                        uint32_t hi[16], lo[16];
                        const float *bj = s_bias + it.n0 + 32 * j;
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const float x0 = fmaxf(__uint_as_float(v[2 * i]) + bj[2 * i], 0.0f);
                            const float x1 = fmaxf(__uint_as_float(v[2 * i + 1]) + bj[2 * i + 1], 0.0f);
                            const __half2 hh = __floats2half2_rn(x0, x1);
                            hi[i] = *reinterpret_cast<const uint32_t *>(&hh);
                            const float2 back = __half22float2(hh);
                            lo[i] = pack_half2(x0 - back.x, x1 - back.y);
                        }
                        if (valid) {
                            const size_t o = pix * p.Cout + it.n0 + 32 * j;
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
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_tempty + 8 * acc);
            if (++acc == 2) { acc = 0; phacc ^= 1; }
        }
#ifdef HALO_STATS
        if (threadIdx.x == 128) { atomicAdd(&g_halo_stats[4], (unsigned long long)st[4]); atomicAdd(&g_halo_stats[5], (unsigned long long)(clock64() - tstart)); }
#endif
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// ---- host side --------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn halo_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *q = nullptr;
        cudaDriverEntryPointQueryResult r;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &q, cudaEnableDefault, &r) == cudaSuccess && r == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)q;
    }
    return fn;
}

static int halo_encode(CUtensorMap *m, const __half *ptr, int rank, const cuuint64_t *dims, const cuuint64_t *strides,
                       const cuuint32_t *box, int RB, const char *what) {
    EncodeTiledFn enc = halo_encode_fn();
    if (!enc) return fail(GND_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, (cuuint32_t)rank, (void *)ptr, dims, strides, box, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, RB == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(GND_ERR_CUDA, "cuTensorMapEncodeTiled(%s) failed: %d", what, (int)r);
    return GND_OK;
}

static inline uint32_t up1024(uint32_t x) { return (x + 1023u) & ~1023u; }

// Picks the strip width: fewest 128-position tiles over the image, then the widest strip, among the
// geometries whose two halo stages leave room for at least `min_sb` weight stages.
static bool halo_geometry(int H, int W, int RB, int planes, int MT, int NT, int min_sb, HaloParams *p) {
    long best_tiles = -1;
    const int L = 128 * MT;
    for (int TW = (W < 254 ? W : 254); TW >= 6; --TW) {
        const int Wp = TW + 2;
        const int strips = (W + TW - 1) / TW;
        const int tps = (H * Wp + 127) / 128;
        const int NR = (L + Wp - 2) / Wp + 1 + 2;
        if (NR > 256) continue;
        const uint32_t a_plane = up1024(1024u + (uint32_t)(NR * Wp * RB) + (uint32_t)RB);
        const uint32_t a_stage = a_plane * planes;
        const uint32_t b_plane = (uint32_t)(NT * RB), b_stage = b_plane * planes;
        const long room = (long)HALO_RING_BUDGET - 2L * a_stage;
        if (room < (long)min_sb * b_stage) continue;
        const long tiles = (long)strips * tps;
        if (best_tiles < 0 || tiles < best_tiles) {
            best_tiles = tiles;
            int sb = (int)(room / b_stage);
            if (sb > HALO_MAX_SB) sb = HALO_MAX_SB;
            p->Wp = Wp; p->TW = TW; p->strips = strips; p->tiles_per_strip = tps;
            p->groups = (tps + MT - 1) / MT;
            p->SA = 2; p->SB = sb;
            p->a_plane = a_plane; p->a_stage = a_stage; p->b_plane = b_plane; p->b_stage = b_stage;
            p->a_tx = (uint32_t)(NR * Wp * RB) * planes; p->b_tx = b_stage;
            p->H = NR;   // stashed; overwritten by the caller
        }
    }
    return best_tiles >= 0;
}

int halo_build(gnd_ctx *ctx, const Act *const inputs[8]) {
    if (!ctx->hmaps) ctx->hmaps = new HaloMaps();
    for (int l = 0; l < 8; ++l) {
        HaloLayer &hl = ctx->hmaps->layer[l];
        hl.usable = false;
        const ConvLayer &L = ctx->conv[l];
        const Act &in = *inputs[l];
        if (L.cin % 64 != 0 || L.cout % 64 != 0 || !in.hi || (L.split && !in.lo)) continue;
        // Cout >= 256: the tap-per-load kernel issues N=256 MMAs (96 B/clk of operand reads from shared memory
        // instead of 128 at N=128) and measures faster on the B200; the halo kernel takes the N <= 128 layers.
        if (L.cout > 128 && !getenv("GNDNET_HALO_ALL")) continue;
        HaloParams p = {};
        p.split = L.split;
        p.RB = L.split ? 64 : 128;
        p.kc = p.RB / 2; p.ksteps = p.RB / 32;
        p.cblocks = L.cin / p.kc;
        p.NT = L.cout < 128 ? 64 : 128;
        if (L.cout % p.NT != 0) continue;
        p.MT = 256 / p.NT;
        p.n_ntiles = L.cout / p.NT;
        const int planes = L.split ? 2 : 1;
        if (!halo_geometry(in.H, in.W, p.RB, planes, p.MT, p.NT, 5, &p) &&
            !halo_geometry(in.H, in.W, p.RB, planes, p.MT, p.NT, 3, &p))
            continue;
        const int NR = p.H;
        p.H = in.H; p.W = in.W; p.Cout = L.cout; p.B = ctx->B;
        p.bias = L.bias;
        hl.smem = (size_t)p.SA * p.a_stage + (size_t)p.SB * p.b_stage + 1024 + HALO_NBARS * 8 + 16 + 256 * 4 + 768 * 4;
        {
            cuuint64_t dims[4] = {(cuuint64_t)in.C, (cuuint64_t)in.W, (cuuint64_t)in.H, (cuuint64_t)ctx->B};
            cuuint64_t strides[3] = {(cuuint64_t)in.C * 2, (cuuint64_t)in.W * in.C * 2, (cuuint64_t)in.H * in.W * in.C * 2};
            cuuint32_t box[4] = {(cuuint32_t)p.kc, (cuuint32_t)p.Wp, (cuuint32_t)NR, 1};
            int rc = halo_encode(&hl.a_hi, in.hi, 4, dims, strides, box, p.RB, "halo activation");
            if (rc != GND_OK) return rc;
            rc = halo_encode(&hl.a_lo, L.split ? in.lo : in.hi, 4, dims, strides, box, p.RB, "halo activation lo");
            if (rc != GND_OK) return rc;
        }
        {
            cuuint64_t dims[3] = {(cuuint64_t)L.cin, (cuuint64_t)L.cout, 9};
            cuuint64_t strides[2] = {(cuuint64_t)L.cin * 2, (cuuint64_t)L.cout * L.cin * 2};
            cuuint32_t box[3] = {(cuuint32_t)p.kc, (cuuint32_t)p.NT, 1};
            int rc = halo_encode(&hl.b_hi, L.whi, 3, dims, strides, box, p.RB, "halo weights");
            if (rc != GND_OK) return rc;
            rc = halo_encode(&hl.b_lo, L.wlo, 3, dims, strides, box, p.RB, "halo weights lo");
            if (rc != GND_OK) return rc;
        }
        hl.p = p;
        hl.usable = true;
        if (getenv("GNDNET_DEBUG"))
            fprintf(stderr, "halo layer %d: %dx%d cin %d cout %d split %d | Wp %d strips %d tiles/strip %d MT %d NT %d NR %d SA %d SB %d smem %zu\n",
                    l, in.H, in.W, L.cin, L.cout, L.split, p.Wp, p.strips, p.tiles_per_strip, p.MT, p.NT, NR, p.SA, p.SB, hl.smem);
    }
    return GND_OK;
}

void halo_free(gnd_ctx *ctx) {
    delete ctx->hmaps;
    ctx->hmaps = nullptr;
}

bool halo_usable(const gnd_ctx *ctx, int l) {
    static const char *env = getenv("GNDNET_CONV");        // "tap" forces the tap-per-load kernel (A/B timing)
    if (env && env[0] == 't') return false;
    return ctx->hmaps && ctx->hmaps->layer[l].usable;
}

int halo_conv_run(gnd_ctx *ctx, int l, const Act &out, int B, bool fuse_regressor, cudaStream_t st) {
    const HaloLayer &hl = ctx->hmaps->layer[l];
    HaloParams p = hl.p;
    p.B = B;
    p.total_items = B * p.strips * p.groups * p.n_ntiles;
    p.out_hi = fuse_regressor ? nullptr : out.hi;
    p.out_lo = fuse_regressor ? nullptr : out.lo;
    p.reg_w = ctx->reg_w;
    p.regpart = fuse_regressor ? ctx->regpart : nullptr;
    if (fuse_regressor && p.Cout != 64) return fail(GND_ERR_UNSUPPORTED, "regressor fusion needs a 64-channel last block");
    if (!fuse_regressor && !out.hi) return fail(GND_ERR_STATE, "layer %d: output activation missing", l);
    GND_CUDA_CHECK(cudaFuncSetAttribute(conv3x3_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hl.smem));
    const int grid = p.total_items < 148 ? p.total_items : 148;
    conv3x3_halo_kernel<<<grid, HALO_THREADS, hl.smem, st>>>(hl.a_hi, hl.a_lo, hl.b_hi, hl.b_lo, p);
    GND_LAUNCH_CHECK("conv3x3_halo_kernel");
#ifdef HALO_STATS
    if (getenv("GNDNET_DEBUG")) {
        unsigned long long h[8];
        cudaStreamSynchronize(st);
        cudaMemcpyFromSymbol(h, g_halo_stats, sizeof(h));
        const double g = (double)grid;
        fprintf(stderr, "halo layer %d: per-CTA clks: issuer total %.0f wait_tempty %.0f wait_afull %.0f wait_bfull %.0f | epilogue total %.0f wait_tfull %.0f | items/CTA %.1f\n",
                l, h[3] / g, h[0] / g, h[1] / g, h[2] / g, h[5] / g, h[4] / g, p.total_items / g);
        unsigned long long z[8] = {0};
        cudaMemcpyToSymbol(g_halo_stats, z, sizeof(z));
    }
#endif
    return GND_OK;
}

}  // namespace gnd
