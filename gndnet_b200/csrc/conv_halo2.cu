// conv_halo2.cu -- 3x3 convolution + folded BatchNorm + ReLU (conv2DBatchNormRelu, modules/segnet.py:11-44) as an
// implicit GEMM on tcgen05 with the input HALO staged once per channel block, on a CTA PAIR
// (tcgen05.mma.cta_group::2, M = 256 = 128 positions from each CTA, the weight tile split between the two CTAs'
// shared memories).  The last block's epilogue also carries the regressor's channel contraction (segnet.py:131,140).
//
// Halo scheme.  A work item covers a run of consecutive positions of a vertical STRIP of the image:
//   strip    = TW image columns plus one halo column on each side (padded width Wp = TW + 2)
//   position q of a strip <-> (row q / Wp, padded column q % Wp); columns 0 and Wp-1 are halo: their outputs are
//            computed and discarded (TW / Wp of the MMA rows are useful)
//   half-item = MT tiles of 128 consecutive positions x NT output channels (MT * NT = 256 TMEM columns)
// For one channel block ONE TMA box {kc channels, Wp columns, NR rows, 1 scan} of the NHWC activation lands in
// shared memory (coordinates outside the image are zero-filled by the TMA unit: the conv's zero padding) as NR*Wp
// rows of RB bytes with the RB-byte swizzle.  The A operand of tap (kh, kw) for tile t is then simply the 128 rows
// that start (q0 - r_lo*Wp) + 128 t + kh*Wp + kw - 1 rows into the box: nine UMMA descriptors into the same bytes,
// their start address moved by whole rows (the swizzle is a function of the absolute shared-memory address --
// tools/umma_shift_probe.cu checks this on the hardware for both swizzle widths).  Activations come from L2 once
// per channel block instead of nine times.  "split" layers (hi + lo fp16 planes, three products per K step:
// DESIGN.md section 3) use 32-channel blocks (64-byte rows, SWIZZLE_64B) so two halo stages of both planes fit.
//
// Why a pair: an M128 x N128 x K16 MMA reads 4 KB of A and 4 KB of B from shared memory in its 64 clocks --
// the whole 128 B/clk of the SM -- so every byte TMA writes into shared memory slows the tensor pipe down
// (measured on the single-CTA predecessor of this kernel: 84-97 clk per MMA against 64.6 in isolation).  With
// cta_group::2 each SM reads its own 4 KB of A but only HALF of B (2 KB), 96 B/clk, and loads half as many weight
// bytes from L2.
//
// The two CTAs of a pair work on two INDEPENDENT half-items (any scan / strip / position range, same output-
// channel tile): the only things the pair instruction shares are the weight tile and the operand OFFSETS, so
// every CTA places its halo box such that the operand of (tile 0, tap (0,0)) starts at the same fixed row
// `row0` of the stage (the TMA destination moves by whole 128-byte units; its swizzle, like UMMA's, is a
// function of the absolute shared-memory address -- tools/umma_shift_probe.cu -DDSTK=k).
//
// Barriers (all in shared memory, same offsets in both CTAs):
//   a_full / b_full   in the LEADER (cluster rank 0): both CTAs' TMA loads complete_tx there; count 1 (the leader's
//                     arrive.expect_tx covers both boxes; the peer only issues its load -- it cannot run a phase
//                     ahead because its slot is freed by the same commit that follows the phase's MMAs.  A remote
//                     arrive.release.cluster per stage from the peer cost a cluster-scope fence each, ~1300 clk)
//   a_empty / b_empty one per CTA, arrived by the leader's tcgen05.commit multicast to both CTAs
//   t_full            one per CTA (commit multicast);  t_empty in the leader, count 8 (4 epilogue warps x 2 CTAs)
//
// Warp roles per CTA (256 threads, one persistent CTA per SM): warp 0 lane 0 halo (A) producer, warp 3 lane 0 weight (B)
// producer, warp 1 lane 0 of the leader MMA issuer (two accumulator sets of 256 TMEM columns, ping-pong), warp 2 TMEM
// allocator, warps 4-7 epilogue (tcgen05.ld -> bias + ReLU -> NHWC fp16 hi/lo stores, or the regressor head).
#include "model.cuh"
#include "tc_ptx.cuh"
#include <stdlib.h>

namespace gnd {

constexpr int H2_THREADS = 256;
constexpr int H2_SA = 2;
constexpr int H2_MAX_SB = 8;
constexpr int H2_NBARS = 2 * H2_SA + 2 * H2_MAX_SB + 4;
#ifndef H2_RING_KB
#define H2_RING_KB 222
#endif
#ifndef H2_MINBLOCKS
#define H2_MINBLOCKS 1          // keep 1: at 255 registers x 256 threads a CTA owns its SM, so two pair kernels from
                                // different streams can never be co-resident.  With 2 (128 registers) they can, and then
                                // deadlock: each holds the 512 TMEM columns of one SM of the other's CTA pair while
                                // blocking in tcgen05.alloc for its own (measured: a two-context run hung; the convs
                                // were also 12 % slower with the 120 KB ring that co-residency would need)
#endif
constexpr int H2_RING_BUDGET = H2_RING_KB * 1024;

struct Halo2Params {
    int H, W, Cout, B;
    int Wp, TW, strips, MT, NT, n_ntiles;
    int tiles_per_strip, groups;
    int cblocks, kc, split, RB, ksteps;
    int SB, row0;
    uint32_t a_plane, a_stage, b_plane, b_stage, a_tx, b_tx;
    int half_items, pair_items;
    int rev;                 // 1: items are walked from the last scan to the first
    const float *bias;
    __half *out_hi, *out_lo;
    const float *reg_w;
    float *regpart;
};

struct Halo2Layer {
    bool usable = false;
    Halo2Params p;
    size_t smem = 0;
    CUtensorMap a_hi, a_lo, b_hi, b_lo;
};

struct Halo2Maps {
    Halo2Layer layer[8];     // throughput geometry: the widest N tile, 512 TMEM columns in use
    Halo2Layer small[8];     // latency geometry for small batches: N tile of 64 (or $GNDNET_H2_SMALL_NT), one M tile per item,
                             // so that a single scan already spreads over most of the 74 CTA pairs
};

struct Half { int b, sx0, q0, r_lo, tiles; };

#ifdef HALO_STATS
__device__ unsigned long long g_h2_stats[8];
#define H2T0() long long _t0 = clock64()
#define H2ADD(i) do { st[i] += clock64() - _t0; } while (0)
#else
#define H2T0()
#define H2ADD(i)
#endif

__device__ __forceinline__ Half h2_half(const Halo2Params &p, int h) {
    Half it;
    const int g = h % p.groups;
    int t = h / p.groups;
    const int s = t % p.strips;
    it.b = t / p.strips;
    it.sx0 = s * p.TW;
    it.q0 = g * p.MT * 128;
    it.r_lo = it.q0 / p.Wp;
    it.tiles = min(p.MT, p.tiles_per_strip - g * p.MT);
    return it;
}

// ---- cluster / pair PTX ----------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}
// TMA loads of a CTA pair: destination in the executing CTA, completion on the given (cluster) mbarrier address
__device__ __forceinline__ void tma2_load_4d(uint32_t dst, const CUtensorMap *map, uint32_t bar_cluster, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma2_load_3d(uint32_t dst, const CUtensorMap *map, uint32_t bar_cluster, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tc2_commit(uint32_t bar) {      // arrives on `bar` of BOTH CTAs when the prior MMAs retire
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void tc2_mma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc2_mma_f16_on(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.eq.u32 p, 0, 0;\n"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc) : "memory");
}

// All pair MMAs of one (tap, channel block): `tiles` accumulators x KSTEPS K=16 steps (x3 products when SPLIT).
template <int KSTEPS, bool SPLIT>
__device__ __forceinline__ void h2_issue_tap(uint32_t d0, int tiles, uint32_t NT, uint32_t a_lo, uint32_t a_tile_step,
                                             uint32_t a_plane16, uint32_t b_lo, uint32_t b_plane16, uint32_t dhi32,
                                             uint32_t idesc, uint32_t acc_on) {
    const uint64_t hi = (uint64_t)dhi32 << 32;
    for (int t = 0; t < tiles; ++t) {
        const uint32_t al = a_lo + (uint32_t)t * a_tile_step;
        const uint32_t d = d0 + (uint32_t)t * NT;
#pragma unroll
        for (int k = 0; k < KSTEPS; ++k) {
            const uint64_t da = hi | (uint64_t)(al + 2 * k), db = hi | (uint64_t)(b_lo + 2 * k);
            if (k == 0) tc2_mma_f16(d, da, db, idesc, acc_on); else tc2_mma_f16_on(d, da, db, idesc);
            if (SPLIT) {
                tc2_mma_f16_on(d, da, hi | (uint64_t)(b_lo + b_plane16 + 2 * k), idesc);
                tc2_mma_f16_on(d, hi | (uint64_t)(al + a_plane16 + 2 * k), db, idesc);
            }
        }
    }
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(H2_THREADS, H2_MINBLOCKS)
conv3x3_halo2_kernel(const __grid_constant__ CUtensorMap tm_a_hi, const __grid_constant__ CUtensorMap tm_a_lo,
                     const __grid_constant__ CUtensorMap tm_b_hi, const __grid_constant__ CUtensorMap tm_b_lo,
                     const Halo2Params p) {
#ifdef HALO_STATS
    const long long t_entry = clock64();
#endif
    extern __shared__ uint8_t smem_raw[];
    // pointer arithmetic on the shared array (not an integer round trip): the compiler keeps the shared address space,
    // so s_bias / s_regw reads are LDS and not generic loads
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t *ring_a = smem;
    uint8_t *ring_b = smem + (size_t)H2_SA * p.a_stage;
    uint64_t *bars = reinterpret_cast<uint64_t *>(ring_b + (size_t)p.SB * p.b_stage);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + H2_NBARS);
    float *s_bias = reinterpret_cast<float *>(tmem_slot + 4);           // [256]
    float *s_regw = s_bias + 256;                                        // [32 channel pairs][20] (fused regressor only)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_rank();
    const bool leader = rank == 0;
    const uint32_t bar_afull = smem_u32(bars), bar_aempty = smem_u32(bars + H2_SA);
    const uint32_t bar_bfull = smem_u32(bars + 2 * H2_SA), bar_bempty = smem_u32(bars + 2 * H2_SA + H2_MAX_SB);
    const uint32_t bar_tfull = smem_u32(bars + 2 * H2_SA + 2 * H2_MAX_SB), bar_tempty = bar_tfull + 16;
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_a_hi) : "memory");
        if (p.split) asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_a_lo) : "memory");
    }
    if (warp == 3 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_b_hi) : "memory");
        if (p.split) asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_b_lo) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < H2_SA; ++s) { mbar_init(bar_afull + 8 * s, 1); mbar_init(bar_aempty + 8 * s, 1); }
        for (int s = 0; s < p.SB; ++s) { mbar_init(bar_bfull + 8 * s, 1); mbar_init(bar_bempty + 8 * s, 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(bar_tfull + 8 * a, 1); mbar_init(bar_tempty + 8 * a, 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    if (threadIdx.x >= 128) {
        for (int i = threadIdx.x - 128; i < p.Cout && i < 256; i += 128) s_bias[i] = p.bias[i];
        if (p.regpart)
            // [tap][c] -> per channel PAIR (c, c + 1): nine (w[tap][c], w[tap][c + 1]) pairs, then the pair of conv biases:
            // five 16-byte rows whose halves are the register pairs of the packed fma.rn.f32x2 below
            for (int i = threadIdx.x - 128; i < 640; i += 128) {
                const int c = (i < 576) ? (i & 63) : i - 576, slot = (i < 576) ? (i >> 6) : 9;
                s_regw[(c >> 1) * 20 + slot * 2 + (c & 1)] = (i < 576) ? p.reg_w[i] : p.bias[c];
            }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();          // barrier inits of both CTAs visible before any remote arrive / multicast commit
    tc_fence_after();
    // each CTA owns its SM and all 512 columns: the allocation starts at column 0 (keeps tcgen05 operands uniform)
    if (*tmem_slot != 0u) __trap();
    constexpr uint32_t tmem_base = 0u;

    if (warp == 0) {
        if (elect_one()) {
            // ===== halo producer (both CTAs): one box per (pair item, channel block), completion on the leader =====
            int sa = 0; uint32_t ph = 0;
            for (int it0 = pair; it0 < p.pair_items; it0 += n_pairs) {
                const int item = p.rev ? p.pair_items - 1 - it0 : it0;     // scan order of this launch (see halo2_conv_run)
                const Half it = h2_half(p, min(2 * (item / p.n_ntiles) + (int)rank, p.half_items - 1));
                const int off = it.q0 - it.r_lo * p.Wp - 1;
                for (int cb = 0; cb < p.cblocks; ++cb) {
                    mbar_wait(bar_aempty + 8 * sa, ph ^ 1);
                    const uint32_t full = map_to_cta(bar_afull + 8 * sa, 0);
                    const uint32_t dst = smem_u32(ring_a + (size_t)sa * p.a_stage) + (uint32_t)((p.row0 - off) * p.RB);
                    if (leader) mbar_expect_tx(bar_afull + 8 * sa, 2 * p.a_tx);
                    tma2_load_4d(dst, &tm_a_hi, full, cb * p.kc, it.sx0 - 1, it.r_lo - 1, it.b);
                    if (p.split) tma2_load_4d(dst + p.a_plane, &tm_a_lo, full, cb * p.kc, it.sx0 - 1, it.r_lo - 1, it.b);
                    if (++sa == H2_SA) { sa = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 3) {
        if (elect_one()) {
            // ===== weight producer (both CTAs): this CTA's half of the NT rows per (pair item, channel block, tap) =====
            int sb = 0; uint32_t ph = 0;
            for (int it0 = pair; it0 < p.pair_items; it0 += n_pairs) {
                const int item = p.rev ? p.pair_items - 1 - it0 : it0;     // scan order of this launch (see halo2_conv_run)
                const int n0 = (item % p.n_ntiles) * p.NT + (int)rank * (p.NT / 2);
                for (int cb = 0; cb < p.cblocks; ++cb)
                    for (int tap = 0; tap < 9; ++tap) {
                        mbar_wait(bar_bempty + 8 * sb, ph ^ 1);
                        const uint32_t full = map_to_cta(bar_bfull + 8 * sb, 0);
                        const uint32_t dst = smem_u32(ring_b + (size_t)sb * p.b_stage);
                        if (leader) mbar_expect_tx(bar_bfull + 8 * sb, 2 * p.b_tx);
                        tma2_load_3d(dst, &tm_b_hi, full, cb * p.kc, n0, tap);
                        if (p.split) tma2_load_3d(dst + p.b_plane, &tm_b_lo, full, cb * p.kc, n0, tap);
                            if (++sb == p.SB) { sb = 0; ph ^= 1; }
                    }
            }
        }
    } else if (warp == 1) {
        if (leader && elect_one()) {
            // ===== MMA issuer (leader only): M = 256 = 128 positions of each CTA =====
            const uint32_t idesc = (1u << 4) | ((uint32_t)(p.NT >> 3) << 17) | ((256u >> 4) << 24);   // F16 x F16 -> F32, K-major A and B
            const uint32_t dhi32 = (uint32_t)(((uint64_t)((8 * p.RB) >> 4)) | ((uint64_t)1 << 14) | ((uint64_t)(p.RB == 128 ? 2 : 4) << 29));
            int sa = 0, sb = 0, acc = 0;
            uint32_t pha = 0, phb = 0, phacc = 0;
#ifdef HALO_STATS
            long long st[4] = {0, 0, 0, 0}; const long long tstart = clock64();
            atomicAdd(&g_h2_stats[4], (unsigned long long)(tstart - t_entry));      // prologue: entry -> issuer ready
#endif
            for (int it0 = pair; it0 < p.pair_items; it0 += n_pairs) {
                const int item = p.rev ? p.pair_items - 1 - it0 : it0;     // scan order of this launch (see halo2_conv_run)
                const int j = item / p.n_ntiles;
                const int tiles = max(h2_half(p, min(2 * j, p.half_items - 1)).tiles, h2_half(p, min(2 * j + 1, p.half_items - 1)).tiles);
                { H2T0(); mbar_wait_cluster(bar_tempty + 8 * acc, phacc ^ 1); H2ADD(0); }
                tc_fence_after();
                const uint32_t d0 = tmem_base + (uint32_t)(acc * 256);
                for (int cb = 0; cb < p.cblocks; ++cb) {
                    { H2T0(); mbar_wait(bar_afull + 8 * sa, pha); H2ADD(1); }
                    tc_fence_after();
                    const uint32_t a_base = smem_u32(ring_a + (size_t)sa * p.a_stage);
                    for (int tap = 0; tap < 9; ++tap) {
                        const int kh = tap / 3, kw = tap - 3 * kh;
                        { H2T0(); mbar_wait(bar_bfull + 8 * sb, phb); H2ADD(2); }
                        tc_fence_after();
                        const uint32_t b_base = smem_u32(ring_b + (size_t)sb * p.b_stage);
                        const uint32_t acc_on = (cb | tap) != 0;
                        const uint32_t a_lo = ((a_base + (uint32_t)((p.row0 + kh * p.Wp + kw) * p.RB)) >> 4) | 0x10000u;
                        const uint32_t b_lo = (b_base >> 4) | 0x10000u;
                        if (p.split) h2_issue_tap<2, true>(d0, tiles, (uint32_t)p.NT, a_lo, 128u * 64u / 16u, p.a_plane >> 4, b_lo, p.b_plane >> 4, dhi32, idesc, acc_on);
                        else h2_issue_tap<4, false>(d0, tiles, (uint32_t)p.NT, a_lo, 128u * 128u / 16u, 0u, b_lo, 0u, dhi32, idesc, acc_on);
                        tc2_commit(bar_bempty + 8 * sb);         // frees the weight slot in both CTAs
                        if (++sb == p.SB) { sb = 0; phb ^= 1; }
                    }
                    tc2_commit(bar_aempty + 8 * sa);             // frees the halo slot in both CTAs
                    if (++sa == H2_SA) { sa = 0; pha ^= 1; }
                }
                tc2_commit(bar_tfull + 8 * acc);                 // accumulators complete -> both epilogues
                if (++acc == 2) { acc = 0; phacc ^= 1; }
            }
#ifdef HALO_STATS
            st[3] = clock64() - tstart;
            for (int i = 0; i < 4; ++i) atomicAdd(&g_h2_stats[i], (unsigned long long)st[i]);
#endif
        }
    } else if (warp >= 4) {
        // ===== epilogue (both CTAs): own TMEM rows -> bias + ReLU -> NHWC fp16 (hi, lo) or regressor partials =====
        const int qd = warp & 3;
        const int row = qd * 32 + lane;
        int acc = 0; uint32_t phacc = 0;
        for (int it0 = pair; it0 < p.pair_items; it0 += n_pairs) {
                const int item = p.rev ? p.pair_items - 1 - it0 : it0;     // scan order of this launch (see halo2_conv_run)
            const Half it = h2_half(p, min(2 * (item / p.n_ntiles) + (int)rank, p.half_items - 1));
            const int n0 = (item % p.n_ntiles) * p.NT;
            mbar_wait(bar_tfull + 8 * acc, phacc);
            tc_fence_after();
            for (int t = 0; t < it.tiles; ++t) {
                const int q = it.q0 + t * 128 + row;
                const int y = q / p.Wp, col = q - y * p.Wp;
                const int x = it.sx0 - 1 + col;
                const bool valid = (col >= 1) && (col <= p.TW) && (x < p.W) && (y < p.H);
                const size_t pix = ((size_t)it.b * p.H + y) * p.W + x;
                const uint32_t taddr = tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)(acc * 256 + t * p.NT);
                if (p.regpart) {
                    // two channels per step through the packed FP32 pipe: 9 FFMA2 + 1 FADD2 + 2 FMNMX + 5 LDS.128 per channel PAIR
                    // (was 18 FFMA + 2 FADD + 2 FMNMX + 6 LDS.128); even and odd channels accumulate apart and meet at the end
                    float2 acc[9];
#pragma unroll
                    for (int k = 0; k < 9; ++k) acc[k] = make_float2(0.0f, 0.0f);
#pragma unroll 1
                    for (int j = 0; j < 2; ++j) {
                        uint32_t v[32];
                        tc_ld_32x32(taddr + 32 * j, v);
                        tc_ld_wait();
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const float4 *w = reinterpret_cast<const float4 *>(s_regw + (16 * j + i) * 20);
                            const float4 q0 = w[0], q1 = w[1], q2 = w[2], q3 = w[3], q4 = w[4];       // q4 = (tap 8 pair, bias pair)
                            float2 x = __fadd2_rn(make_float2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1])), make_float2(q4.z, q4.w));
                            x.x = fmaxf(x.x, 0.0f); x.y = fmaxf(x.y, 0.0f);
                            acc[0] = __ffma2_rn(x, make_float2(q0.x, q0.y), acc[0]); acc[1] = __ffma2_rn(x, make_float2(q0.z, q0.w), acc[1]);
                            acc[2] = __ffma2_rn(x, make_float2(q1.x, q1.y), acc[2]); acc[3] = __ffma2_rn(x, make_float2(q1.z, q1.w), acc[3]);
                            acc[4] = __ffma2_rn(x, make_float2(q2.x, q2.y), acc[4]); acc[5] = __ffma2_rn(x, make_float2(q2.z, q2.w), acc[5]);
                            acc[6] = __ffma2_rn(x, make_float2(q3.x, q3.y), acc[6]); acc[7] = __ffma2_rn(x, make_float2(q3.z, q3.w), acc[7]);
                            acc[8] = __ffma2_rn(x, make_float2(q4.x, q4.y), acc[8]);
                        }
                    }
                    if (valid) {
#pragma unroll
                        for (int k = 0; k < 9; ++k) p.regpart[pix * 9 + k] = acc[k].x + acc[k].y;
                    }
                } else {
#pragma unroll 1
                    for (int j = 0; j < p.NT / 32; ++j) {
                        uint32_t v[32];
                        tc_ld_32x32(taddr + 32 * j, v);
                        tc_ld_wait();
                        uint32_t hi[16], lo[16];
                        const float4 *bj = reinterpret_cast<const float4 *>(s_bias + n0 + 32 * j);       // 16-byte aligned
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const float4 b4 = bj[i >> 1];
                            const float x0 = fmaxf(__uint_as_float(v[2 * i]) + ((i & 1) ? b4.z : b4.x), 0.0f);
                            const float x1 = fmaxf(__uint_as_float(v[2 * i + 1]) + ((i & 1) ? b4.w : b4.y), 0.0f);
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
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(map_to_cta(bar_tempty + 8 * acc, 0));
            if (++acc == 2) { acc = 0; phacc ^= 1; }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();          // the peer's shared memory and barriers stay alive until the leader's last MMA / commit
#ifdef HALO_STATS
    if (leader && threadIdx.x == 0) atomicAdd(&g_h2_stats[5], (unsigned long long)(clock64() - t_entry));   // whole CTA lifetime
#endif
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// ---- host side --------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn2)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                   const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int h2_encode(CUtensorMap *m, const __half *ptr, int rank, const cuuint64_t *dims, const cuuint64_t *strides,
                     const cuuint32_t *box, int RB, const char *what) {
    static EncodeTiledFn2 enc = nullptr;
    if (!enc) {
        void *q = nullptr;
        cudaDriverEntryPointQueryResult r;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &q, cudaEnableDefault, &r) == cudaSuccess && r == cudaDriverEntryPointSuccess)
            enc = (EncodeTiledFn2)q;
    }
    if (!enc) return fail(GND_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, (cuuint32_t)rank, (void *)ptr, dims, strides, box, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, RB == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(GND_ERR_CUDA, "cuTensorMapEncodeTiled(%s) failed: %d", what, (int)r);
    return GND_OK;
}

static inline uint32_t h2_up1024(uint32_t x) { return (x + 1023u) & ~1023u; }

// Strip width: fewest 128-position tiles over the image (counting the half-item granularity of MT tiles),
// then the widest strip, among the geometries that leave room for `min_sb` weight stages.
static bool h2_geometry(int H, int W, int RB, int planes, int MT, int NT, int min_sb, Halo2Params *p, int *NR_out) {
    long best = -1;
    const int L = 128 * MT;
    for (int TW = (W < 254 ? W : 254); TW >= 6; --TW) {
        const int Wp = TW + 2;
        if (RB == 64 && (Wp & 1)) continue;                   // 64-byte rows: the box must move by even row counts
        const int strips = (W + TW - 1) / TW;
        const int tps = (H * Wp + 127) / 128;
        const int groups = (tps + MT - 1) / MT;
        const int NR = (L + Wp - 2) / Wp + 1 + 2;
        if (NR > 256 || Wp > 256) continue;
        const int row0 = (RB == 64) ? Wp + 1 : Wp;
        const uint32_t a_plane = h2_up1024((uint32_t)((row0 + NR * Wp + 4) * RB));
        const uint32_t a_stage = a_plane * planes;
        const uint32_t b_plane = (uint32_t)((NT / 2) * RB), b_stage = b_plane * planes;
        const long room = (long)H2_RING_BUDGET - (long)H2_SA * a_stage;
        if (room < (long)min_sb * b_stage) continue;
        const long tiles = (long)strips * tps;
        if (best < 0 || tiles < best) {
            best = tiles;
            int sb = (int)(room / b_stage);
            if (sb > H2_MAX_SB) sb = H2_MAX_SB;
            p->Wp = Wp; p->TW = TW; p->strips = strips; p->tiles_per_strip = tps; p->groups = groups;
            p->SB = sb; p->row0 = row0;
            p->a_plane = a_plane; p->a_stage = a_stage; p->b_plane = b_plane; p->b_stage = b_stage;
            p->a_tx = (uint32_t)(NR * Wp * RB) * planes; p->b_tx = b_stage;
            *NR_out = NR;
        }
    }
    return best >= 0;
}

int halo2_build(gnd_ctx *ctx, const Act *const inputs[8]) {
    if (!ctx->h2maps) ctx->h2maps = new Halo2Maps();
    // once per build, not per launch: the opt-in limit covers every layer's ring (<= H2_RING_BUDGET + barriers + bias)
    GND_CUDA_CHECK(cudaFuncSetAttribute(conv3x3_halo2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
    const int small_nt = tuning().h2_small_nt;
    for (int l2 = 0; l2 < 16; ++l2) {
        const int l = l2 >> 1;
        const bool small = (l2 & 1) != 0;
        Halo2Layer &hl = small ? ctx->h2maps->small[l] : ctx->h2maps->layer[l];
        hl.usable = false;
        const ConvLayer &L = ctx->conv[l];
        const Act &in = *inputs[l];
        if (L.cin % 64 != 0 || L.cout % 64 != 0 || !in.hi || (L.split && !in.lo)) continue;
        Halo2Params p = {};
        p.split = L.split;
        p.RB = L.split ? 64 : 128;
        p.kc = p.RB / 2; p.ksteps = p.RB / 32;
        p.cblocks = L.cin / p.kc;
        p.NT = L.cout < 128 ? 64 : (L.cout % 256 == 0 ? 256 : 128);
        if (small) p.NT = small_nt < p.NT ? small_nt : p.NT;
        if (p.NT < 32 || p.NT % 32 != 0 || L.cout % p.NT != 0) continue;
        p.MT = small ? 1 : 256 / p.NT;
        p.n_ntiles = L.cout / p.NT;
        const int planes = L.split ? 2 : 1;
        int NR = 0;
        const int want_sb = tuning().h2_minsb;
        if (!h2_geometry(in.H, in.W, p.RB, planes, p.MT, p.NT, want_sb, &p, &NR) &&
            !h2_geometry(in.H, in.W, p.RB, planes, p.MT, p.NT, 4, &p, &NR))
            continue;
        p.H = in.H; p.W = in.W; p.Cout = L.cout; p.B = ctx->B;
        p.bias = L.bias;
        hl.smem = (size_t)H2_SA * p.a_stage + (size_t)p.SB * p.b_stage + 1024 + H2_NBARS * 8 + 16 + 256 * 4 + 768 * 4;
        {
            cuuint64_t dims[4] = {(cuuint64_t)in.C, (cuuint64_t)in.W, (cuuint64_t)in.H, (cuuint64_t)ctx->B};
            cuuint64_t strides[3] = {(cuuint64_t)in.C * 2, (cuuint64_t)in.W * in.C * 2, (cuuint64_t)in.H * in.W * in.C * 2};
            cuuint32_t box[4] = {(cuuint32_t)p.kc, (cuuint32_t)p.Wp, (cuuint32_t)NR, 1};
            int rc = h2_encode(&hl.a_hi, in.hi, 4, dims, strides, box, p.RB, "pair halo activation");
            if (rc != GND_OK) return rc;
            rc = h2_encode(&hl.a_lo, L.split ? in.lo : in.hi, 4, dims, strides, box, p.RB, "pair halo activation lo");
            if (rc != GND_OK) return rc;
        }
        {
            cuuint64_t dims[3] = {(cuuint64_t)L.cin, (cuuint64_t)L.cout, 9};
            cuuint64_t strides[2] = {(cuuint64_t)L.cin * 2, (cuuint64_t)L.cout * L.cin * 2};
            cuuint32_t box[3] = {(cuuint32_t)p.kc, (cuuint32_t)(p.NT / 2), 1};
            int rc = h2_encode(&hl.b_hi, L.whi, 3, dims, strides, box, p.RB, "pair halo weights");
            if (rc != GND_OK) return rc;
            rc = h2_encode(&hl.b_lo, L.wlo, 3, dims, strides, box, p.RB, "pair halo weights lo");
            if (rc != GND_OK) return rc;
        }
        hl.p = p;
        hl.usable = true;
        if (tuning().debug)
            fprintf(stderr, "halo2 layer %d%s: %dx%d cin %d cout %d split %d | Wp %d strips %d tiles/strip %d MT %d NT %d NR %d SB %d row0 %d smem %zu\n",
                    l, small ? " (small-batch)" : "", in.H, in.W, L.cin, L.cout, L.split, p.Wp, p.strips, p.tiles_per_strip, p.MT, p.NT, NR, p.SB, p.row0, hl.smem);
    }
    return GND_OK;
}

void halo2_free(gnd_ctx *ctx) {
    delete ctx->h2maps;
    ctx->h2maps = nullptr;
}

bool halo2_usable(const gnd_ctx *ctx, int l) {
    if (ctx->opt_conv_path == 1) return false;             // gnd_set_option("conv_path", 1): the single-CTA tap kernel
    return ctx->h2maps && ctx->h2maps->layer[l].usable;
}

int halo2_conv_run(gnd_ctx *ctx, int l, const Act &out, int B, bool fuse_regressor, cudaStream_t st) {
    // geometry: the throughput one unless it leaves most of the 74 pairs without work
    const Halo2Layer *pick = &ctx->h2maps->layer[l];
    {
        const Halo2Layer &sm = ctx->h2maps->small[l];
        auto items = [&](const Halo2Params &q) { return ((B * q.strips * q.groups + 1) / 2) * q.n_ntiles; };
        const int force = ctx->opt_conv_geom;                      // gnd_set_option("conv_geometry"): 1 big, 2 small (tests)
        // (an N tile of 64 re-reads the halo once per N tile and runs the MMAs at the shared-memory limit: it pays only
        //  while the wide geometry leaves about half of the pairs idle -- measured at batch 1, 2 and 4)
        if (sm.usable && ((items(pick->p) <= 40 && items(sm.p) > items(pick->p)) || force == 2) && force != 1)
            pick = &sm;
    }
    const Halo2Layer &hl = *pick;
    Halo2Params p = hl.p;
    p.B = B;
    p.half_items = B * p.strips * p.groups;
    p.pair_items = ((p.half_items + 1) / 2) * p.n_ntiles;
    // L2 reuse across launches: a kernel that walks the scans in the REVERSE order of its producer finds the producer's
    // last-written scans still in the 126 MB L2 (write-back) instead of fetching them from HBM.  The chain alternates:
    // down1.conv1 forward, down1.conv2 reverse, pool forward, down2.conv1 reverse, down2.conv2 forward, pool+mask
    // reverse, up2.conv1 forward, up2.conv2 reverse, unpool forward, up1.conv1 reverse, up1.conv2 forward
    p.rev = (ctx->opt_l2_order && (l == 1 || l == 2 || l == 5 || l == 6)) ? 1 : 0;
    p.out_hi = fuse_regressor ? nullptr : out.hi;
    p.out_lo = fuse_regressor ? nullptr : out.lo;
    p.reg_w = ctx->reg_w;
    p.regpart = fuse_regressor ? ctx->regpart : nullptr;
    if (fuse_regressor && p.Cout != 64) return fail(GND_ERR_UNSUPPORTED, "regressor fusion needs a 64-channel last block");
    if (!fuse_regressor && !out.hi) return fail(GND_ERR_STATE, "layer %d: output activation missing", l);
    const int pairs = p.pair_items < 74 ? p.pair_items : 74;
    conv3x3_halo2_kernel<<<2 * pairs, H2_THREADS, hl.smem, st>>>(hl.a_hi, hl.a_lo, hl.b_hi, hl.b_lo, p);
    GND_LAUNCH_CHECK("conv3x3_halo2_kernel");
#ifdef HALO_STATS
    if (tuning().debug) {
        unsigned long long h[8];
        cudaStreamSynchronize(st);
        cudaMemcpyFromSymbol(h, g_h2_stats, sizeof(h));
        const double g = (double)pairs;
        fprintf(stderr, "halo2 layer %d: per-pair clks: issuer total %.0f wait_tempty %.0f wait_afull %.0f wait_bfull %.0f | items/pair %.1f | prologue %.0f cta lifetime %.0f\n",
                l, h[3] / g, h[0] / g, h[1] / g, h[2] / g, p.pair_items / g, h[4] / g, h[5] / g);
        unsigned long long z[8] = {0};
        cudaMemcpyToSymbol(g_h2_stats, z, sizeof(z));
    }
#endif
    return GND_OK;
}

}  // namespace gnd
