// pfn.cu -- PillarFeatureNet + PointPillarsScatter as one kernel.
//
// Replaces modules/pointpillars.py:115-146 (PillarFeatureNet.forward), :52-65 (PFNLayer.forward),
// :202-223 (get_paddings_indicator) and :168-199 (PointPillarsScatter.forward).
//
//   features = [x, y, z, (r), x-mean_x, y-mean_y, z-mean_z, x-cx, y-cy, (|xyz|)]        (:120-133)
//   mean     = sum over the pillar's slots / num_points                                   (:120)
//   cx, cy   = coors_x * vx + (vx/2 + pc_range[0]),  coors_y * vy + (vy/2 + pc_range[1])  (:110-113, :125-126)
//   padded slots are zeroed BEFORE the Linear, so they contribute relu(bias') to the max whenever
//   num_points < max_points                                                              (:137-140, :54-58)
//   out[ch]  = max over slots of relu(W' f + b')   (BatchNorm1d folded into W', b' when use_norm)
//   canvas[b, :, y, x] = out                                                             (:174-189)
//
// One warp per BEV cell: lane l owns channels 2l, 2l+1 (18-20 weights in registers); the pillar's
// points are gathered through the voxelizer's pillar list, decorated by one lane each, staged in
// shared memory and broadcast to all lanes; the 64-channel result is written as one coalesced
// 128/256-byte NHWC row -- and empty cells get their zeros from the same kernel, so the canvas needs
// no separate memset.  HBM-bound: <= 16 B read per kept point + 64*e B written per cell.
#include "model.cuh"
#include <math.h>
#include <stdlib.h>
#include <algorithm>
#include <utility>
#include <vector>

namespace gnd {

struct PfnArgs {
    const float *points;        // fused: [B][N][F] raw scans    dense: voxels [M][P][F]
    int32_t B;
    int64_t N;
    int32_t F;
    int32_t nfeat;
    int32_t with_distance;
    int32_t P;                  // max_points_voxel
    int32_t max_voxels;
    int32_t nx, ny;
    float vx, vy, xoff, yoff;
    int32_t apply_shift;
    float z_shift;
    const float *w;             // [nfeat][64]
    const float *bias;          // [64]
    // fused inputs: kept points grouped by cell (voxelize.cu)
    const int32_t *offc;
    const float4 *sorted;
    float zmid, q_xy, iq_xy, q_z, iq_z;   // fixed-point scales of the pillar-mean reduction
    // dense inputs
    const float *coors4, *num_f;
    int64_t M;
    int32_t gshift;             // fused: a work unit is 2^gshift consecutive cells (<= 32)
    const int32_t *order;       // fused: work units sorted by distance from the sensor's cell (handed out nearest first)
    int32_t *ctr;               // fused: [0] next group of cells to hand out, [1] CTAs finished (last one clears both)
    // output canvas [B][ny][nx][64]
    float *out32;
    __half *out_hi, *out_lo;
};

constexpr int PFN_WARPS = 8;
constexpr int FEAT_STRIDE = 12;
constexpr int PFN_RQ = 128;          // reciprocal table: 1 / n for n < PFN_RQ (max_points_voxel is 100 in every shipped config)

__device__ __forceinline__ void store_cell(const PfnArgs &a, int64_t cell_global, int lane, float m0, float m1) {
    const int64_t o = cell_global * PFN_OUT + 2 * lane;
    if (a.out32) {
        *reinterpret_cast<float2 *>(a.out32 + o) = make_float2(m0, m1);
    } else {
        __half h0 = __float2half_rn(m0), h1 = __float2half_rn(m1);
        *reinterpret_cast<__half2 *>(a.out_hi + o) = __halves2half2(h0, h1);
        if (a.out_lo) {
            __half l0 = __float2half_rn(m0 - __half2float(h0)), l1 = __float2half_rn(m1 - __half2float(h1));
            *reinterpret_cast<__half2 *>(a.out_lo + o) = __halves2half2(l0, l1);
        }
    }
}

__device__ __forceinline__ void load_point(const float *row, int F, bool vec4, float &x, float &y, float &z, float &r) {
    if (vec4) {
        float4 v = __ldg(reinterpret_cast<const float4 *>(row));
        x = v.x; y = v.y; z = v.z; r = v.w;
    } else {
        x = __ldg(row); y = __ldg(row + 1); z = __ldg(row + 2);
        r = (F > 3) ? __ldg(row + 3) : 0.0f;
    }
}

// Shared by both kernels: given a way to fetch slot j's point, run the pillar through the PFN layer.
// fetch(j, x, y, z, r) must be callable by any lane for j < n_real.
template <typename Fetch>
__device__ __forceinline__ void pfn_pillar(const PfnArgs &a, int lane, float (*s_feat)[FEAT_STRIDE], const float *wreg0,
                                           const float *wreg1, float b0, float b1, int n_real, float sumx, float sumy,
                                           float sumz, float n_div, float cxf, float cyf, Fetch fetch, float &m0, float &m1) {
    // :120 points_mean = sum / num_voxels (float32 division)
    const float mx = __fdiv_rn(sumx, n_div), my = __fdiv_rn(sumy, n_div), mz = __fdiv_rn(sumz, n_div);
    // :125-126 pillar centre: coors * v + offset, two float32 roundings like torch's mul then add
    const float pcx = __fadd_rn(__fmul_rn(cxf, a.vx), a.xoff);
    const float pcy = __fadd_rn(__fmul_rn(cyf, a.vy), a.yoff);
    // padded slots give relu(b') (:137-140 then :54-58); with none padded the relu floor is 0
    m0 = (n_real < a.P) ? fmaxf(b0, 0.0f) : 0.0f;
    m1 = (n_real < a.P) ? fmaxf(b1, 0.0f) : 0.0f;
    for (int j0 = 0; j0 < n_real; j0 += 32) {
        const int j = j0 + lane;
        if (j < n_real) {
            float x, y, z, r;
            fetch(j, x, y, z, r);
            float *f = s_feat[lane];
            int k = 0;
            f[k++] = x; f[k++] = y; f[k++] = z;
            if (a.F > 3) f[k++] = r;
            f[k++] = __fsub_rn(x, mx); f[k++] = __fsub_rn(y, my); f[k++] = __fsub_rn(z, mz);
            f[k++] = __fsub_rn(x, pcx); f[k++] = __fsub_rn(y, pcy);
            if (a.with_distance) f[k++] = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z)));
        }
        __syncwarp();
        const int cnt = min(32, n_real - j0);
        for (int p = 0; p < cnt; ++p) {
            const float *f = s_feat[p];
            float y0 = b0, y1 = b1;
#pragma unroll
            for (int k = 0; k < PFN_MAX_FEAT; ++k)
                if (k < a.nfeat) { y0 = fmaf(wreg0[k], f[k], y0); y1 = fmaf(wreg1[k], f[k], y1); }
            m0 = fmaxf(m0, y0);
            m1 = fmaxf(m1, y1);
        }
        __syncwarp();
    }
}

// Zeros for the empty cells of a work unit (about half of a LiDAR BEV grid), written by the whole warp: the unit's
// cells are contiguous in every plane, so lane l stores the 16-byte piece l (+ 32, + 64, ...) of the unit's rows when the
// cell that piece belongs to is empty -- a handful of fully coalesced store instructions per unit instead of 16 per
// empty cell from single lanes.
__device__ __forceinline__ void zero_empty_cells(const PfnArgs &a, int64_t cell0, int ncell, unsigned empty, int lane) {
    if (!empty) return;
    const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
    if (a.out32) {
        const int per = PFN_OUT * 4 / 16;                       // 16-byte pieces per cell
        uint4 *dst = reinterpret_cast<uint4 *>(a.out32 + cell0 * PFN_OUT);
        for (int i = lane; i < ncell * per; i += 32)
            if ((empty >> (i / per)) & 1u) dst[i] = z4;
    } else {
        const int per = PFN_OUT * 2 / 16;
        uint4 *dh = reinterpret_cast<uint4 *>(a.out_hi + cell0 * PFN_OUT);
        uint4 *dl = a.out_lo ? reinterpret_cast<uint4 *>(a.out_lo + cell0 * PFN_OUT) : nullptr;
        for (int i = lane; i < ncell * per; i += 32)
            if ((empty >> (i / per)) & 1u) {
                dh[i] = z4;
                if (dl) dl[i] = z4;
            }
    }
}

// Fused path: the voxelizer left the kept points grouped by CELL (`sorted`, CSR `offc`), z already shifted.
// One warp walks 32 consecutive cells; lane l owns channels 2l, 2l+1.  The Linear is regrouped around the
// pillar centre so that a point costs 4 multiply-adds per channel instead of 9:
//   W f + b = A x' + B y' + C z + D r (+ Wd |p|) + E,   x' = x - cx,  y' = y - cy,
//   A = w_x + w_dmx + w_dcx,  B = w_y + w_dmy + w_dcy,  C = w_z + w_dmz,  D = w_r,
//   E = b + w_x cx + w_y cy - w_dmx mean(x') - w_dmy mean(y') - w_dmz mean(z)            (per pillar)
// and relu(max_p(t_p + E)) = relu(max_p t_p + E).  All products stay below |w| in magnitude except E, whose
// rounding (|w| * 50 m * 2^-24) is the same as the reference's own w_x * x term.
// The pillar mean (:120) is taken over cell-relative coordinates in fixed point (2^-k units chosen on the
// host so that max_points * half a voxel fits 31 bits): one integer warp reduction (redux.sync) per axis
// instead of a 5-level float shuffle tree, and an error (<= 2^-25 of a voxel per point) far below the float32
// summation error of the reference's own mean.
template <bool DIST, bool HAS_R>
__global__ void __launch_bounds__(PFN_WARPS * 32, 4) pfn_scatter_kernel(const PfnArgs a) {
    __shared__ float4 s_pt[PFN_WARPS][32];
    // without the |p| feature the points are staged in PAIRS, (x'a, x'b, y'a, y'b) (za, zb, ra, rb): two LDS.128 deliver, for
    // two points at once, the aligned register pairs of sm_100's packed fma.rn.f32x2 -- per point and channel pair 4 packed
    // FMAs + 1 FMNMX3 + 1 LDS instead of 8 FFMA + 2 FMNMX + 1 LDS, same arithmetic per point
    __shared__ float4 s_pp[PFN_WARPS][16][2];
    __shared__ float s_dist[PFN_WARPS][32];
    __shared__ float2 s_rq[PFN_RQ];                // (1/n) * iq_xy, (1/n) * iq_z: one LDS instead of an IEEE reciprocal per cell
    for (int i = threadIdx.x; i < PFN_RQ; i += PFN_WARPS * 32) {
        const float rn = __frcp_rn((float)max(i, 1));
        s_rq[i] = make_float2(rn * a.iq_xy, rn * a.iq_z);
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int kf = HAS_R ? 4 : 3;
    auto W = [&](int k, int ch) { return __ldg(a.w + k * PFN_OUT + ch); };
    const int ch0 = 2 * lane, ch1 = 2 * lane + 1;
    const float wx0 = W(0, ch0), wx1 = W(0, ch1), wy0 = W(1, ch0), wy1 = W(1, ch1);
    const float wmx0 = W(kf, ch0), wmx1 = W(kf, ch1), wmy0 = W(kf + 1, ch0), wmy1 = W(kf + 1, ch1);
    const float wmz0 = W(kf + 2, ch0), wmz1 = W(kf + 2, ch1);
    const float A0 = wx0 + wmx0 + W(kf + 3, ch0), A1 = wx1 + wmx1 + W(kf + 3, ch1);
    const float B0 = wy0 + wmy0 + W(kf + 4, ch0), B1 = wy1 + wmy1 + W(kf + 4, ch1);
    const float C0 = W(2, ch0) + wmz0, C1 = W(2, ch1) + wmz1;
    const float D0 = HAS_R ? W(3, ch0) : 0.0f, D1 = HAS_R ? W(3, ch1) : 0.0f;
    const float G0 = DIST ? W(kf + 5, ch0) : 0.0f, G1 = DIST ? W(kf + 5, ch1) : 0.0f;
    const float b0 = __ldg(a.bias + ch0), b1 = __ldg(a.bias + ch1);
    const float rb0 = fmaxf(b0, 0.0f), rb1 = fmaxf(b1, 0.0f);
    float *const pr = reinterpret_cast<float *>(&s_pp[warp][lane >> 1][0]) + (lane & 1);      // this lane's slot of its pair
    const float4 *const pp = &s_pp[warp][0][0];
    const int cells = a.nx * a.ny;
    const int gsize = 1 << a.gshift;
    const int groups = (cells + gsize - 1) >> a.gshift;
    const int total = a.B * groups;
    // groups are handed out through a device counter: the points per group vary from 0 to 32 * max_points,
    // so a static split leaves most warps idle behind the few that drew the dense cells next to the sensor
    for (;;) {
        int g = 0;
        if (lane == 0) g = atomicAdd(a.ctr, 1);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= total) break;
        // units are handed out nearest-to-the-sensor first, for every scan of the batch in turn: the dense pillars next to
        // the sensor are drawn early and the kernel's tail is made of the sparse far cells (0.207 -> 0.187 ms per 64 scans
        // against index order, where ~9 % of the warp time was spent waiting at the final barrier)
        const int rank = g / a.B, b = g - rank * a.B;
        const int c0 = __ldg(a.order + rank) << a.gshift;
        const int32_t *offc = a.offc + (int64_t)b * (cells + 1);
        const float4 *pts = a.sorted + (int64_t)b * a.N;
        const int cl = min(c0 + lane, cells);
        const int my_o = __ldg(offc + cl);
        const int my_n = __ldg(offc + min(cl + 1, cells)) - my_o;
        const int ncell = min(gsize, cells - c0);
        // empty cells (about half of a LiDAR BEV grid): the lane that owns the cell writes its 64 zeros with 16-byte
        // stores, once per group, instead of the whole warp walking through the cell
        zero_empty_cells(a, (int64_t)b * cells + c0, ncell, __ballot_sync(0xffffffffu, lane < ncell && my_n == 0), lane);
        // :125-126 pillar centre of the lane's own cell: coors * v + offset, two float32 roundings like torch's mul then add
        const int my_cy = cl / a.nx, my_cx = cl - my_cy * a.nx;
        const float my_pcx = __fadd_rn(__fmul_rn((float)my_cx, a.vx), a.xoff);
        const float my_pcy = __fadd_rn(__fmul_rn((float)my_cy, a.vy), a.yoff);
        const int64_t out0 = ((int64_t)b * cells + c0) * PFN_OUT + 2 * lane;
        unsigned occ = __ballot_sync(0xffffffffu, lane < ncell && my_n > 0);
        while (occ) {
            const int j = __ffs(occ) - 1;
            occ &= occ - 1;
            const int n = __shfl_sync(0xffffffffu, my_n, j);
            const int64_t oidx = out0 + j * PFN_OUT;
            float m0 = 0.0f, m1 = 0.0f;
            {
                const int o = __shfl_sync(0xffffffffu, my_o, j);
                const float pcx = __shfl_sync(0xffffffffu, my_pcx, j), pcy = __shfl_sync(0xffffffffu, my_pcy, j);
                float4 p0 = make_float4(pcx, pcy, a.zmid, 0.f);
                if (lane < n) p0 = __ldg(pts + o + lane);
                const float4 f0 = make_float4(p0.x - pcx, p0.y - pcy, p0.z, p0.w);
                int qx = __float2int_rn(f0.x * a.q_xy), qy = __float2int_rn(f0.y * a.q_xy), qz = __float2int_rn((p0.z - a.zmid) * a.q_z);
                for (int q = 32 + lane; q < n; q += 32) {
                    const float4 p = __ldg(pts + o + q);
                    qx += __float2int_rn((p.x - pcx) * a.q_xy); qy += __float2int_rn((p.y - pcy) * a.q_xy);
                    qz += __float2int_rn((p.z - a.zmid) * a.q_z);
                }
                qx = __reduce_add_sync(0xffffffffu, qx);
                qy = __reduce_add_sync(0xffffffffu, qy);
                qz = __reduce_add_sync(0xffffffffu, qz);
                float2 rq;
                if (n < PFN_RQ) rq = s_rq[n];
                else { const float rn = __frcp_rn((float)n); rq = make_float2(rn * a.iq_xy, rn * a.iq_z); }
                const float mxr = (float)qx * rq.x, myr = (float)qy * rq.x;   // mean(x') , mean(y')
                const float mz = a.zmid + (float)qz * rq.y;
                const float E0 = fmaf(wx0, pcx, fmaf(wy0, pcy, fmaf(-wmx0, mxr, fmaf(-wmy0, myr, fmaf(-wmz0, mz, b0)))));
                const float E1 = fmaf(wx1, pcx, fmaf(wy1, pcy, fmaf(-wmx1, mxr, fmaf(-wmy1, myr, fmaf(-wmz1, mz, b1)))));
                float t0 = -INFINITY, t1 = -INFINITY;
                for (int q0 = 0; q0 < n; q0 += 32) {
                    if (q0 + lane < n) {
                        float4 f = f0;
                        float4 p = p0;
                        if (q0 != 0) { p = __ldg(pts + o + q0 + lane); f = make_float4(p.x - pcx, p.y - pcy, p.z, p.w); }
                        if (DIST) s_pt[warp][lane] = f;
                        else {
                            pr[0] = f.x; pr[2] = f.y; pr[4] = f.z; pr[6] = f.w;
                            if (q0 + lane == n - 1 && !(lane & 1)) { pr[1] = f.x; pr[3] = f.y; pr[5] = f.z; pr[7] = f.w; }   // odd count: the last point twice
                        }
                        if (DIST) s_dist[warp][lane] = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(p.x, p.x), __fmul_rn(p.y, p.y)), __fmul_rn(p.z, p.z)));
                    }
                    __syncwarp();
                    const int cnt = min(32, n - q0);
                    if (DIST) {
                        for (int q = 0; q < cnt; ++q) {
                            const float4 f = s_pt[warp][q];
                            const float dq = s_dist[warp][q];
                            t0 = fmaxf(t0, fmaf(G0, dq, fmaf(D0, f.w, fmaf(C0, f.z, fmaf(B0, f.y, A0 * f.x)))));
                            t1 = fmaxf(t1, fmaf(G1, dq, fmaf(D1, f.w, fmaf(C1, f.z, fmaf(B1, f.y, A1 * f.x)))));
                        }
                    } else {
                        const float2 A00 = make_float2(A0, A0), A11 = make_float2(A1, A1), B00 = make_float2(B0, B0), B11 = make_float2(B1, B1);
                        const float2 C00 = make_float2(C0, C0), C11 = make_float2(C1, C1), D00 = make_float2(D0, D0), D11 = make_float2(D1, D1);
                        const int npair = (cnt + 1) >> 1;
#pragma unroll 2
                        for (int q = 0; q < npair; ++q) {
                            const float4 fa = pp[2 * q], fb = pp[2 * q + 1];
                            const float2 xx = make_float2(fa.x, fa.y), yy = make_float2(fa.z, fa.w), zz = make_float2(fb.x, fb.y), rr = make_float2(fb.z, fb.w);
                            // per point and channel the same order as the scalar form: mul, then two (three) fused multiply-adds
                            float2 u = __ffma2_rn(C00, zz, __ffma2_rn(B00, yy, __fmul2_rn(A00, xx)));
                            float2 v = __ffma2_rn(C11, zz, __ffma2_rn(B11, yy, __fmul2_rn(A11, xx)));
                            if (HAS_R) { u = __ffma2_rn(D00, rr, u); v = __ffma2_rn(D11, rr, v); }
                            t0 = fmaxf(t0, fmaxf(u.x, u.y));
                            t1 = fmaxf(t1, fmaxf(v.x, v.y));
                        }
                    }
                    __syncwarp();
                }
                // padded slots give relu(b') (:137-140 then :54-58)
                m0 = fmaxf(t0 + E0, n < a.P ? rb0 : 0.0f);
                m1 = fmaxf(t1 + E1, n < a.P ? rb1 : 0.0f);
            }
            if (a.out32) {
                *reinterpret_cast<float2 *>(a.out32 + oidx) = make_float2(m0, m1);
            } else {
                const __half2 hh = __floats2half2_rn(m0, m1);
                *reinterpret_cast<__half2 *>(a.out_hi + oidx) = hh;
                if (a.out_lo) {
                    const float2 back = __half22float2(hh);
                    *reinterpret_cast<__half2 *>(a.out_lo + oidx) = __floats2half2_rn(m0 - back.x, m1 - back.y);
                }
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(a.ctr + 1, 1) == (int)gridDim.x - 1) { a.ctr[0] = 0; a.ctr[1] = 0; __threadfence(); }
    }
}

// ---- the same, with the per-point Linear on the (legacy-path) tensor cores --------------------------------------
// ncu on pfn_scatter_kernel: 135 M warp instructions per 64 scans, 31 per kept point, issue-bound (71 % busy); the
// multiply-adds themselves are a third of that.  Here a warp still owns a cell, but its points go through
// mma.sync.m16n8k8 (TF32 operands, fp32 accumulate) EIGHT at a time, with the CHANNELS on the M side and the points
// on the N side:
//   A1 [16 ch x 8] = [ W'hi | W'hi ],  A2 = [ W'lo | 0 ]      W' = the regrouped weights A, B, C, D above (constant
//                                                              fragments, built once per warp)
//   B  [8 x 8 pts] = [ hi(x', y', z, r) ; lo(x', y', z, r) ]   hi = tf32(v), lo = tf32(v - hi)
//   A1 B + A2 B    = W'hi hi + W'hi lo + W'lo hi               (3xTF32: ~2^-21 relative, fp32 class)
// Lane (g, t) of the warp loads feature t of point g -- the 8 points of a step are ONE coalesced 128-byte row of the
// cell-ordered buffer and ARE the B fragment, no shared memory, no shuffles -- and gets back channels x points
// (2t, 2t+1): the maximum over a step's points is one FMNMX3 per accumulator pair, the running maximum of the cell
// stays in 8 registers, and only the end of the cell pays a 2-round reduce-scatter over t (6 shuffles) that leaves lane
// (g, t) with one adjacent channel pair.  (The first version of this kernel had the points on the M side: 16 per step,
// but the maximum then runs over the row groups g, a 3-round 14-shuffle exchange per cell, and the kernel executed
// MORE instructions than the CUDA-core one: 149.5 M per 64 scans, 263 us against 186 us.)
// The pillar mean keeps its order-independent fixed-point form (identical sums, so an identical E term).
__device__ __forceinline__ uint32_t to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ void mma_tf32_16x8x8(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                                uint32_t b0, uint32_t b1, const float (&c)[4]) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
                 : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "f"(c[0]), "f"(c[1]), "f"(c[2]), "f"(c[3]));
}

__device__ __forceinline__ void mma_tf32_16x8x4(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm volatile("mma.sync.aligned.m16n8k4.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%7,%7,%7,%7};"
                 : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3]) : "r"(a0), "r"(a1), "r"(b0), "f"(0.0f));
}
__device__ __forceinline__ void mma_tf32_16x8x8_acc(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0, uint32_t b1) {
    // A = [a0 a1 | a0 a1]: the same 16 x 4 block for both halves of K
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%4,%5}, {%6,%7}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a0), "r"(a1), "r"(b0), "r"(b1));
}

template <bool HAS_R>
__global__ void __launch_bounds__(PFN_WARPS * 32, 4) pfn_scatter_mma_kernel(const PfnArgs a) {
    __shared__ float4 s_ew[32][3];                 // per lane (g, t): the 12 weights of its channel pair's E term
    __shared__ float2 s_rq[PFN_RQ];                // (1/n) * iq_xy, (1/n) * iq_z
    const int lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    constexpr int kf = HAS_R ? 4 : 3;
    auto W = [&](int k, int ch) { return __ldg(a.w + k * PFN_OUT + ch); };
    // A fragments.  Tile mt: row g <-> channel 2 (8 mt + g), row g + 8 <-> the channel after it (so that a lane ends up
    // with an adjacent pair); column t (and t + 4) <-> regrouped weight of input t
    uint32_t ahi[4][2], alo[4][2];
#pragma unroll
    for (int mt = 0; mt < 4; ++mt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int ch = 2 * (8 * mt + g) + h;
            float w;
            if (t == 0) w = W(0, ch) + W(kf, ch) + W(kf + 3, ch);
            else if (t == 1) w = W(1, ch) + W(kf + 1, ch) + W(kf + 4, ch);
            else if (t == 2) w = W(2, ch) + W(kf + 2, ch);
            else w = HAS_R ? W(3, ch) : 0.0f;
            ahi[mt][h] = to_tf32(w);
            alo[mt][h] = to_tf32(w - __uint_as_float(ahi[mt][h]));
        }
    // the adjacent channel pair this lane finalises: tile 2 (t & 1) + (t >> 1), rows g / g + 8
    const int ch0 = 2 * (8 * (2 * (t & 1) + (t >> 1)) + g), ch1 = ch0 + 1;
    if (threadIdx.x < 32) {
        s_ew[lane][0] = make_float4(W(0, ch0), W(1, ch0), -W(kf, ch0), -W(kf + 1, ch0));
        s_ew[lane][1] = make_float4(-W(kf + 2, ch0), __ldg(a.bias + ch0), W(0, ch1), W(1, ch1));
        s_ew[lane][2] = make_float4(-W(kf, ch1), -W(kf + 1, ch1), -W(kf + 2, ch1), __ldg(a.bias + ch1));
    }
    for (int i = threadIdx.x; i < PFN_RQ; i += PFN_WARPS * 32) {
        const float rn = __frcp_rn((float)max(i, 1));
        s_rq[i] = make_float2(rn * a.iq_xy, rn * a.iq_z);
    }
    __syncthreads();
    const float qscale = (t < 2) ? a.q_xy : (t == 2 ? a.q_z : 0.0f);
    const float qbias = (t == 2) ? -a.zmid * a.q_z : 0.0f;      // (v - zmid) * q_z as one FMA: zmid * q_z is exact (q_z = 2^k)
    const int cells = a.nx * a.ny;
    const int gsize = 1 << a.gshift;
    const int groups = (cells + gsize - 1) >> a.gshift;
    const int total = a.B * groups;
    for (;;) {
        int grp = 0;
        if (lane == 0) grp = atomicAdd(a.ctr, 1);
        grp = __shfl_sync(0xffffffffu, grp, 0);
        if (grp >= total) break;
        const int rank = grp / a.B, b = grp - rank * a.B;                     // nearest-to-the-sensor units first, as above
        const int c0 = __ldg(a.order + rank) << a.gshift;
        const int32_t *offc = a.offc + (int64_t)b * (cells + 1);
        const float *pts = reinterpret_cast<const float *>(a.sorted + (int64_t)b * a.N) + t;
        const int cl = min(c0 + lane, cells);
        const int my_o = __ldg(offc + cl);
        const int my_n = __ldg(offc + min(cl + 1, cells)) - my_o;
        const int ncell = min(gsize, cells - c0);
        unsigned occ = __ballot_sync(0xffffffffu, lane < ncell && my_n > 0);
        // first step of the first occupied cell: in flight while the empty cells get their zeros
        int jc = occ ? __ffs(occ) - 1 : 0;
        int n = __shfl_sync(0xffffffffu, my_n, jc), o = __shfl_sync(0xffffffffu, my_o, jc);
        float vnext = occ ? __ldg(pts + 4 * (o + min(g, n - 1))) : 0.0f;
        zero_empty_cells(a, (int64_t)b * cells + c0, ncell, __ballot_sync(0xffffffffu, lane < ncell && my_n == 0), lane);
        // :125-126 pillar centre of the lane's own cell: coors * v + offset, two float32 roundings like torch's mul then add
        const int my_cy = cl / a.nx, my_cx = cl - my_cy * a.nx;
        const float my_pcx = __fadd_rn(__fmul_rn((float)my_cx, a.vx), a.xoff);
        const float my_pcy = __fadd_rn(__fmul_rn((float)my_cy, a.vy), a.yoff);
        const int64_t out0 = ((int64_t)b * cells + c0) * PFN_OUT + ch0;
        while (occ) {
            occ &= occ - 1;
            const int cur = jc, n_cur = n;
            const float *cell_pts = pts + 4 * o;
            float vraw = vnext;
            if (occ) {                              // the next occupied cell's first step, ahead of this cell's arithmetic
                jc = __ffs(occ) - 1;
                n = __shfl_sync(0xffffffffu, my_n, jc);
                o = __shfl_sync(0xffffffffu, my_o, jc);
                vnext = __ldg(pts + 4 * (o + min(g, n - 1)));
            }
            const float pcx = __shfl_sync(0xffffffffu, my_pcx, cur), pcy = __shfl_sync(0xffffffffu, my_pcy, cur);
            const float sub = (t == 0) ? pcx : (t == 1 ? pcy : 0.0f);
            float m[4][2];
#pragma unroll
            for (int mt = 0; mt < 4; ++mt) { m[mt][0] = -INFINITY; m[mt][1] = -INFINITY; }
            int qsum = 0;
#pragma unroll 1
            for (int r0 = 0; r0 < n_cur; r0 += 8) {
                const int row = r0 + g;
                // points past the end repeat the last one: the maximum does not change, the mean skips them.  The next
                // step's row is requested before this step's arithmetic (one load in flight across the MMA chain)
                const float vahead = (r0 + 8 < n_cur) ? __ldg(cell_pts + 4 * min(row + 8, n_cur - 1)) : 0.0f;
                const float v = vraw - sub;
                vraw = vahead;
                const int q = __float2int_rn(fmaf(v, qscale, qbias));
                if (row < n_cur) qsum += q;
                // hi = the value with its low 13 mantissa bits cleared (what the tensor core reads of a 32-bit operand),
                // lo = the exact remainder (<= 13 significant bits, of which the tensor core keeps 11)
                const uint32_t bh = __float_as_uint(v) & 0xffffe000u;
                const uint32_t bl = __float_as_uint(v - __uint_as_float(bh));
                float c[4][4];
#pragma unroll
                for (int mt = 0; mt < 4; ++mt) mma_tf32_16x8x4(c[mt], alo[mt][0], alo[mt][1], bh);                 // W'lo hi (small term first)
#pragma unroll
                for (int mt = 0; mt < 4; ++mt) mma_tf32_16x8x8_acc(c[mt], ahi[mt][0], ahi[mt][1], bh, bl);         // + W'hi hi + W'hi lo
#pragma unroll
                for (int mt = 0; mt < 4; ++mt) {
                    m[mt][0] = fmaxf(m[mt][0], fmaxf(c[mt][0], c[mt][1]));
                    m[mt][1] = fmaxf(m[mt][1], fmaxf(c[mt][2], c[mt][3]));
                }
            }
            // feature sums: the 8 lanes with the same t hold the pieces of feature t
            qsum += __shfl_xor_sync(0xffffffffu, qsum, 4);
            qsum += __shfl_xor_sync(0xffffffffu, qsum, 8);
            qsum += __shfl_xor_sync(0xffffffffu, qsum, 16);
            const int qx = __shfl_sync(0xffffffffu, qsum, 0), qy = __shfl_sync(0xffffffffu, qsum, 1);
            const int qz = __shfl_sync(0xffffffffu, qsum, 2);
            // maxima over the point pairs of the four t lanes: reduce-scatter, lane (g, t) ends with tile 2 (t & 1) + (t >> 1)
            float r4[2][2], r2[2];
            {
                const bool low = (t & 1) == 0;
#pragma unroll
                for (int i = 0; i < 2; ++i)
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const float send = low ? m[2 + i][h] : m[i][h];
                        const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
                        r4[i][h] = fmaxf(low ? m[i][h] : m[2 + i][h], recv);
                    }
            }
            {
                const bool low = (t & 2) == 0;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const float send = low ? r4[1][h] : r4[0][h];
                    const float recv = __shfl_xor_sync(0xffffffffu, send, 2);
                    r2[h] = fmaxf(low ? r4[0][h] : r4[1][h], recv);
                }
            }
            float2 rq;
            if (n_cur < PFN_RQ) rq = s_rq[n_cur];
            else { const float rn = __frcp_rn((float)n_cur); rq = make_float2(rn * a.iq_xy, rn * a.iq_z); }
            const float mxr = (float)qx * rq.x, myr = (float)qy * rq.x;      // mean(x'), mean(y')
            const float mz = fmaf((float)qz, rq.y, a.zmid);
            const float4 e0 = s_ew[lane][0], e1 = s_ew[lane][1], e2 = s_ew[lane][2];
            // e0 = (wx0, wy0, -wmx0, -wmy0)  e1 = (-wmz0, b0, wx1, wy1)  e2 = (-wmx1, -wmy1, -wmz1, b1)
            const float E0 = fmaf(e0.x, pcx, fmaf(e0.y, pcy, fmaf(e0.z, mxr, fmaf(e0.w, myr, fmaf(e1.x, mz, e1.y)))));
            const float E1 = fmaf(e1.z, pcx, fmaf(e1.w, pcy, fmaf(e2.x, mxr, fmaf(e2.y, myr, fmaf(e2.z, mz, e2.w)))));
            // padded slots give relu(b') (:137-140 then :54-58)
            const bool padded = n_cur < a.P;
            const float o0 = fmaxf(r2[0] + E0, padded ? fmaxf(e1.y, 0.0f) : 0.0f);
            const float o1 = fmaxf(r2[1] + E1, padded ? fmaxf(e2.w, 0.0f) : 0.0f);
            const int64_t oidx = out0 + cur * PFN_OUT;
            if (a.out32) {
                *reinterpret_cast<float2 *>(a.out32 + oidx) = make_float2(o0, o1);
            } else {
                const __half2 hh = __floats2half2_rn(o0, o1);
                *reinterpret_cast<__half2 *>(a.out_hi + oidx) = hh;
                if (a.out_lo) {
                    const float2 back = __half22float2(hh);
                    *reinterpret_cast<__half2 *>(a.out_lo + oidx) = __floats2half2_rn(o0 - back.x, o1 - back.y);
                }
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(a.ctr + 1, 1) == (int)gridDim.x - 1) { a.ctr[0] = 0; a.ctr[1] = 0; __threadfence(); }
    }
}

// Drop-in tensors (model.py:36-42): voxels [M][P][F], coors [M][4] float (b,z,y,x), num_points [M] float.
// The canvas must have been zero-filled (pointpillars.py:174).
__global__ void __launch_bounds__(PFN_WARPS * 32) pfn_scatter_dense_kernel(const PfnArgs a) {
    __shared__ float s_feat[PFN_WARPS][32][FEAT_STRIDE];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float w0[PFN_MAX_FEAT], w1[PFN_MAX_FEAT];
#pragma unroll
    for (int k = 0; k < PFN_MAX_FEAT; ++k) {
        w0[k] = (k < a.nfeat) ? __ldg(a.w + k * PFN_OUT + 2 * lane) : 0.0f;
        w1[k] = (k < a.nfeat) ? __ldg(a.w + k * PFN_OUT + 2 * lane + 1) : 0.0f;
    }
    const float b0 = __ldg(a.bias + 2 * lane), b1 = __ldg(a.bias + 2 * lane + 1);
    const bool vec4 = (a.F == 4) && ((((uintptr_t)a.points) & 15) == 0);
    for (int64_t m = (int64_t)blockIdx.x * PFN_WARPS + warp; m < a.M; m += (int64_t)gridDim.x * PFN_WARPS) {
        const float *vox = a.points + m * a.P * a.F;
        const float nf = __ldg(a.num_f + m);
        const int n = max(0, min(a.P, (int)nf));                    // num_voxels.int() > arange (:202-223)
        const float bf = __ldg(a.coors4 + m * 4 + 0), cyf = __ldg(a.coors4 + m * 4 + 2), cxf = __ldg(a.coors4 + m * 4 + 3);
        const int b = (int)bf;
        const int64_t cell = (int64_t)(__fadd_rn(__fmul_rn(cyf, (float)a.nx), cxf));   // y*nx + x (:180)
        auto fetch = [&](int j, float &x, float &y, float &z, float &r) {
            load_point(vox + (int64_t)j * a.F, a.F, vec4, x, y, z, r);
        };
        float sx = 0.f, sy = 0.f, sz = 0.f;
        for (int j = lane; j < a.P; j += 32) {      // :120 sums every slot, padded ones included
            float x, y, z, r;
            fetch(j, x, y, z, r);
            sx += x; sy += y; sz += z;
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            sx += __shfl_xor_sync(0xffffffffu, sx, d);
            sy += __shfl_xor_sync(0xffffffffu, sy, d);
            sz += __shfl_xor_sync(0xffffffffu, sz, d);
        }
        float m0, m1;
        pfn_pillar(a, lane, s_feat[warp], w0, w1, b0, b1, n, sx, sy, sz, nf, cxf, cyf, fetch, m0, m1);
        if (b >= 0 && b < a.B && cell >= 0 && cell < (int64_t)a.nx * a.ny)
            store_cell(a, (int64_t)b * a.nx * a.ny + cell, lane, m0, m1);
    }
}

// ---- multi-layer PillarFeatureNet (cfg.vfe_filters with more than one entry, modules/pointpillars.py:57-65, :96-107) ----
// No shipped config uses it; this is the plain form: one warp per pillar, the per-slot feature rows of the current and
// the next layer in shared memory.  A non-last PFNLayer has out / 2 units and hands concat(x, max over ALL max_points
// slots) to the next layer; padded slots are zeroed only before the FIRST layer (:137-140), afterwards they carry
// concat(relu(b'), max) like the reference's -- they are all equal, so one representative slot stands for them.
struct PfnMulti {
    int n_layers, cmax;
    int cin[4], units[4];
    const float *w[4], *b[4];
};

template <bool FUSED>
__global__ void __launch_bounds__(32) pfn_multi_kernel(const PfnArgs a, const PfnMulti ml) {
    extern __shared__ float s_rows[];                    // [2][P + 1][cmax]
    const int lane = threadIdx.x;
    const int cells = a.nx * a.ny;
    const int64_t total = FUSED ? (int64_t)a.B * cells : a.M;
    const bool vec4 = !FUSED && (a.F == 4) && ((((uintptr_t)a.points) & 15) == 0);
    float *buf0 = s_rows, *buf1 = s_rows + (size_t)(a.P + 1) * ml.cmax;
    for (int64_t m = blockIdx.x; m < total; m += gridDim.x) {
        int b, n_real;
        int64_t cell;
        float nf, cxf, cyf;
        const float *src;                                // rows of F (dense) or 4 (fused) floats
        int stride;
        if (FUSED) {
            b = (int)(m / cells); cell = m - (int64_t)b * cells;
            const int32_t *offc = a.offc + (int64_t)b * (cells + 1);
            const int o = __ldg(offc + cell);
            n_real = __ldg(offc + cell + 1) - o;
            nf = (float)n_real;
            cyf = (float)(cell / a.nx); cxf = (float)(cell - (cell / a.nx) * a.nx);
            src = reinterpret_cast<const float *>(a.sorted + (int64_t)b * a.N + o); stride = 4;
            if (n_real == 0) {                           // empty cell: zeros (the fused path has no memset)
                for (int ch = lane; ch < PFN_OUT; ch += 32) {
                    const int64_t o2 = ((int64_t)b * cells + cell) * PFN_OUT + ch;
                    if (a.out32) a.out32[o2] = 0.0f;
                    else { a.out_hi[o2] = __float2half_rn(0.0f); if (a.out_lo) a.out_lo[o2] = __float2half_rn(0.0f); }
                }
                continue;
            }
        } else {
            nf = __ldg(a.num_f + m);
            n_real = max(0, min(a.P, (int)nf));          // num_voxels.int() > arange (:202-223)
            const float bf = __ldg(a.coors4 + m * 4 + 0);
            cyf = __ldg(a.coors4 + m * 4 + 2); cxf = __ldg(a.coors4 + m * 4 + 3);
            b = (int)bf;
            cell = (int64_t)(__fadd_rn(__fmul_rn(cyf, (float)a.nx), cxf));
            src = a.points + m * a.P * a.F; stride = a.F;
        }
        auto fetch = [&](int j, float &x, float &y, float &z, float &r) { load_point(src + (int64_t)j * stride, FUSED ? 4 : a.F, FUSED || vec4, x, y, z, r); };
        // :120 points_mean: the dense tensor sums every slot (padded ones are zeros there), the fused path its kept points
        float sx = 0.f, sy = 0.f, sz = 0.f;
        for (int j = lane; j < (FUSED ? n_real : a.P); j += 32) {
            float x, y, z, r;
            fetch(j, x, y, z, r);
            sx += x; sy += y; sz += z;
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            sx += __shfl_xor_sync(0xffffffffu, sx, d); sy += __shfl_xor_sync(0xffffffffu, sy, d); sz += __shfl_xor_sync(0xffffffffu, sz, d);
        }
        const float mx = __fdiv_rn(sx, nf), my = __fdiv_rn(sy, nf), mz = __fdiv_rn(sz, nf);
        const float pcx = __fadd_rn(__fmul_rn(cxf, a.vx), a.xoff), pcy = __fadd_rn(__fmul_rn(cyf, a.vy), a.yoff);
        const int slots = n_real + (n_real < a.P ? 1 : 0);
        __syncwarp();
        for (int j = lane; j < slots; j += 32) {
            float *f = buf0 + (size_t)j * ml.cmax;
            if (j < n_real) {
                float x, y, z, r;
                fetch(j, x, y, z, r);
                int k = 0;
                f[k++] = x; f[k++] = y; f[k++] = z;
                if (a.F > 3) f[k++] = r;
                f[k++] = __fsub_rn(x, mx); f[k++] = __fsub_rn(y, my); f[k++] = __fsub_rn(z, mz);
                f[k++] = __fsub_rn(x, pcx); f[k++] = __fsub_rn(y, pcy);
                if (a.with_distance) f[k++] = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z)));
            } else {
                for (int k = 0; k < a.nfeat; ++k) f[k] = 0.0f;      // the representative padded slot
            }
        }
        __syncwarp();
        float *in = buf0, *out = buf1;
        for (int l = 0; l < ml.n_layers; ++l) {
            const int C = ml.cin[l], U = ml.units[l];
            const bool last = l == ml.n_layers - 1;
            for (int ch = lane; ch < U; ch += 32) {
                const float bias = __ldg(ml.b[l] + ch);
                float vmax = -INFINITY;
                for (int sidx = 0; sidx < slots; ++sidx) {
                    const float *f = in + (size_t)sidx * ml.cmax;
                    float y = bias;
                    for (int k = 0; k < C; ++k) y = fmaf(__ldg(ml.w[l] + k * U + ch), f[k], y);
                    y = fmaxf(y, 0.0f);
                    if (!last) out[(size_t)sidx * ml.cmax + ch] = y;
                    vmax = fmaxf(vmax, y);
                }
                if (last) {
                    if (b >= 0 && b < a.B && cell >= 0 && cell < cells) {
                        const int64_t o2 = ((int64_t)b * cells + cell) * PFN_OUT + ch;
                        if (a.out32) a.out32[o2] = vmax;
                        else {
                            const __half h = __float2half_rn(vmax);
                            a.out_hi[o2] = h;
                            if (a.out_lo) a.out_lo[o2] = __float2half_rn(vmax - __half2float(h));
                        }
                    }
                } else {
                    for (int sidx = 0; sidx < slots; ++sidx) out[(size_t)sidx * ml.cmax + U + ch] = vmax;
                }
            }
            __syncwarp();
            float *t = in; in = out; out = t;
        }
    }
}

static int pfn_multi_run(gnd_ctx *ctx, PfnArgs &a, bool fused, int64_t total, cudaStream_t st) {
    PfnMulti ml = {};
    ml.n_layers = ctx->n_vfe;
    ml.cmax = ctx->nfeat;
    for (int l = 0; l < ctx->n_vfe; ++l) {
        ml.cin[l] = ctx->vfe_in[l]; ml.units[l] = ctx->vfe_units[l];
        ml.w[l] = ctx->vfe_w[l]; ml.b[l] = ctx->vfe_b[l];
        if (ml.cin[l] > ml.cmax) ml.cmax = ml.cin[l];
        if (l < ctx->n_vfe - 1 && 2 * ml.units[l] > ml.cmax) ml.cmax = 2 * ml.units[l];
    }
    const size_t smem = (size_t)2 * (a.P + 1) * ml.cmax * sizeof(float);
    if (smem > 200 * 1024) return fail(GND_ERR_UNSUPPORTED, "multi-layer pillar encoder: %zu bytes of shared memory per pillar (max_points_voxel %d x %d features)", smem, a.P, ml.cmax);
    if (fused) GND_CUDA_CHECK(cudaFuncSetAttribute(pfn_multi_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else GND_CUDA_CHECK(cudaFuncSetAttribute(pfn_multi_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t blocks = total < 148 * 16 ? total : 148 * 16;
    if (blocks < 1) return GND_OK;
    if (fused) pfn_multi_kernel<true><<<(int)blocks, 32, smem, st>>>(a, ml);
    else pfn_multi_kernel<false><<<(int)blocks, 32, smem, st>>>(a, ml);
    GND_LAUNCH_CHECK("pfn_multi_kernel");
    return GND_OK;
}

static void fill_common(const gnd_ctx *ctx, PfnArgs &a, int32_t batch) {
    const gnd_config &c = ctx->cfg;
    a.B = batch;
    a.F = c.input_features;
    a.nfeat = ctx->nfeat;
    a.with_distance = c.with_distance;
    a.P = c.max_points_voxel;
    a.max_voxels = c.max_voxels;
    a.nx = ctx->nx; a.ny = ctx->ny;
    // pointpillars.py:110-113: Python-float arithmetic, then float32 when it meets the tensors
    a.vx = (float)c.voxel_size_f64[0];
    a.vy = (float)c.voxel_size_f64[1];
    a.xoff = (float)(c.voxel_size_f64[0] / 2 + c.pc_range_f64[0]);
    a.yoff = (float)(c.voxel_size_f64[1] / 2 + c.pc_range_f64[1]);
    a.w = ctx->pfn_w; a.bias = ctx->pfn_b;
    a.out32 = ctx->canvas.f32; a.out_hi = ctx->canvas.hi; a.out_lo = ctx->canvas.lo;
}

// Set-up time (gnd_bind_workspace, synchronous): for every unit size 2^s the list of units sorted by the distance of
// their middle cell from the cell under the sensor (x = y = 0 clipped into the grid).
int pfn_build_order(gnd_ctx *ctx) {
    if (!ctx->pfn_order || ctx->nz != 1) return GND_OK;
    const gnd_config &c = ctx->cfg;
    const int nx = ctx->nx, ny = ctx->ny, cells = nx * ny;
    double sx = floor((0.0 - c.pc_range_f64[0]) / c.voxel_size_f64[0]), sy = floor((0.0 - c.pc_range_f64[1]) / c.voxel_size_f64[1]);
    sx = sx < 0 ? 0 : (sx > nx - 1 ? nx - 1 : sx);
    sy = sy < 0 ? 0 : (sy > ny - 1 ? ny - 1 : sy);
    std::vector<int32_t> table((size_t)6 * cells, 0);
    std::vector<std::pair<double, int32_t>> key;
    for (int s = 0; s < 6; ++s) {
        const int units = (cells + (1 << s) - 1) >> s;
        key.resize(units);
        for (int u = 0; u < units; ++u) {
            const int mid = (u << s) + ((1 << s) >> 1) < cells ? (u << s) + ((1 << s) >> 1) : cells - 1;
            const double dx = (double)(mid % nx) - sx, dy = (double)(mid / nx) - sy;
            key[u] = std::make_pair(dx * dx + dy * dy, (int32_t)u);
        }
        std::sort(key.begin(), key.end());
        for (int u = 0; u < units; ++u) table[(size_t)s * cells + u] = key[u].second;
    }
    GND_CUDA_CHECK(cudaMemcpy(ctx->pfn_order, table.data(), table.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    return GND_OK;
}

int pfn_scatter_run(gnd_ctx *ctx, const float *points, int32_t batch, int64_t n_points, cudaStream_t st) {
    if (ctx->nz != 1) return fail(GND_ERR_UNSUPPORTED, "pillar scatter needs a single z layer (nz=%d)", ctx->nz);
    PfnArgs a = {};
    fill_common(ctx, a, batch);
    a.points = points; a.N = n_points;
    a.apply_shift = 1; a.z_shift = ctx->cfg.lidar_height;
    if (!ctx->vws.sorted) return fail(GND_ERR_STATE, "workspace has no cell-ordered point buffer");
    a.offc = ctx->vws.offc; a.sorted = ctx->vws.sorted; a.ctr = ctx->pfn_ctr;
    {
        // kept points lie inside their voxel: |x'|, |y'| <= voxel/2 (+ rounding), |z - zmid| <= z extent / 2
        const gnd_config &c = ctx->cfg;
        const double hxy = 0.5 * (c.voxel_size_f64[0] > c.voxel_size_f64[1] ? c.voxel_size_f64[0] : c.voxel_size_f64[1]) * 1.01 + 1e-3;
        const double hz = 0.5 * (c.pc_range_f64[5] - c.pc_range_f64[2]) * 1.01 + 1e-3;
        const double lim = 1073741824.0 / (double)(c.max_points_voxel > 32 ? c.max_points_voxel : 32);
        int e;
        frexp(lim / hxy, &e); a.q_xy = (float)ldexp(1.0, e - 1); a.iq_xy = (float)ldexp(1.0, 1 - e);
        frexp(lim / hz, &e); a.q_z = (float)ldexp(1.0, e - 1); a.iq_z = (float)ldexp(1.0, 1 - e);
        a.zmid = (float)(0.5 * (c.pc_range_f64[5] + c.pc_range_f64[2]));
    }
    // unit size: 8 cells (measured: 32-cell units leave a 40-60 us tail behind the dense cells next to the sensor --
    // 0.245 ms against 0.206 ms at batch 64 -- and 2-cell units pay for their counter traffic), fewer when that is
    // needed to give every resident warp about four units, down to one cell for a single scan
    const int64_t cells_all = (int64_t)batch * ctx->nx * ctx->ny;
    a.gshift = 3;
    while (a.gshift > 0 && (cells_all >> a.gshift) < 4 * 148 * 4 * PFN_WARPS) --a.gshift;
    if (tuning().pfn_gshift >= 0) a.gshift = tuning().pfn_gshift > 5 ? 5 : tuning().pfn_gshift;
    if (ctx->n_vfe > 1) return pfn_multi_run(ctx, a, true, (int64_t)batch * ctx->nx * ctx->ny, st);
    if (!ctx->pfn_order) return fail(GND_ERR_STATE, "workspace has no work-unit order table");
    a.order = ctx->pfn_order + (size_t)a.gshift * ctx->nx * ctx->ny;
    int64_t warps = (cells_all + (1 << a.gshift) - 1) >> a.gshift;
    int blocks = (int)((warps + PFN_WARPS - 1) / PFN_WARPS);
    if (blocks > 148 * 4) blocks = 148 * 4;      // the resident set: work is handed out dynamically
    const bool has_r = ctx->cfg.input_features > 3;
    if (!ctx->cfg.with_distance && ctx->opt_pfn_path == 2) {      // tensor-core variant (no |p| feature: four regrouped inputs)
        if (has_r) pfn_scatter_mma_kernel<true><<<blocks, PFN_WARPS * 32, 0, st>>>(a);
        else pfn_scatter_mma_kernel<false><<<blocks, PFN_WARPS * 32, 0, st>>>(a);
        GND_LAUNCH_CHECK("pfn_scatter_mma_kernel");
        return GND_OK;
    }
    if (ctx->cfg.with_distance) {
        if (has_r) pfn_scatter_kernel<true, true><<<blocks, PFN_WARPS * 32, 0, st>>>(a);
        else pfn_scatter_kernel<true, false><<<blocks, PFN_WARPS * 32, 0, st>>>(a);
    } else {
        if (has_r) pfn_scatter_kernel<false, true><<<blocks, PFN_WARPS * 32, 0, st>>>(a);
        else pfn_scatter_kernel<false, false><<<blocks, PFN_WARPS * 32, 0, st>>>(a);
    }
    GND_LAUNCH_CHECK("pfn_scatter_kernel");
    return GND_OK;
}

int pfn_scatter_dense_run(gnd_ctx *ctx, const float *voxels, const float *coors4, const float *num_points,
                          int64_t n_pillars, int32_t batch, cudaStream_t st) {
    PfnArgs a = {};
    fill_common(ctx, a, batch);
    a.points = voxels; a.coors4 = coors4; a.num_f = num_points; a.M = n_pillars;
    const size_t n = ctx->canvas.elems(batch);
    if (a.out32) GND_CUDA_CHECK(cudaMemsetAsync(a.out32, 0, n * sizeof(float), st));
    if (a.out_hi) GND_CUDA_CHECK(cudaMemsetAsync(a.out_hi, 0, n * sizeof(__half), st));
    if (a.out_lo) GND_CUDA_CHECK(cudaMemsetAsync(a.out_lo, 0, n * sizeof(__half), st));
    if (n_pillars == 0) return GND_OK;
    if (ctx->n_vfe > 1) return pfn_multi_run(ctx, a, false, n_pillars, st);
    int blocks = (int)((n_pillars + PFN_WARPS - 1) / PFN_WARPS);
    if (blocks > 148 * 8) blocks = 148 * 8;
    pfn_scatter_dense_kernel<<<blocks, PFN_WARPS * 32, 0, st>>>(a);
    GND_LAUNCH_CHECK("pfn_scatter_dense_kernel");
    return GND_OK;
}

}  // namespace gnd
