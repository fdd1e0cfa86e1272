// voxelize.cu -- order-exact parallel restatement of the reference's sequential voxelizer.
//
// Replaces utils/point_cloud_ops.py:7-53 (_points_to_voxel_reverse_kernel), :55-104 and the wrapper
// :107-167.  The reference is one sequential loop whose outputs depend on input order three ways:
//   (1) voxel id   = rank of the cell's FIRST in-range point among all cells' first points (:42-48)
//   (2) point slot = rank of the point among the points of its cell, kept while < max_points (:49-52)
//   (3) overflow   = at the first point that would open voxel number max_voxels the loop BREAKS
//                    (:44-45): that point and every later point is dropped.
// Parallel formulation (three launches, all bit-exact, no unordered slot assignment anywhere):
//   vox_rank_kernel    one CTA per (scan, chunk of T consecutive points): all warps compute the float32
//                      cell of each point (coalesced loads); ONE warp then walks the chunk in input
//                      order, 32 points per step, ranking equal cells inside the step with match.any and
//                      across steps with a chunk-local per-cell counter table in shared memory.  Emits
//                      cell[i], the chunk-local rank, the chunk histogram and atomicMin(first[cell]).
//   vox_assign_kernel  one CTA per scan: a bitmap over point indices of all `first` values gives every
//                      occupied cell its voxel id by popcount-prefix (= order of first appearance);
//                      the cell with rank max_voxels yields the break index; chunk histograms become
//                      exclusive prefixes; num_points and the CSR offsets of the pillar lists follow.
//   vox_scatter_kernel one thread per point: slot = chunk prefix + chunk-local rank; kept if
//                      slot < max_points and index < break index; writes the pillar list (and, for the
//                      drop-in API, the dense voxels tensor / per-point slot).
#include "voxelize.cuh"
#include <limits.h>
#include <math.h>
#include <stdlib.h>

namespace gnd {

struct VoxArgs {
    const float *points;
    const int32_t *counts;
    int32_t B;
    int32_t N;
    int32_t F;
    int32_t vec4;          // rows are 16-byte aligned float4
    float lo[3], vs[3];
    float inv[3];          // exact reciprocal when voxel_size is a power of two (then x / vs == x * inv bit for bit), else 0
    int32_t grid[3];       // nx ny nz
    int32_t cells;
    int32_t K, T, G;
    int32_t kc_shift;      // set-exact path: chunk of point i = i >> kc_shift (< VL_KC); 0 = counts are not chunked
    int32_t apply_shift;
    float z_shift;
    int32_t max_points, max_voxels, reverse;
    VoxWs ws;
    int32_t *coors, *num_points, *voxel_count, *offsets, *list;
    float *voxels;
    int32_t *slot;
};

// utils/point_cloud_ops.py:34-37: c = floor((p - lo) / vs) in IEEE float32 (subtract, then true divide),
// rejected when c < 0 or c >= grid.  NaN compares false on both of the reference's tests (UB there);
// here it is rejected.
__device__ __forceinline__ int axis_cell(float p, float lo, float vs, float inv, int grid) {
    const float d = __fsub_rn(p, lo);
    float c = floorf(inv != 0.0f ? __fmul_rn(d, inv) : __fdiv_rn(d, vs));
    if (!(c >= 0.0f) || !(c < (float)grid)) return -1;
    return (int)c;
}

__device__ __forceinline__ void load_xyz(const VoxArgs &a, const float *row, float &x, float &y, float &z) {
    if (a.vec4) {
        float4 v = __ldg(reinterpret_cast<const float4 *>(row));
        x = v.x; y = v.y; z = v.z;
    } else {
        x = __ldg(row); y = __ldg(row + 1); z = __ldg(row + 2);
    }
}

__device__ __forceinline__ int xyz_cell(const VoxArgs &a, float x, float y, float z) {
    if (a.apply_shift) z = __fadd_rn(z, a.z_shift);   // predict_ground.py:104-107
    int cx = axis_cell(x, a.lo[0], a.vs[0], a.inv[0], a.grid[0]);
    int cy = axis_cell(y, a.lo[1], a.vs[1], a.inv[1], a.grid[1]);
    int cz = axis_cell(z, a.lo[2], a.vs[2], a.inv[2], a.grid[2]);
    if ((cx | cy | cz) < 0) return -1;
    return (cz * a.grid[1] + cy) * a.grid[0] + cx;
}

__device__ __forceinline__ int point_cell(const VoxArgs &a, const float *row) {
    float x, y, z;
    load_xyz(a, row, x, y, z);
    return xyz_cell(a, x, y, z);
}

__global__ void __launch_bounds__(256) vox_rank_kernel(const VoxArgs a) {
    extern __shared__ uint16_t s_tbl[];            // [cells] points of each cell seen so far in this chunk
    __shared__ int32_t s_cell[VOX_SUB];
    __shared__ uint16_t s_rank[VOX_SUB];
    const int b = blockIdx.x / a.K, k = blockIdx.x % a.K;
    const int n = a.counts ? max(0, min(a.counts[b], a.N)) : a.N;
    const int start = k * a.T;
    const int end = min(start + a.T, n);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t base = (int64_t)b * a.N;

    for (int c = tid; c < a.cells; c += 256) s_tbl[c] = 0;
    __syncthreads();

    for (int sub = start; sub < end; sub += VOX_SUB) {
        const int cnt = min(VOX_SUB, end - sub);
        // phase A: cells of this sub-tile (all warps, coalesced)
        for (int j = tid; j < cnt; j += 256) {
            const int64_t gi = base + sub + j;
            int c = point_cell(a, a.points + gi * a.F);
            s_cell[j] = c;
            a.ws.cell[gi] = c;
        }
        __syncthreads();
        // phase B: in-order ranking by one warp
        if (warp == 0) {
            int32_t *first = a.ws.first + (int64_t)b * a.cells;
            for (int j0 = 0; j0 < cnt; j0 += 32) {
                const int j = j0 + lane;
                const int c = (j < cnt) ? s_cell[j] : -1;
                const unsigned act = __ballot_sync(0xffffffffu, c >= 0);
                if (c >= 0) {
                    const unsigned peers = __match_any_sync(act, c);
                    const int leader = __ffs(peers) - 1;
                    const int pre = __popc(peers & lanemask_lt());
                    unsigned seen = 0;
                    if (lane == leader) {
                        seen = s_tbl[c];
                        s_tbl[c] = (uint16_t)(seen + __popc(peers));
                        if (seen == 0) atomicMin(&first[c], sub + j);   // first point of the cell in this chunk
                    }
                    seen = __shfl_sync(peers, seen, leader);
                    s_rank[j] = (uint16_t)(seen + pre);
                }
                __syncwarp();
            }
        }
        __syncthreads();
        // phase C: chunk-local ranks out (coalesced)
        for (int j = tid; j < cnt; j += 256) a.ws.lrank[base + sub + j] = s_rank[j];
    }
    __syncthreads();
    uint32_t *hist = a.ws.hist + ((int64_t)b * a.K + k) * a.cells;
    for (int c = tid; c < a.cells; c += 256) hist[c] = s_tbl[c];
}

// Tile-parallel exact ranking (grids up to VOX_RT_MAX_CELLS cells).  A tile is 8 warps x 224 consecutive
// points; every warp ranks ITS 224 points in input order (7 steps of 32: match.any inside the step, a
// private byte counter per cell across steps), all warps at once.  The eight byte counters of a cell sit in
// one 8-byte word, so a single shared-memory load gives a point the counts of the warps in front of it
// (its rank inside the tile) and the tile total; `s_run` carries the per-cell count over the chunk's tiles.
// Outputs are identical to vox_rank_kernel's: cell[i], chunk-local rank, chunk histogram, atomicMin(first).
__device__ __forceinline__ unsigned byte_sum64(uint2 v) { return __vsadu4(v.x, 0u) + __vsadu4(v.y, 0u); }

__global__ void __launch_bounds__(VOX_RT_WARPS * 32) vox_rank_tiles_kernel(const VoxArgs a) {
    extern __shared__ __align__(16) unsigned char s_rt[];
    uint2 *s_cnt = reinterpret_cast<uint2 *>(s_rt);                         // [cells] byte w = points warp w saw in this tile
    uint16_t *s_run = reinterpret_cast<uint16_t *>(s_cnt + a.cells);        // [cells] points seen in earlier tiles of the chunk
    const int b = blockIdx.x / a.K, k = blockIdx.x % a.K;
    const int n = a.counts ? max(0, min(a.counts[b], a.N)) : a.N;
    const int start = k * a.T;
    const int end = min(start + a.T, n);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t base = (int64_t)b * a.N;
    int32_t *first = a.ws.first + (int64_t)b * a.cells;

    for (int c = tid; c < a.cells; c += VOX_RT_WARPS * 32) { s_cnt[c] = make_uint2(0u, 0u); s_run[c] = 0; }
    __syncthreads();

    const uint2 below = warp == 0 ? make_uint2(0u, 0u)
                      : warp < 4  ? make_uint2(0xffffffffu >> (32 - 8 * warp), 0u)
                      : warp == 4 ? make_uint2(0xffffffffu, 0u)
                                  : make_uint2(0xffffffffu, 0xffffffffu >> (32 - 8 * (warp - 4)));
    // the next tile's coordinates are fetched while this tile is ranked
    float px[VOX_RT_STEPS], py[VOX_RT_STEPS], pz[VOX_RT_STEPS];
#pragma unroll
    for (int s = 0; s < VOX_RT_STEPS; ++s) {
        const int i = start + warp * (VOX_RT_STEPS * 32) + s * 32 + lane;
        px[s] = py[s] = pz[s] = 0.0f;
        if (i < end) load_xyz(a, a.points + (base + i) * a.F, px[s], py[s], pz[s]);
    }
    for (int t0 = start; t0 < end; t0 += VOX_RT_TILE) {
        const int wbase = t0 + warp * (VOX_RT_STEPS * 32);
        int c[VOX_RT_STEPS], r[VOX_RT_STEPS];
#pragma unroll
        for (int s = 0; s < VOX_RT_STEPS; ++s) {
            const int i = wbase + s * 32 + lane;
            c[s] = (i < end) ? xyz_cell(a, px[s], py[s], pz[s]) : -1;
        }
#pragma unroll
        for (int s = 0; s < VOX_RT_STEPS; ++s) {
            const int i = wbase + VOX_RT_TILE + s * 32 + lane;
            if (i < end) load_xyz(a, a.points + (base + i) * a.F, px[s], py[s], pz[s]);
        }
        // phase A: in-order ranking inside the warp's 224 points.  The seven match.any are independent of the
        // counters, so they are issued first (their latency overlaps); only the counter read-modify-write of the
        // step leaders is serial across steps.
        unsigned peers[VOX_RT_STEPS];
#pragma unroll
        for (int s = 0; s < VOX_RT_STEPS; ++s) {
            const unsigned act = __ballot_sync(0xffffffffu, c[s] >= 0);
            peers[s] = (c[s] >= 0) ? __match_any_sync(act, c[s]) : 0u;
        }
#pragma unroll
        for (int s = 0; s < VOX_RT_STEPS; ++s) {
            r[s] = 0;
            if (c[s] >= 0) {
                const int leader = __ffs(peers[s]) - 1;
                unsigned seen = 0;
                unsigned char *cnt = reinterpret_cast<unsigned char *>(s_cnt + c[s]) + warp;
                if (lane == leader) { seen = *cnt; *cnt = (unsigned char)(seen + __popc(peers[s])); }
                seen = __shfl_sync(peers[s], seen, leader);
                r[s] = (int)seen + __popc(peers[s] & lanemask_lt());
            }
            __syncwarp();
        }
        __syncthreads();
        // phase B: rank inside the tile = counts of the warps in front + rank inside the warp
        int run[VOX_RT_STEPS], tot[VOX_RT_STEPS];
#pragma unroll
        for (int s = 0; s < VOX_RT_STEPS; ++s) {
            const int i = wbase + s * 32 + lane;
            run[s] = -1; tot[s] = 0;
            if (c[s] >= 0) {
                const uint2 v = s_cnt[c[s]];
                const int pre = (int)byte_sum64(make_uint2(v.x & below.x, v.y & below.y));
                const int rs = (int)s_run[c[s]];
                if (pre + r[s] == 0) { run[s] = rs; tot[s] = (int)byte_sum64(v); }    // first point of its cell in the tile
                r[s] += pre + rs;
            }
            if (i < end) {
                a.ws.cell[base + i] = c[s];
                a.ws.lrank[base + i] = (uint16_t)r[s];
            }
        }
        __syncthreads();
        // phase C: the cell's first point of the tile folds the tile into the chunk state and clears the counters
#pragma unroll
        for (int s = 0; s < VOX_RT_STEPS; ++s) {
            if (run[s] >= 0) {
                if (run[s] == 0) atomicMin(&first[c[s]], wbase + s * 32 + lane);
                s_run[c[s]] = (uint16_t)(run[s] + tot[s]);
                s_cnt[c[s]] = make_uint2(0u, 0u);
            }
        }
        __syncthreads();
    }
    uint32_t *hist = a.ws.hist + ((int64_t)b * a.K + k) * a.cells;
    for (int c = tid; c < a.cells; c += VOX_RT_WARPS * 32) hist[c] = s_run[c];
}


// Exclusive scan of a global int array by ONE CTA of 1024 threads: every thread owns four consecutive elements (four
// independent loads in flight), one block scan per 4096 elements.  (A serial per-thread walk over n / 1024 consecutive
// elements costs one L2 round trip per element: measured 28 k clocks for 10 000 cells.)  `in` may equal `out`.
// Returns the total.
__device__ __forceinline__ int cta_exclusive_scan_global(const int32_t *in, int32_t *out, int n, int *s_warp) {
    const int tid = threadIdx.x;
    int carry = 0;
    for (int base = 0; base < n; base += 4096) {
        const int i0 = base + 4 * tid;
        int v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = (i0 + j < n) ? __ldcg(in + i0 + j) : 0;
        int tot;
        int run = carry + block_exclusive_scan(v[0] + v[1] + v[2] + v[3], s_warp, &tot);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (i0 + j < n) out[i0 + j] = run;
            run += v[j];
        }
        carry += tot;
    }
    return carry;
}

// G CTAs per scan.  Every CTA rebuilds the first-point bitmap and its popcount prefix (cheap: one load per
// cell), then ranks / prefixes / truncates only ITS slice of the cells; the CTA that finishes last for a scan
// turns the per-cell and per-voxel kept counts into the two CSR offset arrays.
constexpr int VOX_DONE_BASE = 0x7f7f7f7f;     // `done` shares the 0x7f memset of `first`

__global__ void __launch_bounds__(1024) vox_assign_kernel(const VoxArgs a) {
    extern __shared__ uint32_t s_bm[];             // bit i set <=> point i is the first point of a cell
    __shared__ int s_tpre[1024];
    __shared__ int s_warp[33];
    __shared__ int s_cut, s_last;
    const int G = a.G;
    const int b = blockIdx.x / G, part = blockIdx.x - b * G, tid = threadIdx.x;
    const int n = a.counts ? max(0, min(a.counts[b], a.N)) : a.N;
    const int nW = (n + 31) >> 5;
    const int32_t *first = a.ws.first + (int64_t)b * a.cells;
    int32_t *vid = a.ws.vid + (int64_t)b * a.cells;
    uint32_t *hist = a.ws.hist + (int64_t)b * a.K * a.cells;
    int32_t *coors = a.coors + (int64_t)b * a.max_voxels * 3;
    int32_t *num = a.num_points + (int64_t)b * a.max_voxels;
    int32_t *off = a.offsets + (int64_t)b * (a.max_voxels + 1);
    int32_t *offc = a.ws.offc + (int64_t)b * (a.cells + 1);

    for (int w = tid; w < nW; w += 1024) s_bm[w] = 0;
    if (tid == 0) s_cut = INT_MAX;
    __syncthreads();
    for (int c0 = tid; c0 < a.cells; c0 += 4096) {            // four independent loads in flight
        int f[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) f[j] = (c0 + j * 1024 < a.cells) ? __ldcg(first + c0 + j * 1024) : INT_MAX;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (f[j] < n) atomicOr(&s_bm[f[j] >> 5], 1u << (f[j] & 31));
    }
    __syncthreads();
    const int seg = (nW + 1023) / 1024;
    int local = 0;
    for (int w = tid * seg; w < min((tid + 1) * seg, nW); ++w) local += __popc(s_bm[w]);
    int n_occ;
    int pre = block_exclusive_scan(local, s_warp, &n_occ);
    s_tpre[tid] = pre;
    const int M = min(n_occ, a.max_voxels);
    // break index (:44-45): the first point of the cell whose rank is max_voxels, i.e. set bit number max_voxels
    if (n_occ > a.max_voxels && pre <= a.max_voxels && a.max_voxels < pre + local) {
        int r = pre;
        for (int w = tid * seg; w < min((tid + 1) * seg, nW); ++w) {
            uint32_t m = s_bm[w];
            const int pc = __popc(m);
            if (r + pc > a.max_voxels) {
                for (int k = a.max_voxels - r; k > 0; --k) m &= m - 1;
                s_cut = w * 32 + __ffs(m) - 1;
                break;
            }
            r += pc;
        }
    }
    __syncthreads();
    const int cut = s_cut;
    const int kcut = (cut < n) ? cut / a.T : a.K;
    const bool overflow = cut < n;

    // this CTA's slice of the cells
    const int per = (a.cells + G - 1) / G;
    const int c_lo = min(part * per, a.cells), c_hi = min(c_lo + per, a.cells);
    const int nx = a.grid[0], ny = a.grid[1];
    int cx = (c_lo + tid) % nx, cyz = (c_lo + tid) / nx;          // c = cyz * nx + cx, advanced by 1024 per iteration
    const int step_x = 1024 % nx, step_yz = 1024 / nx;
    for (int c = c_lo + tid; c < c_hi; c += 1024) {
        // voxel id = number of cells whose first point comes earlier (:42-48)
        const int f = first[c];
        int v = -1;
        if (f < n) {
            const int w = f >> 5, t = w / seg;
            int r = s_tpre[t];
            for (int w2 = t * seg; w2 < w; ++w2) r += __popc(s_bm[w2]);
            r += __popc(s_bm[w] & ((1u << (f & 31)) - 1u));
            if (r < a.max_voxels) {
                v = r;
                const int cy = (a.grid[2] == 1) ? cyz : cyz % ny, cz = (a.grid[2] == 1) ? 0 : cyz / ny;
                if (a.reverse) { coors[v * 3 + 0] = cz; coors[v * 3 + 1] = cy; coors[v * 3 + 2] = cx; }
                else           { coors[v * 3 + 0] = cx; coors[v * 3 + 1] = cy; coors[v * 3 + 2] = cz; }
            }
        }
        vid[c] = v;
        cx += step_x; cyz += step_yz;
        if (cx >= nx) { cx -= nx; ++cyz; }
        // chunk histograms -> exclusive prefixes over chunks; points of the cell before the break index.
        // (loads first, stores after: the K loads of a cell are independent and stay in flight together)
        uint32_t run = 0, before_cut = 0;
        for (int k0 = 0; k0 < a.K; k0 += 8) {
            uint32_t h[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) h[j] = (k0 + j < a.K) ? hist[(int64_t)(k0 + j) * a.cells + c] : 0u;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (k0 + j < a.K) {
                    hist[(int64_t)(k0 + j) * a.cells + c] = run;
                    if (k0 + j == kcut) before_cut = run;
                    run += h[j];
                }
            }
        }
        if (!overflow) {          // truncate (:49-52) right here: kept points of the cell / voxel
            const int kept = (v >= 0) ? min((int)run, a.max_points) : 0;
            if (v >= 0) num[v] = kept;
            offc[c] = kept;
        } else if (v >= 0) {
            num[v] = (int)before_cut;
        }
    }
    if (overflow) {   // add the points of the break chunk that precede the break index, then truncate
        __syncthreads();
        const int32_t *cell = a.ws.cell + (int64_t)b * a.N;
        for (int j = kcut * a.T + tid; j < cut; j += 1024) {
            const int c = cell[j];
            if (c >= c_lo && c < c_hi) { const int v = vid[c]; if (v >= 0) atomicAdd(&num[v], 1); }
        }
        __syncthreads();
        for (int c = c_lo + tid; c < c_hi; c += 1024) {
            const int v = vid[c];
            int kept = 0;
            if (v >= 0) { kept = min(num[v], a.max_points); num[v] = kept; }
            offc[c] = kept;
        }
    }
    if (G > 1) {
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            int32_t *done = a.ws.done + b;
            const int t = atomicAdd(done, 1);
            s_last = (t == VOX_DONE_BASE + G - 1);
            if (s_last) *done = VOX_DONE_BASE;
            __threadfence();
        }
        __syncthreads();
        if (!s_last) return;
    } else {
        __syncthreads();
    }
    // CSR offsets of the kept points in CELL order (what the pillar encoder walks), in place over the kept counts.
    // Each thread owns a run of consecutive cells: one block scan instead of one per 1024 cells.
    {
        const int S = cta_exclusive_scan_global(offc, offc, a.cells, s_warp);
        if (tid == 0) offc[a.cells] = S;
        // ... and in VOXEL order (the pillar lists of the drop-in API)
        cta_exclusive_scan_global(num, off, M, s_warp);
        if (tid == 0) {
            off[M] = S;
            a.voxel_count[b] = M;
            a.ws.cutoff[b] = cut;
        }
    }
}

// grid.y = (scan, chunk): no per-point divisions; grid.x strides over the chunk's points
__global__ void __launch_bounds__(256) vox_scatter_kernel(const VoxArgs a) {
    const int b = blockIdx.y / a.K, k = blockIdx.y - b * a.K;
    const int n = a.counts ? max(0, min(a.counts[b], a.N)) : a.N;
    const int cstart = k * a.T, cend = min(cstart + a.T, a.N);
    const int cut = a.ws.cutoff[b];
    const int64_t base = (int64_t)b * a.N;
    const int32_t *vid = a.ws.vid + (int64_t)b * a.cells;
    const uint32_t *pre = a.ws.hist + ((int64_t)b * a.K + k) * a.cells;
    const int32_t *off = a.offsets + (int64_t)b * (a.max_voxels + 1);
    const int32_t *offc = a.ws.offc + (int64_t)b * (a.cells + 1);
    for (int i = cstart + blockIdx.x * 256 + threadIdx.x; i < cend; i += gridDim.x * 256) {
        const int64_t gi = base + i;
        if (i >= n) { if (a.slot) a.slot[gi] = -1; continue; }
        const int c = a.ws.cell[gi];
        int slot = -1;
        if (c >= 0 && i < cut) {
            const int v = __ldg(vid + c);
            if (v >= 0) {
                const int r = (int)a.ws.lrank[gi] + (int)__ldg(pre + c);
                if (r < a.max_points) {
                    slot = v * a.max_points + r;
                    a.list[base + __ldg(off + v) + r] = i;
                    if (a.ws.sorted) {
                        const float *src = a.points + gi * a.F;
                        float4 pt;
                        if (a.vec4) pt = __ldg(reinterpret_cast<const float4 *>(src));
                        else pt = make_float4(__ldg(src), __ldg(src + 1), __ldg(src + 2), a.F > 3 ? __ldg(src + 3) : 0.0f);
                        if (a.apply_shift) pt.z = __fadd_rn(pt.z, a.z_shift);
                        a.ws.sorted[base + __ldg(offc + c) + r] = pt;
                    }
                    if (a.voxels) {
                        const float *src = a.points + gi * a.F;
                        float *dst = a.voxels + (((int64_t)b * a.max_voxels + v) * a.max_points + r) * a.F;
                        for (int f = 0; f < a.F; ++f) {
                            float val = src[f];   // _shift_cloud adds [0,0,h,0] to the first 4 columns
                            if (a.apply_shift && f < 4) val = __fadd_rn(val, f == 2 ? a.z_shift : 0.0f);
                            dst[f] = val;
                        }
                    }
                }
            }
        }
        if (a.slot) a.slot[gi] = slot;
    }
}

// =============================================================================================================
// Set-exact ("lean") voxelization for the fused pipeline (gnd_infer).
//
// What the rest of the path consumes is, per cell, the SET of its first max_points points in input order (the pillar
// encoder's max and its fixed-point mean are order-independent inside a pillar), plus the bookkeeping the reference
// returns: voxel ids in order of first appearance (coors), num_points, and the max_voxels break.  None of that needs the
// rank of EVERY point inside its cell -- only, for the few over-full cells, the index of their max_points-th point:
//   vl_count     every point: cell (16 bit) out, RED.ADD on the cell's count, RED.MIN on its first index (one atomic
//                pair per run of equal cells inside a warp: LiDAR rings stay in a cell for many consecutive points).
//                The count is kept per CHUNK of the index range (VL_KC = 8 chunks of 2^kc_shift points): for an
//                over-full cell the chunk that holds its max_points-th point follows from a prefix over 8 numbers, every
//                point in an earlier chunk is kept outright, and only the points of THAT chunk are candidates -- an
//                eighth of the collect / select work on a LiDAR scan, and what makes the path pay on dense camera
//                frames where every cell is over-full
//   vl_cells     per cell: voxel id by popcount-prefix over the first-point bitmap, the break index, kept counts and
//                the two CSR offset arrays (as vox_assign_kernel); over-full cells go on a compact list
//   vl_collect   points of over-full cells (before the break) -> the cell's candidate list
//   vl_select    one warp per over-full cell: the max_points-th smallest candidate index by a two-level radix select
//                (1024-index buckets, then a 1024-bit bitmap) = the cell's keep threshold
//   vl_fill      every point: kept iff its cell has a voxel, index < break and index <= threshold; slot by atomicAdd
//                (any order); writes the cell-ordered float4 the pillar encoder reads
// Same coors / num_points / voxel_count / cell offsets / kept sets as the order-exact kernels above (tests compare both paths
// bit for bit, heights included); the per-voxel point index lists in slot order are produced only by the exact path
// (gnd_voxelize, gnd_points_to_voxel).  All per-cell state is private to this path and restored by the kernels themselves
// (vox_lean_init sets it once at bind time).
constexpr int VL_NONE = 0x7f7f7f7f;
constexpr int VL_SHIFT = 10;                    // select buckets of 1024 indices; N <= 1.6 M -> <= 1563 buckets

__global__ void __launch_bounds__(256) vl_count_kernel(const VoxArgs a) {
    const int b = blockIdx.y;
    const int n = a.counts ? max(0, min(a.counts[b], a.N)) : a.N;
    const int64_t base = (int64_t)b * a.N;
    uint16_t *cell16 = a.ws.lrank + base;
    int32_t *first = a.ws.lfirst + (int64_t)b * a.cells;
    int32_t *cnt = a.kc_shift ? a.ws.lcntk + (int64_t)b * a.cells * VL_KC : a.ws.lcnt + (int64_t)b * a.cells;
    const int lane = threadIdx.x & 31;
    if (blockIdx.x == 0 && threadIdx.x < 4) a.ws.lstat[b * 4 + threadIdx.x] = 0;      // consumed by the kernels that follow
    const int nround = (n + 31) & ~31;                                                   // whole warps stay converged
    const int stride = gridDim.x * 256;
    for (int i0 = blockIdx.x * 256 + threadIdx.x; i0 < nround; i0 += 2 * stride) {
        // two 32-point groups per warp and iteration: both rows are loaded before either is used
        const int i1 = i0 + stride;
        float x0 = 0.f, y0 = 0.f, z0 = 0.f, x1 = 0.f, y1 = 0.f, z1 = 0.f;
        if (i0 < n) load_xyz(a, a.points + (base + i0) * a.F, x0, y0, z0);
        if (i1 < n) load_xyz(a, a.points + (base + i1) * a.F, x1, y1, z1);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int i = h ? i1 : i0;
            if (i - lane >= nround) break;                // warp-uniform
            int c = -1;
            if (i < n) {
                c = h ? xyz_cell(a, x1, y1, z1) : xyz_cell(a, x0, y0, z0);
                cell16[i] = (uint16_t)(c < 0 ? 0xffff : c);
            }
            // runs of equal cells among the 32 consecutive points of the warp: the run's first lane speaks for all of it
            const int prev = __shfl_up_sync(0xffffffffu, c, 1);
            const bool head = (lane == 0) || (c != prev);
            const unsigned heads = __ballot_sync(0xffffffffu, head);
            if (head && c >= 0) {
                const unsigned later = heads & ~((2u << lane) - 1u);
                const int len = (later ? __ffs(later) - 1 : 32) - lane;
                // (the 32 points of a warp step share a chunk: chunks are multiples of 32 points)
                atomicAdd(a.kc_shift ? cnt + c * VL_KC + (i >> a.kc_shift) : cnt + c, len);
                atomicMin(first + c, i);
            }
        }
    }
}

// Slot hand-out for a run of equal cells inside a warp: the run's first lane takes `take` consecutive slots with ONE
// atomicAdd and every member lane gets base + its position in the run (same-address atomics with a return value are
// served one at a time; a ring of a LiDAR stays in one cell for many consecutive points).
struct RunSlot { int raw, head_lane; };
// phase 1: run analysis + the atomic (its result is not consumed here, so several groups' atomics overlap)
__device__ __forceinline__ RunSlot run_slots_issue(int32_t *counter_of_cell, int c, bool member, int lane) {
    const int key = member ? c : -1 - lane;               // non-members never join a run
    const int prev = __shfl_up_sync(0xffffffffu, key, 1);
    const bool head = (lane == 0) || (key != prev);
    const unsigned heads = __ballot_sync(0xffffffffu, head);
    const unsigned upto = heads & ((2u << lane) - 1u);    // heads at or before this lane
    const unsigned later = heads & ~((2u << lane) - 1u);
    const int len = (later ? __ffs(later) - 1 : 32) - lane;      // valid in head lanes
    RunSlot r;
    r.head_lane = 31 - __clz(upto);
    r.raw = (head && member) ? atomicAdd(counter_of_cell, len) : 0;
    return r;
}
// phase 2: this lane's slot
__device__ __forceinline__ int run_slots_take(const RunSlot &r, int lane) {
    return __shfl_sync(0xffffffffu, r.raw, r.head_lane) + (lane - r.head_lane);
}

__global__ void __launch_bounds__(1024) vl_cells_kernel(const VoxArgs a) {
    extern __shared__ uint32_t s_bm[];             // bit i set <=> point i is the first point of a cell
    __shared__ int s_tpre[1024];
    __shared__ int s_warp[33];
    __shared__ int s_cut, s_last;
    const int G = a.G;
    const int b = blockIdx.x / G, part = blockIdx.x - b * G, tid = threadIdx.x;
    const int n = a.counts ? max(0, min(a.counts[b], a.N)) : a.N;
    const int nW = (n + 31) >> 5;
    int32_t *first = a.ws.lfirst + (int64_t)b * a.cells;
    int32_t *cnt = a.ws.lcnt + (int64_t)b * a.cells;
    int32_t *fill = a.ws.lfill + (int64_t)b * a.cells;
    int32_t *thr = a.ws.lthr + (int64_t)b * a.cells;
    int32_t *hoff = a.ws.lhoff + (int64_t)b * a.cells;
    int32_t *hot = a.ws.lhot + (int64_t)b * a.cells;
    int32_t *stat = a.ws.lstat + b * 4;
    int32_t *vid = a.ws.vid + (int64_t)b * a.cells;
    int32_t *coors = a.coors + (int64_t)b * a.max_voxels * 3;
    int32_t *num = a.num_points + (int64_t)b * a.max_voxels;
    int32_t *off = a.offsets + (int64_t)b * (a.max_voxels + 1);
    int32_t *offc = a.ws.offc + (int64_t)b * (a.cells + 1);

    for (int w = tid; w < nW; w += 1024) s_bm[w] = 0;
    if (tid == 0) s_cut = INT_MAX;
    __syncthreads();
    for (int c0 = tid; c0 < a.cells; c0 += 4096) {            // four independent loads in flight
        int f[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) f[j] = (c0 + j * 1024 < a.cells) ? __ldcg(first + c0 + j * 1024) : INT_MAX;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (f[j] < n) atomicOr(&s_bm[f[j] >> 5], 1u << (f[j] & 31));
    }
    __syncthreads();
    const int seg = (nW + 1023) / 1024;
    int local = 0;
    for (int w = tid * seg; w < min((tid + 1) * seg, nW); ++w) local += __popc(s_bm[w]);
    int n_occ;
    const int pre = block_exclusive_scan(local, s_warp, &n_occ);
    s_tpre[tid] = pre;
    const int M = min(n_occ, a.max_voxels);
    // break index (point_cloud_ops.py:44-45): the first point of the cell whose rank is max_voxels
    if (n_occ > a.max_voxels && pre <= a.max_voxels && a.max_voxels < pre + local) {
        int r = pre;
        for (int w = tid * seg; w < min((tid + 1) * seg, nW); ++w) {
            uint32_t m = s_bm[w];
            const int pc = __popc(m);
            if (r + pc > a.max_voxels) {
                for (int k = a.max_voxels - r; k > 0; --k) m &= m - 1;
                s_cut = w * 32 + __ffs(m) - 1;
                break;
            }
            r += pc;
        }
    }
    __syncthreads();
    const int cut = s_cut;
    const bool overflow = cut < n;

    const int per = (a.cells + G - 1) / G;
    const int c_lo = min(part * per, a.cells), c_hi = min(c_lo + per, a.cells);
    const int nx = a.grid[0], ny = a.grid[1];
    // pass 1 over the slice: voxel ids and coors; with the break, counts are re-taken over the points before it
    for (int c = c_lo + tid; c < c_hi; c += 1024) {
        const int f = __ldcg(first + c);
        int v = -1;
        if (f < n) {
            const int w = f >> 5, t = w / seg;
            int r = s_tpre[t];
            for (int w2 = t * seg; w2 < w; ++w2) r += __popc(s_bm[w2]);
            r += __popc(s_bm[w] & ((1u << (f & 31)) - 1u));
            if (r < a.max_voxels) {
                v = r;
                const int cx = c % nx, cyz = c / nx;
                const int cy = (a.grid[2] == 1) ? cyz : cyz % ny, cz = (a.grid[2] == 1) ? 0 : cyz / ny;
                if (a.reverse) { coors[v * 3 + 0] = cz; coors[v * 3 + 1] = cy; coors[v * 3 + 2] = cx; }
                else           { coors[v * 3 + 0] = cx; coors[v * 3 + 1] = cy; coors[v * 3 + 2] = cz; }
            }
        }
        vid[c] = v;
        fill[c] = 0;
    }
    if (overflow) {
        __syncthreads();
        const uint16_t *cell16 = a.ws.lrank + (int64_t)b * a.N;
        for (int j = tid; j < cut; j += 1024) {
            const int c = cell16[j];
            if (c >= c_lo && c < c_hi && vid[c] >= 0) atomicAdd(fill + c, 1);
        }
        __syncthreads();
    }
    // pass 2: kept counts, over-full cells (their list slots / candidate ranges come from one atomicAdd pair per CTA)
    int my_hot = 0, my_cand = 0;
    for (int c = c_lo + tid; c < c_hi; c += 1024) {
        int have = overflow ? fill[c] : __ldcg(cnt + c);
        if (vid[c] < 0) have = 0;
        if (have > a.max_points) { ++my_hot; my_cand += have; }
    }
    int tot_hot, tot_cand;
    int hot_pos = block_exclusive_scan(my_hot, s_warp, &tot_hot);
    int cand_pos = block_exclusive_scan(my_cand, s_warp, &tot_cand);
    __shared__ int s_hot0, s_cand0;
    if (tid == 0) {
        s_hot0 = tot_hot ? atomicAdd(stat + 0, tot_hot) : 0;
        s_cand0 = tot_cand ? atomicAdd(stat + 1, tot_cand) : 0;
    }
    __syncthreads();
    hot_pos += s_hot0; cand_pos += s_cand0;
    for (int c = c_lo + tid; c < c_hi; c += 1024) {
        const int v = vid[c];
        int have = __ldcg(cnt + c);
        cnt[c] = 0;                                       // restored for the next call
        if (overflow) { have = fill[c]; fill[c] = 0; }
        if (v < 0) have = 0;
        const int kept = min(have, a.max_points);
        if (v >= 0) num[v] = kept;
        offc[c] = kept;
        if (have > a.max_points) {
            hot[hot_pos++] = c;
            hoff[c] = cand_pos;
            cand_pos += have;
            thr[c] = -1;                                  // set by vl_select_kernel
        } else {
            hoff[c] = -1;
            thr[c] = (v >= 0) ? INT_MAX : -1;             // no voxel (beyond max_voxels): nothing of the cell is kept
        }
    }
    if (G > 1) {
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            s_last = (atomicAdd(stat + 2, 1) == G - 1);
            __threadfence();
        }
        __syncthreads();
        if (!s_last) return;
    } else {
        __syncthreads();
    }
    // every CTA of the scan has read `first`: restore it, then the CSR offsets (cell order / voxel order)
    for (int c = tid; c < a.cells; c += 1024) first[c] = VL_NONE;
    {
        const int S = cta_exclusive_scan_global(offc, offc, a.cells, s_warp);
        if (tid == 0) offc[a.cells] = S;
        cta_exclusive_scan_global(num, off, M, s_warp);
        if (tid == 0) {
            off[M] = S;
            a.voxel_count[b] = M;
            a.ws.cutoff[b] = cut;
        }
    }
}

// The same per-cell step for grids of up to 1024 * VL_CPT cells (both shipped grids): ONE CTA per scan, every thread
// keeps its cells (c = j * 1024 + tid: coalesced) in registers from the first load to the last store.  The multi-CTA
// kernel above is bound by ~14 dependent L2 round trips (~1.3 k clocks each on this part: re-reading first / counts /
// kept counts that other CTAs wrote); here there are two (the cell loads, and the voxel-order offsets).
constexpr int VL_CPT = 16;

// exclusive scan of one 64-bit value per thread over a 1024-thread CTA (three packed counters per value)
__device__ __forceinline__ unsigned long long block_exclusive_scan64(unsigned long long v, unsigned long long *s_w,
                                                                     unsigned long long *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        const unsigned long long w = s_w[lane];
        unsigned long long winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += t;
        }
        s_w[lane] = winc - w;
        if (lane == 31) s_w[32] = winc;
    }
    __syncthreads();
    const unsigned long long res = s_w[warp] + inc - v;
    *total = s_w[32];
    __syncthreads();
    return res;
}

// Chunked counts of one cell (VL_KC consecutive int32): the total, and for an over-full cell the chunk k* that holds its
// max_points-th point, that point's rank r inside the chunk (1-based) and the chunk's population (the candidates).
struct ChunkSplit { int total, kstar, r, cand; };
__device__ __forceinline__ ChunkSplit chunk_split(const int32_t *ck, int max_points) {
    const int4 lo = __ldcg(reinterpret_cast<const int4 *>(ck)), hi = __ldcg(reinterpret_cast<const int4 *>(ck) + 1);
    const int v[VL_KC] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
    ChunkSplit s;
    s.total = 0; s.kstar = 0; s.r = max_points; s.cand = 0;
    bool found = false;
#pragma unroll
    for (int k = 0; k < VL_KC; ++k) {
        if (!found && s.total + v[k] >= max_points) { found = true; s.kstar = k; s.r = max_points - s.total; s.cand = v[k]; }
        s.total += v[k];
    }
    return s;
}
// keep-threshold slot of an over-full cell between the per-cell kernel and vl_select: which chunk to collect and the rank
// to select, as a negative number (-1 = the whole index range, rank max_points: un-chunked counts / after the break)
__device__ __forceinline__ int thr_encode(int kstar, int r) { return -2 - (kstar | (r << 4)); }
__device__ __forceinline__ void thr_decode(int t, int max_points, int &kstar, int &r) {
    if (t >= -1) { kstar = -1; r = max_points; }
    else { const int code = -2 - t; kstar = code & 15; r = code >> 4; }
}

template <bool CHUNKED>
__global__ void __launch_bounds__(1024) vl_cells1_kernel(const VoxArgs a) {
    extern __shared__ uint32_t s_bm[];             // [nW] first-point bitmap, then (8-byte aligned) [cells] packed counters
    __shared__ int s_tpre[1024];
    __shared__ int s_warp[33];
    __shared__ unsigned long long s_w64[33];
    __shared__ int s_cut, s_last;
    const int G = a.G;
    const int b = blockIdx.x / G, part = blockIdx.x - b * G, tid = threadIdx.x;
    const int n = a.counts ? max(0, min(a.counts[b], a.N)) : a.N;
    const int nW = (n + 31) >> 5;
    const int nWcap = (a.N + 31) >> 5;             // layout by the capacity, not by this scan's count
    unsigned long long *s_pk = reinterpret_cast<unsigned long long *>(s_bm + ((nWcap + 2) & ~1));
    int32_t *first = a.ws.lfirst + (int64_t)b * a.cells;
    int32_t *cnt = a.ws.lcnt + (int64_t)b * a.cells;
    const int32_t *cntk = a.ws.lcntk + (int64_t)b * a.cells * VL_KC;
    int32_t *fill = a.ws.lfill + (int64_t)b * a.cells;
    int32_t *thr = a.ws.lthr + (int64_t)b * a.cells;
    int32_t *hoff = a.ws.lhoff + (int64_t)b * a.cells;
    int32_t *hot = a.ws.lhot + (int64_t)b * a.cells;
    int32_t *stat = a.ws.lstat + b * 4;
    int32_t *vid = a.ws.vid + (int64_t)b * a.cells;
    int32_t *coors = a.coors + (int64_t)b * a.max_voxels * 3;
    int32_t *num = a.num_points + (int64_t)b * a.max_voxels;
    int32_t *off = a.offsets + (int64_t)b * (a.max_voxels + 1);
    int32_t *offc = a.ws.offc + (int64_t)b * (a.cells + 1);
    const int cpt = (a.cells + 1023) / 1024;       // <= VL_CPT (checked by the launcher)
    constexpr bool chunked = CHUNKED;         // compile-time: the un-chunked instance keeps its 2 * cpt plain loads in flight

    // every CTA of the scan looks at ALL cells (first index + count, 2 * cpt coalesced loads in flight): the voxel ids
    // and, when the break does not fire, every offset follow from them without any exchange between the CTAs
    int fa[VL_CPT], ha[VL_CPT];
#pragma unroll
    for (int j = 0; j < VL_CPT; ++j) {
        const int c = j * 1024 + tid;
        const bool ok = j < cpt && c < a.cells;
        fa[j] = ok ? __ldcg(first + c) : INT_MAX;
        // ha: the cell's point count, or for an over-full cell MINUS the number of its candidates (chunked counts: the
        // population of the chunk that holds the max_points-th point; else all of its points)
        if (!ok) ha[j] = 0;
        else if (chunked) {
            const ChunkSplit cs = chunk_split(cntk + (int64_t)c * VL_KC, a.max_points);
            ha[j] = cs.total > a.max_points ? -cs.cand : cs.total;
        } else {
            const int h = __ldcg(cnt + c);
            ha[j] = h > a.max_points ? -h : h;
        }
    }
    for (int w = tid; w < nW; w += 1024) s_bm[w] = 0;
    if (tid == 0) s_cut = INT_MAX;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < VL_CPT; ++j)
        if (fa[j] < n) atomicOr(&s_bm[fa[j] >> 5], 1u << (fa[j] & 31));
    __syncthreads();
    const int seg = (nW + 1023) / 1024;
    int local = 0;
    for (int w = tid * seg; w < min((tid + 1) * seg, nW); ++w) local += __popc(s_bm[w]);
    int n_occ;
    const int pre = block_exclusive_scan(local, s_warp, &n_occ);
    s_tpre[tid] = pre;
    const int M = min(n_occ, a.max_voxels);
    if (n_occ > a.max_voxels && pre <= a.max_voxels && a.max_voxels < pre + local) {      // the break index (:44-45)
        int r = pre;
        for (int w = tid * seg; w < min((tid + 1) * seg, nW); ++w) {
            uint32_t m = s_bm[w];
            const int pc = __popc(m);
            if (r + pc > a.max_voxels) {
                for (int k = a.max_voxels - r; k > 0; --k) m &= m - 1;
                s_cut = w * 32 + __ffs(m) - 1;
                break;
            }
            r += pc;
        }
    }
    __syncthreads();
    const int cut = s_cut;
    const bool overflow = cut < n;                 // the same in every CTA of the scan
    const int per = (a.cells + G - 1) / G;
    const int c_lo = min(part * per, a.cells), c_hi = min(c_lo + per, a.cells);
    const int nx = a.grid[0], ny = a.grid[1];

    auto voxel_of = [&](int c, int f) -> int {     // rank of the cell's first point among all first points (:42-48)
        if (f >= n) return -1;
        const int w = f >> 5, t = w / seg;
        int r = s_tpre[t];
        for (int w2 = t * seg; w2 < w; ++w2) r += __popc(s_bm[w2]);
        r += __popc(s_bm[w] & ((1u << (f & 31)) - 1u));
        if (r >= a.max_voxels) return -1;
        const int cx = c % nx, cyz = c / nx;
        const int cy = (a.grid[2] == 1) ? cyz : cyz % ny, cz = (a.grid[2] == 1) ? 0 : cyz / ny;
        if (a.reverse) { coors[r * 3 + 0] = cz; coors[r * 3 + 1] = cy; coors[r * 3 + 2] = cx; }
        else           { coors[r * 3 + 0] = cx; coors[r * 3 + 1] = cy; coors[r * 3 + 2] = cz; }
        return r;
    };

    if (!overflow) {
        // every occupied cell has a voxel.  Three packed counters per cell -- kept [0,21), over-full cells [21,36),
        // their points [36,57) (N < 2^21) -- staged in shared memory, scanned with thread-consecutive ownership
        // (one block scan), read back by the coalesced owners.
#pragma unroll
        for (int j = 0; j < VL_CPT; ++j) {
            const int c = j * 1024 + tid;
            if (j < cpt && c < a.cells) {
                const int h = (fa[j] < n) ? ha[j] : 0;
                s_pk[c] = h < 0 ? ((unsigned long long)a.max_points | (1ull << 21) | ((unsigned long long)(-h) << 36))
                                : (unsigned long long)h;
            }
        }
        __syncthreads();
        const int q0 = min(tid * cpt, a.cells), q1 = min(q0 + cpt, a.cells);
        unsigned long long mine = 0;
        for (int c = q0; c < q1; ++c) mine += s_pk[c];
        unsigned long long tot;
        unsigned long long run = block_exclusive_scan64(mine, s_w64, &tot);
        for (int c = q0; c < q1; ++c) { const unsigned long long v = s_pk[c]; s_pk[c] = run; run += v; }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < VL_CPT; ++j) {
            const int c = j * 1024 + tid;
            if (!(j < cpt && c >= c_lo && c < c_hi)) continue;          // this CTA's slice
            const int v = voxel_of(c, fa[j]);
            const int h = (v >= 0) ? ha[j] : 0;
            const unsigned long long ex = s_pk[c];
            vid[c] = v;
            offc[c] = (int)(ex & 0x1fffffull);
            if (v >= 0) num[v] = h < 0 ? a.max_points : h;
            if (h < 0) {
                hot[(int)((ex >> 21) & 0x7fffull)] = c;
                hoff[c] = (int)(ex >> 36);
                int code = -1;                            // un-chunked: every point of the cell is a candidate
                if (chunked) {
                    const ChunkSplit cs = chunk_split(cntk + (int64_t)c * VL_KC, a.max_points);
                    code = thr_encode(cs.kstar, cs.r);
                }
                thr[c] = code;                            // replaced by the keep threshold in vl_select_kernel
            } else {
                hoff[c] = -1;
                thr[c] = (v >= 0) ? INT_MAX : -1;
            }
            fill[c] = 0;                                  // vl_collect / vl_fill slot counter
        }
        if (part == 0 && tid == 0) {
            const int S = (int)(tot & 0x1fffffull);
            offc[a.cells] = S;
            off[M] = S;                                   // total kept points (the per-voxel offsets belong to the exact path)
            stat[0] = (int)((tot >> 21) & 0x7fffull);
            stat[1] = (int)(tot >> 36);
            a.voxel_count[b] = M;
            a.ws.cutoff[b] = cut;
        }
        return;
    }

    // ---- the max_voxels break fired (no shipped config): counts are re-taken over the points before the break, slice by
    // slice, and the last CTA of the scan turns them into offsets
    for (int c = c_lo + tid; c < c_hi; c += 1024) {
        vid[c] = voxel_of(c, __ldcg(first + c));
        fill[c] = 0;
    }
    __syncthreads();
    {
        const uint16_t *cell16 = a.ws.lrank + (int64_t)b * a.N;
        for (int i = tid; i < cut; i += 1024) {
            const int c = cell16[i];
            if (c >= c_lo && c < c_hi && __ldcg(vid + c) >= 0) atomicAdd(fill + c, 1);
        }
    }
    __syncthreads();
    int my_hot = 0, my_cand = 0;
    for (int c = c_lo + tid; c < c_hi; c += 1024) {
        const int have = __ldcg(fill + c);
        if (have > a.max_points) { ++my_hot; my_cand += have; }
    }
    int tot_hot, tot_cand;
    int hot_pos = block_exclusive_scan(my_hot, s_warp, &tot_hot);
    int cand_pos = block_exclusive_scan(my_cand, s_warp, &tot_cand);
    __shared__ int s_hot0, s_cand0;
    if (tid == 0) {
        s_hot0 = tot_hot ? atomicAdd(stat + 0, tot_hot) : 0;
        s_cand0 = tot_cand ? atomicAdd(stat + 1, tot_cand) : 0;
    }
    __syncthreads();
    hot_pos += s_hot0; cand_pos += s_cand0;
    for (int c = c_lo + tid; c < c_hi; c += 1024) {
        const int v = __ldcg(vid + c);
        const int have = __ldcg(fill + c);
        fill[c] = 0;
        const int kept = min(have, a.max_points);
        if (v >= 0) num[v] = kept;
        offc[c] = kept;
        if (have > a.max_points) {
            hot[hot_pos++] = c;
            hoff[c] = cand_pos;
            cand_pos += have;
            thr[c] = -1;
        } else {
            hoff[c] = -1;
            thr[c] = (v >= 0) ? INT_MAX : -1;
        }
    }
    if (G > 1) {
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            s_last = (atomicAdd(stat + 2, 1) == G - 1);
            __threadfence();
        }
        __syncthreads();
        if (!s_last) return;
    } else {
        __syncthreads();
    }
    const int S = cta_exclusive_scan_global(offc, offc, a.cells, s_warp);
    if (tid == 0) {
        offc[a.cells] = S;
        off[M] = S;
        a.voxel_count[b] = M;
        a.ws.cutoff[b] = cut;
    }
}

template <bool CHUNKED>
__global__ void __launch_bounds__(256) vl_collect_kernel(const VoxArgs a) {
    const int b = blockIdx.y;
    {   // every CTA of the per-cell kernel has read the first indices / counts by now: restore them for the next call
        int32_t *first = a.ws.lfirst + (int64_t)b * a.cells, *cnt = a.ws.lcnt + (int64_t)b * a.cells;
        for (int c = blockIdx.x * 256 + threadIdx.x; c < a.cells; c += gridDim.x * 256) { first[c] = VL_NONE; cnt[c] = 0; }
        if (CHUNKED) {
            int4 *ck = reinterpret_cast<int4 *>(a.ws.lcntk + (int64_t)b * a.cells * VL_KC);
            for (int q = blockIdx.x * 256 + threadIdx.x; q < a.cells * (VL_KC / 4); q += gridDim.x * 256) ck[q] = make_int4(0, 0, 0, 0);
        }
    }
    if (a.ws.lstat[b * 4] == 0) return;                   // no over-full cell in this scan
    const int n = a.counts ? max(0, min(a.counts[b], a.N)) : a.N;
    const int lim = min(n, a.ws.cutoff[b]);
    const int64_t base = (int64_t)b * a.N;
    const uint16_t *cell16 = a.ws.lrank + base;
    const int32_t *hoff = a.ws.lhoff + (int64_t)b * a.cells;
    const int32_t *thr = a.ws.lthr + (int64_t)b * a.cells;
    int32_t *fill = a.ws.lfill + (int64_t)b * a.cells;
    int32_t *cand = a.ws.cell + base;                     // the exact path's cell buffer doubles as the candidate lists
    const int lane = threadIdx.x & 31;
    const int lround = (lim + 31) & ~31;
    const int stride = gridDim.x * 256;
    constexpr int U = 4;           // 32-point groups per warp and iteration: their cell -> list offset -> slot chains overlap
    for (int i0 = blockIdx.x * 256 + threadIdx.x; i0 < lround; i0 += U * stride) {
        int i[U], c[U], o[U];
        RunSlot rs[U];
#pragma unroll
        for (int h = 0; h < U; ++h) { i[h] = i0 + h * stride; c[h] = (i[h] < lim) ? (int)cell16[i[h]] : 0xffff; }
#pragma unroll
        for (int h = 0; h < U; ++h) {
            o[h] = (c[h] != 0xffff) ? __ldg(hoff + c[h]) : -1;
            if (CHUNKED && o[h] >= 0) {                   // chunked counts: only the chunk that holds the max_points-th point
                int kstar, r;
                thr_decode(__ldcg(thr + c[h]), a.max_points, kstar, r);
                if (kstar >= 0 && (i[h] >> a.kc_shift) != kstar) o[h] = -1;
            }
        }
#pragma unroll
        for (int h = 0; h < U; ++h) rs[h] = run_slots_issue(fill + (o[h] >= 0 ? c[h] : 0), c[h], o[h] >= 0, lane);
#pragma unroll
        for (int h = 0; h < U; ++h) {
            const int slot = run_slots_take(rs[h], lane);
            if (o[h] >= 0) cand[o[h] + slot] = i[h];
        }
    }
}

constexpr int VL_SEL_WARPS = 8;

__global__ void __launch_bounds__(VL_SEL_WARPS * 32) vl_select_kernel(const VoxArgs a, int nbuckets) {
    extern __shared__ uint32_t s_sel[];                   // per warp: [nbuckets] bucket counts + [32] bitmap words
    const int b = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nh = a.ws.lstat[b * 4];
    uint32_t *hist = s_sel + (size_t)warp * (nbuckets + 32);
    uint32_t *bits = hist + nbuckets;
    const int32_t *hot = a.ws.lhot + (int64_t)b * a.cells;
    const int32_t *hoff = a.ws.lhoff + (int64_t)b * a.cells;
    int32_t *fill = a.ws.lfill + (int64_t)b * a.cells;
    int32_t *thr = a.ws.lthr + (int64_t)b * a.cells;
    const int32_t *cand_all = a.ws.cell + (int64_t)b * a.N;
    for (int j = blockIdx.x * VL_SEL_WARPS + warp; j < nh; j += gridDim.x * VL_SEL_WARPS) {
        const int c = hot[j];
        const int m = fill[c];                            // candidates collected
        const int32_t *cand = cand_all + hoff[c];
        int kstar, want;                                  // rank (1-based) of the keep threshold among the candidates
        thr_decode(thr[c], a.max_points, kstar, want);
        if (m < want) {                                   // cannot happen (the counts say the rank exists); stay safe
            if (lane == 0) { thr[c] = INT_MAX; fill[c] = 0; }
            continue;
        }
        for (int k = lane; k < nbuckets; k += 32) hist[k] = 0;
        bits[lane] = 0;
        __syncwarp();
        for (int k = lane; k < m; k += 32) atomicAdd(&hist[cand[k] >> VL_SHIFT], 1u);
        __syncwarp();
        // the bucket that holds the max_points-th smallest index
        const int per = (nbuckets + 31) / 32;
        const int k0 = min(lane * per, nbuckets), k1 = min(k0 + per, nbuckets);
        int mine = 0;
        for (int k = k0; k < k1; ++k) mine += (int)hist[k];
        const int incl = warp_inclusive_scan(mine);
        const int before_lane = incl - mine;
        const bool holds = before_lane < want && want <= incl;
        int bstar = 0, before = 0;
        if (holds) {
            int run = before_lane;
            for (int k = k0; k < k1; ++k) {
                if (run + (int)hist[k] >= want) { bstar = k; before = run; break; }
                run += (int)hist[k];
            }
        }
        const unsigned who = __ballot_sync(0xffffffffu, holds);
        const int src = __ffs(who) - 1;
        bstar = __shfl_sync(0xffffffffu, bstar, src);
        before = __shfl_sync(0xffffffffu, before, src);
        for (int k = lane; k < m; k += 32) {
            const int i = cand[k];
            if ((i >> VL_SHIFT) == bstar) atomicOr(&bits[(i >> 5) & 31], 1u << (i & 31));
        }
        __syncwarp();
        // the r-th set bit (1-based) of the 1024-bit map, r = wanted rank - candidates in earlier buckets
        const int r = want - before;
        const uint32_t word = bits[lane];
        const int pc = __popc(word);
        const int pincl = warp_inclusive_scan(pc);
        const bool here = (pincl - pc) < r && r <= pincl;
        int pos = 0;
        if (here) {
            uint32_t w = word;
            for (int k = r - (pincl - pc) - 1; k > 0; --k) w &= w - 1;
            pos = lane * 32 + __ffs(w) - 1;
        }
        const unsigned who2 = __ballot_sync(0xffffffffu, here);
        pos = __shfl_sync(0xffffffffu, pos, __ffs(who2) - 1);
        if (lane == 0) {
            thr[c] = (bstar << VL_SHIFT) + pos;
            fill[c] = 0;                                  // becomes the slot counter of vl_fill_kernel
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(256) vl_fill_kernel(const VoxArgs a) {
    const int b = blockIdx.y;
    const int n = a.counts ? max(0, min(a.counts[b], a.N)) : a.N;
    const int lim = min(n, a.ws.cutoff[b]);
    const int64_t base = (int64_t)b * a.N;
    const uint16_t *cell16 = a.ws.lrank + base;
    const int32_t *thr = a.ws.lthr + (int64_t)b * a.cells;
    const int32_t *offc = a.ws.offc + (int64_t)b * (a.cells + 1);
    int32_t *fill = a.ws.lfill + (int64_t)b * a.cells;
    const int lane = threadIdx.x & 31;
    const int lround = (lim + 31) & ~31;
    const int stride = gridDim.x * 256;
    // four 32-point groups per warp and iteration; per group the chain is cell -> threshold -> slot atomic -> store
    // (three dependent memory round trips: ncu shows nothing but long-scoreboard stalls), and the four chains and the
    // 16-byte point loads overlap
    constexpr int U = 4;
    for (int i0 = blockIdx.x * 256 + threadIdx.x; i0 < lround; i0 += U * stride) {
        int i[U], c[U], ofs[U];
        bool keep[U];
        float4 pt[U];
        RunSlot rs[U];
#pragma unroll
        for (int h = 0; h < U; ++h) { i[h] = i0 + h * stride; c[h] = (i[h] < lim) ? (int)cell16[i[h]] : 0xffff; }
#pragma unroll
        for (int h = 0; h < U; ++h) {
            pt[h] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c[h] != 0xffff) {                         // in-range point: its row is most likely kept, fetch it now
                const float *src = a.points + (base + i[h]) * a.F;
                if (a.vec4) pt[h] = __ldg(reinterpret_cast<const float4 *>(src));
                else pt[h] = make_float4(__ldg(src), __ldg(src + 1), __ldg(src + 2), a.F > 3 ? __ldg(src + 3) : 0.0f);
            }
        }
#pragma unroll
        for (int h = 0; h < U; ++h) {
            keep[h] = (c[h] != 0xffff) && i[h] <= __ldg(thr + c[h]);      // thr = -1: the cell has no voxel
            ofs[h] = (c[h] != 0xffff) ? __ldg(offc + c[h]) : 0;
        }
#pragma unroll
        for (int h = 0; h < U; ++h) rs[h] = run_slots_issue(fill + (keep[h] ? c[h] : 0), c[h], keep[h], lane);
#pragma unroll
        for (int h = 0; h < U; ++h) {
            const int slot = run_slots_take(rs[h], lane);
            if (keep[h]) {
                if (a.apply_shift) pt[h].z = __fadd_rn(pt[h].z, a.z_shift);
                a.ws.sorted[base + ofs[h] + slot] = pt[h];
            }
        }
    }
}

// predict_ground.py:104-107: cloud[:, :4] += float32([0, 0, h, 0]) (so -0.0 becomes +0.0 in every column)
__global__ void shift_cloud_kernel(float *points, int64_t rows, int F, float h) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < rows)
        for (int f = 0; f < min(F, 4); ++f) points[i * F + f] = __fadd_rn(points[i * F + f], f == 2 ? h : 0.0f);
}

// ---- host side ----------------------------------------------------------------------------
static void host_grid(const gnd_voxel_params *p, int32_t g[3]) {
    for (int j = 0; j < 3; ++j) {
        volatile float d = p->coors_range[3 + j] - p->coors_range[j];
        volatile float q = d / p->voxel_size[j];
        g[j] = (int32_t)rintf(q);            // np.round == round-half-even (:24-27)
    }
}

int vox_make_plan(const gnd_voxel_params *p, int32_t batch, int64_t n_points, VoxPlan *plan) {
    if (!p || !plan) return fail(GND_ERR_INVALID, "null voxel params");
    if (batch < 1 || n_points < 0) return fail(GND_ERR_INVALID, "batch %d / n_points %lld", batch, (long long)n_points);
    if (p->num_features < 3) return fail(GND_ERR_INVALID, "points need >= 3 columns, got %d", p->num_features);
    if (p->max_points < 1 || p->max_voxels < 1) return fail(GND_ERR_INVALID, "max_points/max_voxels must be >= 1");
    if (p->max_points > 65535) return fail(GND_ERR_UNSUPPORTED, "max_points %d > 65535", p->max_points);
    for (int j = 0; j < 3; ++j)
        if (!(p->voxel_size[j] > 0.0f)) return fail(GND_ERR_INVALID, "voxel_size[%d] must be > 0", j);
    host_grid(p, plan->grid);
    int64_t cells = 1;
    for (int j = 0; j < 3; ++j) {
        if (plan->grid[j] < 1) return fail(GND_ERR_INVALID, "empty voxel grid on axis %d", j);
        cells *= plan->grid[j];
    }
    if (cells > 100000) return fail(GND_ERR_UNSUPPORTED, "voxel grid of %lld cells > 100000 (shared-memory cell table)", (long long)cells);
    if (n_points > 1600000) return fail(GND_ERR_UNSUPPORTED, "%lld points per scan > 1600000 (shared-memory first-point bitmap)", (long long)n_points);
    plan->cells = (int32_t)cells;
    plan->tiled = (cells <= VOX_RT_MAX_CELLS && !tuning().vox_serial) ? 1 : 0;
    const int64_t sub = plan->tiled ? VOX_RT_TILE : VOX_SUB;
    int64_t n = n_points < 1 ? 1 : n_points;
    int64_t kmin = (n + VOX_MAX_T - 1) / VOX_MAX_T;
    int64_t kmax = (n + sub - 1) / sub;
    // tiled: two CTAs per SM in ONE wave (floor); serial: about two CTAs per SM over the batch
    int64_t k = plan->tiled ? (2 * 148) / batch : (2 * 148 + batch - 1) / batch;
    if (k < 1) k = 1;
    {   // the chunk-histogram table is K x cells: small grids (the 2 500-cell camera grid) afford more, smaller chunks,
        // so that one dense frame still spreads over every SM
        int64_t kcap = 655360 / cells;
        if (kcap < VOX_MAX_K) kcap = VOX_MAX_K;
        if (kcap > 2 * 148) kcap = 2 * 148;
        if (k > kcap) k = kcap;
    }
    if (tuning().vox_k > 0 && tuning().vox_k < k) k = tuning().vox_k;
    if (k > kmax) k = kmax;
    if (k < kmin) k = kmin;
    int64_t t = (n + k - 1) / k;
    t = (t + sub - 1) / sub * sub;
    plan->T = (int32_t)t;
    plan->K = (int32_t)((n + t - 1) / t);
    {   // assign CTAs per scan: about two per SM over the batch, at most 16
        int g = (2 * 148) / batch;
        if (tuning().vox_g > 0) g = tuning().vox_g;
        plan->G = g < 1 ? 1 : (g > 16 ? 16 : g);
    }
    plan->rank_smem = plan->tiled ? (size_t)plan->cells * (sizeof(uint2) + sizeof(uint16_t)) : (size_t)plan->cells * sizeof(uint16_t);
    plan->assign_smem = (size_t)((n_points + 31) / 32 + 1) * sizeof(uint32_t);
    plan->lean_ok = (cells <= 65534) ? 1 : 0;
    return GND_OK;
}

size_t vox_carve(const VoxPlan &plan, int32_t batch, int64_t n_points, void *ws, VoxWs *out, bool with_sorted) {
    Carver c(ws);
    VoxWs w;
    size_t bn = (size_t)batch * (size_t)(n_points < 1 ? 1 : n_points);
    w.cell = c.take<int32_t>(bn);
    w.lrank = c.take<uint16_t>(bn);
    w.first = c.take<int32_t>((size_t)batch * plan.cells + (size_t)batch);    // + done[B], same 0x7f fill
    w.done = w.first ? w.first + (size_t)batch * plan.cells : nullptr;
    w.hist = c.take<uint32_t>((size_t)batch * plan.K * plan.cells);
    w.vid = c.take<int32_t>((size_t)batch * plan.cells);
    w.cutoff = c.take<int32_t>((size_t)batch);
    w.offc = c.take<int32_t>((size_t)batch * (plan.cells + 1));
    w.sorted = with_sorted ? c.take<float4>(bn) : nullptr;
    const bool lean = with_sorted && plan.lean_ok;
    const size_t bc = (size_t)batch * plan.cells;
    w.lfirst = lean ? c.take<int32_t>(bc) : nullptr;
    w.lcnt = lean ? c.take<int32_t>(bc) : nullptr;
    w.lcntk = lean ? c.take<int32_t>(bc * VL_KC) : nullptr;
    w.lfill = lean ? c.take<int32_t>(bc) : nullptr;
    w.lthr = lean ? c.take<int32_t>(bc) : nullptr;
    w.lhoff = lean ? c.take<int32_t>(bc) : nullptr;
    w.lhot = lean ? c.take<int32_t>(bc) : nullptr;
    w.lstat = lean ? c.take<int32_t>((size_t)batch * 4) : nullptr;
    if (out) *out = w;
    return c.bytes();
}

static void fill_args(VoxArgs &a, const gnd_voxel_params *p, const VoxPlan &plan, const VoxWs &ws, const float *points,
                      const int32_t *counts, int32_t batch, int64_t n_points) {
    a.points = points; a.counts = counts; a.B = batch; a.N = (int32_t)n_points; a.F = p->num_features;
    a.vec4 = (p->num_features == 4 && ((uintptr_t)points % 16) == 0) ? 1 : 0;
    for (int j = 0; j < 3; ++j) {
        a.lo[j] = p->coors_range[j]; a.vs[j] = p->voxel_size[j]; a.grid[j] = plan.grid[j];
        int e = 0;
        const float m = frexpf(a.vs[j], &e);
        a.inv[j] = (m == 0.5f && e > -100 && e < 100) ? 1.0f / a.vs[j] : 0.0f;
    }
    a.cells = plan.cells; a.K = plan.K; a.T = plan.T; a.G = plan.G;
    a.kc_shift = 0;
    a.apply_shift = p->apply_z_shift; a.z_shift = p->z_shift;
    a.max_points = p->max_points; a.max_voxels = p->max_voxels; a.reverse = p->reverse_index;
    a.ws = ws;
}

static int vl_attrs() {
    static bool done[64] = {};
    int dev = 0;
    GND_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !done[dev]) {
        GND_CUDA_CHECK(cudaFuncSetAttribute(vl_cells_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024 + 8));
        GND_CUDA_CHECK(cudaFuncSetAttribute(vl_cells1_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024 + 8));
        GND_CUDA_CHECK(cudaFuncSetAttribute(vl_cells1_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024 + 8));
        GND_CUDA_CHECK(cudaFuncSetAttribute(vl_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
        if (dev >= 0 && dev < 64) done[dev] = true;
    }
    return GND_OK;
}

int vox_lean_init(const VoxPlan &plan, const VoxWs &ws, int32_t batch) {      // set-up time (gnd_bind_workspace): synchronous
    if (!ws.lfirst) return GND_OK;
    const size_t bc = (size_t)batch * plan.cells;
    GND_CUDA_CHECK(cudaMemset(ws.lfirst, 0x7f, bc * sizeof(int32_t)));
    GND_CUDA_CHECK(cudaMemset(ws.lcnt, 0, bc * sizeof(int32_t)));
    GND_CUDA_CHECK(cudaMemset(ws.lcntk, 0, bc * VL_KC * sizeof(int32_t)));
    GND_CUDA_CHECK(cudaMemset(ws.lfill, 0, bc * sizeof(int32_t)));
    GND_CUDA_CHECK(cudaMemset(ws.lstat, 0, (size_t)batch * 4 * sizeof(int32_t)));
    return GND_OK;
}

int vox_run_lean(const gnd_voxel_params *p, const VoxPlan &plan, const VoxWs &ws, const float *points,
                 const int32_t *counts, int32_t batch, int64_t n_points, int32_t *coors, int32_t *num_points,
                 int32_t *voxel_count, int32_t *point_offsets, cudaStream_t stream) {
    if (!plan.lean_ok || !ws.lfirst || !ws.sorted) return fail(GND_ERR_STATE, "workspace has no set-exact voxelizer state");
    if (n_points > 0x1fffff) return fail(GND_ERR_UNSUPPORTED, "set-exact voxelizer: more than 2^21 points per scan");
    if (!coors || !num_points || !voxel_count || !point_offsets) return fail(GND_ERR_INVALID, "voxelizer output pointer is null");
    if (n_points > 0 && !points) return fail(GND_ERR_INVALID, "points is null");
    VoxArgs a;
    fill_args(a, p, plan, ws, points, counts, batch, n_points);
    a.coors = coors; a.num_points = num_points; a.voxel_count = voxel_count; a.offsets = point_offsets;
    a.list = nullptr; a.voxels = nullptr; a.slot = nullptr;
    int rc = vl_attrs();
    if (rc != GND_OK) return rc;
    const int64_t n_cap = n_points;
    const size_t smem1 = (size_t)(((n_cap + 31) / 32 + 2) & ~(size_t)1) * sizeof(uint32_t) + (size_t)plan.cells * sizeof(unsigned long long) + 16;
    const bool single = plan.cells <= 1024 * VL_CPT && smem1 <= 200 * 1024;
    // chunked counts pay on DENSE clouds only (camera frames: every cell over-full, ~400 points each): measured at
    // 64 x 1 M points 1.61 ms against 2.03 ms for the order-exact kernels (un-chunked set-exact: slower than both), but
    // on LiDAR scans +0.02 ms per 64 scans (8 counters per cell to read and to clear, one more dependent gather per
    // candidate) against the 0.01 ms the smaller candidate lists save
    if (single && n_points > (int64_t)32 * plan.cells) {     // VL_KC chunks of 2^kc_shift points
        int sh = 5;
        while (((int64_t)VL_KC << sh) < n_points) ++sh;
        a.kc_shift = sh;
    }
    // point-parallel kernels: about eight CTAs per SM over the whole batch
    int gx = (int)((n_points + 255) / 256);
    const int want = (148 * 8 + batch - 1) / batch;
    if (gx > want) gx = want;
    if (gx < 1) gx = 1;
    vl_count_kernel<<<dim3(gx, batch), 256, 0, stream>>>(a);
    GND_LAUNCH_CHECK("vl_count_kernel");
    if (single) {
        // every CTA of a scan redoes the all-cells part: as few CTAs per scan as still fill the SMs
        int g = 148 / batch;
        if (tuning().vox_g > 0) g = tuning().vox_g;
        a.G = g < 1 ? 1 : (g > 16 ? 16 : g);
        if (a.kc_shift) vl_cells1_kernel<true><<<batch * a.G, 1024, smem1, stream>>>(a);
        else vl_cells1_kernel<false><<<batch * a.G, 1024, smem1, stream>>>(a);
        GND_LAUNCH_CHECK("vl_cells1_kernel");
    } else {
        vl_cells_kernel<<<batch * plan.G, 1024, plan.assign_smem, stream>>>(a);
        GND_LAUNCH_CHECK("vl_cells_kernel");
    }
    if (n_points > 0) {
        if (a.kc_shift) vl_collect_kernel<true><<<dim3(gx, batch), 256, 0, stream>>>(a);
        else vl_collect_kernel<false><<<dim3(gx, batch), 256, 0, stream>>>(a);
        GND_LAUNCH_CHECK("vl_collect_kernel");
        const int nbuckets = (int)(n_points >> VL_SHIFT) + 1;
        int sx = (148 * 4 + batch - 1) / batch;
        if (sx < 1) sx = 1;
        vl_select_kernel<<<dim3(sx, batch), VL_SEL_WARPS * 32, (size_t)VL_SEL_WARPS * (nbuckets + 32) * sizeof(uint32_t), stream>>>(a, nbuckets);
        GND_LAUNCH_CHECK("vl_select_kernel");
        vl_fill_kernel<<<dim3(gx, batch), 256, 0, stream>>>(a);
        GND_LAUNCH_CHECK("vl_fill_kernel");
    }
    return GND_OK;
}

int vox_run(const gnd_voxel_params *p, const VoxPlan &plan, const VoxWs &ws, const float *points,
            const int32_t *counts, int32_t batch, int64_t n_points, int32_t *coors, int32_t *num_points,
            int32_t *voxel_count, int32_t *point_offsets, int32_t *point_list, float *voxels,
            int32_t *point_slot, cudaStream_t stream) {
    if (!coors || !num_points || !voxel_count || !point_offsets || !point_list)
        return fail(GND_ERR_INVALID, "voxelizer output pointer is null");
    if (n_points > 0 && !points) return fail(GND_ERR_INVALID, "points is null");
    VoxArgs a;
    fill_args(a, p, plan, ws, points, counts, batch, n_points);
    a.ws = ws; a.coors = coors; a.num_points = num_points; a.voxel_count = voxel_count;
    a.offsets = point_offsets; a.list = point_list; a.voxels = voxels; a.slot = point_slot;

    {   // opt-in shared-memory limits: once per device, not per call
        static bool done[64] = {};
        int dev = 0;
        GND_CUDA_CHECK(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 64 || !done[dev]) {
            GND_CUDA_CHECK(cudaFuncSetAttribute(vox_rank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            GND_CUDA_CHECK(cudaFuncSetAttribute(vox_assign_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024 + 8));
            GND_CUDA_CHECK(cudaFuncSetAttribute(vox_rank_tiles_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            if (dev >= 0 && dev < 64) done[dev] = true;
        }
    }
    GND_CUDA_CHECK(cudaMemsetAsync(ws.first, 0x7f, ((size_t)batch * plan.cells + (size_t)batch) * sizeof(int32_t), stream));
    if (n_points > 0) {
        if (plan.tiled) {
            vox_rank_tiles_kernel<<<batch * plan.K, VOX_RT_WARPS * 32, plan.rank_smem, stream>>>(a);
            GND_LAUNCH_CHECK("vox_rank_tiles_kernel");
        } else {
            vox_rank_kernel<<<batch * plan.K, 256, plan.rank_smem, stream>>>(a);
            GND_LAUNCH_CHECK("vox_rank_kernel");
        }
    } else {
        GND_CUDA_CHECK(cudaMemsetAsync(ws.hist, 0, (size_t)batch * plan.K * plan.cells * sizeof(uint32_t), stream));
    }
    vox_assign_kernel<<<batch * plan.G, 1024, plan.assign_smem, stream>>>(a);
    GND_LAUNCH_CHECK("vox_assign_kernel");
    if (n_points > 0) {
        int gx = (plan.T + 255) / 256;
        const int want = (148 * 16 + batch * plan.K - 1) / (batch * plan.K);     // ~16 CTAs per SM over the grid
        if (gx > want) gx = want;
        if (gx < 1) gx = 1;
        vox_scatter_kernel<<<dim3(gx, batch * plan.K), 256, 0, stream>>>(a);
        GND_LAUNCH_CHECK("vox_scatter_kernel");
    }
    return GND_OK;
}

}  // namespace gnd

// ---- C ABI ----------------------------------------------------------------------------------
using namespace gnd;

extern "C" int gnd_voxel_grid_size(const gnd_voxel_params *p, int32_t *h_grid_xyz) {
    if (!p || !h_grid_xyz) return fail(GND_ERR_INVALID, "null argument");
    host_grid(p, h_grid_xyz);
    return GND_OK;
}

extern "C" int gnd_voxel_workspace_bytes(const gnd_voxel_params *p, int32_t batch, int64_t n_points, size_t *h_bytes) {
    if (!h_bytes) return fail(GND_ERR_INVALID, "null argument");
    VoxPlan plan;
    int rc = vox_make_plan(p, batch, n_points, &plan);
    if (rc != GND_OK) return rc;
    *h_bytes = vox_carve(plan, batch, n_points, nullptr, nullptr);
    return GND_OK;
}

extern "C" int gnd_points_to_voxel(const gnd_voxel_params *p, const float *points, const int32_t *counts,
                                   int32_t batch, int64_t n_points, void *workspace, size_t workspace_bytes,
                                   int32_t *coors, int32_t *num_points, int32_t *voxel_count,
                                   int32_t *point_offsets, int32_t *point_list, float *voxels,
                                   int32_t *point_slot, void *stream) {
    g_launches = 0;
    VoxPlan plan;
    int rc = vox_make_plan(p, batch, n_points, &plan);
    if (rc != GND_OK) return rc;
    VoxWs ws;
    size_t need = vox_carve(plan, batch, n_points, workspace, &ws);
    if (!workspace || workspace_bytes < need)
        return fail(GND_ERR_WORKSPACE, "voxel workspace: need %zu bytes, got %zu", need, workspace_bytes);
    return vox_run(p, plan, ws, points, counts, batch, n_points, coors, num_points, voxel_count,
                   point_offsets, point_list, voxels, point_slot, (cudaStream_t)stream);
}

extern "C" int gnd_shift_cloud(float *points, int64_t n_rows, int32_t num_features, float height, void *stream) {
    g_launches = 0;
    if (n_rows < 0 || num_features < 3) return fail(GND_ERR_INVALID, "bad shape");
    if (n_rows == 0) return GND_OK;
    if (!points) return fail(GND_ERR_INVALID, "points is null");
    shift_cloud_kernel<<<(unsigned)((n_rows + 255) / 256), 256, 0, (cudaStream_t)stream>>>(points, n_rows, num_features, height);
    GND_LAUNCH_CHECK("shift_cloud_kernel");
    return GND_OK;
}
