// segment.cu -- per-point ground / obstacle labelling against the estimated height map.
//
// Replaces utils/utils.py:198-220 (segment_cloud) and :222-248 (segment_cloud_noground).
// Arithmetic restated exactly:
//   :204  xy -= grid_size[:2]     float32 array minus int64 array: float64 subtract rounded back to float32
//   :205  / voxel_size            float64 division by the Python scalar
//   :206-207 floor, astype(int32)
//   :213  0 < x < nx and 0 < y < ny      (index 0 is OUT of range)
//   :214  z > elevation_map[x, y] + threshold    in float64
// HBM-bound gather: 12-16 B read + 1-8 B written per point, the 40 KB map stays in L1/L2.
#include "common.cuh"
#include <limits.h>
#include <math.h>
#include <atomic>

namespace gnd {

struct SegArgs {
    gnd_segment_params p;
    float *points;
    const int32_t *counts;
    int32_t B;
    int64_t N;
    const float *elev;
    double *lab64;
    int8_t *lab8;
    double inv;          // exact reciprocal of voxel_size when it is a power of two, else 0
    int32_t quad;        // 16-byte rows, N % 4 == 0 and aligned outputs: four points per thread
};

// `inv` != 0: voxel_size is a power of two, so s / vs == s * inv bit for bit and the float64 divide is skipped
__device__ __forceinline__ int seg_axis(float v, double lo, double vs, double inv, float *shifted) {
    float s = __double2float_rn((double)v - lo);          // :204 (in-place float32 update)
    *shifted = s;
    double q = floor(inv != 0.0 ? (double)s * inv : (double)s / vs);   // :205-206
    if (!(q >= -2147483648.0 && q <= 2147483647.0)) return INT_MIN;   // :207 cvttsd2si semantics
    return (int)q;
}

// label of one point whose row has been loaded (x, y, z): writes the shifted x, y back when asked (:204)
__device__ __forceinline__ int seg_label(const SegArgs &a, int b, float *row, float x, float y, float z) {
    if (a.p.apply_z_shift) z = __fadd_rn(z, a.p.z_shift);
    float sx, sy;
    const int cx = seg_axis(x, a.p.grid_lo[0], a.p.voxel_size, a.inv, &sx);
    const int cy = seg_axis(y, a.p.grid_lo[1], a.p.voxel_size, a.inv, &sy);
    if (a.p.mutate_points) { row[0] = sx; row[1] = sy; }
    if (0 < cx && cx < a.p.map_nx && 0 < cy && cy < a.p.map_ny) {   // :213
        const float h = __ldg(a.elev + (int64_t)b * a.p.map_batch_stride + cx * a.p.map_sx + cy * a.p.map_sy);
        return ((double)z > (double)h + a.p.threshold) ? 1 : 0;     // :214-217
    }
    return -1;
}

__global__ void __launch_bounds__(256) segment_kernel(const SegArgs a) {
    const int F = a.p.num_features;
    const int b = blockIdx.y;                              // one grid row per scan: no per-point division
    const int n = a.counts ? max(0, min(a.counts[b], (int)a.N)) : (int)a.N;
    if (a.quad) {
        // four consecutive points per thread: the four 16-byte row loads are in flight together, the four int8
        // labels leave as one 4-byte store
        for (int64_t i = ((int64_t)blockIdx.x * 256 + threadIdx.x) * 4; i < a.N; i += (int64_t)gridDim.x * 1024) {
            const int64_t gi = (int64_t)b * a.N + i;
            float4 v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                v[k] = (i + k < n) ? *reinterpret_cast<const float4 *>(a.points + (gi + k) * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
            int lab[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                lab[k] = (i + k < n) ? seg_label(a, b, a.points + (gi + k) * 4, v[k].x, v[k].y, v[k].z) : -1;
            if (a.lab8)
                *reinterpret_cast<uint32_t *>(a.lab8 + gi) = (uint32_t)(uint8_t)lab[0] | ((uint32_t)(uint8_t)lab[1] << 8) |
                                                             ((uint32_t)(uint8_t)lab[2] << 16) | ((uint32_t)(uint8_t)lab[3] << 24);
            if (a.lab64) {
                *reinterpret_cast<double2 *>(a.lab64 + gi) = make_double2((double)lab[0], (double)lab[1]);
                *reinterpret_cast<double2 *>(a.lab64 + gi + 2) = make_double2((double)lab[2], (double)lab[3]);
            }
        }
        return;
    }
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < a.N; i += (int64_t)gridDim.x * 256) {
        const int64_t gi = (int64_t)b * a.N + i;
        int label = -1;
        if (i < n) {
            float *row = a.points + gi * F;
            label = seg_label(a, b, row, row[0], row[1], row[2]);
        }
        if (a.lab8) a.lab8[gi] = (int8_t)label;
        if (a.lab64) a.lab64[gi] = (double)label;
    }
}

// ---- stable obstacle compaction (:242-243, :248) ----------------------------------------------
constexpr int CMP_BLOCK = 1024;

__global__ void __launch_bounds__(CMP_BLOCK) compact_count_kernel(const int8_t *lab, int64_t n, int32_t *block_counts) {
    __shared__ int s_warp[33];
    int64_t i = (int64_t)blockIdx.x * CMP_BLOCK + threadIdx.x;
    int flag = (i < n && lab[i] == 1) ? 1 : 0;
    int total;
    block_exclusive_scan(flag, s_warp, &total);
    if (threadIdx.x == 0) block_counts[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024) compact_scan_kernel(int32_t *block_counts, int nblocks, int32_t *obstacle_count) {
    __shared__ int s_warp[33];
    int carry = 0;
    for (int base = 0; base < nblocks; base += 1024) {
        int idx = base + threadIdx.x;
        int v = idx < nblocks ? block_counts[idx] : 0;
        int total;
        int ex = block_exclusive_scan(v, s_warp, &total);
        if (idx < nblocks) block_counts[idx] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) *obstacle_count = carry;
}

__global__ void __launch_bounds__(CMP_BLOCK) compact_write_kernel(const int8_t *lab, const float *src, int64_t n, int F,
                                                                 const int32_t *block_offsets, float *dst) {
    __shared__ int s_warp[33];
    int64_t i = (int64_t)blockIdx.x * CMP_BLOCK + threadIdx.x;
    int flag = (i < n && lab[i] == 1) ? 1 : 0;
    int total;
    int ex = block_exclusive_scan(flag, s_warp, &total);
    if (flag) {
        int64_t o = (int64_t)block_offsets[blockIdx.x] + ex;
        for (int f = 0; f < F; ++f) dst[o * F + f] = src[i * F + f];
    }
}

int segment_run(const gnd_segment_params *p, float *points, const int32_t *counts, int32_t batch,
                int64_t n_points, const float *elevation, double *labels_f64, int8_t *labels_i8,
                cudaStream_t stream) {
    if (!p) return fail(GND_ERR_INVALID, "null segment params");
    if (batch < 1 || n_points < 0) return fail(GND_ERR_INVALID, "bad batch / n_points");
    if (p->num_features < 3) return fail(GND_ERR_INVALID, "points need >= 3 columns");
    if (!(p->voxel_size > 0.0)) return fail(GND_ERR_INVALID, "voxel_size must be > 0");
    if (p->map_nx < 1 || p->map_ny < 1) return fail(GND_ERR_INVALID, "empty elevation map");
    if (n_points == 0) return GND_OK;
    if (!points || !elevation) return fail(GND_ERR_INVALID, "null points / elevation");
    if (!labels_f64 && !labels_i8) return fail(GND_ERR_INVALID, "no label output given");
    SegArgs a;
    a.p = *p; a.points = points; a.counts = counts; a.B = batch; a.N = n_points;
    a.elev = elevation; a.lab64 = labels_f64; a.lab8 = labels_i8;
    {
        int e = 0;
        const double m = frexp(p->voxel_size, &e);
        a.inv = (m == 0.5 && e > -500 && e < 500) ? 1.0 / p->voxel_size : 0.0;
    }
    if (batch > 65535) return fail(GND_ERR_UNSUPPORTED, "batch %d > 65535", batch);
    a.quad = (p->num_features == 4 && (n_points & 3) == 0 && (((uintptr_t)points) & 15) == 0 &&
              (!labels_i8 || (((uintptr_t)labels_i8) & 3) == 0) && (!labels_f64 || (((uintptr_t)labels_f64) & 15) == 0)) ? 1 : 0;
    int gx = (int)((n_points + (a.quad ? 1023 : 255)) / (a.quad ? 1024 : 256));
    const int want = (148 * 16 + batch - 1) / batch;          // ~16 CTAs per SM over the whole grid
    if (gx > want) gx = want;
    segment_kernel<<<dim3(gx, batch), 256, 0, stream>>>(a);
    GND_LAUNCH_CHECK("segment_kernel");
    return GND_OK;
}

}  // namespace gnd

using namespace gnd;

extern "C" int gnd_segment_cloud(const gnd_segment_params *p, float *points, const int32_t *counts,
                                 int32_t batch, int64_t n_points, const float *elevation,
                                 double *labels_f64, int8_t *labels_i8, void *stream) {
    g_launches = 0;
    return segment_run(p, points, counts, batch, n_points, elevation, labels_f64, labels_i8, (cudaStream_t)stream);
}

extern "C" int gnd_compact_workspace_bytes(int64_t n_points, size_t *h_bytes) {
    if (!h_bytes || n_points < 0) return fail(GND_ERR_INVALID, "bad argument");
    *h_bytes = align_up((size_t)((n_points + CMP_BLOCK - 1) / CMP_BLOCK + 1) * sizeof(int32_t), 256);
    return GND_OK;
}

extern "C" int gnd_compact_obstacles(const int8_t *labels_i8, const float *points_obstacle, int64_t n_points,
                                     int32_t obstacle_features, void *workspace, size_t workspace_bytes,
                                     float *out_obstacle, int32_t *obstacle_count, void *stream) {
    g_launches = 0;
    if (n_points < 0 || obstacle_features < 1) return fail(GND_ERR_INVALID, "bad shape");
    if (!obstacle_count) return fail(GND_ERR_INVALID, "obstacle_count is null");
    cudaStream_t st = (cudaStream_t)stream;
    if (n_points == 0) { GND_CUDA_CHECK(cudaMemsetAsync(obstacle_count, 0, sizeof(int32_t), st)); return GND_OK; }
    if (!labels_i8 || !points_obstacle || !out_obstacle) return fail(GND_ERR_INVALID, "null pointer");
    size_t need;
    gnd_compact_workspace_bytes(n_points, &need);
    if (!workspace || workspace_bytes < need) return fail(GND_ERR_WORKSPACE, "compaction workspace: need %zu bytes", need);
    int nblocks = (int)((n_points + CMP_BLOCK - 1) / CMP_BLOCK);
    int32_t *bc = (int32_t *)workspace;
    compact_count_kernel<<<nblocks, CMP_BLOCK, 0, st>>>(labels_i8, n_points, bc);
    GND_LAUNCH_CHECK("compact_count_kernel");
    compact_scan_kernel<<<1, 1024, 0, st>>>(bc, nblocks, obstacle_count);
    GND_LAUNCH_CHECK("compact_scan_kernel");
    compact_write_kernel<<<nblocks, CMP_BLOCK, 0, st>>>(labels_i8, points_obstacle, n_points, obstacle_features, bc, out_obstacle);
    GND_LAUNCH_CHECK("compact_write_kernel");
    return GND_OK;
}

// ---- obstacle compaction for callers that cannot allocate device memory -------------------------------------------
// gnd_compact_obstacles_static is gnd_compact_obstacles without the caller-provided workspace.  Who needs it: code that
// is handed device buffers it does not own and has no allocator of its own to ask for the few KB of block counters --
// a publisher callback running inside someone else's memory pool, a plug-in loaded into a process whose device memory
// is fully pre-partitioned at start-up, a real-time loop whose rules forbid cudaMalloc (it synchronises the device and
// can take milliseconds).  The per-block counters live in the library image instead: one fixed array per GPU, there from
// the moment the library is loaded on that GPU until the process exits.
//
// g_compact_static is a DEVICE array.  In the host functions below its name is only a handle for the runtime's symbol
// table, never a pointer a kernel may be given; the device address is asked for once per GPU with cudaGetSymbolAddress
// (while that GPU is current) and cached in g_compact_static_addr[device].
constexpr int64_t CMP_STATIC_MAX_POINTS = GND_COMPACT_STATIC_MAX_POINTS;
constexpr int CMP_STATIC_SLOTS = (int)(CMP_STATIC_MAX_POINTS / CMP_BLOCK) + 1;    // one counter per 1024 labels (+1 spare)
static_assert(CMP_STATIC_MAX_POINTS % CMP_BLOCK == 0, "the limit is a whole number of compaction blocks");
static __device__ __align__(256) int32_t g_compact_static[CMP_STATIC_SLOTS];       // 16 KB + 4 B per GPU

constexpr int CMP_STATIC_MAX_DEVICES = 64;
static std::atomic<int32_t *> g_compact_static_addr[CMP_STATIC_MAX_DEVICES];

// Device address of g_compact_static on the CURRENT device (looked up on the first call per device, then cached).
static int compact_static_scratch(int32_t **out) {
    int dev = -1;
    GND_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= CMP_STATIC_MAX_DEVICES) return gnd::fail(GND_ERR_UNSUPPORTED, "device ordinal %d >= %d", dev, CMP_STATIC_MAX_DEVICES);
    int32_t *p = g_compact_static_addr[dev].load(std::memory_order_acquire);
    if (!p) {
        void *sym = nullptr;
        GND_CUDA_CHECK(cudaGetSymbolAddress(&sym, g_compact_static));
        if (!sym) return gnd::fail(GND_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the compaction scratch");
        p = (int32_t *)sym;
        g_compact_static_addr[dev].store(p, std::memory_order_release);   // two threads racing here store the same value
    }
    *out = p;
    return GND_OK;
}

extern "C" int gnd_compact_obstacles_static(const int8_t *labels_i8, const float *points_obstacle, int64_t n_points,
                                            int32_t obstacle_features, float *out_obstacle, int32_t *obstacle_count,
                                            void *stream) {
    g_launches = 0;
    if (n_points < 0 || obstacle_features < 1) return fail(GND_ERR_INVALID, "bad shape");
    if (n_points > CMP_STATIC_MAX_POINTS)
        return fail(GND_ERR_UNSUPPORTED,
                    "gnd_compact_obstacles_static handles at most %lld points per call (got %lld): split the cloud or use "
                    "gnd_compact_obstacles with a workspace of your own",
                    (long long)CMP_STATIC_MAX_POINTS, (long long)n_points);
    if (!obstacle_count) return fail(GND_ERR_INVALID, "obstacle_count is null");
    cudaStream_t st = (cudaStream_t)stream;
    if (n_points == 0) { GND_CUDA_CHECK(cudaMemsetAsync(obstacle_count, 0, sizeof(int32_t), st)); return GND_OK; }
    if (!labels_i8 || !points_obstacle || !out_obstacle) return fail(GND_ERR_INVALID, "null pointer");
    int32_t *bc = nullptr;
    if (int rc = compact_static_scratch(&bc)) return rc;
    // Nothing in the scratch has to start at zero: the count kernel overwrites counters [0, nblocks) and the two
    // kernels after it read only those, so whatever an earlier call left behind is never seen.
    const int nblocks = (int)((n_points + CMP_BLOCK - 1) / CMP_BLOCK);
    compact_count_kernel<<<nblocks, CMP_BLOCK, 0, st>>>(labels_i8, n_points, bc);
    GND_LAUNCH_CHECK("compact_count_kernel");
    compact_scan_kernel<<<1, 1024, 0, st>>>(bc, nblocks, obstacle_count);
    GND_LAUNCH_CHECK("compact_scan_kernel");
    compact_write_kernel<<<nblocks, CMP_BLOCK, 0, st>>>(labels_i8, points_obstacle, n_points, obstacle_features, bc, out_obstacle);
    GND_LAUNCH_CHECK("compact_write_kernel");
    return GND_OK;
}
