// cloud_io.cu -- the steps either side of the hot path in the streaming node.
//
// Upstream (ros_node.py:242-263):
//   utils/utils.py:148-195    cloud_msg_to_numpy: xyz = bytes 0..11 of every point_step-byte row of msg.data,
//                             three little-endian float32 (:179-180)
//   utils/transform.py:148-165 transformCloud: res = dot(trafo, [x, y, z, 1.0]) in float64 (row sums in k order,
//                             no contraction), stored back as float32
//   ros_node.py:263           cloud[~np.isnan(cloud).any(axis=1)]: rows with a NaN dropped, order kept
// Downstream (ros_node.py:280-288):
//   utils/ros_utils.py:171-191 np2ros_pub_2: 16-byte wire rows (x, y, z, intensity = label as float32)
//   utils/ros_utils.py:210-248 array_to_pointcloud2: 12-byte rows x, y, z -- exactly what gnd_compact_obstacles writes
//
// All HBM-bound streaming kernels: 12..point_step bytes read + 12 bytes written per point (x2 reads when NaN rows
// are dropped: the order-preserving compaction counts first, then writes).
#include "common.cuh"

namespace gnd {

constexpr int IO_BLOCK = 1024;

struct UnpackArgs {
    const uint8_t *data;
    int64_t n;
    int32_t step;
    int32_t aligned;       // rows start on 4-byte boundaries
    int32_t has_t;
    double t[12];          // first three rows of the 4x4
    int32_t drop_nan;
    int32_t F;
    float *out;
    int32_t *block_counts;
    int32_t *count;
};

__device__ __forceinline__ bool unpack_row(const UnpackArgs &a, int64_t i, float &x, float &y, float &z) {
    const uint8_t *row = a.data + i * (int64_t)a.step;
    uint32_t w[3];
    if (a.aligned) {
        const uint32_t *r = reinterpret_cast<const uint32_t *>(row);
        w[0] = __ldg(r); w[1] = __ldg(r + 1); w[2] = __ldg(r + 2);
    } else {
#pragma unroll
        for (int j = 0; j < 3; ++j)
            w[j] = (uint32_t)__ldg(row + 4 * j) | ((uint32_t)__ldg(row + 4 * j + 1) << 8) |
                   ((uint32_t)__ldg(row + 4 * j + 2) << 16) | ((uint32_t)__ldg(row + 4 * j + 3) << 24);
    }
    x = __uint_as_float(w[0]); y = __uint_as_float(w[1]); z = __uint_as_float(w[2]);
    if (a.has_t) {
        const double v0 = (double)x, v1 = (double)y, v2 = (double)z;
        float o[3];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            double acc = __dadd_rn(0.0, __dmul_rn(a.t[4 * r + 0], v0));
            acc = __dadd_rn(acc, __dmul_rn(a.t[4 * r + 1], v1));
            acc = __dadd_rn(acc, __dmul_rn(a.t[4 * r + 2], v2));
            acc = __dadd_rn(acc, __dmul_rn(a.t[4 * r + 3], 1.0));
            o[r] = __double2float_rn(acc);
        }
        x = o[0]; y = o[1]; z = o[2];
    }
    return !(a.drop_nan && (isnan(x) || isnan(y) || isnan(z)));
}

__global__ void __launch_bounds__(IO_BLOCK) unpack_count_kernel(const UnpackArgs a) {
    __shared__ int s_warp[33];
    const int64_t i = (int64_t)blockIdx.x * IO_BLOCK + threadIdx.x;
    float x, y, z;
    const int flag = (i < a.n && unpack_row(a, i, x, y, z)) ? 1 : 0;
    int total;
    block_exclusive_scan(flag, s_warp, &total);
    if (threadIdx.x == 0) a.block_counts[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024) unpack_scan_kernel(int32_t *block_counts, int nblocks, int32_t *count) {
    __shared__ int s_warp[33];
    int carry = 0;
    for (int base = 0; base < nblocks; base += 1024) {
        const int idx = base + threadIdx.x;
        const int v = idx < nblocks ? block_counts[idx] : 0;
        int total;
        const int ex = block_exclusive_scan(v, s_warp, &total);
        if (idx < nblocks) block_counts[idx] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) *count = carry;
}

__global__ void __launch_bounds__(IO_BLOCK) unpack_write_kernel(const UnpackArgs a) {
    __shared__ int s_warp[33];
    const int64_t i = (int64_t)blockIdx.x * IO_BLOCK + threadIdx.x;
    float x = 0.f, y = 0.f, z = 0.f;
    const int flag = (i < a.n && unpack_row(a, i, x, y, z)) ? 1 : 0;
    int ex = 0;
    int64_t o = i;
    if (a.drop_nan) {
        int total;
        ex = block_exclusive_scan(flag, s_warp, &total);
        o = (int64_t)a.block_counts[blockIdx.x] + ex;
    }
    if (flag) {
        float *dst = a.out + o * a.F;
        if (a.F == 4) {
            *reinterpret_cast<float4 *>(dst) = make_float4(x, y, z, 0.0f);
        } else {
            dst[0] = x; dst[1] = y; dst[2] = z;
            for (int f = 3; f < a.F; ++f) dst[f] = 0.0f;
        }
    }
    if (!a.drop_nan && i == 0) *a.count = (int32_t)a.n;
}

__global__ void __launch_bounds__(256) pack_segmented_kernel(const float *points, int64_t n, int F, const int8_t *lab8,
                                                             const double *lab64, float4 *out) {
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
        const float *p = points + i * F;
        const float l = lab8 ? (float)lab8[i] : __double2float_rn(lab64[i]);
        out[i] = make_float4(__ldg(p), __ldg(p + 1), __ldg(p + 2), l);
    }
}

}  // namespace gnd

using namespace gnd;

extern "C" int gnd_unpack_workspace_bytes(int64_t n_points, size_t *h_bytes) {
    if (!h_bytes || n_points < 0) return fail(GND_ERR_INVALID, "bad argument");
    *h_bytes = align_up((size_t)((n_points + IO_BLOCK - 1) / IO_BLOCK + 1) * sizeof(int32_t), 256);
    return GND_OK;
}

extern "C" int gnd_unpack_cloud(const gnd_unpack_params *p, const uint8_t *msg_data, int64_t n_points, void *workspace,
                                size_t workspace_bytes, float *out_points, int32_t *out_count, void *stream) {
    g_launches = 0;
    if (!p || !out_count) return fail(GND_ERR_INVALID, "null argument");
    if (n_points < 0 || n_points > 0x7fffffffLL) return fail(GND_ERR_INVALID, "n_points %lld", (long long)n_points);
    if (p->point_step < 12) return fail(GND_ERR_INVALID, "point_step %d < 12 (x, y, z float32)", p->point_step);
    if (p->out_features < 3) return fail(GND_ERR_INVALID, "out_features must be >= 3");
    cudaStream_t st = (cudaStream_t)stream;
    if (n_points == 0) { GND_CUDA_CHECK(cudaMemsetAsync(out_count, 0, sizeof(int32_t), st)); return GND_OK; }
    if (!msg_data || !out_points) return fail(GND_ERR_INVALID, "null pointer");
    if (p->out_features == 4 && ((uintptr_t)out_points & 15)) return fail(GND_ERR_INVALID, "out_points must be 16-byte aligned");
    UnpackArgs a;
    a.data = msg_data; a.n = n_points; a.step = p->point_step;
    a.aligned = (p->point_step % 4 == 0 && ((uintptr_t)msg_data & 3) == 0) ? 1 : 0;
    a.has_t = p->has_transform;
    for (int i = 0; i < 12; ++i) a.t[i] = p->transform[i];
    a.drop_nan = p->drop_nan; a.F = p->out_features; a.out = out_points; a.count = out_count;
    a.block_counts = nullptr;
    const int nblocks = (int)((n_points + IO_BLOCK - 1) / IO_BLOCK);
    if (p->drop_nan) {
        size_t need;
        gnd_unpack_workspace_bytes(n_points, &need);
        if (!workspace || workspace_bytes < need) return fail(GND_ERR_WORKSPACE, "unpack workspace: need %zu bytes", need);
        a.block_counts = (int32_t *)workspace;
        unpack_count_kernel<<<nblocks, IO_BLOCK, 0, st>>>(a);
        GND_LAUNCH_CHECK("unpack_count_kernel");
        unpack_scan_kernel<<<1, 1024, 0, st>>>(a.block_counts, nblocks, out_count);
        GND_LAUNCH_CHECK("unpack_scan_kernel");
    }
    unpack_write_kernel<<<nblocks, IO_BLOCK, 0, st>>>(a);
    GND_LAUNCH_CHECK("unpack_write_kernel");
    return GND_OK;
}

extern "C" int gnd_pack_segmented(const float *points, int64_t n_points, int32_t num_features, const int8_t *labels_i8,
                                  const double *labels_f64, float *out_rows, void *stream) {
    g_launches = 0;
    if (n_points < 0 || num_features < 3) return fail(GND_ERR_INVALID, "bad shape");
    if (n_points == 0) return GND_OK;
    if (!points || !out_rows || (!labels_i8 && !labels_f64)) return fail(GND_ERR_INVALID, "null pointer");
    if ((uintptr_t)out_rows & 15) return fail(GND_ERR_INVALID, "out_rows must be 16-byte aligned");
    int blocks = (int)((n_points + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    pack_segmented_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(points, n_points, num_features, labels_i8, labels_f64,
                                                                     reinterpret_cast<float4 *>(out_rows));
    GND_LAUNCH_CHECK("pack_segmented_kernel");
    return GND_OK;
}

// ---- the evaluation reductions of evaluate_SemanticKITTI.py on the device ----------------
//   :94-100  get_GndSeg: 0 for the ground classes, -1 for sem_label in {0, 1} (unlabelled / outlier), else 1
//   :102-111 remove_outliers: keep points with pred >= 0 and GndSeg >= 0, then ground = 1 - value
//   :198-204 intersection / union / pred sum / GT sum over the kept points (integers: exact, order-free)
//   :224-226 mse = sum(square(target - pred.T) * mask) / sum(mask) in float64
namespace gnd {

struct EvalArgs {
    const int8_t *lab8;
    const double *lab64;
    const uint32_t *sem;
    int64_t n;
    int32_t n_classes;
    uint32_t classes[16];
    unsigned long long *counts;      // [4]: intersection, union, predicted ground, GT ground
};

__device__ __forceinline__ void eval_point(const EvalArgs &a, int pred, uint32_t s, unsigned &c_i, unsigned &c_u,
                                           unsigned &c_p, unsigned &c_g) {
    bool is_gnd = false;
    for (int k = 0; k < a.n_classes; ++k) is_gnd |= (s == a.classes[k]);
    const int gt = (s == 0u || s == 1u) ? -1 : (is_gnd ? 0 : 1);       // the {0,1} assignment comes second (:98-99)
    if (pred >= 0 && gt >= 0) {
        const bool pg = pred == 0, gg = gt == 0;
        c_i += (pg && gg); c_u += (pg || gg); c_p += pg; c_g += gg;
    }
}

__global__ void __launch_bounds__(256) eval_counts_kernel(const EvalArgs a) {
    unsigned c_i = 0, c_u = 0, c_p = 0, c_g = 0;
    const int64_t tid = (int64_t)blockIdx.x * 256 + threadIdx.x, nthr = (int64_t)gridDim.x * 256;
    int64_t done = 0;
    if (a.lab8 && (((uintptr_t)a.lab8) & 3) == 0 && (((uintptr_t)a.sem) & 15) == 0) {
        // four points per thread and iteration: one 4-byte label load, one 16-byte class load
        const int64_t n4 = a.n >> 2;
        const uint32_t *l4 = reinterpret_cast<const uint32_t *>(a.lab8);
        const uint4 *s4 = reinterpret_cast<const uint4 *>(a.sem);
        for (int64_t i = tid; i < n4; i += nthr) {
            const uint32_t l = __ldg(l4 + i);
            const uint4 sv = __ldg(s4 + i);
            eval_point(a, (int)(int8_t)(l & 0xffu), sv.x, c_i, c_u, c_p, c_g);
            eval_point(a, (int)(int8_t)((l >> 8) & 0xffu), sv.y, c_i, c_u, c_p, c_g);
            eval_point(a, (int)(int8_t)((l >> 16) & 0xffu), sv.z, c_i, c_u, c_p, c_g);
            eval_point(a, (int)(int8_t)(l >> 24), sv.w, c_i, c_u, c_p, c_g);
        }
        done = n4 << 2;
    }
    for (int64_t i = done + tid; i < a.n; i += nthr) {
        const int pred = a.lab8 ? (int)a.lab8[i] : (int)a.lab64[i];
        eval_point(a, pred, __ldg(a.sem + i), c_i, c_u, c_p, c_g);
    }
    c_i = __reduce_add_sync(0xffffffffu, c_i); c_u = __reduce_add_sync(0xffffffffu, c_u);
    c_p = __reduce_add_sync(0xffffffffu, c_p); c_g = __reduce_add_sync(0xffffffffu, c_g);
    // one set of four global atomics per BLOCK: same-address atomics serialise in L2
    __shared__ unsigned s_c[4];
    if (threadIdx.x < 4) s_c[threadIdx.x] = 0;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&s_c[0], c_i); atomicAdd(&s_c[1], c_u); atomicAdd(&s_c[2], c_p); atomicAdd(&s_c[3], c_g);
    }
    __syncthreads();
    if (threadIdx.x < 4 && s_c[threadIdx.x]) atomicAdd(a.counts + threadIdx.x, (unsigned long long)s_c[threadIdx.x]);
}

// one block, fixed reduction tree: the float64 sums are reproducible run to run
__global__ void __launch_bounds__(1024) eval_sqerr_kernel(const float *pred, const double *target, const double *mask,
                                                          int nx, int ny, double *sums) {
    __shared__ double s_a[1024], s_b[1024];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < nx * ny; i += 1024) {
        const int x = i / ny, y = i - x * ny;                 // target / mask are [nx][ny]; pred.T[x][y] = pred[y][x]
        const double d = target[i] - (double)pred[(size_t)y * nx + x];
        a += d * d * mask[i];
        b += mask[i];
    }
    s_a[threadIdx.x] = a; s_b[threadIdx.x] = b;
    __syncthreads();
    for (int s = 512; s > 0; s >>= 1) {
        if (threadIdx.x < s) { s_a[threadIdx.x] += s_a[threadIdx.x + s]; s_b[threadIdx.x] += s_b[threadIdx.x + s]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { sums[0] = s_a[0]; sums[1] = s_b[0]; }
}

}  // namespace gnd

extern "C" int gnd_eval_ground_counts(const int8_t *labels_i8, const double *labels_f64, const uint32_t *sem_label,
                                      int64_t n_points, const uint32_t *h_ground_classes, int32_t n_classes,
                                      uint64_t *counts4, void *stream) {
    g_launches = 0;
    if (n_points < 0 || n_classes < 0 || n_classes > 16) return fail(GND_ERR_INVALID, "bad n_points / n_classes (<= 16)");
    if (!counts4 || (n_classes && !h_ground_classes)) return fail(GND_ERR_INVALID, "null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    GND_CUDA_CHECK(cudaMemsetAsync(counts4, 0, 4 * sizeof(uint64_t), st));
    if (n_points == 0) return GND_OK;
    if ((!labels_i8 && !labels_f64) || !sem_label) return fail(GND_ERR_INVALID, "null pointer");
    EvalArgs a;
    a.lab8 = labels_i8; a.lab64 = labels_f64; a.sem = sem_label; a.n = n_points; a.n_classes = n_classes;
    for (int k = 0; k < 16; ++k) a.classes[k] = k < n_classes ? h_ground_classes[k] : 0u;
    a.counts = reinterpret_cast<unsigned long long *>(counts4);
    int blocks = (int)((n_points + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    eval_counts_kernel<<<blocks, 256, 0, st>>>(a);
    GND_LAUNCH_CHECK("eval_counts_kernel");
    return GND_OK;
}

extern "C" int gnd_eval_masked_sqerr(const float *pred, const double *target, const double *mask, int32_t nx, int32_t ny,
                                     double *sums2, void *stream) {
    g_launches = 0;
    if (nx < 1 || ny < 1) return fail(GND_ERR_INVALID, "empty map");
    if (!pred || !target || !mask || !sums2) return fail(GND_ERR_INVALID, "null pointer");
    eval_sqerr_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(pred, target, mask, nx, ny, sums2);
    GND_LAUNCH_CHECK("eval_sqerr_kernel");
    return GND_OK;
}
