// raster.cu -- ground-truth rasterisation of the evaluation / label tooling:
//   utils/utils.py:250-270  lidar_to_img        mask[x, y] = fill where float64(z) + 1.732 < 10
//   utils/utils.py:273-295  lidar_to_heightmap  per-cell mean of the first points in INPUT ORDER, reference quirks kept:
//       divisor = count + 1 (num_points starts at 1); the guard `k - 1 <= max_points` accepts max_points + 1 points and
//       writes the last one to slot 0 of the NEXT cell in C order (numba does not bounds-check), where the later write
//       in input order wins; for the very last cell that write leaves the array (undefined in the reference): dropped.
// Cell arithmetic of both (:253-257, :276-281) = segment_cloud's: float32(p - lo), float64 divide, floor, 0 < c < n.
//
// The height map needs "the first max_points + 1 points of every cell in input order": that is the voxelizer's
// problem, so it is handed to the voxelizer.  A pre-pass writes, per point, a surrogate (y_cell + 0.5, x_cell + 0.5,
// 0 | out of range, z): on a unit grid the order-exact voxelizer (voxelize.cu) then returns per-cell point lists in
// input order, and one thread per cell walks its list.
#include "voxelize.cuh"
#include <limits.h>
#include <math.h>

namespace gnd {

__device__ __forceinline__ int raster_axis(float v, double lo, double vs, int n_axis) {
    const float s = __double2float_rn((double)v - lo);
    const double q = floor((double)s / vs);
    if (!(q > 0.0 && q < (double)n_axis)) return -1;
    return (int)q;
}

__global__ void __launch_bounds__(256) raster_img_kernel(const float *points, int64_t n, int F, double lo_x, double lo_y,
                                                         double vs, int nx, int ny, double fill, double *img) {
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
        const float *p = points + i * F;
        if (!((double)__ldg(p + 2) + 1.732 < 10.0)) continue;                    // :252, :264
        const int x = raster_axis(__ldg(p), lo_x, vs, nx), y = raster_axis(__ldg(p + 1), lo_y, vs, ny);
        if (x < 0 || y < 0) continue;
        img[(int64_t)x * ny + y] = fill;                                         // same value from every writer
    }
}

__global__ void __launch_bounds__(256) raster_surrogate_kernel(const float *points, int64_t n, int F, double lo_x, double lo_y,
                                                               double vs, int nx, int ny, float4 *sur) {
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
        const float *p = points + i * F;
        const float z = __ldg(p + 2);
        int x = -1, y = -1;
        if (z < 10.0f) { x = raster_axis(__ldg(p), lo_x, vs, nx); y = raster_axis(__ldg(p + 1), lo_y, vs, ny); }   // :290
        const bool ok = x >= 0 && y >= 0;
        sur[i] = make_float4(ok ? (float)y + 0.5f : -8.0f, ok ? (float)x + 0.5f : -8.0f, ok ? 0.0f : 8.0f, z);
    }
}

__global__ void raster_cellvox_kernel(const int32_t *coors, const int32_t *voxel_count, int ny, int32_t *cellvox) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < *voxel_count) cellvox[coors[v * 3 + 1] * ny + coors[v * 3 + 2]] = v;      // coors are (z, y', x') = (0, x, y)
}

__global__ void raster_height_kernel(const float4 *sur, const int32_t *cellvox, const int32_t *num, const int32_t *off,
                                     const int32_t *list, int cells, int max_points, double *out) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= cells) return;
    const int v = cellvox[c];
    const int cnt = v >= 0 ? num[v] : 0;                       // <= max_points + 1
    const int32_t *mine = v >= 0 ? list + off[v] : nullptr;
    // slot 0: the cell's own first point, or the (max_points + 1)-th point of the previous cell if that came later
    double s = 0.0;
    int i1 = -1;
    if (cnt > 0) { i1 = mine[0]; s = (double)sur[i1].w; }
    if (c > 0) {
        const int pv = cellvox[c - 1];
        if (pv >= 0 && num[pv] == max_points + 1) {
            const int i2 = list[off[pv] + max_points];
            if (i2 > i1) s = (double)sur[i2].w;
        }
    }
    const int own = min(cnt, max_points);
    for (int r = 1; r < own; ++r) s += (double)sur[mine[r]].w;         // left to right, like the reference's slots
    out[c] = s / (double)(cnt + 1);                                    // :295: num_points = accepted points + 1
}

struct RasterPlan {
    int nx, ny;
    gnd_voxel_params vp;
    VoxPlan plan;
    size_t o_sur, o_vox, o_coors, o_num, o_cnt, o_off, o_list, o_cellvox, bytes;
};

static int raster_plan(int64_t n, const double *gs, double vs, int max_points, RasterPlan *rp) {
    if (!gs || !(vs > 0.0)) return fail(GND_ERR_INVALID, "bad grid_size / voxel_size");
    if (n < 0 || n > 1600000) return fail(GND_ERR_UNSUPPORTED, "%lld points > 1600000", (long long)n);
    rp->nx = (int)((gs[2] - gs[0]) / vs);                      // :261-262 int() truncation of the float64 quotient
    rp->ny = (int)((gs[3] - gs[1]) / vs);
    if (rp->nx < 1 || rp->ny < 1) return fail(GND_ERR_INVALID, "empty map");
    if ((int64_t)rp->nx * rp->ny > 100000) return fail(GND_ERR_UNSUPPORTED, "map of %d x %d cells > 100000", rp->nx, rp->ny);
    if (max_points < 1 || max_points > 65534) return fail(GND_ERR_UNSUPPORTED, "max_points %d", max_points);
    gnd_voxel_params &p = rp->vp;
    p.voxel_size[0] = p.voxel_size[1] = 1.0f; p.voxel_size[2] = 1.0f;
    p.coors_range[0] = 0.0f; p.coors_range[1] = 0.0f; p.coors_range[2] = -0.5f;
    p.coors_range[3] = (float)rp->ny; p.coors_range[4] = (float)rp->nx; p.coors_range[5] = 0.5f;
    p.max_points = max_points + 1; p.max_voxels = rp->nx * rp->ny; p.reverse_index = 1; p.num_features = 4;
    p.apply_z_shift = 0; p.z_shift = 0.0f;
    int rc = vox_make_plan(&p, 1, n, &rp->plan);
    if (rc != GND_OK) return rc;
    const size_t cells = (size_t)rp->nx * rp->ny, nn = (size_t)(n < 1 ? 1 : n);
    size_t cur = 0;
    auto take = [&cur](size_t b) { cur = align_up(cur, 256); size_t r = cur; cur += b; return r; };
    rp->o_sur = take(nn * sizeof(float4));
    rp->o_vox = take(vox_carve(rp->plan, 1, n, nullptr, nullptr));
    rp->o_coors = take(cells * 3 * sizeof(int32_t));
    rp->o_num = take(cells * sizeof(int32_t));
    rp->o_cnt = take(sizeof(int32_t));
    rp->o_off = take((cells + 1) * sizeof(int32_t));
    rp->o_list = take(nn * sizeof(int32_t));
    rp->o_cellvox = take(cells * sizeof(int32_t));
    rp->bytes = align_up(cur, 256);
    return GND_OK;
}

}  // namespace gnd

using namespace gnd;

extern "C" int gnd_raster_workspace_bytes(int64_t n_points, const double *h_grid_size4, double voxel_size, int32_t max_points,
                                          size_t *h_bytes) {
    if (!h_bytes) return fail(GND_ERR_INVALID, "null argument");
    RasterPlan rp;
    int rc = raster_plan(n_points, h_grid_size4, voxel_size, max_points, &rp);
    if (rc != GND_OK) return rc;
    *h_bytes = rp.bytes;
    return GND_OK;
}

extern "C" int gnd_lidar_to_img(const float *points, int64_t n_points, int32_t num_features, const double *h_grid_size4,
                                double voxel_size, double fill, double *img, void *stream) {
    g_launches = 0;
    RasterPlan rp;
    int rc = raster_plan(n_points, h_grid_size4, voxel_size, 1, &rp);
    if (rc != GND_OK) return rc;
    if (num_features < 3 || !img || (n_points > 0 && !points)) return fail(GND_ERR_INVALID, "bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    GND_CUDA_CHECK(cudaMemsetAsync(img, 0, (size_t)rp.nx * rp.ny * sizeof(double), st));
    if (n_points == 0) return GND_OK;
    int blocks = (int)((n_points + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    raster_img_kernel<<<blocks, 256, 0, st>>>(points, n_points, num_features, h_grid_size4[0], h_grid_size4[1], voxel_size,
                                              rp.nx, rp.ny, fill, img);
    GND_LAUNCH_CHECK("raster_img_kernel");
    return GND_OK;
}

extern "C" int gnd_lidar_to_heightmap(const float *points, int64_t n_points, int32_t num_features, const double *h_grid_size4,
                                      double voxel_size, int32_t max_points, void *workspace, size_t workspace_bytes,
                                      double *heightmap, void *stream) {
    g_launches = 0;
    RasterPlan rp;
    int rc = raster_plan(n_points, h_grid_size4, voxel_size, max_points, &rp);
    if (rc != GND_OK) return rc;
    if (num_features < 3 || !heightmap || (n_points > 0 && !points)) return fail(GND_ERR_INVALID, "bad argument");
    if (!workspace || workspace_bytes < rp.bytes) return fail(GND_ERR_WORKSPACE, "raster workspace: need %zu bytes", rp.bytes);
    if ((uintptr_t)workspace & 255) return fail(GND_ERR_INVALID, "workspace must be 256-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    char *w = (char *)workspace;
    float4 *sur = (float4 *)(w + rp.o_sur);
    int32_t *coors = (int32_t *)(w + rp.o_coors), *num = (int32_t *)(w + rp.o_num), *cnt = (int32_t *)(w + rp.o_cnt);
    int32_t *off = (int32_t *)(w + rp.o_off), *list = (int32_t *)(w + rp.o_list), *cellvox = (int32_t *)(w + rp.o_cellvox);
    const int cells = rp.nx * rp.ny;
    if (n_points > 0) {
        int blocks = (int)((n_points + 255) / 256);
        if (blocks > 148 * 8) blocks = 148 * 8;
        raster_surrogate_kernel<<<blocks, 256, 0, st>>>(points, n_points, num_features, h_grid_size4[0], h_grid_size4[1],
                                                        voxel_size, rp.nx, rp.ny, sur);
        GND_LAUNCH_CHECK("raster_surrogate_kernel");
    }
    VoxWs vws;
    vox_carve(rp.plan, 1, n_points, w + rp.o_vox, &vws);
    rc = vox_run(&rp.vp, rp.plan, vws, reinterpret_cast<const float *>(sur), nullptr, 1, n_points, coors, num, cnt, off, list,
                 nullptr, nullptr, st);
    if (rc != GND_OK) return rc;
    GND_CUDA_CHECK(cudaMemsetAsync(cellvox, 0xff, (size_t)cells * sizeof(int32_t), st));
    raster_cellvox_kernel<<<(cells + 255) / 256, 256, 0, st>>>(coors, cnt, rp.ny, cellvox);
    GND_LAUNCH_CHECK("raster_cellvox_kernel");
    raster_height_kernel<<<(cells + 127) / 128, 128, 0, st>>>(sur, cellvox, num, off, list, cells, max_points, heightmap);
    GND_LAUNCH_CHECK("raster_height_kernel");
    return GND_OK;
}
