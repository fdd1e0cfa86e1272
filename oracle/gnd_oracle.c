/*
 * gnd_oracle.c -- TEST INFRASTRUCTURE ONLY (CPU oracle).
 *
 * Plain-C restatement of the integer / index stages of the GndNet inference
 * hot path.  It exists so that tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs have a checker that travels to the GPU
 * box (where /root/reference does not exist).  Nothing under gndnet_b200/ may
 * import, link or call this file.
 *
 * Each function cites the reference lines it restates (paths relative to the
 * reference repo root).  Parity of this restatement is pinned by
 * oracle/ref_harness.py, which runs the unmodified reference (numba) next to
 * it in the build container and writes the golden fixtures in tests/golden/.
 *
 * Build: see oracle/Makefile  (gcc -O2 -ffp-contract=off, no -ffast-math:
 * the voxel index arithmetic must be IEEE float32 subtract + divide).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__GNUC__)
#define GND_EXPORT __attribute__((visibility("default")))
#else
#define GND_EXPORT
#endif

/* predict_ground.py:104-107  _shift_cloud: cloud += float32([0,0,h,0]).
 * Every column gets a float32 add (x+0, y+0, z+h, r+0), so -0.0 becomes +0.0
 * exactly as in the reference.  cols may be 3 (camera) or 4. */
GND_EXPORT void gnd_oracle_shift_cloud(float *cloud, int64_t n, int cols, float height)
{
    for (int64_t i = 0; i < n; ++i) {
        volatile float *p = cloud + i * cols;
        for (int j = 0; j < cols; ++j) {
            float add = (j == 2) ? height : 0.0f;
            p[j] = p[j] + add;
        }
    }
}

/* utils/point_cloud_ops.py:24-27 (kernel) and :142-143 (wrapper):
 * grid = round((hi - lo) / vs) evaluated in the points' dtype (float32),
 * numpy round == round-half-even == rintf in the default rounding mode. */
GND_EXPORT void gnd_oracle_grid_size(const float voxel_size[3], const float coors_range[6],
                                     int32_t grid_xyz[3])
{
    for (int j = 0; j < 3; ++j) {
        volatile float d = coors_range[3 + j] - coors_range[j];
        volatile float q = d / voxel_size[j];
        grid_xyz[j] = (int32_t)rintf(q);
    }
}

/* utils/point_cloud_ops.py:7-53 (_points_to_voxel_reverse_kernel) and :55-104
 * (_points_to_voxel_kernel, reverse_index = 0), plus the wrapper's buffer
 * contract (:147-151): voxels/coors/num_points_per_voxel arrive zero-filled,
 * coor_to_voxelidx arrives filled with -1 and is shaped (nz,ny,nx) when
 * reverse_index else (nx,ny,nz).
 *
 * point_slot (optional, may be NULL): for every input point, the flat slot
 * voxelidx*max_points + num it was stored at, or -1 if it was dropped.  Not a
 * reference output; used by the tests to check the fused CUDA path, which
 * never materialises the dense voxels tensor.
 *
 * Returns voxel_num.  NaN coordinates are undefined behaviour in the
 * reference (they pass both range tests and index garbage); here, as in the
 * CUDA path, a point with a NaN coordinate is dropped. */
GND_EXPORT int32_t gnd_oracle_points_to_voxel(
    const float *points, int64_t n, int cols,
    const float voxel_size[3], const float coors_range[6],
    int32_t max_points, int32_t max_voxels, int reverse_index,
    float *voxels, int32_t *coors, int32_t *num_points_per_voxel,
    int32_t *coor_to_voxelidx, int64_t *point_slot)
{
    int32_t grid[3];
    gnd_oracle_grid_size(voxel_size, coors_range, grid);
    /* extents of coor_to_voxelidx in storage order */
    int64_t d1, d2;
    if (reverse_index) { d1 = grid[1]; d2 = grid[0]; }   /* [z][y][x] */
    else               { d1 = grid[1]; d2 = grid[2]; }   /* [x][y][z] */
    int32_t voxel_num = 0;
    int64_t i = 0;
    for (; i < n; ++i) {
        const float *p = points + i * cols;
        int32_t c3[3];
        int failed = 0;
        for (int j = 0; j < 3; ++j) {
            /* :34  c = floor((points[i,j] - coors_range[j]) / voxel_size[j]) in float32 */
            volatile float d = p[j] - coors_range[j];
            volatile float q = d / voxel_size[j];
            float c = floorf(q);
            /* :35  if c < 0 or c >= grid_size[j] */
            if (!(c >= 0.0f) || (double)c >= (double)grid[j]) { failed = 1; break; }
            c3[j] = (int32_t)c;
        }
        if (point_slot) point_slot[i] = -1;
        if (failed) continue;
        int32_t coor[3];
        if (reverse_index) { coor[0] = c3[2]; coor[1] = c3[1]; coor[2] = c3[0]; }
        else               { coor[0] = c3[0]; coor[1] = c3[1]; coor[2] = c3[2]; }
        int64_t cell = ((int64_t)coor[0] * d1 + coor[1]) * d2 + coor[2];
        int32_t voxelidx = coor_to_voxelidx[cell];
        if (voxelidx == -1) {
            voxelidx = voxel_num;
            if (voxel_num >= max_voxels) break;            /* :44-45 */
            voxel_num += 1;
            coor_to_voxelidx[cell] = voxelidx;
            coors[voxelidx * 3 + 0] = coor[0];
            coors[voxelidx * 3 + 1] = coor[1];
            coors[voxelidx * 3 + 2] = coor[2];
        }
        int32_t num = num_points_per_voxel[voxelidx];
        if (num < max_points) {                            /* :49-52 */
            if (voxels)
                memcpy(voxels + ((int64_t)voxelidx * max_points + num) * cols, p,
                       sizeof(float) * (size_t)cols);
            if (point_slot) point_slot[i] = (int64_t)voxelidx * max_points + num;
            num_points_per_voxel[voxelidx] = num + 1;
        }
    }
    if (point_slot) for (int64_t k = i; k < n; ++k) point_slot[k] = -1;
    return voxel_num;
}

/* utils/utils.py:198-220  segment_cloud and :222-248 segment_cloud_noground.
 *
 * points        [n, cols] float32, C order; columns 0,1 are MUTATED like the
 *               reference does (:204  lidar_data -= grid_size[:2], rounded to
 *               float32 because it is an in-place update of a float32 view).
 * grid_lo       grid_size[0], grid_size[1] as the doubles numpy/numba see.
 * voxel_size    Python scalar -> double (:205 divides in float64).
 * emap          elevation_map as the reference indexes it: emap[x, y], given
 *               with explicit element strides (the callers pass pred.T, an
 *               F-ordered view, predict_ground.py:168).
 * labels        float64 [n]  (np.zeros default dtype, :201).
 * points_obstacle  NULL for segment_cloud; else [n, ocols] compacted in place
 *               (:242-243), return value = obs_cnt.
 */
GND_EXPORT int64_t gnd_oracle_segment_cloud(
    float *points, int64_t n, int cols,
    const double grid_lo[2], double voxel_size,
    const float *emap, int64_t emap_nx, int64_t emap_ny, int64_t emap_sx, int64_t emap_sy,
    double threshold, double *labels,
    float *points_obstacle, int ocols)
{
    int64_t obs_cnt = 0;
    for (int64_t i = 0; i < n; ++i) {
        float *p = points + i * cols;
        int32_t cell[2];
        for (int j = 0; j < 2; ++j) {
            float shifted = (float)((double)p[j] - grid_lo[j]);   /* :204 */
            p[j] = shifted;
            double q = floor((double)shifted / voxel_size);        /* :205-206 */
            /* :207 astype(int32): x86 cvttsd2si semantics for out-of-range / NaN */
            int32_t ci;
            if (!(q >= -2147483648.0 && q <= 2147483647.0)) ci = INT32_MIN;
            else ci = (int32_t)q;
            cell[j] = ci;
        }
        int32_t x = cell[0], y = cell[1];
        if (0 < x && x < emap_nx && 0 < y && y < emap_ny) {        /* :213 */
            double h = (double)emap[x * emap_sx + y * emap_sy];
            if ((double)p[2] > h + threshold) {                    /* :214 */
                labels[i] = 1.0;
                if (points_obstacle) {
                    memmove(points_obstacle + obs_cnt * ocols, points_obstacle + i * ocols,
                            sizeof(float) * (size_t)ocols);
                    obs_cnt += 1;
                }
            } else {
                labels[i] = 0.0;
            }
        } else {
            labels[i] = -1.0;                                      /* :219 */
        }
    }
    return obs_cnt;
}

/* ---------------------------------------------------------------------------
 * SURVEY.md 8(f) row 1 -- the steps immediately upstream of the voxelizer in the streaming node:
 *   utils/utils.py:148-195  cloud_msg_to_numpy: xyz = the first 12 bytes of every point_step-byte row of
 *                           msg.data viewed as three little-endian float32 (:179-180)
 *   utils/transform.py:148-165 transformCloud: v = append(xyz, 1.0) (float64), res = dot(trafo, v) (float64,
 *                           numba np.dot on a 4x4 @ 4 -> row sums in k order), points[i] = res[:3] as float32
 *   ros_node.py:263         cloud = cloud[~np.isnan(cloud).any(axis=1)]  (stable, after the transform)
 * out is [n][out_cols] float32 (out_cols >= 3, extra columns zero); returns the number of rows kept.
 * ------------------------------------------------------------------------- */
GND_EXPORT int64_t gnd_oracle_unpack_cloud(const uint8_t *data, int64_t n, int32_t point_step, const double *trafo,
                                           int drop_nan, int out_cols, float *out)
{
    int64_t m = 0;
    for (int64_t i = 0; i < n; ++i) {
        float p[3];
        memcpy(p, data + i * (int64_t)point_step, 12);
        if (trafo) {
            const double v[4] = {(double)p[0], (double)p[1], (double)p[2], 1.0};
            for (int r = 0; r < 3; ++r) {
                volatile double acc = 0.0;
                for (int k = 0; k < 4; ++k) acc = acc + trafo[4 * r + k] * v[k];
                p[r] = (float)acc;
            }
        }
        if (drop_nan && (isnan(p[0]) || isnan(p[1]) || isnan(p[2]))) continue;
        float *o = out + m * out_cols;
        o[0] = p[0]; o[1] = p[1]; o[2] = p[2];
        for (int j = 3; j < out_cols; ++j) o[j] = 0.0f;
        ++m;
    }
    return m;
}

/* SURVEY.md 8(f) row 2 -- the wire layouts immediately downstream:
 *   utils/ros_utils.py:171-191 np2ros_pub_2: structured rows (x, y, z, intensity) float32, 16 bytes per point,
 *                           intensity = the segment_cloud label (float64 -> float32)
 *   utils/ros_utils.py:210-248 array_to_pointcloud2 without colours: 12-byte rows x, y, z (the compacted obstacle
 *                           rows of segment_cloud_noground are already in this layout)
 * points is [n][cols] float32; out is [n][4] float32. */
GND_EXPORT void gnd_oracle_pack_segmented(const float *points, int64_t n, int cols, const double *labels, float *out)
{
    for (int64_t i = 0; i < n; ++i) {
        out[4 * i + 0] = points[i * cols + 0];
        out[4 * i + 1] = points[i * cols + 1];
        out[4 * i + 2] = points[i * cols + 2];
        out[4 * i + 3] = (float)labels[i];
    }
}

/* ---------------------------------------------------------------------------
 * SURVEY.md 8(f) row 3, ground-truth rasterisation (evaluate_SemanticKITTI.py:118-127 get_target_gnd):
 *   utils/utils.py:250-270 lidar_to_img        mask[x, y] = fill for points with float64(z) + 1.732 < 10
 *   utils/utils.py:273-295 lidar_to_heightmap  mean of the first points of each cell, with the reference's quirks:
 *     - num_points starts at 1 ("to avoid division by zero") and the divisor is that counter: count + 1;
 *     - the guard is `k - 1 <= max_points`, so a cell accepts max_points + 1 points and the last one is written to
 *       heightmap[x, y, max_points] -- one past the cell's slots.  numba does not bounds-check: in C order that is
 *       slot 0 of the NEXT cell (x, y + 1) (or (x + 1, 0)), which keeps whichever write came last in input order.
 *       For the very last cell the write leaves the array (real undefined behaviour): dropped here.
 * Cell arithmetic as in segment_cloud (:204-207): float32(p - lo), float64 divide by the Python scalar, floor, and the
 * exclusive test 0 < x < nx, 0 < y < ny.  points [n][cols] float32 (not modified); out maps [nx][ny] float64.
 * ------------------------------------------------------------------------- */
static int gnd_oracle_cell(float v, double lo, double vs, int n_axis)
{
    volatile float s = (float)((double)v - lo);
    double q = floor((double)s / vs);
    if (!(q > 0.0 && q < (double)n_axis)) return -1;
    return (int)q;
}

GND_EXPORT void gnd_oracle_lidar_to_img(const float *points, int64_t n, int cols, const double *grid_size4,
                                        double voxel_size, double fill, int nx, int ny, double *img)
{
    for (int64_t c = 0; c < (int64_t)nx * ny; ++c) img[c] = 0.0;
    for (int64_t i = 0; i < n; ++i) {
        const float *p = points + i * cols;
        if (!((double)p[2] + 1.732 < 10.0)) continue;
        int x = gnd_oracle_cell(p[0], grid_size4[0], voxel_size, nx);
        int y = gnd_oracle_cell(p[1], grid_size4[1], voxel_size, ny);
        if (x < 0 || y < 0) continue;
        img[(int64_t)x * ny + y] = fill;
    }
}

GND_EXPORT int gnd_oracle_lidar_to_heightmap(const float *points, int64_t n, int cols, const double *grid_size4,
                                             double voxel_size, int max_points, int nx, int ny, double *slots /* scratch
                                             [nx*ny*max_points + 1] */, int32_t *num /* scratch [nx*ny] */, double *out)
{
    const int64_t cells = (int64_t)nx * ny;
    for (int64_t k = 0; k < cells * max_points + 1; ++k) slots[k] = 0.0;
    for (int64_t c = 0; c < cells; ++c) num[c] = 1;
    for (int64_t i = 0; i < n; ++i) {
        const float *p = points + i * cols;
        if (!(p[2] < 10.0f)) continue;
        int x = gnd_oracle_cell(p[0], grid_size4[0], voxel_size, nx);
        int y = gnd_oracle_cell(p[1], grid_size4[1], voxel_size, ny);
        if (x < 0 || y < 0) continue;
        const int64_t c = (int64_t)x * ny + y;
        const int k = num[c];
        if (k - 1 <= max_points) {
            slots[c * max_points + (k - 1)] = (double)p[2];      /* k - 1 == max_points lands in the next cell's slot 0 */
            num[c] += 1;
        }
    }
    for (int64_t c = 0; c < cells; ++c) {
        double s = 0.0;
        for (int k = 0; k < max_points; ++k) s += slots[c * max_points + k];
        out[c] = s / (double)num[c];
    }
    return 0;
}
