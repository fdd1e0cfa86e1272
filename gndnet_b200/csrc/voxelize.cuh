// voxelize.cuh -- internal interface of the order-exact parallel voxelizer (see voxelize.cu).
#pragma once
#include "common.cuh"

namespace gnd {

constexpr int VOX_SUB = 2048;        // points ranked per sub-tile of a chunk
constexpr int VOX_MAX_T = 61440;     // chunk size limit: chunk-local counts are uint16
constexpr int VOX_MAX_K = 64;        // chunks per scan on big grids (K x cells <= 655 360 lifts it up to 296)
// tile-parallel ranker: 8 warps x 7 steps x 32 points per tile (per-warp byte counters hold <= 224)
constexpr int VOX_RT_WARPS = 8;
constexpr int VOX_RT_STEPS = 7;
constexpr int VOX_RT_TILE = VOX_RT_WARPS * VOX_RT_STEPS * 32;
constexpr int VOX_RT_MAX_CELLS = 11264;   // 10 bytes of shared memory per cell, two CTAs per SM
constexpr int VL_KC = 8;             // set-exact path: chunks of the index range that the per-cell counts are kept for

struct VoxPlan {
    int32_t grid[3];     // nx, ny, nz   (utils/point_cloud_ops.py:24-27)
    int32_t cells;       // nx*ny*nz
    int32_t K;           // chunks per scan
    int32_t T;           // points per chunk (multiple of VOX_SUB)
    int32_t G;           // vox_assign_kernel CTAs per scan
    int32_t tiled;       // 1: vox_rank_tiles_kernel (grids <= VOX_RT_MAX_CELLS cells), 0: vox_rank_kernel
    size_t rank_smem;    // dynamic smem of the rank kernel
    size_t assign_smem;  // dynamic smem of the assign kernel
    int32_t lean_ok;     // the set-exact path applies (cells fit 16 bits)
};

// scratch carved from the caller's workspace
struct VoxWs {
    int32_t *cell;       // [B][N]      linear cell of every point, -1 if rejected
    uint16_t *lrank;     // [B][N]      rank of the point among its cell's points inside its chunk
    int32_t *first;      // [B][cells]  index of the first point of the cell (INT_MAX-like if none)
    int32_t *done;       // [B]         assign CTAs finished per scan (counts up from the 0x7f fill, self-resetting)
    uint32_t *hist;      // [B][K][cells] chunk histogram -> exclusive prefix over chunks (in place)
    int32_t *vid;        // [B][cells]  voxel id of the cell, -1 if none
    int32_t *cutoff;     // [B]         first dropped index when max_voxels overflows, else INT_MAX
    int32_t *offc;       // [B][cells+1] CSR offsets of the kept points in CELL order (pillar-encode input)
    float4 *sorted;      // [B][N]      kept points (x, y, z [+shift], r) grouped by cell, slot order inside; may be null
    // ---- set-exact ("lean") path of the fused pipeline: private, self-resetting state (see voxelize.cu) ----
    int32_t *lfirst;     // [B][cells]  index of the cell's first in-range point (INT_MAX-like between calls)
    int32_t *lcnt;       // [B][cells]  in-range points of the cell (0 between calls)
    int32_t *lcntk;      // [B][cells][VL_KC] the same count per CHUNK of the index range (single-CTA per-cell kernel only)
    int32_t *lfill;      // [B][cells]  slot counter / collected-candidate counter
    int32_t *lthr;       // [B][cells]  largest kept point index of the cell (INT_MAX: keep all)
    int32_t *lhoff;      // [B][cells]  offset of the cell's candidate list in hot_idx, -1 if the cell is not over-full
    int32_t *lhot;       // [B][cells]  compact list of over-full cells
    int32_t *lstat;      // [B][4]      over-full cells, total candidates, cells CTAs finished, spare
};

int vox_make_plan(const gnd_voxel_params *p, int32_t batch, int64_t n_points, VoxPlan *plan);
size_t vox_carve(const VoxPlan &plan, int32_t batch, int64_t n_points, void *ws, VoxWs *out, bool with_sorted = false);
// Fused pipeline only: the same voxel ids / coors / num_points / voxel_count and, per cell, the same SET of kept points
// (cell-ordered `sorted` + CSR `offc`) as vox_run, without ranking every point in input order; the per-voxel point index
// lists (`point_list`) and their offsets are NOT produced (only point_offsets[M] = the number of kept points).
int vox_run_lean(const gnd_voxel_params *p, const VoxPlan &plan, const VoxWs &ws, const float *points,
                 const int32_t *counts, int32_t batch, int64_t n_points, int32_t *coors, int32_t *num_points,
                 int32_t *voxel_count, int32_t *point_offsets, cudaStream_t stream);
int vox_lean_init(const VoxPlan &plan, const VoxWs &ws, int32_t batch);
int vox_run(const gnd_voxel_params *p, const VoxPlan &plan, const VoxWs &ws, const float *points,
            const int32_t *counts, int32_t batch, int64_t n_points, int32_t *coors, int32_t *num_points,
            int32_t *voxel_count, int32_t *point_offsets, int32_t *point_list, float *voxels,
            int32_t *point_slot, cudaStream_t stream);

}  // namespace gnd
