/*
 * gndnet_b200.h -- C ABI of libgndnet_b200.so (sm_100a CUDA kernels for the GndNet
 * inference hot path: voxelize -> pillar-encode -> scatter -> segnet -> point labels).
 *
 * Boundary rules:
 *   - extern "C", plain pointers and sizes, no torch / C++ types in any signature;
 *   - every data pointer is a DEVICE pointer owned by the caller unless its name starts
 *     with h_ (host); every call takes an explicit cudaStream_t (passed as void*);
 *   - calls only enqueue work: no allocation, no synchronisation, CUDA-graph capturable
 *     (gnd_create / gnd_load_weights / gnd_bind_workspace are the set-up exceptions);
 *   - every call returns an int status (GND_OK or a negative gnd_status); the message
 *     of the last failure on the calling thread is gnd_last_error_string().
 *
 * Each entry point names the reference interface it replaces (paths relative to the
 * pem-gini/GndNet repo root).  The reference has no FFI of its own -- it is pure Python
 * (numba + torch) -- so the reference-side binding is a ctypes stub; see INTEGRATION.md.
 */
#ifndef GNDNET_B200_H
#define GNDNET_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define GND_API __attribute__((visibility("default")))
#else
#define GND_API
#endif

typedef enum gnd_status {
    GND_OK = 0,
    GND_ERR_INVALID = -1,      /* bad argument */
    GND_ERR_WORKSPACE = -2,    /* workspace missing or too small */
    GND_ERR_CUDA = -3,         /* a CUDA runtime / driver call failed */
    GND_ERR_UNSUPPORTED = -4,  /* outside the supported envelope (see DESIGN.md limits) */
    GND_ERR_STATE = -5         /* weights not loaded / workspace not bound */
} gnd_status;

/* arithmetic used by the segnet convolutions */
typedef enum gnd_precision {
    GND_PREC_FP32 = 0,   /* CUDA-core fp32 implicit GEMM (strict mode, 1e-3 m bar)          */
    GND_PREC_MIXED = 1,  /* tcgen05: split-fp16 (hi+lo, 3 products) encoder, fp16 decoder    */
    GND_PREC_SPLIT = 2   /* tcgen05: split-fp16 everywhere (fp32-class accuracy)             */
} gnd_precision;

GND_API const char *gnd_last_error_string(void);
GND_API int gnd_version(void);

/* ------------------------------------------------------------------------------------------
 * Voxelizer.  Replaces utils/point_cloud_ops.py:107-167 points_to_voxel and the numba loops
 * :7-53 (_points_to_voxel_reverse_kernel) / :55-104 (_points_to_voxel_kernel), batched over
 * independent scans, optionally fused with the caller's z-shift (predict_ground.py:104-107).
 * ---------------------------------------------------------------------------------------- */
typedef struct gnd_voxel_params {
    float voxel_size[3];     /* xyz, already in the points' dtype (float32), :138-139        */
    float coors_range[6];    /* xyzxyz min/max, float32, :140-141                             */
    int32_t max_points;      /* per voxel, :50                                                */
    int32_t max_voxels;      /* per scan, :44-45 (break on overflow)                          */
    int32_t reverse_index;   /* 1: coors are (z,y,x) (:38), 0: (x,y,z) (:86)                  */
    int32_t num_features;    /* columns of the point array (row stride in floats), >= 3       */
    int32_t apply_z_shift;   /* 1: use z + z_shift (float32 add) as the point's z everywhere  */
    float z_shift;           /* cfg.lidar_height                                              */
} gnd_voxel_params;

/* grid = round((hi - lo) / vs) in float32 (:24-27, :142-143); h_grid_xyz[3] on the host. */
GND_API int gnd_voxel_grid_size(const gnd_voxel_params *p, int32_t *h_grid_xyz);

/* Bytes of scratch gnd_points_to_voxel needs for `batch` scans of at most n_points each. */
GND_API int gnd_voxel_workspace_bytes(const gnd_voxel_params *p, int32_t batch, int64_t n_points,
                                      size_t *h_bytes);

/*
 * points        [batch][n_points][num_features] float32
 * counts        [batch] int32 real points per scan (<= n_points), or NULL for "all n_points"
 * coors         [batch][max_voxels][3] int32     out: first voxel_count[b] rows valid
 * num_points    [batch][max_voxels]    int32     out
 * voxel_count   [batch]                int32     out: M per scan (the reference's voxel_num)
 * point_offsets [batch][max_voxels+1]  int32     out: CSR offsets into point_list (row M = kept points)
 * point_list    [batch][n_points]      int32     out: kept point indices, voxel-major, input order inside
 * voxels        [batch][max_voxels][max_points][num_features] float32, ZERO-FILLED by the caller
 *               (:149-150), or NULL: the fused path never materialises it
 * point_slot    [batch][n_points] int32, or NULL: voxel*max_points + slot the point went to, -1 if dropped
 * NaN coordinates are undefined behaviour in the reference; here such points are dropped.
 */
GND_API int gnd_points_to_voxel(const gnd_voxel_params *p, const float *points, const int32_t *counts,
                                int32_t batch, int64_t n_points, void *workspace, size_t workspace_bytes,
                                int32_t *coors, int32_t *num_points, int32_t *voxel_count,
                                int32_t *point_offsets, int32_t *point_list, float *voxels,
                                int32_t *point_slot, void *stream);

/* predict_ground.py:104-107 _shift_cloud / utils/utils.py:137-139: cloud[:, 2] += height in place. */
GND_API int gnd_shift_cloud(float *points, int64_t n_rows, int32_t num_features, float height, void *stream);

/* ------------------------------------------------------------------------------------------
 * Per-point labelling.  Replaces utils/utils.py:198-220 segment_cloud and :222-248
 * segment_cloud_noground.
 * ---------------------------------------------------------------------------------------- */
typedef struct gnd_segment_params {
    double grid_lo[2];       /* grid_size[0], grid_size[1] as float64 (:204)                  */
    double voxel_size;       /* Python scalar -> float64 division (:205)                      */
    double threshold;        /* :214                                                          */
    int32_t map_nx, map_ny;  /* elevation_map.shape: the map is indexed [x, y] (:214)         */
    int64_t map_sx, map_sy;  /* element strides of that [x, y] view (callers pass pred.T)      */
    int64_t map_batch_stride;/* elements between consecutive scans' maps                      */
    int32_t num_features;    /* row stride of points in floats, >= 3                          */
    int32_t apply_z_shift;   /* 1: compare z + z_shift (fused path: points are the raw scan)  */
    float z_shift;
    int32_t mutate_points;   /* 1: write the shifted x,y back like the reference does (:204)  */
} gnd_segment_params;

/*
 * points      [batch][n_points][num_features] float32 (columns 0,1 rewritten if mutate_points)
 * counts      [batch] int32 or NULL
 * elevation   float32 height maps
 * labels_f64  [batch][n_points] float64 in {1,0,-1} (the reference's return dtype), or NULL
 * labels_i8   [batch][n_points] int8, or NULL
 */
GND_API int gnd_segment_cloud(const gnd_segment_params *p, float *points, const int32_t *counts,
                              int32_t batch, int64_t n_points, const float *elevation,
                              double *labels_f64, int8_t *labels_i8, void *stream);

GND_API int gnd_compact_workspace_bytes(int64_t n_points, size_t *h_bytes);
/*
 * segment_cloud_noground's obstacle compaction (:242-243, :248), one scan: rows of points_obstacle
 * [n_points][obstacle_features] whose label is 1 are moved to the front in input order (stable).
 * out_obstacle may alias nothing (separate output buffer); obstacle_count is a device int32.
 */
GND_API int gnd_compact_obstacles(const int8_t *labels_i8, const float *points_obstacle, int64_t n_points,
                                  int32_t obstacle_features, void *workspace, size_t workspace_bytes,
                                  float *out_obstacle, int32_t *obstacle_count, void *stream);

/*
 * gnd_compact_obstacles for callers that have no way to provide the workspace.  (No counterpart in the reference:
 * numpy allocates wherever it likes.  This exists for the deployment side only.)
 *
 * Some users of this library are handed device buffers but cannot create one: a callback running inside another
 * component's memory pool, a plug-in in a process whose GPU memory was carved up completely at start-up, a real-time
 * loop in which cudaMalloc is banned because it stalls the whole device.  For them the block counters the compaction
 * needs sit in a fixed array inside the library image (16 KB, present on every GPU that loads the library, for the
 * life of the process).  The call itself allocates and frees nothing, on the host or the device.
 *
 * Same arguments and results as gnd_compact_obstacles minus workspace / workspace_bytes.  Differences to know about:
 *   - n_points may not exceed GND_COMPACT_STATIC_MAX_POINTS; a larger cloud is refused with GND_ERR_UNSUPPORTED (split
 *     it, or use gnd_compact_obstacles);
 *   - the first call on each GPU looks up the array's device address (a one-off runtime query, made with that GPU
 *     current).  Make that first call during set-up, outside any stream capture;
 *   - every call on a GPU uses the SAME array.  Two calls whose kernels can overlap on that GPU (different streams,
 *     different threads) corrupt each other's counters: run them on one stream or order them with events.  Calls on
 *     different GPUs are independent.  The array keeps the previous call's values; the call does not depend on them.
 */
#define GND_COMPACT_STATIC_MAX_POINTS 4194304   /* 4 Mi points per call */
GND_API int gnd_compact_obstacles_static(const int8_t *labels_i8, const float *points_obstacle, int64_t n_points,
                                         int32_t obstacle_features, float *out_obstacle, int32_t *obstacle_count,
                                         void *stream);

/* ------------------------------------------------------------------------------------------
 * Model context.  Replaces model.py:13-42 GroundEstimatorNet (+ modules/pointpillars.py,
 * modules/segnet.py) in eval mode, and predict_ground.py:110-121,165-168 InferGround +
 * segmentation as one enqueue.
 * ---------------------------------------------------------------------------------------- */
typedef struct gnd_ctx gnd_ctx;

typedef struct gnd_config {
    float pc_range[6];          /* cfg.pc_range                                               */
    float voxel_size[3];        /* cfg.voxel_size                                             */
    double pc_range_f64[6];     /* same, as the Python floats the model's offsets are made of */
    double voxel_size_f64[3];
    double grid_range[4];       /* cfg.grid_range (segmentation frame)                        */
    float lidar_height;         /* cfg.lidar_height                                           */
    int32_t max_points_voxel;   /* cfg.max_points_voxel                                       */
    int32_t max_voxels;         /* cfg.max_voxels                                             */
    int32_t input_features;     /* cfg.input_features (3 or 4)                                */
    int32_t use_norm;           /* cfg.use_norm: Linear(no bias)+BatchNorm1d vs Linear(bias)   */
    int32_t with_distance;      /* cfg.with_distance                                          */
    int32_t precision;          /* gnd_precision                                              */
    int32_t bn_batch_stats;     /* 0: BatchNorm2d with running statistics (eval mode, ros_node.py:227), folded into
                                   the convolutions.  1: the module's TRAIN-mode arithmetic -- per-channel mean and
                                   biased variance of the batch being processed -- which is what predict_ground.py
                                   computes because it never calls .eval() (GND_PREC_FP32 only, use_norm = 0)   */
    int32_t n_vfe;              /* len(cfg.vfe_filters), 1..GND_MAX_VFE; 0 is read as 1 layer of 64 (every shipped config) */
    int32_t vfe_filters[4];     /* cfg.vfe_filters (modules/pointpillars.py:96-107): the last one must be 64 (the scatter
                                   and the segnet take 64 channels), the others even and <= 128; non-last PFN layers have
                                   out / 2 units and hand concat(x, max over the pillar) to the next layer              */
} gnd_config;
#define GND_MAX_VFE 4

GND_API int gnd_create(const gnd_config *cfg, int32_t device, gnd_ctx **out);
GND_API void gnd_destroy(gnd_ctx *ctx);

/* One state_dict entry: the reference's key, float32 HOST data, element count. */
typedef struct gnd_weight {
    const char *name;
    const float *h_data;
    int64_t numel;
} gnd_weight;

/* model.load_state_dict(checkpoint['state_dict']) (predict_ground.py:196): folds BatchNorm
 * (eval mode), repacks to the kernels' layouts and uploads.  Synchronous set-up call. */
GND_API int gnd_load_weights(gnd_ctx *ctx, const gnd_weight *table, int32_t n_entries);

GND_API int gnd_workspace_bytes(const gnd_ctx *ctx, int32_t batch, int64_t n_points, size_t *h_bytes);
/* Carves the caller's device buffer for the CAPACITY (batch, n_points) and builds the TMA descriptors.  Every later
 * stage call may pass any batch <= the bound batch and any n_points <= the bound n_points (the row count of the
 * caller's [batch][n_points][F] buffer); clouds of different sizes -- ros_node.py's NaN-filtered clouds (:263),
 * predict_ground.py's per-file sizes -- therefore run without re-binding.  Rows beyond counts[b] are ignored.
 *
 * Threading contract (the reference drives the path from a daemon thread while a callback thread segments the
 * latest output, ros_node.py:51-60,96-126,270-274): a gnd_ctx owns mutable device scratch (work-queue counters,
 * voxelizer tables, activations), so AT MOST ONE stage call per ctx may be in flight at a time -- serialise calls on
 * one ctx (gndnet_b200.model takes a lock) and enqueue them on one stream, or use one ctx per thread / stream.
 * The ctx-free entry points (gnd_points_to_voxel, gnd_segment_cloud, gnd_compact_obstacles, gnd_unpack_cloud, ...)
 * keep all state in caller-owned buffers and are re-entrant.  Two ctxs may run concurrently on different streams:
 * the convolution kernel takes a whole SM per CTA (all 512 TMEM columns), so kernels of different ctxs never
 * co-reside on an SM and cannot starve each other's TMEM allocation. */
GND_API int gnd_bind_workspace(gnd_ctx *ctx, void *workspace, size_t bytes, int32_t batch, int64_t n_points);

GND_API int gnd_grid_shape(const gnd_ctx *ctx, int32_t *h_nz_ny_nx);

/* Stage 1 (fused path): voxelize `batch` raw scans (z-shift fused) into the bound workspace. */
GND_API int gnd_voxelize(gnd_ctx *ctx, const float *points, const int32_t *counts, int32_t batch,
                         int64_t n_points, void *stream);
/* Stage 2: PillarFeatureNet + PointPillarsScatter (pointpillars.py:115-146, :168-199) from the
 * workspace's pillar lists straight into the BEV canvas. */
GND_API int gnd_pfn_scatter(gnd_ctx *ctx, const float *points, int32_t batch, int64_t n_points, void *stream);
/* Stage 2, drop-in tensors: forward(voxels [M,P,F], coors [M,4] float (b,z,y,x), num_points [M] float)
 * as model.py:36-42 receives them. */
GND_API int gnd_pfn_scatter_dense(gnd_ctx *ctx, const float *voxels, const float *coors4,
                                  const float *num_points, int64_t n_pillars, int32_t batch, void *stream);
/* Stage 3: segnetGndEst.forward (segnet.py:133-142) canvas -> height [batch][ny][nx] float32. */
GND_API int gnd_segnet(gnd_ctx *ctx, int32_t batch, float *height, void *stream);
/* Profiling aid (bench.py roofline): gnd_segnet with CUDA events around each of the 8 convolution
 * launches; SYNCHRONOUS; h_conv_ms[8] per-layer device milliseconds, *h_total_ms the whole stage. */
GND_API int gnd_segnet_profile(gnd_ctx *ctx, int32_t batch, float *height, float *h_conv_ms, float *h_total_ms,
                               void *stream);
/* Stage 4: labels for the raw scans against `height` (z-shift fused; grid_range frame). */
GND_API int gnd_segment(gnd_ctx *ctx, const float *points, const int32_t *counts, int32_t batch,
                        int64_t n_points, const float *height, double threshold, int8_t *labels_i8,
                        double *labels_f64, void *stream);
/* Stages 1-4 back to back on `stream`.  The voxelization inside is set-exact (see gnd_set_option "vox_path"): pillar
 * ids, coors, num_points and every pillar's kept points equal the reference's; the per-pillar point index lists in
 * slot order ("point_list") and their offsets ("point_offsets" rows [0, M)) are only materialised by gnd_voxelize /
 * gnd_points_to_voxel.  After a set-exact run the workspace's "point_offsets" holds ONE valid entry per scan, row
 * M = voxel_count[b] (the number of kept points); the rows below it are stale or uninitialised. */
GND_API int gnd_infer(gnd_ctx *ctx, const float *points, const int32_t *counts, int32_t batch,
                      int64_t n_points, double threshold, float *height, int8_t *labels_i8,
                      double *labels_f64, void *stream);

/* ------------------------------------------------------------------------------------------
 * Streaming-node I/O either side of the path.
 * ---------------------------------------------------------------------------------------- */
typedef struct gnd_unpack_params {
    int32_t point_step;      /* msg.point_step: x, y, z float32 at byte offsets 0, 4, 8 of each row
                                (utils/utils.py:179-180)                                        */
    int32_t has_transform;   /* 1: apply `transform` (utils/transform.py:148-165 transformCloud) */
    double transform[16];    /* row-major 4x4 (transform.py:107-111, :99-105), float64           */
    int32_t drop_nan;        /* 1: drop rows with a NaN after the transform (ros_node.py:263)    */
    int32_t out_features;    /* columns of the output rows, >= 3 (extra columns are zero)        */
} gnd_unpack_params;

GND_API int gnd_unpack_workspace_bytes(int64_t n_points, size_t *h_bytes);
/*
 * Replaces ros_node.py:242-263: cloud_msg_to_numpy(msg)["xyz"] (utils/utils.py:148-195), transformCloud
 * (utils/transform.py:148-165) and the NaN filter, on the device.
 * msg_data   [n_points][point_step] bytes of sensor_msgs/PointCloud2.data (device)
 * out_points [n_points][out_features] float32; the first *out_count rows are valid, input order kept
 * out_count  device int32 -- usable directly as gnd_infer's `counts`
 */
GND_API int gnd_unpack_cloud(const gnd_unpack_params *p, const uint8_t *msg_data, int64_t n_points,
                             void *workspace, size_t workspace_bytes, float *out_points,
                             int32_t *out_count, void *stream);
/*
 * Replaces utils/ros_utils.py:171-191 np2ros_pub_2 up to the msgify call: the 16-byte wire rows
 * (x, y, z, intensity = label as float32) of the segmented cloud.  One of labels_i8 / labels_f64.
 * (utils/ros_utils.py:210-248 array_to_pointcloud2's 12-byte rows are what gnd_compact_obstacles
 * writes for obstacle_features == 3.)
 */
GND_API int gnd_pack_segmented(const float *points, int64_t n_points, int32_t num_features,
                               const int8_t *labels_i8, const double *labels_f64, float *out_rows,
                               void *stream);

/*
 * Evaluation reductions.  Replaces evaluate_SemanticKITTI.py:94-111 (get_GndSeg,
 * remove_outliers) and :198-204: counts4 (device uint64[4]) = intersection, union, predicted-ground and
 * GT-ground counts over the points with pred >= 0 and a usable SemanticKITTI label; iou = c[0]/c[1],
 * prec = c[0]/c[2], recall = c[0]/c[3].  h_ground_classes is a HOST list (<= 16) of ground class ids.
 */
GND_API int gnd_eval_ground_counts(const int8_t *labels_i8, const double *labels_f64,
                                   const uint32_t *sem_label, int64_t n_points,
                                   const uint32_t *h_ground_classes, int32_t n_classes,
                                   uint64_t *counts4, void *stream);
/* evaluate_SemanticKITTI.py:224-226: sums2 (device double[2]) = sum(square(target - pred.T) * mask), sum(mask).
 * pred [ny][nx] float32 (the model output); target, mask [nx][ny] float64 (lidar_to_heightmap / lidar_to_img). */
GND_API int gnd_eval_masked_sqerr(const float *pred, const double *target, const double *mask,
                                  int32_t nx, int32_t ny, double *sums2, void *stream);

/*
 * Ground-truth rasterisation of the evaluation tooling (evaluate_SemanticKITTI.py:118-127).
 * Maps are [nx][ny] float64 with nx = int((gs[2]-gs[0])/vs), ny = int((gs[3]-gs[1])/vs); h_grid_size4 is a HOST array
 * (cfg.grid_range as float64).  points [n][num_features] float32 are not modified.
 * gnd_lidar_to_img        replaces utils/utils.py:250-270 (lidar_to_img).
 * gnd_lidar_to_heightmap  replaces utils/utils.py:273-295 (lidar_to_heightmap), input-order semantics and the
 *                         reference's count+1 divisor / one-past-the-end slot write included (see csrc/raster.cu).
 */
GND_API int gnd_raster_workspace_bytes(int64_t n_points, const double *h_grid_size4, double voxel_size,
                                       int32_t max_points, size_t *h_bytes);
GND_API int gnd_lidar_to_img(const float *points, int64_t n_points, int32_t num_features,
                             const double *h_grid_size4, double voxel_size, double fill, double *img,
                             void *stream);
GND_API int gnd_lidar_to_heightmap(const float *points, int64_t n_points, int32_t num_features,
                                   const double *h_grid_size4, double voxel_size, int32_t max_points,
                                   void *workspace, size_t workspace_bytes, double *heightmap, void *stream);

/* Test / profiling hook: device pointer + bytes of a named workspace buffer ("canvas", "coors",
 * "num_points", "voxel_count", "point_offsets", "point_list", "act:<stage>", ...). */
GND_API int gnd_get_buffer(gnd_ctx *ctx, const char *name, void **ptr, size_t *bytes);
/* Number of kernels the last gnd_* call on this ctx enqueued (bench.py's gpu_launches). */
GND_API int gnd_last_launch_count(const gnd_ctx *ctx);
/* Per-ctx switches (tests, A/B timing; never read from the environment at launch time):
 *   "conv_path"      0 = CTA-pair halo kernel where the shape fits (default), 1 = single-CTA tap kernel (the fallback)
 *   "conv_geometry"  0 = automatic, 1 = throughput geometry, 2 = small-batch geometry of the pair kernel
 *   "l2_order"       0 (default) = every segnet kernel walks the scans of a batch forward, 1 = consecutive kernels walk
 *                    them in alternating directions, so that a kernel starts on the scans its producer wrote last
 *                    (A/B switch: no step-time difference was measured).  Same results bit for bit.
 *   "pfn_path"       fused pillar encoder: 0 = automatic (default; today the CUDA-core kernel: 0.21 ms against 0.24 ms per
 *                    64 scans), 1 = CUDA-core kernel, 2 = mma.sync TF32x3 kernel (cfg.with_distance == 0 only; else the
 *                    CUDA-core kernel runs)
 *   "vox_path"       gnd_infer only: 0 = automatic (default): set-exact voxelization -- same coors / num_points /
 *                    voxel_count and the same kept points per pillar, but no slot-ordered point_list / offsets -- unless
 *                    the cloud is dense (more than 32 points per cell on average) and the batch small (< 12 M points
 *                    in all), 1 = always the order-exact kernels gnd_voxelize uses (point_list filled), 2 = always
 *                    the set-exact kernels */
GND_API int gnd_set_option(gnd_ctx *ctx, const char *name, int32_t value);
/* Asynchronous per-convolution timing for bench.py's roofline: after _begin, each of the next `slots` gnd_segnet /
 * gnd_infer calls records library-owned CUDA events around its 8 convolution launches (nothing is read back, the
 * stream is not synchronised); _end waits for them, writes h_conv_ms[recorded][8] and h_segnet_ms[recorded] (device
 * milliseconds) and frees the events.  h_stage_ms (may be NULL) [recorded][4]: for slots filled by gnd_infer, the device
 * time of its four stages as that call ran them -- voxelize, pillar-encode + scatter, segnet, labels -- else -1. */
GND_API int gnd_conv_timing_begin(gnd_ctx *ctx, int32_t slots);
GND_API int gnd_conv_timing_end(gnd_ctx *ctx, float *h_conv_ms, float *h_segnet_ms, float *h_stage_ms, int32_t *h_recorded);

#ifdef __cplusplus
}
#endif
#endif /* GNDNET_B200_H */
