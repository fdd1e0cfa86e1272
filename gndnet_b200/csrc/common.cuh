// common.cuh -- shared helpers for libgndnet_b200 (error reporting, launch accounting, scans).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/gndnet_b200.h"

namespace gnd {

// thread-local last-error text (gnd_last_error_string)
char *err_buf();
int fail(int code, const char *fmt, ...);
// kernels enqueued since the last reset (gnd_last_launch_count)
extern thread_local int g_launches;
inline void count_launch(int n = 1) { g_launches += n; }

#define GND_CUDA_CHECK(expr)                                                                      \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess)                                                                    \
            return ::gnd::fail(GND_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                               __FILE__, __LINE__);                                               \
    } while (0)

#define GND_LAUNCH_CHECK(name)                                                                    \
    do {                                                                                          \
        cudaError_t _e = cudaPeekAtLastError();                                                   \
        if (_e != cudaSuccess)                                                                    \
            return ::gnd::fail(GND_ERR_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(_e)); \
        ::gnd::count_launch();                                                                    \
    } while (0)

// Debug / tuning knobs from the environment, read ONCE per process (first use) -- never at launch time, so a
// running service cannot change kernel geometry under a captured CUDA graph or between two threads' calls.
//   GNDNET_DEBUG            print per-layer geometry when tensor maps are built
//   GNDNET_VOX_SERIAL       force the one-ranking-warp voxel kernel        GNDNET_VOX_K / GNDNET_VOX_G  chunking overrides
//   GNDNET_PFN_GSHIFT       pillar-encode work-unit size (log2 cells)       GNDNET_H2_SMALL_NT / GNDNET_H2_MINSB  conv ring
struct Tuning {
    int debug, vox_serial, vox_k, vox_g, pfn_gshift, h2_small_nt, h2_minsb;
};
const Tuning &tuning();

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// carve a caller-owned buffer
struct Carver {
    char *base;
    size_t off;
    explicit Carver(void *p) : base((char *)p), off(0) {}
    template <typename T>
    T *take(size_t n) {
        off = align_up(off, 256);
        T *r = base ? (T *)(base + off) : nullptr;
        off += n * sizeof(T);
        return r;
    }
    size_t bytes() const { return align_up(off, 256); }
};

// ---- device helpers -------------------------------------------------------------------------
__device__ __forceinline__ unsigned lanemask_lt() {
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

__device__ __forceinline__ int warp_inclusive_scan(int v) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
    }
    return v;
}

// Exclusive scan of one int per thread over the whole block (blockDim.x multiple of 32, <= 1024).
// s_warp must hold 33 ints.  Returns the exclusive prefix; *total = block sum.  Ends with a barrier,
// so s_warp may be reused right after.
__device__ __forceinline__ int block_exclusive_scan(int v, int *s_warp, int *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    int inc = warp_inclusive_scan(v);
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int w = (lane < nwarp) ? s_warp[lane] : 0;
        int winc = warp_inclusive_scan(w);
        s_warp[lane] = winc - w;          // exclusive prefix of warp sums
        if (lane == 31) s_warp[32] = winc;
    }
    __syncthreads();
    int res = s_warp[warp] + inc - v;
    *total = s_warp[32];
    __syncthreads();
    return res;
}

}  // namespace gnd
