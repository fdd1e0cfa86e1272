// umma_shift_probe.cu -- hardware probe (not product code): does a tcgen05.mma shared-memory descriptor
// (K-major, SWIZZLE_128B, SBO 1024) read the right rows when its start address is offset by a number of
// 128-byte rows that is NOT a multiple of 8 from the 1024-byte-aligned address the TMA box was written to?
// The halo-reuse conv (one TMA box per channel block, nine shifted operand views) depends on the answer.
// mode 0: base_offset field = 0;  mode 1: base_offset = (start >> 7) & 7.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o umma_shift_probe umma_shift_probe.cu -lcuda
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

#ifndef RB
#define RB 128
#endif
#ifndef DSTK
#define DSTK 0
#endif
#ifndef NBV
#define NBV 64
#endif
constexpr int ROWS = 192, KC = RB / 2, NB = NBV, NSHIFT = 40;
constexpr int KSTEPS = KC / 16;
constexpr uint64_t LAYOUT = (RB == 128) ? 2 : 4;   // SWIZZLE_128B / SWIZZLE_64B
constexpr uint32_t SBO = 8 * RB;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void tc_mma_f16(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tc_ld_32x32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, int mode) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(SBO >> 4) << 32;
    d |= (uint64_t)1 << 46;
    if (mode == 1) d |= (uint64_t)((saddr >> 7) & 7) << 49;
    d |= LAYOUT << 61;
    return d;
}

__global__ void __launch_bounds__(128, 1) probe(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, float *out, long long *clk) {
    extern __shared__ uint8_t raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
    // [0, 1024) slack | A: ROWS*128 B | slack 8 KB | B: NB*128 | barriers
    uint8_t *sA = smem + 1024 + DSTK * RB;   // TMA destination: DSTK rows past a 1024-byte boundary
    uint8_t *sB = sA + ROWS * RB + 8192;
    uint64_t *bars = reinterpret_cast<uint64_t *>(sB + NB * RB);
    uint32_t *tslot = reinterpret_cast<uint32_t *>(bars + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar_full = smem_u32(bars), bar_done = smem_u32(bars + 1);
    if (threadIdx.x == 0) {
        mbar_init(bar_full, 1); mbar_init(bar_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tslot)), "r"((uint32_t)(NB < 64 ? 64 : NB)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tslot;
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar_full, ROWS * RB + NB * RB);
        tma_load_2d(smem_u32(sA), &tm_a, bar_full, 0, 0);
        tma_load_2d(smem_u32(sB), &tm_b, bar_full, 0, 0);
    }
    mbar_wait(bar_full, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t idesc = (1u << 4) | ((uint32_t)(NB >> 3) << 17) | ((128u >> 4) << 24);
    uint32_t ph = 0;
    for (int mode = 0; mode < 2; ++mode)
        for (int s = 0; s < NSHIFT; ++s) {
            if (threadIdx.x == 0) {
                const uint32_t a0 = smem_u32(sA) + (uint32_t)s * RB, b0 = smem_u32(sB);
                for (int k = 0; k < KSTEPS; ++k) tc_mma_f16(tmem, umma_desc(a0 + 32 * k, mode), umma_desc(b0 + 32 * k, 0), idesc, k != 0);
                tc_commit(bar_done);
            }
            mbar_wait(bar_done, ph); ph ^= 1;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int row = warp * 32 + lane;
            for (int j = 0; j < NB / 32; ++j) {
                uint32_t v[32];
                tc_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + 32 * j, v);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                for (int i = 0; i < 32; ++i) out[(((size_t)mode * NSHIFT + s) * 128 + row) * NB + 32 * j + i] = __uint_as_float(v[i]);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncthreads();
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
    // timing: NREP back-to-back MMAs per start shift (same accumulator), clocks per MMA
    for (int s = 0; s < NSHIFT; ++s) {
        long long t0 = clock64();
        if (threadIdx.x == 0) {
            const uint32_t a0 = smem_u32(sA) + (uint32_t)s * RB, b0 = smem_u32(sB);
            for (int r = 0; r < 64; ++r)
                for (int k = 0; k < KSTEPS; ++k) tc_mma_f16(tmem, umma_desc(a0 + 32 * k, 0), umma_desc(b0 + 32 * k, 0), idesc, 1);
            tc_commit(bar_done);
        }
        mbar_wait(bar_done, ph); ph ^= 1;
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        long long t1 = clock64();
        if (threadIdx.x == 0) clk[s] = (t1 - t0);
        __syncthreads();
    }
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)(NB < 64 ? 64 : NB)) : "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    void *fp = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
    EncodeTiledFn enc = (EncodeTiledFn)fp;
    std::vector<__half> hA((size_t)ROWS * KC), hB((size_t)NB * KC);
    std::vector<float> fA(hA.size()), fB(hB.size());
    srand(1);
    for (size_t i = 0; i < hA.size(); ++i) { fA[i] = (float)(rand() % 17 - 8); hA[i] = __float2half(fA[i]); }
    for (size_t i = 0; i < hB.size(); ++i) { fB[i] = (float)(rand() % 9 - 4); hB[i] = __float2half(fB[i]); }
    __half *dA, *dB; float *dO;
    const size_t on = (size_t)2 * NSHIFT * 128 * NB;
    CK(cudaMalloc(&dA, hA.size() * 2)); CK(cudaMalloc(&dB, hB.size() * 2)); CK(cudaMalloc(&dO, on * 4));
    CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
    CUtensorMap ta, tb;
    cuuint64_t da[2] = {KC, ROWS}, sa[1] = {KC * 2}; cuuint32_t ba[2] = {KC, ROWS}, es[2] = {1, 1};
    cuuint64_t db[2] = {KC, NB}; cuuint32_t bb[2] = {KC, NB};
    if (enc(&ta, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, dA, da, sa, ba, es, CU_TENSOR_MAP_INTERLEAVE_NONE, (RB == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B), CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) { printf("encode A failed\n"); return 1; }
    if (enc(&tb, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, dB, db, sa, bb, es, CU_TENSOR_MAP_INTERLEAVE_NONE, (RB == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B), CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) { printf("encode B failed\n"); return 1; }
    const int smem = 1024 + 1024 + ROWS * RB + 8192 + NB * RB + 64;
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    long long *dC; CK(cudaMalloc(&dC, NSHIFT * 8));
    probe<<<1, 128, smem>>>(ta, tb, dO, dC);
    CK(cudaDeviceSynchronize());
    std::vector<float> hO(on);
    CK(cudaMemcpy(hO.data(), dO, on * 4, cudaMemcpyDeviceToHost));
    for (int mode = 0; mode < 2; ++mode) {
        printf("mode %d (base_offset %s): ", mode, mode ? "=(addr>>7)&7" : "=0");
        for (int s = 0; s < NSHIFT; ++s) {
            long bad = 0;
            for (int r = 0; r < 128; ++r)
                for (int n = 0; n < NB; ++n) {
                    float ref = 0;
                    for (int k = 0; k < KC; ++k) ref += fA[(size_t)(s + r) * KC + k] * fB[(size_t)n * KC + k];
                    if (hO[(((size_t)mode * NSHIFT + s) * 128 + r) * NB + n] != ref) ++bad;
                }
            printf("%s", bad ? "x" : ".");
        }
        printf("\n");
    }
    long long hc[NSHIFT]; CK(cudaMemcpy(hc, dC, NSHIFT * 8, cudaMemcpyDeviceToHost));
    printf("clk per MMA (N=%d, %d MMAs) by row shift:", NB, 64 * KSTEPS);
    for (int s = 0; s < NSHIFT; ++s) printf(" %.1f", (double)hc[s] / (64 * KSTEPS));
    printf("\n");
    return 0;
}
