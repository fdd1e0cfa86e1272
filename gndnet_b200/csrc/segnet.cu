// segnet.cu -- segnetGndEst.forward (modules/segnet.py:133-142): schedule of the encoder/decoder,
// the pooling / unpooling kernels, the regression head, and the strict-fp32 CUDA-core convolution.
//
//   down1 = conv(64->128), conv(128->128), maxpool 2x2 + indices          (segnet.py:49-61)
//   down2 = conv(128->256), conv(256->256), maxpool 2x2 + indices
//   up2   = unpool(indices2), conv(256->256), conv(256->128)               (segnet.py:81-92)
//   up1   = unpool(indices1), conv(128->128), conv(128->64)
//   regressor conv(64->1), no BatchNorm / ReLU                             (segnet.py:131,140)
// every conv is 3x3, stride 1, zero pad 1, bias, BatchNorm2d(eval, eps 1e-5) folded at load, ReLU
// (segnet.py:11-44).  MaxPool2d(2,2) floors odd sizes and keeps the FIRST maximum in row-major window
// order; MaxUnpool2d writes the value back at that position of a zero tensor of the pre-pool shape.
// pool(down2) followed at once by unpool is a mask: keep the window's first maximum, zero the rest.
//
// All activations are NHWC.  The tcgen05 implicit-GEMM convolutions live in conv_tc.cu; this file
// holds the fp32 mode (GND_PREC_FP32), which is the 1e-3 m parity path.
#include "model.cuh"

namespace gnd {

// ---- element access for either activation format ------------------------------------------
__device__ __forceinline__ float act_load(const Act &a, size_t i) {
    if (a.f32) return a.f32[i];
    float v = __half2float(a.hi[i]);
    if (a.lo) v += __half2float(a.lo[i]);     // exact: hi and lo are adjacent 11-bit pieces
    return v;
}
__device__ __forceinline__ void act_store(const Act &a, size_t i, float v) {
    if (a.f32) { a.f32[i] = v; return; }
    __half h = __float2half_rn(v);
    a.hi[i] = h;
    if (a.lo) a.lo[i] = __float2half_rn(v - __half2float(h));
}

// first maximum of the 2x2 window in row-major order, strict '>' (ATen max_pool2d)
__device__ __forceinline__ int window_argmax(float v00, float v01, float v10, float v11, float *vmax) {
    int k = 0; float m = v00;
    if (v01 > m) { m = v01; k = 1; }
    if (v10 > m) { m = v10; k = 2; }
    if (v11 > m) { m = v11; k = 3; }
    *vmax = m;
    return k;
}

// MaxPool2d(2,2,return_indices) (segnet.py:54,60): in [B][H][W][C] -> out [B][H/2][W/2][C] + argmax 0..3
__global__ void __launch_bounds__(256) pool_kernel(const Act in, const Act out, uint8_t *idx, int B) {
    const int C = in.C, Ho = out.H, Wo = out.W;
    const size_t total = (size_t)B * Ho * Wo * C;
    for (size_t g = (size_t)blockIdx.x * 256 + threadIdx.x; g < total; g += (size_t)gridDim.x * 256) {
        const int c = (int)(g % C);
        size_t p = g / C;
        const int wo = (int)(p % Wo); p /= Wo;
        const int ho = (int)(p % Ho);
        const int b = (int)(p / Ho);
        const size_t r0 = (((size_t)b * in.H + 2 * ho) * in.W + 2 * wo) * C + c;
        const size_t r1 = r0 + (size_t)in.W * C;
        float m;
        int k = window_argmax(act_load(in, r0), act_load(in, r0 + C), act_load(in, r1), act_load(in, r1 + C), &m);
        act_store(out, g, m);
        idx[g] = (uint8_t)k;
    }
}

// pool + unpool back to the same shape (down2 -> up2): out[h][w] = in[h][w] if it is its window's first max else 0
__global__ void __launch_bounds__(256) pool_mask_kernel(const Act in, const Act out, int B) {
    const int C = in.C, H = in.H, W = in.W, Hp = H / 2, Wp = W / 2;
    const size_t total = (size_t)B * H * W * C;
    for (size_t g = (size_t)blockIdx.x * 256 + threadIdx.x; g < total; g += (size_t)gridDim.x * 256) {
        const int c = (int)(g % C);
        size_t p = g / C;
        const int w = (int)(p % W); p /= W;
        const int h = (int)(p % H);
        const int b = (int)(p / H);
        float v = 0.0f;
        if ((h >> 1) < Hp && (w >> 1) < Wp) {
            const size_t r0 = (((size_t)b * H + (h & ~1)) * W + (w & ~1)) * C + c;
            const size_t r1 = r0 + (size_t)W * C;
            float m;
            int k = window_argmax(act_load(in, r0), act_load(in, r0 + C), act_load(in, r1), act_load(in, r1 + C), &m);
            if (k == ((h & 1) * 2 + (w & 1))) v = m;
        }
        act_store(out, g, v);
    }
}

// MaxUnpool2d(2,2) to the recorded shape (segnet.py:84,89): in [B][Hp][Wp][C] + idx -> out [B][H][W][C]
__global__ void __launch_bounds__(256) unpool_kernel(const Act in, const uint8_t *idx, const Act out, int B) {
    const int C = out.C, H = out.H, W = out.W, Hp = in.H, Wp = in.W;
    const size_t total = (size_t)B * H * W * C;
    for (size_t g = (size_t)blockIdx.x * 256 + threadIdx.x; g < total; g += (size_t)gridDim.x * 256) {
        const int c = (int)(g % C);
        size_t p = g / C;
        const int w = (int)(p % W); p /= W;
        const int h = (int)(p % H);
        const int b = (int)(p / H);
        float v = 0.0f;
        if ((h >> 1) < Hp && (w >> 1) < Wp) {
            const size_t s = (((size_t)b * Hp + (h >> 1)) * Wp + (w >> 1)) * C + c;
            if (idx[s] == ((h & 1) * 2 + (w & 1))) v = act_load(in, s);
        }
        act_store(out, g, v);
    }
}

// ---- vectorised fp16 (hi [+ lo]) versions: 8 channels (one 16-byte vector per plane) per thread --------
struct H8 { float v[8]; };

__device__ __forceinline__ H8 load_h8(const Act &a, size_t i) {
    H8 r;
    const uint4 h = __ldg(reinterpret_cast<const uint4 *>(a.hi + i));
    const __half2 *hp = reinterpret_cast<const __half2 *>(&h);
#pragma unroll
    for (int k = 0; k < 4; ++k) { float2 f = __half22float2(hp[k]); r.v[2 * k] = f.x; r.v[2 * k + 1] = f.y; }
    if (a.lo) {
        const uint4 l = __ldg(reinterpret_cast<const uint4 *>(a.lo + i));
        const __half2 *lp = reinterpret_cast<const __half2 *>(&l);
#pragma unroll
        for (int k = 0; k < 4; ++k) { float2 f = __half22float2(lp[k]); r.v[2 * k] += f.x; r.v[2 * k + 1] += f.y; }
    }
    return r;
}

__device__ __forceinline__ void store_h8(const Act &a, size_t i, const H8 &x) {
    uint4 h, l;
    __half2 *hp = reinterpret_cast<__half2 *>(&h), *lp = reinterpret_cast<__half2 *>(&l);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        hp[k] = __floats2half2_rn(x.v[2 * k], x.v[2 * k + 1]);
        const float2 b = __half22float2(hp[k]);
        lp[k] = __floats2half2_rn(x.v[2 * k] - b.x, x.v[2 * k + 1] - b.y);
    }
    *reinterpret_cast<uint4 *>(a.hi + i) = h;
    if (a.lo) *reinterpret_cast<uint4 *>(a.lo + i) = l;
}

// one thread per (pooled pixel, 8 channels)
__global__ void __launch_bounds__(256) pool_h8_kernel(const Act in, const Act out, uint8_t *idx, int B) {
    const int C8 = in.C / 8, Ho = out.H, Wo = out.W;
    const size_t total = (size_t)B * Ho * Wo * C8;
    for (size_t g = (size_t)blockIdx.x * 256 + threadIdx.x; g < total; g += (size_t)gridDim.x * 256) {
        const int c = (int)(g % C8) * 8;
        size_t p = g / C8;
        const int wo = (int)(p % Wo); p /= Wo;
        const int ho = (int)(p % Ho);
        const int b = (int)(p / Ho);
        const size_t r0 = (((size_t)b * in.H + 2 * ho) * in.W + 2 * wo) * in.C + c;
        const size_t r1 = r0 + (size_t)in.W * in.C;
        const H8 a = load_h8(in, r0), bb = load_h8(in, r0 + in.C), cc = load_h8(in, r1), d = load_h8(in, r1 + in.C);
        H8 m;
        uint32_t klo = 0, khi = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int w = window_argmax(a.v[k], bb.v[k], cc.v[k], d.v[k], &m.v[k]);
            if (k < 4) klo |= (uint32_t)w << (8 * k); else khi |= (uint32_t)w << (8 * (k - 4));
        }
        const size_t o = g * 8;
        store_h8(out, o, m);
        *reinterpret_cast<uint2 *>(idx + o) = make_uint2(klo, khi);
    }
}

// one thread per (2x2 window incl. partial border windows, 8 channels)
__global__ void __launch_bounds__(256) pool_mask_h8_kernel(const Act in, const Act out, int B, int rev) {
    const int C8 = in.C / 8, H = in.H, W = in.W, Hp = H / 2, Wp = W / 2, Hc = (H + 1) / 2, Wc = (W + 1) / 2;
    const size_t total = (size_t)B * Hc * Wc * C8;
    H8 zero;
#pragma unroll
    for (int k = 0; k < 8; ++k) zero.v[k] = 0.0f;
    for (size_t g0 = (size_t)blockIdx.x * 256 + threadIdx.x; g0 < total; g0 += (size_t)gridDim.x * 256) {
        const size_t g = rev ? total - 1 - g0 : g0;        // last scan first: its rows are the ones the conv before left in L2
        const int c = (int)(g % C8) * 8;
        size_t p = g / C8;
        const int wc = (int)(p % Wc); p /= Wc;
        const int hc = (int)(p % Hc);
        const int b = (int)(p / Hc);
        const size_t r0 = (((size_t)b * H + 2 * hc) * W + 2 * wc) * in.C + c;
        const size_t r1 = r0 + (size_t)W * in.C;
        if (hc < Hp && wc < Wp) {
            const H8 a = load_h8(in, r0), bb = load_h8(in, r0 + in.C), cc = load_h8(in, r1), d = load_h8(in, r1 + in.C);
            H8 o0 = zero, o1 = zero, o2 = zero, o3 = zero;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float m;
                const int w = window_argmax(a.v[k], bb.v[k], cc.v[k], d.v[k], &m);
                if (w == 0) o0.v[k] = m; else if (w == 1) o1.v[k] = m; else if (w == 2) o2.v[k] = m; else o3.v[k] = m;
            }
            store_h8(out, r0, o0); store_h8(out, r0 + in.C, o1); store_h8(out, r1, o2); store_h8(out, r1 + in.C, o3);
        } else {                                   // floor mode: the odd last row / column is never pooled -> zeros
            const bool h1 = 2 * hc + 1 < H, w1 = 2 * wc + 1 < W;
            store_h8(out, r0, zero);
            if (w1) store_h8(out, r0 + in.C, zero);
            if (h1) store_h8(out, r1, zero);
            if (h1 && w1) store_h8(out, r1 + in.C, zero);
        }
    }
}

// one thread per (2x2 output window incl. partial border windows, 8 channels).  MaxUnpool2d only MOVES values
// (segnet.py:84,89), so the planes are handled as raw 16-bit patterns: the four argmax bytes of a word give two byte
// masks (bit 0 / bit 1 of every byte, spread by a multiply), their four boolean combinations select the window
// position, PRMT widens byte masks to half masks.  ~100 instructions per thread against ~520 through float
// conversions (ncu: the kernel was issue-bound at 3 TB/s).
__device__ __forceinline__ void unpool_plane(const __half *__restrict__ src, __half *__restrict__ dst, bool have, size_t s,
                                             size_t r0, size_t r1, int C, uint2 kk, bool w1, bool h1) {
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (have) v = __ldg(reinterpret_cast<const uint4 *>(src + s));
    const uint32_t vv[4] = {v.x, v.y, v.z, v.w};
    const uint32_t kw[2] = {kk.x, kk.y};
    uint32_t o[4][4];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint32_t m0 = (kw[h] & 0x01010101u) * 0xffu;            // 0xff in every byte whose argmax has bit 0 set
        const uint32_t m1 = ((kw[h] >> 1) & 0x01010101u) * 0xffu;     // ... bit 1
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            const uint32_t bm = ((w & 2) ? m1 : ~m1) & ((w & 1) ? m0 : ~m0);
            o[w][2 * h] = vv[2 * h] & __byte_perm(bm, 0u, 0x1100);
            o[w][2 * h + 1] = vv[2 * h + 1] & __byte_perm(bm, 0u, 0x3322);
        }
    }
    *reinterpret_cast<uint4 *>(dst + r0) = make_uint4(o[0][0], o[0][1], o[0][2], o[0][3]);
    if (w1) *reinterpret_cast<uint4 *>(dst + r0 + C) = make_uint4(o[1][0], o[1][1], o[1][2], o[1][3]);
    if (h1) *reinterpret_cast<uint4 *>(dst + r1) = make_uint4(o[2][0], o[2][1], o[2][2], o[2][3]);
    if (h1 && w1) *reinterpret_cast<uint4 *>(dst + r1 + C) = make_uint4(o[3][0], o[3][1], o[3][2], o[3][3]);
}

__global__ void __launch_bounds__(256) unpool_h8_kernel(const Act in, const uint8_t *idx, const Act out, int B) {
    const int C8 = out.C / 8, H = out.H, W = out.W, Hp = in.H, Wp = in.W, Hc = (H + 1) / 2, Wc = (W + 1) / 2;
    const size_t total = (size_t)B * Hc * Wc * C8;
    for (size_t g = (size_t)blockIdx.x * 256 + threadIdx.x; g < total; g += (size_t)gridDim.x * 256) {
        const int c = (int)(g % C8) * 8;
        size_t p = g / C8;
        const int wc = (int)(p % Wc); p /= Wc;
        const int hc = (int)(p % Hc);
        const int b = (int)(p / Hc);
        const size_t r0 = (((size_t)b * H + 2 * hc) * W + 2 * wc) * out.C + c;
        const size_t r1 = r0 + (size_t)W * out.C;
        const bool h1 = 2 * hc + 1 < H, w1 = 2 * wc + 1 < W;
        const bool have = hc < Hp && wc < Wp;      // floor mode: an odd last row / column is never pooled -> zeros
        const size_t s = (((size_t)b * Hp + hc) * Wp + wc) * out.C + c;
        uint2 kk = make_uint2(0u, 0u);
        if (have) kk = __ldg(reinterpret_cast<const uint2 *>(idx + s));
        unpool_plane(in.hi, out.hi, have, s, r0, r1, out.C, kk, w1, h1);
        if (out.lo) {
            if (in.lo) unpool_plane(in.lo, out.lo, have, s, r0, r1, out.C, kk, w1, h1);
            else unpool_plane(in.hi, out.lo, false, s, r0, r1, out.C, kk, w1, h1);          // no low part: zeros
        }
    }
}


// ---- strict fp32 convolution: 8x8 pixels x 64 output channels per CTA, 4x4 outputs per thread ----
__global__ void __launch_bounds__(256) conv3x3_f32_kernel(const float *__restrict__ in, const float *__restrict__ w,
                                                          const float *__restrict__ bias, float *__restrict__ out,
                                                          int H, int W, int Cin, int Cout, int relu) {
    __shared__ __align__(16) float s_in[100][16];
    __shared__ __align__(16) float s_w[9][16][64];
    const int tiles_w = (W + 7) / 8;
    const int h0 = (blockIdx.x / tiles_w) * 8, w0 = (blockIdx.x % tiles_w) * 8;
    const int n0 = blockIdx.y * 64, b = blockIdx.z;
    const int tid = threadIdx.x, tc = tid & 15, tp = tid >> 4;
    const int prow = tp >> 1, pcol0 = (tp & 1) * 4;
    float acc[4][4];
#pragma unroll
    for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[p][k] = 0.0f;

    for (int c0 = 0; c0 < Cin; c0 += 16) {
        for (int q = tid; q < 400; q += 256) {
            const int pix = q >> 2, c4 = (q & 3) * 4;
            const int gh = h0 + pix / 10 - 1, gw = w0 + pix % 10 - 1;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gh >= 0 && gh < H && gw >= 0 && gw < W)
                v = __ldg(reinterpret_cast<const float4 *>(in + (((size_t)b * H + gh) * W + gw) * Cin + c0 + c4));
            *reinterpret_cast<float4 *>(&s_in[pix][c4]) = v;
        }
        for (int q = tid; q < 2304; q += 256) {
            const int row = q >> 4, c4 = (q & 15) * 4;
            const int tap = row >> 4, ci = row & 15;
            *reinterpret_cast<float4 *>(&s_w[tap][ci][c4]) =
                __ldg(reinterpret_cast<const float4 *>(w + ((size_t)tap * Cin + c0 + ci) * Cout + n0 + c4));
        }
        __syncthreads();
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
            const int kh = tap / 3, kw = tap % 3;
#pragma unroll
            for (int ci = 0; ci < 16; ci += 4) {
                float4 a[4];
#pragma unroll
                for (int p = 0; p < 4; ++p)
                    a[p] = *reinterpret_cast<const float4 *>(&s_in[(prow + kh) * 10 + pcol0 + p + kw][ci]);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float4 bw = *reinterpret_cast<const float4 *>(&s_w[tap][ci + k][tc * 4]);
#pragma unroll
                    for (int p = 0; p < 4; ++p) {
                        const float av = (k == 0) ? a[p].x : (k == 1) ? a[p].y : (k == 2) ? a[p].z : a[p].w;
                        acc[p][0] = fmaf(av, bw.x, acc[p][0]);
                        acc[p][1] = fmaf(av, bw.y, acc[p][1]);
                        acc[p][2] = fmaf(av, bw.z, acc[p][2]);
                        acc[p][3] = fmaf(av, bw.w, acc[p][3]);
                    }
                }
            }
        }
        __syncthreads();
    }
    const float4 bv = __ldg(reinterpret_cast<const float4 *>(bias + n0 + tc * 4));
    const int gh = h0 + prow;
    if (gh < H) {
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int gw = w0 + pcol0 + p;
            if (gw < W) {
                float4 o = make_float4(acc[p][0] + bv.x, acc[p][1] + bv.y, acc[p][2] + bv.z, acc[p][3] + bv.w);
                if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
                *reinterpret_cast<float4 *>(out + (((size_t)b * H + gh) * W + gw) * Cout + n0 + tc * 4) = o;
            }
        }
    }
}

// regressor (segnet.py:131,140): 3x3 conv 64->1 + bias on fp32 NHWC; one warp per output pixel
__global__ void __launch_bounds__(256) regressor_f32_kernel(const float *__restrict__ in, const float *__restrict__ w,
                                                            const float *__restrict__ bias, float *__restrict__ out,
                                                            int B, int H, int W) {
    const int lane = threadIdx.x & 31;
    const size_t total = (size_t)B * H * W;
    for (size_t g = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5); g < total; g += (size_t)gridDim.x * 8) {
        const int wq = (int)(g % W);
        const int hq = (int)((g / W) % H);
        const int b = (int)(g / ((size_t)W * H));
        float s = 0.0f;
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
            const int gh = hq + tap / 3 - 1, gw = wq + tap % 3 - 1;
            if (gh >= 0 && gh < H && gw >= 0 && gw < W) {
                const float2 v = __ldg(reinterpret_cast<const float2 *>(in + (((size_t)b * H + gh) * W + gw) * 64 + 2 * lane));
                const float2 ww = __ldg(reinterpret_cast<const float2 *>(w + tap * 64 + 2 * lane));
                s = fmaf(v.x, ww.x, s);
                s = fmaf(v.y, ww.y, s);
            }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
        if (lane == 0) out[g] = s + __ldg(bias);
    }
}

// tcgen05 modes: the last conv's epilogue leaves per-pixel, per-tap partial sums of the regressor
// (regpart [B][H][W][9]); this gathers the 9 shifted partials and adds the bias.
__global__ void __launch_bounds__(256) regressor_gather_kernel(const float *__restrict__ part, const float *__restrict__ bias,
                                                               float *__restrict__ out, int B, int H, int W) {
    const size_t total = (size_t)B * H * W;
    for (size_t g = (size_t)blockIdx.x * 256 + threadIdx.x; g < total; g += (size_t)gridDim.x * 256) {
        const int wq = (int)(g % W);
        const int hq = (int)((g / W) % H);
        const int b = (int)(g / ((size_t)W * H));
        float s = 0.0f;
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
            const int gh = hq + tap / 3 - 1, gw = wq + tap % 3 - 1;
            if (gh >= 0 && gh < H && gw >= 0 && gw < W) s += __ldg(part + (((size_t)b * H + gh) * W + gw) * 9 + tap);
        }
        out[g] = s + __ldg(bias);
    }
}

static inline int ew_blocks(size_t total) {
    size_t b = (total + 255) / 256;
    return (int)(b > 148 * 16 ? 148 * 16 : (b < 1 ? 1 : b));
}

// ---- BatchNorm2d with the statistics of the batch in flight (the module's train-mode arithmetic, segnet.py:35-38
// left in training mode by predict_ground.py): per-channel mean and biased variance over B*H*W in float64, then
// y = relu(x * alpha + beta'), alpha = gamma / sqrt(var + eps), beta' = beta - mean * alpha  (eps = 1e-5).
// Two-stage, fixed-order reduction: reproducible run to run.
__global__ void __launch_bounds__(256) bn_partial_kernel(const float *__restrict__ x, int64_t npix, int C, double *partial) {
    __shared__ double s_s[256], s_q[256];
    const int R = 256 / C;                                   // pixel lanes per channel (C = 64, 128, 256)
    const int c = threadIdx.x % C, r = threadIdx.x / C;
    const int64_t per = (npix + gridDim.x - 1) / gridDim.x;
    const int64_t p0 = (int64_t)blockIdx.x * per, p1 = min(p0 + per, npix);
    double s = 0.0, q = 0.0;
    for (int64_t p = p0 + r; p < p1; p += R) {
        const double v = (double)__ldg(x + p * C + c);
        s += v; q += v * v;
    }
    s_s[threadIdx.x] = s; s_q[threadIdx.x] = q;
    __syncthreads();
    if (r == 0) {
        for (int k = 1; k < R; ++k) { s += s_s[k * C + c]; q += s_q[k * C + c]; }
        partial[((size_t)blockIdx.x * 256 + c) * 2 + 0] = s;
        partial[((size_t)blockIdx.x * 256 + c) * 2 + 1] = q;
    }
}

// `stats` [C][2]: the batch mean and the UNBIASED variance, what nn.BatchNorm2d in training mode blends into its
// running buffers (momentum 0.1) -- the host side applies that update (gndnet_b200.model, bn_mode="batch").
__global__ void bn_final_kernel(const double *partial, int nblk, int64_t npix, const float *gamma, const float *beta,
                                float *ab, float *stats) {
    const int c = threadIdx.x;
    double s = 0.0, q = 0.0;
    for (int b = 0; b < nblk; ++b) { s += partial[((size_t)b * 256 + c) * 2]; q += partial[((size_t)b * 256 + c) * 2 + 1]; }
    const double mean = s / (double)npix;
    double var = q / (double)npix - mean * mean;             // biased variance, as the normalisation uses
    if (var < 0.0) var = 0.0;
    const double alpha = (double)gamma[c] / sqrt(var + 1e-5);
    ab[c] = (float)alpha;
    ab[256 + c] = (float)((double)beta[c] - mean * alpha);
    if (stats) {
        stats[2 * c] = (float)mean;
        stats[2 * c + 1] = (float)(npix > 1 ? var * ((double)npix / (double)(npix - 1)) : var);
    }
}

__global__ void __launch_bounds__(256) bn_apply_relu_kernel(float *x, int64_t n4, int C, const float *__restrict__ ab) {
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (int64_t)gridDim.x * 256) {
        const int c = (int)((i * 4) % C);
        float4 v = reinterpret_cast<float4 *>(x)[i];
        v.x = fmaxf(fmaf(v.x, ab[c], ab[256 + c]), 0.0f);
        v.y = fmaxf(fmaf(v.y, ab[c + 1], ab[256 + c + 1]), 0.0f);
        v.z = fmaxf(fmaf(v.z, ab[c + 2], ab[256 + c + 2]), 0.0f);
        v.w = fmaxf(fmaf(v.w, ab[c + 3], ab[256 + c + 3]), 0.0f);
        reinterpret_cast<float4 *>(x)[i] = v;
    }
}

static int conv_f32(gnd_ctx *ctx, int layer, const Act &in, const Act &out, int B, cudaStream_t st) {
    const ConvLayer &L = ctx->conv[layer];
    const bool batch_bn = ctx->cfg.bn_batch_stats != 0;
    dim3 grid(((in.H + 7) / 8) * ((in.W + 7) / 8), L.cout / 64, B);
    conv3x3_f32_kernel<<<grid, 256, 0, st>>>(in.f32, L.w32, L.bias, out.f32, in.H, in.W, L.cin, L.cout, batch_bn ? 0 : 1);
    GND_LAUNCH_CHECK("conv3x3_f32_kernel");
    if (batch_bn) {
        if (!ctx->bn_partial || !ctx->bn_ab) return fail(GND_ERR_STATE, "workspace has no batch-statistics scratch");
        const int64_t npix = (int64_t)B * out.H * out.W;
        int nblk = (int)((npix + 63) / 64);
        if (nblk > BN_BLOCKS) nblk = BN_BLOCKS;
        bn_partial_kernel<<<nblk, 256, 0, st>>>(out.f32, npix, L.cout, ctx->bn_partial);
        GND_LAUNCH_CHECK("bn_partial_kernel");
        bn_final_kernel<<<1, L.cout, 0, st>>>(ctx->bn_partial, nblk, npix, L.gamma, L.beta, ctx->bn_ab,
                                              ctx->bn_stats ? ctx->bn_stats + (size_t)layer * 512 : nullptr);
        GND_LAUNCH_CHECK("bn_final_kernel");
        const int64_t n4 = npix * L.cout / 4;
        bn_apply_relu_kernel<<<ew_blocks((size_t)n4), 256, 0, st>>>(out.f32, n4, L.cout, ctx->bn_ab);
        GND_LAUNCH_CHECK("bn_apply_relu_kernel");
    }
    return GND_OK;
}

int segnet_run(gnd_ctx *ctx, int32_t B, float *height, cudaStream_t st, cudaEvent_t *ev) {
    const bool tc = ctx->cfg.precision != GND_PREC_FP32;
    int rc;
    bool ring = false;
    if (!ev && ctx->tim_next < ctx->tim_slots) {          // armed by gnd_conv_timing_begin: next slot of the event ring
        ev = &ctx->tim_events[(size_t)ctx->tim_next++ * TIM_STRIDE];
        ring = true;
        cudaEventRecord(ev[16], st);
    }
#define CONV(l, in, out, fuse)                                                          \
    do {                                                                                \
        if (ev) cudaEventRecord(ev[2 * (l)], st);                                        \
        rc = tc ? tc_conv_run(ctx, l, in, out, B, fuse, st) : conv_f32(ctx, l, in, out, B, st); \
        if (rc != GND_OK) return rc;                                                    \
        if (ev) cudaEventRecord(ev[2 * (l) + 1], st);                                    \
    } while (0)
    CONV(0, ctx->canvas, ctx->d1a, false);
    CONV(1, ctx->d1a, ctx->d1b, false);
    if (tc) pool_h8_kernel<<<ew_blocks(ctx->p1.elems(B) / 8), 256, 0, st>>>(ctx->d1b, ctx->p1, ctx->idx1, B);
    else pool_kernel<<<ew_blocks(ctx->p1.elems(B)), 256, 0, st>>>(ctx->d1b, ctx->p1, ctx->idx1, B);
    GND_LAUNCH_CHECK("pool_kernel");
    CONV(2, ctx->p1, ctx->d2a, false);
    CONV(3, ctx->d2a, ctx->d2b, false);
    if (tc) pool_mask_h8_kernel<<<ew_blocks(ctx->u2.elems(B) / 32 + 1), 256, 0, st>>>(ctx->d2b, ctx->u2, B, ctx->opt_l2_order);
    else pool_mask_kernel<<<ew_blocks(ctx->u2.elems(B)), 256, 0, st>>>(ctx->d2b, ctx->u2, B);
    GND_LAUNCH_CHECK("pool_mask_kernel");
    CONV(4, ctx->u2, ctx->u2a, false);
    CONV(5, ctx->u2a, ctx->u2b, false);
    if (tc) unpool_h8_kernel<<<ew_blocks(ctx->u1.elems(B) / 32 + 1), 256, 0, st>>>(ctx->u2b, ctx->idx1, ctx->u1, B);
    else unpool_kernel<<<ew_blocks(ctx->u1.elems(B)), 256, 0, st>>>(ctx->u2b, ctx->idx1, ctx->u1, B);
    GND_LAUNCH_CHECK("unpool_kernel");
    CONV(6, ctx->u1, ctx->u1a, false);
    CONV(7, ctx->u1a, ctx->u1b, tc);
#undef CONV
    const size_t px = (size_t)B * ctx->ny * ctx->nx;
    if (tc) {
        regressor_gather_kernel<<<ew_blocks(px), 256, 0, st>>>(ctx->regpart, ctx->reg_b, height, B, ctx->ny, ctx->nx);
        GND_LAUNCH_CHECK("regressor_gather_kernel");
    } else {
        size_t blocks = (px + 7) / 8;
        regressor_f32_kernel<<<(int)(blocks > 148 * 16 ? 148 * 16 : blocks), 256, 0, st>>>(ctx->u1b.f32, ctx->reg_w, ctx->reg_b,
                                                                                          height, B, ctx->ny, ctx->nx);
        GND_LAUNCH_CHECK("regressor_f32_kernel");
    }
    if (ring) cudaEventRecord(ev[17], st);
    return GND_OK;
}

}  // namespace gnd
