// Pooling and layout kernels (NHWC bf16): 3x3/2 max-pool with saved arg-max, global average pool,
// NCHW fp32 <-> NHWC bf16 conversion, stem weight packing, fp32 -> bf16 casts.
// Reference ops: nn.MaxPool2d(3,2,1) and nn.AdaptiveAvgPool2d(1) of torchvision ResNet
// (R/03 模型并行/03_model_parallel.ipynb#cell5, seq1/seq2), `.to(device)` input
// copies (ddp_gpus.py:49-50).
#include "common.cuh"
#include "ptdb200.h"

namespace {

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
    float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
    uint4 u;
    u.x = pack_bf16x2(f[0], f[1]); u.y = pack_bf16x2(f[2], f[3]);
    u.z = pack_bf16x2(f[4], f[5]); u.w = pack_bf16x2(f[6], f[7]);
    return u;
}

// ---- max pool 3x3 stride 2 pad 1 --------------------------------------------------------------
// One thread = 8 channels of one output pixel.  idx stores the winning tap (0..8, first max in
// scan order like torch) per channel as one byte.
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y,
                                                          uint8_t* __restrict__ idx, int N, int H, int W, int C,
                                                          int OH, int OW) {
    const int cg = C >> 3;
    const long long total = (long long)N * OH * OW * cg;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const int g = (int)(t % cg);
        long long pix = t / cg;
        const int ow = (int)(pix % OW); pix /= OW;
        const int oh = (int)(pix % OH);
        const int n = (int)(pix / OH);
        float best[8];
        int bi[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { best[i] = -INFINITY; bi[i] = 0; }
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int hi = oh * 2 - 1 + r;
            if (hi < 0 || hi >= H) continue;
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                const int wi = ow * 2 - 1 + s;
                if (wi < 0 || wi >= W) continue;
                float f[8];
                unpack8(*reinterpret_cast<const uint4*>(x + ((size_t)((size_t)n * H + hi) * W + wi) * C + g * 8), f);
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    if (f[i] > best[i] || (f[i] != f[i] && best[i] == best[i])) { best[i] = f[i]; bi[i] = r * 3 + s; }
            }
        }
        const size_t off = ((size_t)((size_t)n * OH + oh) * OW + ow) * C + g * 8;
        *reinterpret_cast<uint4*>(y + off) = pack8(best);
        uint2 packed;
        packed.x = bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24);
        packed.y = bi[4] | (bi[5] << 8) | (bi[6] << 16) | (bi[7] << 24);
        *reinterpret_cast<uint2*>(idx + off) = packed;
    }
}

// One thread = 8 channels of one INPUT pixel: sum dy of the (at most 4) windows whose arg-max is here.
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const bf16* __restrict__ dy, const uint8_t* __restrict__ idx,
                                                          bf16* __restrict__ dx, int N, int H, int W, int C, int OH,
                                                          int OW) {
    const int cg = C >> 3;
    const long long total = (long long)N * H * W * cg;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const int g = (int)(t % cg);
        long long pix = t / cg;
        const int w = (int)(pix % W); pix /= W;
        const int h = (int)(pix % H);
        const int n = (int)(pix / H);
        float acc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = 0.f;
        // windows (oh, ow) with oh*2-1+r == h, r in [0,3)
        for (int r = 0; r < 3; ++r) {
            const int th = h + 1 - r;
            if (th < 0 || (th & 1)) continue;
            const int oh = th >> 1;
            if (oh >= OH) continue;
            for (int s = 0; s < 3; ++s) {
                const int tw = w + 1 - s;
                if (tw < 0 || (tw & 1)) continue;
                const int ow = tw >> 1;
                if (ow >= OW) continue;
                const size_t off = ((size_t)((size_t)n * OH + oh) * OW + ow) * C + g * 8;
                const uint2 packed = *reinterpret_cast<const uint2*>(idx + off);
                float d[8];
                unpack8(*reinterpret_cast<const uint4*>(dy + off), d);
                const int tap = r * 3 + s;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const uint32_t word = i < 4 ? packed.x : packed.y;
                    if ((int)((word >> ((i & 3) * 8)) & 0xff) == tap) acc[i] += d[i];
                }
            }
        }
        *reinterpret_cast<uint4*>(dx + ((size_t)((size_t)n * H + h) * W + w) * C + g * 8) = pack8(acc);
    }
}

// ---- global average pool ------------------------------------------------------------------------
__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, int N,
                                                          int HW, int C) {
    const int cg = C >> 3;
    const long long total = (long long)N * cg;
    const float inv = 1.f / (float)HW;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const int g = (int)(t % cg);
        const int n = (int)(t / cg);
        float acc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = 0.f;
        for (int p = 0; p < HW; ++p) {
            float f[8];
            unpack8(*reinterpret_cast<const uint4*>(x + ((size_t)n * HW + p) * C + g * 8), f);
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[i] += f[i];
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] *= inv;
        *reinterpret_cast<uint4*>(y + (size_t)n * C + g * 8) = pack8(acc);
    }
}

__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const bf16* __restrict__ dy, bf16* __restrict__ dx, int N,
                                                          int HW, int C) {
    const int cg = C >> 3;
    const long long total = (long long)N * HW * cg;
    const float inv = 1.f / (float)HW;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const int g = (int)(t % cg);
        const long long pix = t / cg;
        const int n = (int)(pix / HW);
        float f[8];
        unpack8(*reinterpret_cast<const uint4*>(dy + (size_t)n * C + g * 8), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] *= inv;
        *reinterpret_cast<uint4*>(dx + (size_t)pix * C + g * 8) = pack8(f);
    }
}

// ---- layout / dtype -----------------------------------------------------------------------------
// NCHW fp32 [N,3,H,W] -> NHWC bf16 with 4 channels (4th = 0): the stem's input format.
__global__ void __launch_bounds__(256) nchw3_to_nhwc4_kernel(const float* __restrict__ x, bf16* __restrict__ y, int N,
                                                             long long HW) {
    const long long total = (long long)N * HW;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const long long n = t / HW, p = t - n * HW;
        const float* src = x + n * 3 * HW + p;
        uint2 o;
        o.x = pack_bf16x2(src[0], src[HW]);
        o.y = pack_bf16x2(src[2 * HW], 0.f);
        *reinterpret_cast<uint2*>(y + t * 4) = o;
    }
}

// Generic NCHW fp32 -> NHWC bf16 and back, tiled through shared memory (C, HW arbitrary).
__global__ void __launch_bounds__(256) nchw_to_nhwc_kernel(const float* __restrict__ x, bf16* __restrict__ y, int C,
                                                           int HW) {
    __shared__ float tile[32][33];
    const int n = blockIdx.z;
    const int c0 = blockIdx.y * 32, p0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
    for (int i = ty; i < 32; i += 8) {
        const int c = c0 + i, p = p0 + tx;
        tile[i][tx] = (c < C && p < HW) ? x[((size_t)n * C + c) * HW + p] : 0.f;
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int p = p0 + i, c = c0 + tx;
        if (c < C && p < HW) y[((size_t)n * HW + p) * C + c] = __float2bfloat16(tile[tx][i]);
    }
}
__global__ void __launch_bounds__(256) nhwc_to_nchw_kernel(const bf16* __restrict__ x, float* __restrict__ y, int C,
                                                           int HW) {
    __shared__ float tile[32][33];
    const int n = blockIdx.z;
    const int c0 = blockIdx.y * 32, p0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int i = ty; i < 32; i += 8) {
        const int p = p0 + i, c = c0 + tx;
        tile[i][tx] = (c < C && p < HW) ? __bfloat162float(x[((size_t)n * HW + p) * C + c]) : 0.f;
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int c = c0 + i, p = p0 + tx;
        if (c < C && p < HW) y[((size_t)n * C + c) * HW + p] = tile[tx][i];
    }
}

__global__ void __launch_bounds__(256) cast_f32_bf16_kernel(const float* __restrict__ x, bf16* __restrict__ y,
                                                            long long n) {
    const long long n4 = n >> 2;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n4;
         t += (long long)gridDim.x * blockDim.x) {
        const float4 v = reinterpret_cast<const float4*>(x)[t];
        uint2 o;
        o.x = pack_bf16x2(v.x, v.y);
        o.y = pack_bf16x2(v.z, v.w);
        reinterpret_cast<uint2*>(y)[t] = o;
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        const long long i = (n4 << 2) + threadIdx.x;
        y[i] = __float2bfloat16(x[i]);
    }
}

// Stem weights: fp32 logical [64][3][7][7] with arbitrary element strides -> bf16 [64][8][8][4] zero padded.
__global__ void __launch_bounds__(256) stem_pack_kernel(const float* __restrict__ w, bf16* __restrict__ w4,
                                                        long long sk, long long sc, long long sr, long long ss) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 64 * 256) return;
    const int c = t & 3, s = (t >> 2) & 7, r = (t >> 5) & 7, k = t >> 8;
    float v = 0.f;
    if (c < 3 && s < 7 && r < 7) v = w[k * sk + c * sc + r * sr + s * ss];
    w4[t] = __float2bfloat16(v);
}
// dw4 fp32 [64][8][8][4] -> dw fp32 logical [64][3][7][7] (strided), accumulate or overwrite.
__global__ void __launch_bounds__(256) stem_unpack_kernel(const float* __restrict__ dw4, float* __restrict__ dw,
                                                          long long sk, long long sc, long long sr, long long ss,
                                                          int accumulate) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 64 * 256) return;
    const int c = t & 3, s = (t >> 2) & 7, r = (t >> 5) & 7, k = t >> 8;
    if (c < 3 && s < 7 && r < 7) {
        float* d = dw + k * sk + c * sc + r * sr + s * ss;
        *d = accumulate ? *d + dw4[t] : dw4[t];
    }
}

// ---- space-to-depth stem input ---------------------------------------------------------------------
// NCHW fp32 [N,3,H,W] -> bf16 [N][OH+3][OW+3][16]: pixel (hs, ws) holds the 2x2 block of the 3-padded image at
// (2hs-3 .. 2hs-2, 2ws-3 .. 2ws-2), channel index (dy*2 + dx)*3 + ch, entries 12..15 zero.  In this layout the
// 7x7 stride-2 convolution is a 4x4 stride-1 convolution over 16 channels, and the im2col row of output pixel
// (oh, ow) for filter row r' is the 128 CONTIGUOUS bytes starting at pixel (oh + r', ow): a plain tiled TMA
// descriptor with an overlapping pixel stride fetches it (igemm.cu, AM_PATCH).
__global__ void __launch_bounds__(256) nchw3_to_s2d_kernel(const float* __restrict__ x, bf16* __restrict__ y, int N,
                                                           int H, int W, int Hs, int Ws) {
    const long long total = (long long)N * Hs * Ws;
    const long long HW = (long long)H * W;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const int ws = (int)(t % Ws);
        const long long t2 = t / Ws;
        const int hs = (int)(t2 % Hs);
        const long long n = t2 / Hs;
        const float* src = x + n * 3 * HW;
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = 0.f;
#pragma unroll
        for (int dy = 0; dy < 2; ++dy) {
            const int ih = 2 * hs + dy - 3;
#pragma unroll
            for (int dx = 0; dx < 2; ++dx) {
                const int iw = 2 * ws + dx - 3;
                if (ih >= 0 && ih < H && iw >= 0 && iw < W) {
#pragma unroll
                    for (int ch = 0; ch < 3; ++ch) v[(dy * 2 + dx) * 3 + ch] = src[ch * HW + (long long)ih * W + iw];
                }
            }
        }
        uint4 o0, o1;
        o0.x = pack_bf16x2(v[0], v[1]);   o0.y = pack_bf16x2(v[2], v[3]);
        o0.z = pack_bf16x2(v[4], v[5]);   o0.w = pack_bf16x2(v[6], v[7]);
        o1.x = pack_bf16x2(v[8], v[9]);   o1.y = pack_bf16x2(v[10], v[11]);
        o1.z = 0u;                        o1.w = 0u;
        uint4* dst = reinterpret_cast<uint4*>(y + t * 16);
        dst[0] = o0;
        dst[1] = o1;
    }
}

// Stem weights for the space-to-depth form: bf16 [64][4 r'][4 s'][16 c'], c' = (dy*2+dx)*3 + ch, tap (2r'+dy, 2s'+dx).
__device__ __forceinline__ bool s2d_decode(int t, int& k, int& ch, int& r, int& s) {
    const int c = t & 15, sp = (t >> 4) & 3, rp = (t >> 6) & 3;
    k = t >> 8;
    if (c >= 12) return false;
    const int q = c / 3;
    ch = c - q * 3;
    r = 2 * rp + (q >> 1);
    s = 2 * sp + (q & 1);
    return r < 7 && s < 7;
}
__global__ void __launch_bounds__(256) stem_pack_s2d_kernel(const float* __restrict__ w, bf16* __restrict__ w16,
                                                            long long sk, long long sc, long long sr, long long ss) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 64 * 256) return;
    int k, ch, r, s;
    const bool ok = s2d_decode(t, k, ch, r, s);
    w16[t] = __float2bfloat16(ok ? w[k * sk + ch * sc + r * sr + s * ss] : 0.f);
}
__global__ void __launch_bounds__(256) stem_unpack_s2d_kernel(const float* __restrict__ dw16, float* __restrict__ dw,
                                                              long long sk, long long sc, long long sr, long long ss,
                                                              int accumulate) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 64 * 256) return;
    int k, ch, r, s;
    if (s2d_decode(t, k, ch, r, s)) {
        float* d = dw + k * sk + ch * sc + r * sr + s * ss;
        *d = accumulate ? *d + dw16[t] : dw16[t];
    }
}

int ew_grid(long long work_items, int threads) {
    long long blocks = (work_items + threads - 1) / threads;
    const long long cap = (long long)ptd_num_sms() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace

extern "C" int ptd_maxpool3x3s2_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, void* stream) {
    PTD_REQUIRE(C % 8 == 0, "maxpool: C (%d) must be a multiple of 8", C);
    const int OH = (H + 2 - 3) / 2 + 1, OW = (W + 2 - 3) / 2 + 1;
    maxpool_fwd_kernel<<<ew_grid((long long)N * OH * OW * (C / 8), 256), 256, 0, (cudaStream_t)stream>>>(
        (const bf16*)x, (bf16*)y, (uint8_t*)idx, N, H, W, C, OH, OW);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_maxpool3x3s2_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C,
                                    void* stream) {
    PTD_REQUIRE(C % 8 == 0, "maxpool: C (%d) must be a multiple of 8", C);
    const int OH = (H + 2 - 3) / 2 + 1, OW = (W + 2 - 3) / 2 + 1;
    maxpool_bwd_kernel<<<ew_grid((long long)N * H * W * (C / 8), 256), 256, 0, (cudaStream_t)stream>>>(
        (const bf16*)dy, (const uint8_t*)idx, (bf16*)dx, N, H, W, C, OH, OW);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream) {
    PTD_REQUIRE(C % 8 == 0, "avgpool: C (%d) must be a multiple of 8", C);
    avgpool_fwd_kernel<<<ew_grid((long long)N * (C / 8), 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)x, (bf16*)y,
                                                                                                N, HW, C);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream) {
    PTD_REQUIRE(C % 8 == 0, "avgpool: C (%d) must be a multiple of 8", C);
    avgpool_bwd_kernel<<<ew_grid((long long)N * HW * (C / 8), 256), 256, 0, (cudaStream_t)stream>>>(
        (const bf16*)dy, (bf16*)dx, N, HW, C);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_nchw3_f32_to_nhwc4_bf16(const float* x, void* y, int N, int H, int W, void* stream) {
    nchw3_to_nhwc4_kernel<<<ew_grid((long long)N * H * W, 256), 256, 0, (cudaStream_t)stream>>>(x, (bf16*)y, N,
                                                                                                 (long long)H * W);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_nchw_f32_to_nhwc_bf16(const float* x, void* y, int N, int C, int HW, void* stream) {
    dim3 grid((HW + 31) / 32, (C + 31) / 32, N);
    nchw_to_nhwc_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, (bf16*)y, C, HW);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_nhwc_bf16_to_nchw_f32(const void* x, float* y, int N, int C, int HW, void* stream) {
    dim3 grid((HW + 31) / 32, (C + 31) / 32, N);
    nhwc_to_nchw_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)x, y, C, HW);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_cast_f32_to_bf16(const float* x, void* y, long long n, void* stream) {
    PTD_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & 7) == 0,
                "cast: pointers must be 16-/8-byte aligned");
    cast_f32_bf16_kernel<<<ew_grid(n / 4 + 1, 256), 256, 0, (cudaStream_t)stream>>>(x, (bf16*)y, n);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_nchw3_f32_to_s2d_bf16(const float* x, void* y, int N, int H, int W, void* stream) {
    PTD_REQUIRE(H % 2 == 0 && W % 2 == 0, "s2d stem input: H (%d) and W (%d) must be even", H, W);
    const int Hs = H / 2 + 3, Ws = W / 2 + 3;
    nchw3_to_s2d_kernel<<<ew_grid((long long)N * Hs * Ws, 256), 256, 0, (cudaStream_t)stream>>>(x, (bf16*)y, N, H, W, Hs, Ws);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_stem_pack_weight_s2d(const float* w, void* w16, long long sk, long long sc, long long sr,
                                        long long ss, void* stream) {
    stem_pack_s2d_kernel<<<64, 256, 0, (cudaStream_t)stream>>>(w, (bf16*)w16, sk, sc, sr, ss);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_stem_unpack_wgrad_s2d(const float* dw16, float* dw, long long sk, long long sc, long long sr,
                                         long long ss, int accumulate, void* stream) {
    stem_unpack_s2d_kernel<<<64, 256, 0, (cudaStream_t)stream>>>(dw16, dw, sk, sc, sr, ss, accumulate);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_stem_pack_weight(const float* w, void* w4, long long sk, long long sc, long long sr, long long ss,
                                    void* stream) {
    stem_pack_kernel<<<64, 256, 0, (cudaStream_t)stream>>>(w, (bf16*)w4, sk, sc, sr, ss);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_stem_unpack_wgrad(const float* dw4, float* dw, long long sk, long long sc, long long sr,
                                     long long ss, int accumulate, void* stream) {
    stem_unpack_kernel<<<64, 256, 0, (cudaStream_t)stream>>>(dw4, dw, sk, sc, sr, ss, accumulate);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
