// BatchNorm (training) fused with ReLU and the residual add, NHWC bf16, fp32 statistics.
//
// The per-channel sum / sum-of-squares come from the convolution epilogue (igemm.cu), so the
// forward needs exactly one pass over the activation: read y (+ identity), write z.  The backward
// is the usual two passes: reduce (sum g, sum g*xhat), then apply.  The masked gradient
// g = dz * (z > 0) of a residual block's last BN is also the gradient of the identity branch and is
// written out by the apply pass when asked for.
// Reference behaviour: nn.BatchNorm2d in torchvision Bottleneck (momentum 0.1, eps 1e-5),
// reached from R/03 模型并行/03_model_parallel.ipynb#cell5.
#include <mutex>

#include "common.cuh"
#include "ptdb200.h"

namespace {

constexpr int BN_THREADS = 256;

struct alignas(16) Pack8 {
    uint4 u;
};

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


// Block-level reduction of per-thread channel partials (8 channels per thread, threads with equal
// `g = tid % cg` own the same channels): shuffle across the lanes of a warp that share g, then one
// shared-memory atomic per (warp, channel), then one global atomic per (block, channel).
__device__ __forceinline__ void block_channel_reduce(float (&s)[8], float (&q)[8], int cg, int g, float* __restrict__ out_s,
                                                     float* __restrict__ out_q, float* sh /* [2][cg*8] */) {
    const int C = cg * 8;
    for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) sh[i] = 0.f;
    __syncthreads();
    if (cg < 32) {
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            if (o >= cg) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    s[i] += __shfl_xor_sync(0xffffffffu, s[i], o);
                    q[i] += __shfl_xor_sync(0xffffffffu, q[i], o);
                }
            }
        }
    }
    const int lane = threadIdx.x & 31;
    if (cg >= 32 || lane < cg) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            atomicAdd(&sh[g * 8 + i], s[i]);
            atomicAdd(&sh[C + g * 8 + i], q[i]);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < C; i += blockDim.x) {
        atomicAdd(out_s + i, sh[i]);
        atomicAdd(out_q + i, sh[C + i]);
    }
}

// Per-channel sums of a [M][C] bf16 matrix (used when the producer did not fuse them, e.g. tests).
__global__ void __launch_bounds__(BN_THREADS) bn_stats_kernel(const bf16* __restrict__ y, float* __restrict__ sum,
                                                              float* __restrict__ sq, long long M, int C) {
    const int cg = C >> 3;
    const int rows_per_iter = BN_THREADS / cg;
    const int g = threadIdx.x % cg;
    long long row = (long long)blockIdx.x * rows_per_iter + threadIdx.x / cg;
    const long long row_step = (long long)gridDim.x * rows_per_iter;
    float s[8], q[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = q[i] = 0.f;
    for (; row < M; row += row_step) {
        float f[8];
        unpack8(*reinterpret_cast<const uint4*>(y + row * C + g * 8), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] += f[i] * f[i]; }
    }
    extern __shared__ float sh_red[];
    block_channel_reduce(s, q, cg, g, sum, sq, sh_red);
}

// z = act( (y - mean) * invstd * gamma + beta  [+ residual] ), plus running-stat update by block 0.
template <bool RELU, bool RES>
__global__ void __launch_bounds__(BN_THREADS, 3)
bn_apply_kernel(const bf16* __restrict__ y, const bf16* __restrict__ res, bf16* __restrict__ z,
                const float* __restrict__ sum, const float* __restrict__ sq, const float* __restrict__ gamma,
                const float* __restrict__ beta, float* __restrict__ running_mean, float* __restrict__ running_var,
                float* __restrict__ save_mean, float* __restrict__ save_invstd, unsigned char* __restrict__ mask,
                long long M, int C, float eps, float momentum) {
    pdl_wait();
    const int cg = C >> 3;
    const int rows_per_iter = BN_THREADS / cg;
    const int g = threadIdx.x % cg;
    float scale[8], shift[8];
    const float inv_m = 1.f / (float)M;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int c = g * 8 + i;
        const float mean = sum[c] * inv_m;
        const float var = fmaxf(sq[c] * inv_m - mean * mean, 0.f);
        const float invstd = rsqrtf(var + eps);
        scale[i] = gamma[c] * invstd;
        shift[i] = beta[c] - mean * scale[i];
        if (blockIdx.x == 0 && threadIdx.x < cg) {
            save_mean[c] = mean;
            save_invstd[c] = invstd;
            if (running_mean != nullptr) {
                const float unbiased = M > 1 ? var * ((float)M / (float)(M - 1)) : var;
                running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
                running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
            }
        }
    }
    long long row = (long long)blockIdx.x * rows_per_iter + threadIdx.x / cg;
    const long long row_step = (long long)gridDim.x * rows_per_iter;
    // The loop is latency-bound (load -> math -> store per warp), so every thread keeps U independent rows in flight.
    constexpr int U = RES ? 2 : 4;
    for (; row < M; row += U * row_step) {
        uint4 ry[U], rr[RES ? U : 1];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long ru = row + u * row_step;
            if (ru < M) {
                const size_t off = (size_t)ru * C + g * 8;
                ry[u] = *reinterpret_cast<const uint4*>(y + off);
                if (RES) rr[RES ? u : 0] = *reinterpret_cast<const uint4*>(res + off);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long ru = row + u * row_step;
            if (ru >= M) break;
            const size_t off = (size_t)ru * C + g * 8;
            float f[8], r[8];
            unpack8(ry[u], f);
            if (RES) unpack8(rr[RES ? u : 0], r);
            unsigned bits = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float v = fmaf(f[i], scale[i], shift[i]);
                if (RES) v += r[i];
                if (RELU) {
                    if (v > 0.f) bits |= 1u << i;
                    v = fmaxf(v, 0.f);
                }
                f[i] = v;
            }
            *reinterpret_cast<uint4*>(z + off) = pack8(f);
            // one bit per element that survived the ReLU: the backward reads this byte instead of 16 bytes of z
            if (RELU && mask != nullptr) mask[(size_t)ru * cg + g] = (unsigned char)bits;
        }
    }
}

// z = relu(bn(y) + bn2(y2)): the last BatchNorm of a residual block whose shortcut is itself conv + BN.  The
// shortcut's normalised activation is never written: it is applied here, straight from the shortcut convolution's
// output y2, with its own batch statistics (and its running statistics / saved mean, invstd are updated here too).
__global__ void __launch_bounds__(BN_THREADS, 3)
bn_apply_dual_kernel(const bf16* __restrict__ y, const bf16* __restrict__ y2, bf16* __restrict__ z,
                     unsigned char* __restrict__ mask, const float* __restrict__ sum, const float* __restrict__ sq,
                     const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ running_mean,
                     float* __restrict__ running_var, float* __restrict__ save_mean, float* __restrict__ save_invstd,
                     const float* __restrict__ sum2, const float* __restrict__ sq2, const float* __restrict__ gamma2,
                     const float* __restrict__ beta2, float* __restrict__ running_mean2, float* __restrict__ running_var2,
                     float* __restrict__ save_mean2, float* __restrict__ save_invstd2, long long M, int C, float eps,
                     float momentum) {
    pdl_wait();
    const int cg = C >> 3;
    const int rows_per_iter = BN_THREADS / cg;
    const int g = threadIdx.x % cg;
    float scale[8], scale2[8], shift[8];
    const float inv_m = 1.f / (float)M;
    const float unb = M > 1 ? (float)M / (float)(M - 1) : 1.f;
    const bool writer = blockIdx.x == 0 && threadIdx.x < cg;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int c = g * 8 + i;
        const float mean = sum[c] * inv_m, mean2 = sum2[c] * inv_m;
        const float var = fmaxf(sq[c] * inv_m - mean * mean, 0.f), var2 = fmaxf(sq2[c] * inv_m - mean2 * mean2, 0.f);
        const float invstd = rsqrtf(var + eps), invstd2 = rsqrtf(var2 + eps);
        scale[i] = gamma[c] * invstd;
        scale2[i] = gamma2[c] * invstd2;
        shift[i] = (beta[c] - mean * scale[i]) + (beta2[c] - mean2 * scale2[i]);
        if (writer) {
            save_mean[c] = mean; save_invstd[c] = invstd;
            save_mean2[c] = mean2; save_invstd2[c] = invstd2;
            if (running_mean != nullptr) {
                running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
                running_var[c] = (1.f - momentum) * running_var[c] + momentum * var * unb;
            }
            if (running_mean2 != nullptr) {
                running_mean2[c] = (1.f - momentum) * running_mean2[c] + momentum * mean2;
                running_var2[c] = (1.f - momentum) * running_var2[c] + momentum * var2 * unb;
            }
        }
    }
    long long row = (long long)blockIdx.x * rows_per_iter + threadIdx.x / cg;
    const long long row_step = (long long)gridDim.x * rows_per_iter;
    constexpr int U = 2;
    for (; row < M; row += U * row_step) {
        uint4 ry[U], rr[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long ru = row + u * row_step;
            if (ru < M) {
                const size_t off = (size_t)ru * C + g * 8;
                ry[u] = *reinterpret_cast<const uint4*>(y + off);
                rr[u] = *reinterpret_cast<const uint4*>(y2 + off);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long ru = row + u * row_step;
            if (ru >= M) break;
            const size_t off = (size_t)ru * C + g * 8;
            float f[8], r[8];
            unpack8(ry[u], f);
            unpack8(rr[u], r);
            unsigned bits = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float v = fmaf(f[i], scale[i], fmaf(r[i], scale2[i], shift[i]));
                if (v > 0.f) bits |= 1u << i;
                f[i] = fmaxf(v, 0.f);
            }
            *reinterpret_cast<uint4*>(z + off) = pack8(f);
            if (mask != nullptr) mask[(size_t)ru * cg + g] = (unsigned char)bits;
        }
    }
}

// z = act(y * scale[c] + shift[c] [+ residual])  (inference-mode BN, or any per-channel affine)
template <bool RELU, bool RES>
__global__ void __launch_bounds__(BN_THREADS)
affine_act_kernel(const bf16* __restrict__ y, const bf16* __restrict__ res, bf16* __restrict__ z,
                  const float* __restrict__ scale_, const float* __restrict__ shift_, long long M, int C) {
    const int cg = C >> 3;
    const int rows_per_iter = BN_THREADS / cg;
    const int g = threadIdx.x % cg;
    float scale[8], shift[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { scale[i] = scale_[g * 8 + i]; shift[i] = shift_[g * 8 + i]; }
    long long row = (long long)blockIdx.x * rows_per_iter + threadIdx.x / cg;
    const long long row_step = (long long)gridDim.x * rows_per_iter;
    for (; row < M; row += row_step) {
        const size_t off = (size_t)row * C + g * 8;
        float f[8], r[8];
        unpack8(*reinterpret_cast<const uint4*>(y + off), f);
        if (RES) unpack8(*reinterpret_cast<const uint4*>(res + off), r);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float v = fmaf(f[i], scale[i], shift[i]);
            if (RES) v += r[i];
            if (RELU) v = fmaxf(v, 0.f);
            f[i] = v;
        }
        *reinterpret_cast<uint4*>(z + off) = pack8(f);
    }
}

// Backward pass 1: sum_g[c] += g, sum_gx[c] += g * xhat with g = dz * (z > 0 if RELU).
// RELU: 0 = no activation, 1 = mask recomputed from z (> 0), 2 = mask read from the packed bit tensor that
// bn_apply wrote (z then points at [M][C/8] bytes).
template <int RELU>
__global__ void __launch_bounds__(BN_THREADS, 2)
bn_bwd_reduce_kernel(const bf16* __restrict__ dz, const void* __restrict__ zm, const bf16* __restrict__ y,
                     const float* __restrict__ mean, const float* __restrict__ invstd, float* __restrict__ sum_g,
                     float* __restrict__ sum_gx, long long M, int C) {
    pdl_wait();
    const int cg = C >> 3;
    const int rows_per_iter = BN_THREADS / cg;
    const int g = threadIdx.x % cg;
    float mu[8], s[8], q[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { mu[i] = mean[g * 8 + i]; s[i] = q[i] = 0.f; }
    long long row = (long long)blockIdx.x * rows_per_iter + threadIdx.x / cg;
    const long long row_step = (long long)gridDim.x * rows_per_iter;
    const bf16* z = reinterpret_cast<const bf16*>(zm);
    const unsigned char* mask = reinterpret_cast<const unsigned char*>(zm);
    // rows in flight per thread (independent 16-byte loads); two resident blocks per SM keep ~128 KB on the wire
    constexpr int U = RELU == 1 ? 4 : (RELU == 2 ? 6 : 8);
    for (; row < M; row += U * row_step) {
        uint4 rd[U], rx[U], ra[RELU == 1 ? U : 1];
        unsigned rb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long ru = row + u * row_step;
            rb[u] = 0;
            if (ru < M) {
                const size_t off = (size_t)ru * C + g * 8;
                rd[u] = *reinterpret_cast<const uint4*>(dz + off);
                rx[u] = *reinterpret_cast<const uint4*>(y + off);
                if (RELU == 1) ra[RELU == 1 ? u : 0] = *reinterpret_cast<const uint4*>(z + off);
                if (RELU == 2) rb[u] = mask[(size_t)ru * cg + g];
            } else {
                rd[u] = make_uint4(0, 0, 0, 0);
                rx[u] = make_uint4(0, 0, 0, 0);
                if (RELU == 1) ra[RELU == 1 ? u : 0] = make_uint4(0, 0, 0, 0);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float d[8], a[8], x[8];
            unpack8(rd[u], d);
            unpack8(rx[u], x);
            if (RELU == 1) unpack8(ra[RELU == 1 ? u : 0], a);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const bool dead = RELU == 1 ? !(a[i] > 0.f) : (RELU == 2 ? !((rb[u] >> i) & 1u) : false);
                const float gi = dead ? 0.f : d[i];
                s[i] += gi;
                q[i] = fmaf(gi, x[i] - mu[i], q[i]);
            }
        }
    }
    pdl_trigger();      // what is left is the block-level reduction and one atomic per channel
#pragma unroll
    for (int i = 0; i < 8; ++i) q[i] *= invstd[g * 8 + i];       // xhat = (x - mean) * invstd, factored out of the loop
    extern __shared__ float sh_red[];
    block_channel_reduce(s, q, cg, g, sum_g, sum_gx, sh_red);
}

// Backward pass 2: dy = gamma*invstd * (g - sum_g/M - xhat*sum_gx/M); optionally g_out = g.
template <int RELU, bool GOUT>
__global__ void __launch_bounds__(BN_THREADS, 3)
bn_bwd_apply_kernel(const bf16* __restrict__ dz, const void* __restrict__ zm, const bf16* __restrict__ y,
                    const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ gamma,
                    const float* __restrict__ sum_g, const float* __restrict__ sum_gx, bf16* __restrict__ dy,
                    bf16* __restrict__ g_out, long long M, int C) {
    pdl_wait();
    const int cg = C >> 3;
    const int rows_per_iter = BN_THREADS / cg;
    const int g = threadIdx.x % cg;
    // dy = k * (g - mg - xhat * mgx) with xhat = (x - mu) * is, folded to dy = k * g + (a0 + a1 * x)
    float k[8], a0[8], a1[8];
    const float inv_m = 1.f / (float)M;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int c = g * 8 + i;
        const float is = invstd[c], mu = mean[c];
        k[i] = gamma[c] * is;
        const float t = k[i] * sum_gx[c] * inv_m * is;
        a1[i] = -t;
        a0[i] = fmaf(t, mu, -k[i] * sum_g[c] * inv_m);
    }
    const bf16* z = reinterpret_cast<const bf16*>(zm);
    const unsigned char* mask = reinterpret_cast<const unsigned char*>(zm);
    long long row = (long long)blockIdx.x * rows_per_iter + threadIdx.x / cg;
    const long long row_step = (long long)gridDim.x * rows_per_iter;
    constexpr int U = RELU == 1 ? 2 : 3;      // rows in flight per thread
    for (; row < M; row += U * row_step) {
        uint4 rd[U], rx[U], ra[RELU == 1 ? U : 1];
        unsigned rb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long ru = row + u * row_step;
            rb[u] = 0;
            if (ru < M) {
                const size_t off = (size_t)ru * C + g * 8;
                rd[u] = *reinterpret_cast<const uint4*>(dz + off);
                rx[u] = *reinterpret_cast<const uint4*>(y + off);
                if (RELU == 1) ra[RELU == 1 ? u : 0] = *reinterpret_cast<const uint4*>(z + off);
                if (RELU == 2) rb[u] = mask[(size_t)ru * cg + g];
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long ru = row + u * row_step;
            if (ru >= M) break;
            const size_t off = (size_t)ru * C + g * 8;
            float d[8], a[8], x[8], o[8];
            unpack8(rd[u], d);
            unpack8(rx[u], x);
            if (RELU == 1) unpack8(ra[RELU == 1 ? u : 0], a);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const bool dead = RELU == 1 ? !(a[i] > 0.f) : (RELU == 2 ? !((rb[u] >> i) & 1u) : false);
                const float gi = dead ? 0.f : d[i];
                d[i] = gi;
                o[i] = fmaf(k[i], gi, fmaf(a1[i], x[i], a0[i]));
            }
            *reinterpret_cast<uint4*>(dy + off) = pack8(o);
            if (GOUT) *reinterpret_cast<uint4*>(g_out + off) = pack8(d);
        }
    }
}

// =================================================================================================
// Backward of z = relu(bn(y) + bn2(y2)) (bn_apply_dual_kernel): both BatchNorms see the same gradient
// g = dz * mask, so dz and the mask are read once per pass for the two of them.
// =================================================================================================
__global__ void __launch_bounds__(BN_THREADS, 2)
bn_bwd_reduce_dual_kernel(const bf16* __restrict__ dz, const unsigned char* __restrict__ mask, const bf16* __restrict__ y,
                          const bf16* __restrict__ y2, const float* __restrict__ mean, const float* __restrict__ invstd,
                          const float* __restrict__ mean2, const float* __restrict__ invstd2, float* __restrict__ sum_g,
                          float* __restrict__ sum_gx, float* __restrict__ sum_g2, float* __restrict__ sum_gx2, long long M,
                          int C) {
    pdl_wait();
    const int cg = C >> 3;
    const int rows_per_iter = BN_THREADS / cg;
    const int g = threadIdx.x % cg;
    float mu[8], mu2[8], s[8], q[8], q2[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { mu[i] = mean[g * 8 + i]; mu2[i] = mean2[g * 8 + i]; s[i] = q[i] = q2[i] = 0.f; }
    long long row = (long long)blockIdx.x * rows_per_iter + threadIdx.x / cg;
    const long long row_step = (long long)gridDim.x * rows_per_iter;
    constexpr int U = 4;
    for (; row < M; row += U * row_step) {
        uint4 rd[U], rx[U], rx2[U];
        unsigned rb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long ru = row + u * row_step;
            rb[u] = 0;                                  // rows past M: all elements dead
            if (ru < M) {
                const size_t off = (size_t)ru * C + g * 8;
                rd[u] = *reinterpret_cast<const uint4*>(dz + off);
                rx[u] = *reinterpret_cast<const uint4*>(y + off);
                rx2[u] = *reinterpret_cast<const uint4*>(y2 + off);
                rb[u] = mask[(size_t)ru * cg + g];
            } else {
                rd[u] = rx[u] = rx2[u] = make_uint4(0, 0, 0, 0);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float d[8], x[8], x2[8];
            unpack8(rd[u], d);
            unpack8(rx[u], x);
            unpack8(rx2[u], x2);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float gi = ((rb[u] >> i) & 1u) ? d[i] : 0.f;
                s[i] += gi;
                q[i] = fmaf(gi, x[i] - mu[i], q[i]);
                q2[i] = fmaf(gi, x2[i] - mu2[i], q2[i]);
            }
        }
    }
    pdl_trigger();
    float s2[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { q[i] *= invstd[g * 8 + i]; q2[i] *= invstd2[g * 8 + i]; s2[i] = s[i]; }
    extern __shared__ float sh_red[];
    block_channel_reduce(s, q, cg, g, sum_g, sum_gx, sh_red);
    __syncthreads();
    block_channel_reduce(s2, q2, cg, g, sum_g2, sum_gx2, sh_red);
}

__global__ void __launch_bounds__(BN_THREADS, 2)
bn_bwd_apply_dual_kernel(const bf16* __restrict__ dz, const unsigned char* __restrict__ mask, const bf16* __restrict__ y,
                         const bf16* __restrict__ y2, const float* __restrict__ mean, const float* __restrict__ invstd,
                         const float* __restrict__ gamma, const float* __restrict__ mean2, const float* __restrict__ invstd2,
                         const float* __restrict__ gamma2, const float* __restrict__ sum_g, const float* __restrict__ sum_gx,
                         const float* __restrict__ sum_g2, const float* __restrict__ sum_gx2, bf16* __restrict__ dy,
                         bf16* __restrict__ dy2, long long M, int C) {
    pdl_wait();
    const int cg = C >> 3;
    const int rows_per_iter = BN_THREADS / cg;
    const int g = threadIdx.x % cg;
    float k[8], a0[8], a1[8], k2[8], b0[8], b1[8];      // dy = k*g + a0 + a1*y,  dy2 = k2*g + b0 + b1*y2
    const float inv_m = 1.f / (float)M;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int c = g * 8 + i;
        {
            const float is = invstd[c], mu = mean[c];
            k[i] = gamma[c] * is;
            const float t = k[i] * sum_gx[c] * inv_m * is;
            a1[i] = -t;
            a0[i] = fmaf(t, mu, -k[i] * sum_g[c] * inv_m);
        }
        {
            const float is = invstd2[c], mu = mean2[c];
            k2[i] = gamma2[c] * is;
            const float t = k2[i] * sum_gx2[c] * inv_m * is;
            b1[i] = -t;
            b0[i] = fmaf(t, mu, -k2[i] * sum_g2[c] * inv_m);
        }
    }
    long long row = (long long)blockIdx.x * rows_per_iter + threadIdx.x / cg;
    const long long row_step = (long long)gridDim.x * rows_per_iter;
    constexpr int U = 2;
    for (; row < M; row += U * row_step) {
        uint4 rd[U], rx[U], rx2[U];
        unsigned rb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long ru = row + u * row_step;
            rb[u] = 0;
            if (ru < M) {
                const size_t off = (size_t)ru * C + g * 8;
                rd[u] = *reinterpret_cast<const uint4*>(dz + off);
                rx[u] = *reinterpret_cast<const uint4*>(y + off);
                rx2[u] = *reinterpret_cast<const uint4*>(y2 + off);
                rb[u] = mask[(size_t)ru * cg + g];
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long ru = row + u * row_step;
            if (ru >= M) break;
            const size_t off = (size_t)ru * C + g * 8;
            float d[8], x[8], x2[8], o[8], o2[8];
            unpack8(rd[u], d);
            unpack8(rx[u], x);
            unpack8(rx2[u], x2);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float gi = ((rb[u] >> i) & 1u) ? d[i] : 0.f;
                o[i] = fmaf(k[i], gi, fmaf(a1[i], x[i], a0[i]));
                o2[i] = fmaf(k2[i], gi, fmaf(b1[i], x2[i], b0[i]));
            }
            *reinterpret_cast<uint4*>(dy + off) = pack8(o);
            *reinterpret_cast<uint4*>(dy2 + off) = pack8(o2);
        }
    }
}

// =================================================================================================
// BatchNorm + ReLU + 3x3/2 max-pool in one pass each way (the ResNet stem's tail)
// =================================================================================================
// Forward: p = maxpool(relu(bn(y))) straight from the convolution output; the 112x112 activation between BN and
// the pool is never written (or read back by the pool, or by the BN backward).  idx keeps the winning tap per
// element (0..8, first maximum in scan order as torch does), or POOL_DEAD when the maximum is <= 0: the ReLU then
// blocks the gradient, so the backward needs no separate ReLU mask.
// Backward: the gradient that reaches BN output pixel (h, w) is the sum of dp over the (at most four) windows whose
// arg-max is (h, w); it is rebuilt on the fly in both BN backward passes from dp and idx (4x smaller than the
// activation) instead of being materialised by a pooling backward kernel.
constexpr unsigned POOL_DEAD = 15u;

__global__ void __launch_bounds__(BN_THREADS)
bn_relu_pool_fwd_kernel(const bf16* __restrict__ y, bf16* __restrict__ p, uint8_t* __restrict__ idx,
                        const float* __restrict__ sum, const float* __restrict__ sq, const float* __restrict__ gamma,
                        const float* __restrict__ beta, float* __restrict__ running_mean, float* __restrict__ running_var,
                        float* __restrict__ save_mean, float* __restrict__ save_invstd, int N, int H, int W, int C,
                        float eps, float momentum) {
    pdl_wait();
    const int cg = C >> 3;
    const int g = threadIdx.x % cg;
    const int OH = H >> 1, OW = W >> 1;
    const long long M = (long long)N * H * W;
    float scale[8], shift[8];
    const float inv_m = 1.f / (float)M;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int c = g * 8 + i;
        const float mean = sum[c] * inv_m;
        const float var = fmaxf(sq[c] * inv_m - mean * mean, 0.f);
        const float invstd = rsqrtf(var + eps);
        scale[i] = gamma[c] * invstd;
        shift[i] = beta[c] - mean * scale[i];
        if (blockIdx.x == 0 && threadIdx.x < cg) {
            save_mean[c] = mean;
            save_invstd[c] = invstd;
            if (running_mean != nullptr) {
                const float unbiased = M > 1 ? var * ((float)M / (float)(M - 1)) : var;
                running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
                running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
            }
        }
    }
    const int pix_per_iter = BN_THREADS / cg;
    const long long total = (long long)N * OH * OW;
    for (long long pix = (long long)blockIdx.x * pix_per_iter + threadIdx.x / cg; pix < total;
         pix += (long long)gridDim.x * pix_per_iter) {
        const int ow = (int)(pix % OW);
        const long long t2 = pix / OW;
        const int oh = (int)(t2 % OH);
        const long long n = t2 / OH;
        // all nine taps are loaded up front (clamped coordinates), validity is applied afterwards
        uint4 raw[9];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int hi = min(max(oh * 2 - 1 + r, 0), H - 1);
#pragma unroll
            for (int s_ = 0; s_ < 3; ++s_) {
                const int wi = min(max(ow * 2 - 1 + s_, 0), W - 1);
                raw[r * 3 + s_] = *reinterpret_cast<const uint4*>(y + ((size_t)(n * H + hi) * W + wi) * C + g * 8);
            }
        }
        float best[8];
        unsigned bi[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { best[i] = 0.f; bi[i] = POOL_DEAD; }      // relu floor: only a tap > 0 can win
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const bool hv = (oh * 2 - 1 + r) >= 0;          // H, W even: the window never runs past the far edge
#pragma unroll
            for (int s_ = 0; s_ < 3; ++s_) {
                const bool ok = hv && (ow * 2 - 1 + s_) >= 0;
                float f[8];
                unpack8(raw[r * 3 + s_], f);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float v = fmaf(f[i], scale[i], shift[i]);
                    if (ok && v > best[i]) { best[i] = v; bi[i] = r * 3 + s_; }
                }
            }
        }
        const size_t off = (size_t)pix * C + g * 8;
        *reinterpret_cast<uint4*>(p + off) = pack8(best);
        uint2 packed;
        packed.x = bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24);
        packed.y = bi[4] | (bi[5] << 8) | (bi[6] << 16) | (bi[7] << 24);
        *reinterpret_cast<uint2*>(idx + off) = packed;
    }
}

// Gradient reaching the four BN-output pixels of the 2x2 block (2a.., 2b..) from the (up to) four windows that
// cover it: window (a, b) owns taps (1,1) (1,2) (2,1) (2,2) of the block, (a, b+1) taps (1,0) (2,0), (a+1, b) taps
// (0,1) (0,2), (a+1, b+1) tap (0,0).  gq[q][i]: q = 2 * (row in block) + (col in block).
__device__ __forceinline__ void pool_block_grad(const bf16* __restrict__ dp, const uint8_t* __restrict__ idx, long long n,
                                                int a, int b, int OH, int OW, int C, int g, float (&gq)[4][8]) {
    const int a1 = min(a + 1, OH - 1), b1 = min(b + 1, OW - 1);
    const bool va = a + 1 < OH, vb = b + 1 < OW;
    const size_t o00 = ((size_t)(n * OH + a) * OW + b) * C + g * 8, o01 = ((size_t)(n * OH + a) * OW + b1) * C + g * 8;
    const size_t o10 = ((size_t)(n * OH + a1) * OW + b) * C + g * 8, o11 = ((size_t)(n * OH + a1) * OW + b1) * C + g * 8;
    const uint2 i00 = *reinterpret_cast<const uint2*>(idx + o00), i01 = *reinterpret_cast<const uint2*>(idx + o01);
    const uint2 i10 = *reinterpret_cast<const uint2*>(idx + o10), i11 = *reinterpret_cast<const uint2*>(idx + o11);
    const uint4 r00 = *reinterpret_cast<const uint4*>(dp + o00), r01 = *reinterpret_cast<const uint4*>(dp + o01);
    const uint4 r10 = *reinterpret_cast<const uint4*>(dp + o10), r11 = *reinterpret_cast<const uint4*>(dp + o11);
    float d00[8], d01[8], d10[8], d11[8];
    unpack8(r00, d00); unpack8(r01, d01); unpack8(r10, d10); unpack8(r11, d11);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int sh = (i & 3) * 8;
        const unsigned t00 = ((i < 4 ? i00.x : i00.y) >> sh) & 0xffu;
        const unsigned t01 = vb ? ((i < 4 ? i01.x : i01.y) >> sh) & 0xffu : POOL_DEAD;
        const unsigned t10 = va ? ((i < 4 ? i10.x : i10.y) >> sh) & 0xffu : POOL_DEAD;
        const unsigned t11 = (va && vb) ? ((i < 4 ? i11.x : i11.y) >> sh) & 0xffu : POOL_DEAD;
        gq[0][i] = (t00 == 4u ? d00[i] : 0.f);
        gq[1][i] = (t00 == 5u ? d00[i] : 0.f) + (t01 == 3u ? d01[i] : 0.f);
        gq[2][i] = (t00 == 7u ? d00[i] : 0.f) + (t10 == 1u ? d10[i] : 0.f);
        gq[3][i] = (t00 == 8u ? d00[i] : 0.f) + (t01 == 6u ? d01[i] : 0.f) + (t10 == 2u ? d10[i] : 0.f) +
                   (t11 == 0u ? d11[i] : 0.f);
    }
}

// PASS 0: sum_g += g, sum_gx += g * xhat.   PASS 1: dy = k * g + a0 + a1 * y (see bn_bwd_apply_kernel).
template <int PASS>
__global__ void __launch_bounds__(BN_THREADS, 2)
bn_relu_pool_bwd_kernel(const bf16* __restrict__ dp, const uint8_t* __restrict__ idx, const bf16* __restrict__ y,
                        const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ gamma,
                        float* __restrict__ sum_g, float* __restrict__ sum_gx, bf16* __restrict__ dy, int N, int H, int W,
                        int C) {
    pdl_wait();
    const int cg = C >> 3;
    const int g = threadIdx.x % cg;
    const int OH = H >> 1, OW = W >> 1;
    float c0[8], c1[8], c2[8];          // PASS 0: mu, sum, sumx.  PASS 1: k, a0, a1
    const float inv_m = 1.f / ((float)N * H * W);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int c = g * 8 + i;
        if (PASS == 0) {
            c0[i] = mean[c]; c1[i] = 0.f; c2[i] = 0.f;
        } else {
            const float is = invstd[c], mu = mean[c];
            c0[i] = gamma[c] * is;
            const float t = c0[i] * sum_gx[c] * inv_m * is;
            c2[i] = -t;
            c1[i] = fmaf(t, mu, -c0[i] * sum_g[c] * inv_m);
        }
    }
    const int blk_per_iter = BN_THREADS / cg;
    const long long total = (long long)N * OH * OW;          // 2x2 blocks of BN-output pixels
    for (long long blk = (long long)blockIdx.x * blk_per_iter + threadIdx.x / cg; blk < total;
         blk += (long long)gridDim.x * blk_per_iter) {
        const int b = (int)(blk % OW);
        const long long t2 = blk / OW;
        const int a = (int)(t2 % OH);
        const long long n = t2 / OH;
        size_t yo[4];
        uint4 ry[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            yo[q] = ((size_t)(n * H + 2 * a + (q >> 1)) * W + 2 * b + (q & 1)) * C + g * 8;
            ry[q] = *reinterpret_cast<const uint4*>(y + yo[q]);
        }
        float gq[4][8];
        pool_block_grad(dp, idx, n, a, b, OH, OW, C, g, gq);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float x[8];
            unpack8(ry[q], x);
            if (PASS == 0) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    c1[i] += gq[q][i];
                    c2[i] = fmaf(gq[q][i], x[i] - c0[i], c2[i]);
                }
            } else {
                float o[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = fmaf(c0[i], gq[q][i], fmaf(c2[i], x[i], c1[i]));
                *reinterpret_cast<uint4*>(dy + yo[q]) = pack8(o);
            }
        }
    }
    if (PASS == 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) c2[i] *= invstd[g * 8 + i];
        extern __shared__ float sh_red[];
        block_channel_reduce(c1, c2, cg, g, sum_g, sum_gx, sh_red);
    }
}

// `per_sm` = blocks of the kernel that are resident per SM (its launch bound): large tensors get exactly one wave
// of grid-striding blocks.
int bn_grid(long long M, int C, int per_sm = 8) {
    const int rows_per_iter = BN_THREADS / (C >> 3);
    // >= 8 row groups per block: small tensors get few blocks, so the per-block tail (shared-memory
    // reduction + one atomic per channel) and the launch itself stay cheap
    long long blocks = ((M + rows_per_iter - 1) / rows_per_iter + 7) / 8;
    const long long cap = (long long)ptd_num_sms() * per_sm;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

bool bn_shape_ok(int C) { return C >= 64 && C <= 2048 && (C & (C - 1)) == 0; }

}  // namespace

extern "C" int ptd_bn_stats(const void* y, float* sum, float* sq, long long M, int C, void* stream) {
    PTD_REQUIRE(bn_shape_ok(C), "bn: C (%d) must be a power of two in [64, 2048]", C);
    bn_stats_kernel<<<bn_grid(M, C), BN_THREADS, 2 * C * sizeof(float), (cudaStream_t)stream>>>((const bf16*)y, sum, sq, M, C);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

extern "C" int ptd_bn_apply(const void* y, const void* residual, void* z, const float* sum, const float* sq,
                            const float* gamma, const float* beta, float* running_mean, float* running_var,
                            float* save_mean, float* save_invstd, long long M, int C, float eps, float momentum,
                            int relu, void* stream) {
    return ptd_bn_apply_mask(y, residual, z, nullptr, sum, sq, gamma, beta, running_mean, running_var, save_mean,
                             save_invstd, M, C, eps, momentum, relu, stream);
}

// Same, and when `relu` and `mask` are set also writes one bit per element ([M][C/8] bytes, bit i of byte
// (row, g) = element (row, 8g+i) survived the ReLU).  ptd_bn_bwd_reduce / ptd_bn_bwd_apply take that tensor in
// place of z with relu = 2, which cuts a sixth of the bytes the BatchNorm backward has to read.
extern "C" int ptd_bn_apply_mask(const void* y, const void* residual, void* z, void* mask, const float* sum,
                                 const float* sq, const float* gamma, const float* beta, float* running_mean,
                                 float* running_var, float* save_mean, float* save_invstd, long long M, int C, float eps,
                                 float momentum, int relu, void* stream) {
    PTD_REQUIRE(bn_shape_ok(C), "bn: C (%d) must be a power of two in [64, 2048]", C);
    const int grid = bn_grid(M, C, 3);
    cudaStream_t st = (cudaStream_t)stream;
#define PTD_BN_APPLY(RELU, RES)                                                                                     \
    PTD_CUDA_CHECK(ptd_launch(bn_apply_kernel<RELU, RES>, dim3(grid), dim3(BN_THREADS), 0, st, 1, (const bf16*)y,       \
                              (const bf16*)residual, (bf16*)z, sum, sq, gamma, beta, running_mean, running_var,     \
                              save_mean, save_invstd, (unsigned char*)mask, M, C, eps, momentum))
    if (relu && residual) PTD_BN_APPLY(true, true);
    else if (relu) PTD_BN_APPLY(true, false);
    else if (residual) PTD_BN_APPLY(false, true);
    else PTD_BN_APPLY(false, false);
#undef PTD_BN_APPLY
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

// Backward of ptd_bn_apply_dual: dy / dy2 = gradients w.r.t. y / y2; sums / sums2 = [dbeta | dgamma] of the two
// BatchNorms (2*C floats each, zeroed by the caller); mask = the bit tensor the forward wrote.
extern "C" int ptd_bn_bwd_dual(const void* dz, const void* mask, const void* y, const void* y2, const float* mean,
                               const float* invstd, const float* gamma, const float* mean2, const float* invstd2,
                               const float* gamma2, float* sums, float* sums2, void* dy, void* dy2, long long M, int C,
                               void* stream) {
    PTD_REQUIRE(bn_shape_ok(C), "bn: C (%d) must be a power of two in [64, 2048]", C);
    PTD_REQUIRE(mask != nullptr, "bn_bwd_dual: the forward's ReLU bit mask is required");
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = bn_grid(M, C, 2);
    PTD_CUDA_CHECK(ptd_launch(bn_bwd_reduce_dual_kernel, dim3(grid), dim3(BN_THREADS), 2 * C * sizeof(float), st, 1,
                              (const bf16*)dz, (const unsigned char*)mask, (const bf16*)y, (const bf16*)y2, mean, invstd,
                              mean2, invstd2, sums, sums + C, sums2, sums2 + C, M, C));
    PTD_CUDA_CHECK(ptd_launch(bn_bwd_apply_dual_kernel, dim3(grid), dim3(BN_THREADS), 0, st, 1, (const bf16*)dz,
                              (const unsigned char*)mask, (const bf16*)y, (const bf16*)y2, mean, invstd, gamma, mean2,
                              invstd2, gamma2, sums, sums + C, sums2, sums2 + C, (bf16*)dy, (bf16*)dy2, M, C));
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

// z = relu(bn(y) + bn2(y2)) with both sets of batch statistics; see bn_apply_dual_kernel.
extern "C" int ptd_bn_apply_dual(const void* y, const void* y2, void* z, void* mask, const float* sum, const float* sq,
                                 const float* gamma, const float* beta, float* running_mean, float* running_var,
                                 float* save_mean, float* save_invstd, const float* sum2, const float* sq2,
                                 const float* gamma2, const float* beta2, float* running_mean2, float* running_var2,
                                 float* save_mean2, float* save_invstd2, long long M, int C, float eps, float momentum,
                                 void* stream) {
    PTD_REQUIRE(bn_shape_ok(C), "bn: C (%d) must be a power of two in [64, 2048]", C);
    PTD_CUDA_CHECK(ptd_launch(bn_apply_dual_kernel, dim3(bn_grid(M, C, 3)), dim3(BN_THREADS), 0, (cudaStream_t)stream, 1,
                              (const bf16*)y, (const bf16*)y2, (bf16*)z, (unsigned char*)mask, sum, sq, gamma, beta,
                              running_mean, running_var, save_mean, save_invstd, sum2, sq2, gamma2, beta2, running_mean2,
                              running_var2, save_mean2, save_invstd2, M, C, eps, momentum));
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

extern "C" int ptd_affine_act(const void* y, const void* residual, void* z, const float* scale, const float* shift,
                              long long M, int C, int relu, void* stream) {
    PTD_REQUIRE(bn_shape_ok(C), "affine_act: C (%d) must be a power of two in [64, 2048]", C);
    const int grid = bn_grid(M, C);
    cudaStream_t st = (cudaStream_t)stream;
#define PTD_AFF(RELU, RES) \
    affine_act_kernel<RELU, RES><<<grid, BN_THREADS, 0, st>>>((const bf16*)y, (const bf16*)residual, (bf16*)z, scale, shift, M, C)
    if (relu && residual) PTD_AFF(true, true);
    else if (relu) PTD_AFF(true, false);
    else if (residual) PTD_AFF(false, true);
    else PTD_AFF(false, false);
#undef PTD_AFF
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

extern "C" int ptd_bn_bwd_reduce(const void* dz, const void* z, const void* y, const float* mean, const float* invstd,
                                 float* sum_g, float* sum_gx, long long M, int C, int relu, void* stream) {
    PTD_REQUIRE(bn_shape_ok(C), "bn: C (%d) must be a power of two in [64, 2048]", C);
    const int grid = min(bn_grid(M, C), ptd_num_sms() * 2);      // two resident blocks per SM (launch bounds)
    cudaStream_t st = (cudaStream_t)stream;
    const size_t sh = 2 * C * sizeof(float);
#define PTD_RED(RELU)                                                                                                 \
    PTD_CUDA_CHECK(ptd_launch(bn_bwd_reduce_kernel<RELU>, dim3(grid), dim3(BN_THREADS), sh, st, 1, (const bf16*)dz, z, \
                              (const bf16*)y, mean, invstd, sum_g, sum_gx, M, C))
    if (relu == 2) PTD_RED(2);
    else if (relu) PTD_RED(1);
    else PTD_RED(0);
#undef PTD_RED
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

extern "C" int ptd_bn_bwd_apply(const void* dz, const void* z, const void* y, const float* mean, const float* invstd,
                                const float* gamma, const float* sum_g, const float* sum_gx, void* dy, void* g_out,
                                long long M, int C, int relu, void* stream) {
    PTD_REQUIRE(bn_shape_ok(C), "bn: C (%d) must be a power of two in [64, 2048]", C);
    const int grid = bn_grid(M, C, 3);
    cudaStream_t st = (cudaStream_t)stream;
#define PTD_BWD(RELU, GOUT)                                                                                         \
    PTD_CUDA_CHECK(ptd_launch(bn_bwd_apply_kernel<RELU, GOUT>, dim3(grid), dim3(BN_THREADS), 0, st, 1, (const bf16*)dz, \
                              z, (const bf16*)y, mean, invstd, gamma, sum_g, sum_gx, (bf16*)dy, (bf16*)g_out, M, C))
    if (relu == 2 && g_out) PTD_BWD(2, true);
    else if (relu == 2) PTD_BWD(2, false);
    else if (relu && g_out) PTD_BWD(1, true);
    else if (relu) PTD_BWD(1, false);
    else if (g_out) PTD_BWD(0, true);
    else PTD_BWD(0, false);
#undef PTD_BWD
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

// p[N,H/2,W/2,C] = maxpool3x3s2p1(relu(bn(y[N,H,W,C]))), idx = winning tap per element (u8, 15 = gradient blocked).
extern "C" int ptd_bn_relu_maxpool_fwd(const void* y, void* p, void* idx, const float* sum, const float* sq,
                                       const float* gamma, const float* beta, float* running_mean, float* running_var,
                                       float* save_mean, float* save_invstd, int N, int H, int W, int C, float eps,
                                       float momentum, void* stream) {
    PTD_REQUIRE(bn_shape_ok(C), "bn: C (%d) must be a power of two in [64, 2048]", C);
    PTD_REQUIRE(H % 2 == 0 && W % 2 == 0, "bn_relu_maxpool: H (%d) and W (%d) must be even", H, W);
    const int grid = bn_grid((long long)N * (H / 2) * (W / 2), C, 4);
    PTD_CUDA_CHECK(ptd_launch(bn_relu_pool_fwd_kernel, dim3(grid), dim3(BN_THREADS), 0, (cudaStream_t)stream, 1,
                              (const bf16*)y, (bf16*)p, (uint8_t*)idx, sum, sq, gamma, beta, running_mean, running_var,
                              save_mean, save_invstd, N, H, W, C, eps, momentum));
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

// dy[N,H,W,C] = gradient w.r.t. the convolution output y, from dp[N,H/2,W/2,C]; sums (2*C floats, zeroed by the
// caller) receive [dbeta | dgamma].
extern "C" int ptd_bn_relu_maxpool_bwd(const void* dp, const void* idx, const void* y, const float* mean,
                                       const float* invstd, const float* gamma, float* sums, void* dy, int N, int H,
                                       int W, int C, void* stream) {
    PTD_REQUIRE(bn_shape_ok(C), "bn: C (%d) must be a power of two in [64, 2048]", C);
    PTD_REQUIRE(H % 2 == 0 && W % 2 == 0, "bn_relu_maxpool: H (%d) and W (%d) must be even", H, W);
    const int grid = bn_grid((long long)N * (H / 2) * (W / 2), C, 2);
    cudaStream_t st = (cudaStream_t)stream;
    PTD_CUDA_CHECK(ptd_launch(bn_relu_pool_bwd_kernel<0>, dim3(grid), dim3(BN_THREADS), 2 * C * sizeof(float), st, 1,
                              (const bf16*)dp, (const uint8_t*)idx, (const bf16*)y, mean, invstd, gamma, sums, sums + C,
                              (bf16*)nullptr, N, H, W, C));
    PTD_CUDA_CHECK(ptd_launch(bn_relu_pool_bwd_kernel<1>, dim3(grid), dim3(BN_THREADS), 0, st, 1, (const bf16*)dp,
                              (const uint8_t*)idx, (const bf16*)y, mean, invstd, gamma, sums, sums + C, (bf16*)dy, N, H,
                              W, C));
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

// =================================================================================================
// BatchNorm backward for callers that must not allocate device memory
// =================================================================================================
// ptd_bn_bwd_reduce / ptd_bn_bwd_apply above need a caller-provided, caller-zeroed scratch of 2*C floats for
// the per-channel sums.  Some callers cannot provide one: a plug-in running inside another runtime's memory
// plan (every byte of device memory already belongs to that runtime's arena, and a cudaMalloc from the side
// would also stall all of its streams), a fixed-footprint service that sizes its buffers once at start-up,
// or a plain C host that only owns its inputs and outputs.  For them the entry point below keeps the sums in
// a small array that is part of this library's device image: nothing is allocated or freed, here or in any
// routine it calls (cudaGetSymbolAddress, cudaMemsetAsync, cudaMemcpyAsync and kernel launches only).
//
// Limit: 2048 channels (2 * 2048 floats = 16 KiB, held by every GPU that loads the library until it exits).
// A wider layer is refused with an error instead of silently overrunning the array.  Like every BN kernel of this
// library the channel count must also be a power of two, 64 at least (64, 128, ..., 2048).
//
// In host code the array's name is only a host-side handle, not a device address - a kernel handed the name
// would fault (or quietly touch the wrong memory on a system where the GPU can reach host memory).  The real
// address comes from cudaGetSymbolAddress, asked once per device with that device current, and is cached.
// That lookup happens inside the FIRST call on each device; make that first call before starting any stream
// capture on that device.
//
// The array is shared by all calls and keeps whatever the previous call left in it, so every call clears the
// 2*C floats it is about to accumulate into, on the caller's stream, before the reduce pass.  Calls are
// therefore safe back to back on one stream, but two calls whose GPU work can overlap (different streams,
// or threads launching concurrently on one device) race on the array and must be ordered by the caller.
namespace {

constexpr int BN_STATIC_MAX_C = 2048;
static __device__ __align__(256) float g_bn_static_sums[2 * BN_STATIC_MAX_C];

// Device address of g_bn_static_sums for the current device (looked up once per device, then cached).
int bn_static_sums_ptr(float** out) {
    static std::mutex mu;
    static float* cached[64] = {nullptr};
    int dev = 0;
    PTD_CUDA_CHECK(cudaGetDevice(&dev));
    PTD_REQUIRE(dev >= 0 && dev < 64, "bn_backward_static: device index %d not supported", dev);
    std::lock_guard<std::mutex> lk(mu);
    if (cached[dev] == nullptr) {
        float* p = nullptr;
        PTD_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_bn_static_sums));
        PTD_REQUIRE(p != nullptr, "bn_backward_static: cudaGetSymbolAddress returned a null address");
        cached[dev] = p;
    }
    *out = cached[dev];
    return PTD_OK;
}

}  // namespace

// dy = BN-backward(dz) with the ReLU mask taken from z when `relu`; g_out (optional) receives the masked
// incoming gradient; dgamma / dbeta (fp32[C]) receive the parameter gradients.  Same math and kernels as
// ptd_bn_bwd_reduce + ptd_bn_bwd_apply, but no scratch argument: see the block comment above.
extern "C" int ptd_bn_backward_static(const void* dz, const void* z, const void* y, const float* mean,
                                      const float* invstd, const float* gamma, void* dy, void* g_out, float* dgamma,
                                      float* dbeta, long long M, int C, int relu, void* stream) {
    PTD_REQUIRE(C <= BN_STATIC_MAX_C,
                "bn_backward_static: C = %d exceeds the built-in scratch (max %d channels); use ptd_bn_bwd_reduce/"
                "ptd_bn_bwd_apply with your own 2*C-float scratch instead", C, BN_STATIC_MAX_C);
    PTD_REQUIRE(bn_shape_ok(C), "bn: C (%d) must be a power of two in [64, 2048]", C);
    PTD_REQUIRE(dgamma != nullptr && dbeta != nullptr && dy != nullptr, "bn_backward_static: dy, dgamma and dbeta are required");
    PTD_REQUIRE(relu == 0 || relu == 1,
                "bn_backward_static: relu must be 0 or 1 (got %d); the bit-mask mode (2) of ptd_bn_bwd_reduce/apply is not "
                "offered here", relu);
    PTD_REQUIRE(M > 0, "bn_backward_static: M (%lld rows) must be positive", M);
    PTD_REQUIRE(dz != nullptr && y != nullptr && mean != nullptr && invstd != nullptr && gamma != nullptr,
                "bn_backward_static: dz, y, mean, invstd and gamma are required");
    PTD_REQUIRE(!relu || z != nullptr, "bn_backward_static: relu = 1 needs z (the activation the ReLU mask is taken from)");
    float* sums = nullptr;
    int rc = bn_static_sums_ptr(&sums);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    float* sum_g = sums;          // -> dbeta
    float* sum_gx = sums + C;     // -> dgamma
    PTD_CUDA_CHECK(cudaMemsetAsync(sums, 0, 2 * (size_t)C * sizeof(float), st));     // leftovers of the previous call
    const int rgrid = min(bn_grid(M, C), ptd_num_sms() * 2);
    if (relu)
        bn_bwd_reduce_kernel<1><<<rgrid, BN_THREADS, 2 * C * sizeof(float), st>>>(
            (const bf16*)dz, (const bf16*)z, (const bf16*)y, mean, invstd, sum_g, sum_gx, M, C);
    else
        bn_bwd_reduce_kernel<0><<<rgrid, BN_THREADS, 2 * C * sizeof(float), st>>>(
            (const bf16*)dz, (const bf16*)z, (const bf16*)y, mean, invstd, sum_g, sum_gx, M, C);
    PTD_CUDA_CHECK(cudaGetLastError());
    const int agrid = bn_grid(M, C, 3);
#define PTD_BWD_S(RELU, GOUT)                                                                                          \
    bn_bwd_apply_kernel<RELU, GOUT><<<agrid, BN_THREADS, 0, st>>>((const bf16*)dz, (const bf16*)z, (const bf16*)y, mean, \
                                                                 invstd, gamma, sum_g, sum_gx, (bf16*)dy, (bf16*)g_out, M, C)
    if (relu && g_out) PTD_BWD_S(1, true);
    else if (relu) PTD_BWD_S(1, false);
    else if (g_out) PTD_BWD_S(0, true);
    else PTD_BWD_S(0, false);
#undef PTD_BWD_S
    PTD_CUDA_CHECK(cudaGetLastError());
    PTD_CUDA_CHECK(cudaMemcpyAsync(dbeta, sum_g, (size_t)C * sizeof(float), cudaMemcpyDeviceToDevice, st));
    PTD_CUDA_CHECK(cudaMemcpyAsync(dgamma, sum_gx, (size_t)C * sizeof(float), cudaMemcpyDeviceToDevice, st));
    return PTD_OK;
}
