// Loss and optimizer kernels.
//
//  * softmax cross-entropy, forward + backward in one pass, class-index or probability targets
//    (the reference always passes float targets: F.cross_entropy(output, ys) with ys float,
//    R/02 DDP基本概念/ddp_gpus.py:40, and one-hot floats, R/03 模型并行/03_model_parallel.ipynb#cell7).
//  * multi-tensor SGD (lr, momentum, weight decay) and Adam over a table of tensors, one launch for
//    the whole model, optionally refreshing the bf16 shadow copy the tensor-core kernels read
//    (reference: torch.optim.SGD(lr=1e-3) ddp_gpus.py:78; torch.optim.Adam(lr=1e-3) notebook 03 cell7).
#include "common.cuh"
#include "ptdb200.h"

namespace {

// ---- cross entropy ------------------------------------------------------------------------------
// One block per row.  loss_sum += -sum_c t_c * log_softmax(x)_c * scale ; dlogits = (p * sum_t - t) * scale.
template <bool SOFT, typename DOUT>
__global__ void __launch_bounds__(256)
softmax_ce_kernel(const float* __restrict__ logits, const float* __restrict__ soft_t, const long long* __restrict__ idx_t,
                  float* __restrict__ loss_sum, DOUT* __restrict__ dlogits, int C, int ldd, float scale) {
    const int row = blockIdx.x;
    const float* x = logits + (size_t)row * C;
    __shared__ float red[8];
    __shared__ float bcast;
    // max
    float m = -INFINITY;
    for (int c = threadIdx.x; c < C; c += blockDim.x) m = fmaxf(m, x[c]);
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        float v = red[0];
        for (int i = 1; i < (int)(blockDim.x >> 5); ++i) v = fmaxf(v, red[i]);
        bcast = v;
    }
    __syncthreads();
    m = bcast;
    __syncthreads();
    // sum exp, and for soft targets sum_t and sum t*x
    float se = 0.f, st = 0.f, stx = 0.f;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        se += __expf(x[c] - m);
        if (SOFT) {
            const float t = soft_t[(size_t)row * C + c];
            st += t;
            stx += t * x[c];
        }
    }
    float vals[3] = {se, st, stx};
    float tot[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float v = warp_sum(vals[k]);
        __syncthreads();
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
        __syncthreads();
        float a = 0.f;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) a += red[i];
        tot[k] = a;
    }
    const float lse = m + __logf(tot[0]);
    float sum_t, loss;
    long long ti = 0;
    if (SOFT) {
        sum_t = tot[1];
        loss = __fmul_rn(lse, sum_t) - tot[2];      // no FMA: x*t - t*x must cancel exactly (one-class toy loss)
    } else {
        ti = idx_t[row];
        sum_t = 1.f;
        loss = lse - x[ti];
    }
    if (threadIdx.x == 0) atomicAdd(loss_sum, loss * scale);
    const float inv_se = 1.f / tot[0];
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const float p = __expf(x[c] - m) * inv_se;
        const float t = SOFT ? soft_t[(size_t)row * C + c] : (c == ti ? 1.f : 0.f);
        const float g = (p * sum_t - t) * scale;
        if constexpr (sizeof(DOUT) == 2) dlogits[(size_t)row * ldd + c] = __float2bfloat16(g);
        else dlogits[(size_t)row * ldd + c] = g;
    }
    // zero the padding columns (ldd > C) so a TMA box that covers them reads zeros
    for (int c = C + threadIdx.x; c < ldd; c += blockDim.x) {
        if constexpr (sizeof(DOUT) == 2) dlogits[(size_t)row * ldd + c] = __float2bfloat16(0.f);
        else dlogits[(size_t)row * ldd + c] = 0.f;
    }
}

// ---- multi-tensor optimizers ----------------------------------------------------------------------
struct TensorEntry {
    float* p;        // parameter (fp32 master)
    float* g;        // gradient (fp32)
    float* s1;       // momentum buffer / Adam exp_avg (may be null for plain SGD)
    float* s2;       // Adam exp_avg_sq
    bf16* shadow;    // bf16 copy refreshed in the same pass (may be null)
    long long n;
};
constexpr int OPT_CHUNK = 8192;   // elements per block

// lr_dev (optional): the learning rate read from device memory instead of the launch argument, so that a step that
// was captured into a CUDA graph follows a schedule (the host rewrites one float between replays).
__global__ void __launch_bounds__(256)
multi_sgd_kernel(const TensorEntry* __restrict__ tab, const int2* __restrict__ blockmap, float lr, float momentum,
                 float weight_decay, float grad_scale, int first_step, const float* __restrict__ lr_dev) {
    if (lr_dev != nullptr) lr = *lr_dev;
    const int2 bm = blockmap[blockIdx.x];
    const TensorEntry e = tab[bm.x];
    const long long base = (long long)bm.y * OPT_CHUNK;
    const long long end = min(e.n, base + OPT_CHUNK);
    for (long long i = base + threadIdx.x; i < end; i += blockDim.x) {
        float p = e.p[i];
        float g = e.g[i] * grad_scale + weight_decay * p;
        if (e.s1 != nullptr) {
            const float buf = first_step ? g : momentum * e.s1[i] + g;
            e.s1[i] = buf;
            g = buf;
        }
        p -= lr * g;
        e.p[i] = p;
        if (e.shadow != nullptr) e.shadow[i] = __float2bfloat16(p);
    }
}

// Adam's step count lives on the device when step_dev is given: adam_advance_kernel increments it (one thread, before
// the update kernel on the same stream) and every block derives the bias corrections from it.  A host-side count would
// be frozen at capture time inside a CUDA graph - the corrections would never decay (torch: capturable=True).
__global__ void adam_advance_kernel(int* step_dev) { *step_dev += 1; }

__global__ void __launch_bounds__(256)
multi_adam_kernel(const TensorEntry* __restrict__ tab, const int2* __restrict__ blockmap, float lr, float beta1,
                  float beta2, float eps, float weight_decay, float bias1, float bias2_sqrt, float grad_scale,
                  const int* __restrict__ step_dev, const float* __restrict__ lr_dev) {
    if (lr_dev != nullptr) lr = *lr_dev;
    if (step_dev != nullptr) {
        const float t = (float)*step_dev;
        bias1 = 1.f - powf(beta1, t);
        bias2_sqrt = sqrtf(1.f - powf(beta2, t));
    }
    const int2 bm = blockmap[blockIdx.x];
    const TensorEntry e = tab[bm.x];
    const long long base = (long long)bm.y * OPT_CHUNK;
    const long long end = min(e.n, base + OPT_CHUNK);
    const float step = lr / bias1;
    for (long long i = base + threadIdx.x; i < end; i += blockDim.x) {
        float p = e.p[i];
        const float g = e.g[i] * grad_scale + weight_decay * p;
        const float m = beta1 * e.s1[i] + (1.f - beta1) * g;
        const float v = beta2 * e.s2[i] + (1.f - beta2) * g * g;
        e.s1[i] = m;
        e.s2[i] = v;
        p -= step * m / (sqrtf(v) / bias2_sqrt + eps);
        e.p[i] = p;
        if (e.shadow != nullptr) e.shadow[i] = __float2bfloat16(p);
    }
}

__global__ void __launch_bounds__(256) multi_cast_kernel(const TensorEntry* __restrict__ tab,
                                                         const int2* __restrict__ blockmap) {
    const int2 bm = blockmap[blockIdx.x];
    const TensorEntry e = tab[bm.x];
    const long long base = (long long)bm.y * OPT_CHUNK;
    const long long end = min(e.n, base + OPT_CHUNK);
    for (long long i = base + threadIdx.x; i < end; i += blockDim.x) e.shadow[i] = __float2bfloat16(e.p[i]);
}

__global__ void __launch_bounds__(256) multi_zero_kernel(const TensorEntry* __restrict__ tab,
                                                         const int2* __restrict__ blockmap) {
    const int2 bm = blockmap[blockIdx.x];
    const TensorEntry e = tab[bm.x];
    const long long base = (long long)bm.y * OPT_CHUNK;
    const long long end = min(e.n, base + OPT_CHUNK);
    for (long long i = base + threadIdx.x; i < end; i += blockDim.x) e.g[i] = 0.f;
}

}  // namespace

extern "C" int ptd_softmax_ce(const float* logits, const float* soft_targets, const long long* index_targets,
                              float* loss_sum, void* dlogits, int dlogits_is_bf16, int rows, int C, int ldd,
                              float scale, void* stream) {
    PTD_REQUIRE((soft_targets != nullptr) != (index_targets != nullptr), "softmax_ce: give exactly one target kind");
    PTD_REQUIRE(ldd >= C, "softmax_ce: ldd (%d) < C (%d)", ldd, C);
    cudaStream_t st = (cudaStream_t)stream;
    if (soft_targets) {
        if (dlogits_is_bf16)
            softmax_ce_kernel<true, bf16><<<rows, 256, 0, st>>>(logits, soft_targets, nullptr, loss_sum, (bf16*)dlogits, C, ldd, scale);
        else
            softmax_ce_kernel<true, float><<<rows, 256, 0, st>>>(logits, soft_targets, nullptr, loss_sum, (float*)dlogits, C, ldd, scale);
    } else {
        if (dlogits_is_bf16)
            softmax_ce_kernel<false, bf16><<<rows, 256, 0, st>>>(logits, nullptr, index_targets, loss_sum, (bf16*)dlogits, C, ldd, scale);
        else
            softmax_ce_kernel<false, float><<<rows, 256, 0, st>>>(logits, nullptr, index_targets, loss_sum, (float*)dlogits, C, ldd, scale);
    }
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

// `table` is a device array of n_blocks-independent TensorEntry records (6 x 8 bytes each) and
// `blockmap` a device array of (tensor index, chunk index) pairs, one per block; both are built
// once on the host side (pytorchdistributed_b200/ops/optim.py) and reused every step.
extern "C" int ptd_multi_sgd(const void* table, const void* blockmap, int n_blocks, float lr, float momentum,
                             float weight_decay, float grad_scale, int first_step, void* stream) {
    if (n_blocks <= 0) return PTD_OK;
    multi_sgd_kernel<<<n_blocks, 256, 0, (cudaStream_t)stream>>>((const TensorEntry*)table, (const int2*)blockmap, lr,
                                                                 momentum, weight_decay, grad_scale, first_step, nullptr);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
// Same, with the learning rate in device memory (lr_dev: one float) - what a CUDA-graph-captured step uses.
extern "C" int ptd_multi_sgd_dev(const void* table, const void* blockmap, int n_blocks, const float* lr_dev,
                                 float momentum, float weight_decay, float grad_scale, int first_step, void* stream) {
    if (n_blocks <= 0) return PTD_OK;
    PTD_REQUIRE(lr_dev != nullptr, "multi_sgd_dev: lr_dev is required");
    multi_sgd_kernel<<<n_blocks, 256, 0, (cudaStream_t)stream>>>((const TensorEntry*)table, (const int2*)blockmap, 0.f,
                                                                 momentum, weight_decay, grad_scale, first_step, lr_dev);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_multi_adam(const void* table, const void* blockmap, int n_blocks, float lr, float beta1, float beta2,
                              float eps, float weight_decay, int step, float grad_scale, void* stream) {
    if (n_blocks <= 0) return PTD_OK;
    const float bias1 = 1.f - powf(beta1, (float)step);
    const float bias2_sqrt = sqrtf(1.f - powf(beta2, (float)step));
    multi_adam_kernel<<<n_blocks, 256, 0, (cudaStream_t)stream>>>((const TensorEntry*)table, (const int2*)blockmap, lr,
                                                                  beta1, beta2, eps, weight_decay, bias1, bias2_sqrt,
                                                                  grad_scale, nullptr, nullptr);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
// Adam with its step count (step_dev: one int, incremented here before the update) and learning rate (lr_dev: one
// float) in device memory: replaying a captured step advances the bias corrections like an eager step does.
extern "C" int ptd_multi_adam_dev(const void* table, const void* blockmap, int n_blocks, const float* lr_dev, float beta1,
                                  float beta2, float eps, float weight_decay, int* step_dev, float grad_scale,
                                  void* stream) {
    PTD_REQUIRE(lr_dev != nullptr && step_dev != nullptr, "multi_adam_dev: lr_dev and step_dev are required");
    adam_advance_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step_dev);
    PTD_CUDA_CHECK(cudaGetLastError());
    if (n_blocks <= 0) return PTD_OK;
    multi_adam_kernel<<<n_blocks, 256, 0, (cudaStream_t)stream>>>((const TensorEntry*)table, (const int2*)blockmap, 0.f,
                                                                  beta1, beta2, eps, weight_decay, 1.f, 1.f, grad_scale,
                                                                  step_dev, lr_dev);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_multi_cast_bf16(const void* table, const void* blockmap, int n_blocks, void* stream) {
    if (n_blocks <= 0) return PTD_OK;
    multi_cast_kernel<<<n_blocks, 256, 0, (cudaStream_t)stream>>>((const TensorEntry*)table, (const int2*)blockmap);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_multi_zero_grad(const void* table, const void* blockmap, int n_blocks, void* stream) {
    if (n_blocks <= 0) return PTD_OK;
    multi_zero_kernel<<<n_blocks, 256, 0, (cudaStream_t)stream>>>((const TensorEntry*)table, (const int2*)blockmap);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_opt_chunk_elems() { return OPT_CHUNK; }
