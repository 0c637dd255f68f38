// CUDA-core kernels for the tutorial-sized models: Linear(20,1) of the DDP scripts
// (R/02 DDP基本概念/ddp_gpus.py:77) and the 10-20-20-20-5 MLP of the DataParallel notebook
// (R/01 数据并行/01_multi_gpus_data_parallelism.ipynb#cell4).  At K <= 64 a tensor-core tile would
// be almost empty, so these stay on fp32 FMA with the weights held in shared memory.
#include "common.cuh"
#include "ptdb200.h"

namespace {

constexpr int SMALL_MAX = 64;   // max in/out features handled by the small-linear kernels

// y[B][N] = act(x[B][K] * w[N][K]^T + b)
__global__ void __launch_bounds__(128) small_linear_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                               const float* __restrict__ b, float* __restrict__ y,
                                                               int B, int K, int N, int relu) {
    __shared__ float sw[SMALL_MAX * SMALL_MAX];
    __shared__ float sb[SMALL_MAX];
    for (int i = threadIdx.x; i < N * K; i += blockDim.x) sw[i] = w[i];
    for (int i = threadIdx.x; i < N; i += blockDim.x) sb[i] = b ? b[i] : 0.f;
    __syncthreads();
    for (int row = blockIdx.x * blockDim.x + threadIdx.x; row < B; row += gridDim.x * blockDim.x) {
        float xr[SMALL_MAX];
        for (int k = 0; k < K; ++k) xr[k] = x[(size_t)row * K + k];
        for (int n = 0; n < N; ++n) {
            float acc = sb[n];
            for (int k = 0; k < K; ++k) acc = fmaf(xr[k], sw[n * K + k], acc);
            if (relu) acc = fmaxf(acc, 0.f);
            y[(size_t)row * N + n] = acc;
        }
    }
}

// dx[B][K] = dy[B][N] * w[N][K]  (dy already masked by the caller for ReLU), one thread per row.
__global__ void __launch_bounds__(128) small_linear_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ w,
                                                                 float* __restrict__ dx, int B, int K, int N) {
    __shared__ float sw[SMALL_MAX * SMALL_MAX];
    for (int i = threadIdx.x; i < N * K; i += blockDim.x) sw[i] = w[i];
    __syncthreads();
    for (int row = blockIdx.x * blockDim.x + threadIdx.x; row < B; row += gridDim.x * blockDim.x) {
        float acc[SMALL_MAX];
        for (int k = 0; k < K; ++k) acc[k] = 0.f;
        for (int n = 0; n < N; ++n) {
            const float d = dy[(size_t)row * N + n];
            for (int k = 0; k < K; ++k) acc[k] = fmaf(d, sw[n * K + k], acc[k]);
        }
        for (int k = 0; k < K; ++k) dx[(size_t)row * K + k] = acc[k];
    }
}

// dw[N][K] += sum_rows dy[row][n] * x[row][k];  db[N] += sum_rows dy[row][n].  One block handles a slab of
// rows, thread (n,k) pairs accumulate in registers, then one atomic per pair per block.
__global__ void __launch_bounds__(256) small_linear_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                                 float* __restrict__ dw, float* __restrict__ db, int B,
                                                                 int K, int N, int rows_per_block) {
    extern __shared__ float sh[];     // [rows][K] x, then [rows][N] dy
    const int r0 = blockIdx.x * rows_per_block;
    const int nr = min(rows_per_block, B - r0);
    if (nr <= 0) return;
    float* sx = sh;
    float* sdy = sh + rows_per_block * K;
    for (int i = threadIdx.x; i < nr * K; i += blockDim.x) sx[i] = x[(size_t)r0 * K + i];
    for (int i = threadIdx.x; i < nr * N; i += blockDim.x) sdy[i] = dy[(size_t)r0 * N + i];
    __syncthreads();
    for (int pair = threadIdx.x; pair < N * K; pair += blockDim.x) {
        const int n = pair / K, k = pair - n * K;
        float acc = 0.f;
        for (int r = 0; r < nr; ++r) acc = fmaf(sdy[r * N + n], sx[r * K + k], acc);
        atomicAdd(dw + pair, acc);
    }
    if (db != nullptr)
        for (int n = threadIdx.x; n < N; n += blockDim.x) {
            float acc = 0.f;
            for (int r = 0; r < nr; ++r) acc += sdy[r * N + n];
            atomicAdd(db + n, acc);
        }
}

// Whole MLP forward in one kernel: up to 8 layers, every width <= 64, ReLU between layers.
struct MlpDesc {
    const float* w[8];
    const float* b[8];
    int dims[9];
    int n_layers;
};
__global__ void __launch_bounds__(128) mlp_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, MlpDesc d,
                                                      int B) {
    extern __shared__ float sh[];     // all weights and biases back to back
    int off = 0;
    for (int l = 0; l < d.n_layers; ++l) {
        const int nw = d.dims[l] * d.dims[l + 1];
        for (int i = threadIdx.x; i < nw; i += blockDim.x) sh[off + i] = d.w[l][i];
        off += nw;
        for (int i = threadIdx.x; i < d.dims[l + 1]; i += blockDim.x) sh[off + i] = d.b[l] ? d.b[l][i] : 0.f;
        off += d.dims[l + 1];
    }
    __syncthreads();
    for (int row = blockIdx.x * blockDim.x + threadIdx.x; row < B; row += gridDim.x * blockDim.x) {
        float a[SMALL_MAX], t[SMALL_MAX];
        for (int k = 0; k < d.dims[0]; ++k) a[k] = x[(size_t)row * d.dims[0] + k];
        int o = 0;
        for (int l = 0; l < d.n_layers; ++l) {
            const int K = d.dims[l], N = d.dims[l + 1];
            const float* w = sh + o;
            const float* b = sh + o + K * N;
            for (int n = 0; n < N; ++n) {
                float acc = b[n];
                for (int k = 0; k < K; ++k) acc = fmaf(a[k], w[n * K + k], acc);
                t[n] = (l + 1 < d.n_layers) ? fmaxf(acc, 0.f) : acc;
            }
            for (int n = 0; n < N; ++n) a[n] = t[n];
            o += K * N + N;
        }
        const int NO = d.dims[d.n_layers];
        for (int n = 0; n < NO; ++n) y[(size_t)row * NO + n] = a[n];
    }
}

}  // namespace

extern "C" int ptd_small_linear_fwd(const float* x, const float* w, const float* b, float* y, int B, int K, int N,
                                    int relu, void* stream) {
    PTD_REQUIRE(K >= 1 && N >= 1 && K <= SMALL_MAX && N <= SMALL_MAX, "small_linear: K (%d), N (%d) must be in [1,%d]", K, N, SMALL_MAX);
    const int grid = max(1, min((B + 127) / 128, ptd_num_sms() * 4));
    small_linear_fwd_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(x, w, b, y, B, K, N, relu);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
extern "C" int ptd_small_linear_bwd(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db,
                                    int B, int K, int N, void* stream) {
    PTD_REQUIRE(K >= 1 && N >= 1 && K <= SMALL_MAX && N <= SMALL_MAX, "small_linear: K (%d), N (%d) must be in [1,%d]", K, N, SMALL_MAX);
    cudaStream_t st = (cudaStream_t)stream;
    if (dx != nullptr) {
        const int grid = max(1, min((B + 127) / 128, ptd_num_sms() * 4));
        small_linear_dgrad_kernel<<<grid, 128, 0, st>>>(dy, w, dx, B, K, N);
        PTD_CUDA_CHECK(cudaGetLastError());
    }
    if (dw != nullptr) {
        const int rows_per_block = 64;
        const int grid = (B + rows_per_block - 1) / rows_per_block;
        const size_t smem = (size_t)rows_per_block * (K + N) * sizeof(float);
        small_linear_wgrad_kernel<<<grid, 256, smem, st>>>(x, dy, dw, db, B, K, N, rows_per_block);
        PTD_CUDA_CHECK(cudaGetLastError());
    }
    return PTD_OK;
}
// weights/biases: host arrays of device pointers; dims: host array of n_layers+1 widths.
extern "C" int ptd_mlp_fwd(const float* x, float* y, const float* const* weights, const float* const* biases,
                           const int* dims, int n_layers, int B, void* stream) {
    PTD_REQUIRE(n_layers >= 1 && n_layers <= 8, "mlp_fwd: 1..8 layers supported, got %d", n_layers);
    MlpDesc d;
    size_t smem = 0;
    for (int l = 0; l <= n_layers; ++l) {
        PTD_REQUIRE(dims[l] >= 1 && dims[l] <= SMALL_MAX, "mlp_fwd: width %d out of [1,%d]", dims[l], SMALL_MAX);
        d.dims[l] = dims[l];
    }
    for (int l = 0; l < n_layers; ++l) {
        d.w[l] = weights[l];
        d.b[l] = biases ? biases[l] : nullptr;
        smem += (size_t)(dims[l] * dims[l + 1] + dims[l + 1]) * sizeof(float);
    }
    d.n_layers = n_layers;
    PTD_REQUIRE(smem <= 48 * 1024, "mlp_fwd: weights (%zu B) exceed 48 KiB of shared memory", smem);
    const int grid = max(1, min((B + 127) / 128, ptd_num_sms() * 4));
    mlp_fwd_kernel<<<grid, 128, smem, (cudaStream_t)stream>>>(x, y, d, B);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
