// Native executors for a whole residual bottleneck: ONE call launches every kernel of the block's forward
// (4-8 launches) or backward (12-16 launches) on the caller's stream.
//
// Why: at small per-stage work (pipeline micro-batches, late ResNet stages) the step is bound by the host issuing
// launches; driving ~22 kernels of a block from Python costs ~0.6 ms, from here a few microseconds.  The reference
// pays the same overhead through torch's module-by-module eager dispatch (R/03 模型并行/03_model_parallel.ipynb#cell12;
// its split_size=1 run takes 4.87 s) - SURVEY.md 3.5 names launch overhead as the pipeline's real bound.
//
// The caller (ops/functional.py: _Bottleneck) owns all memory: it passes the activations to keep for the backward,
// and scratch / gradient buffers that are zeroed here with cudaMemsetAsync.
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "common.cuh"
#include "ptdb200.h"

namespace {

struct UnitGeom {
    int cin, cout, k, stride, pad, h_in, w_in, h_out, w_out;
    long long m_out;
};

// conv1 1x1/1, conv2 3x3/stride, conv3 1x1/1, downsample 1x1/stride (on the block input)
void geometry(int N, int H, int W, int inplanes, int planes, int stride, UnitGeom g[4]) {
    const int OH = (H + 2 - 3) / stride + 1, OW = (W + 2 - 3) / stride + 1;
    g[0] = {inplanes, planes, 1, 1, 0, H, W, H, W, (long long)N * H * W};
    g[1] = {planes, planes, 3, stride, 1, H, W, OH, OW, (long long)N * OH * OW};
    g[2] = {planes, planes * 4, 1, 1, 0, OH, OW, OH, OW, (long long)N * OH * OW};
    g[3] = {inplanes, planes * 4, 1, stride, 0, H, W, OH, OW, (long long)N * OH * OW};
}

// One side stream + a ring of events per device: the weight-gradient kernels of a block only depend on the BN
// backward of their own unit, so they run beside the data-gradient chain and fill the tails of the small
// late-stage GEMMs (98 tiles on 148 SMs).  Fork/join is event-based, so it is also recorded when the step is
// being captured into a CUDA graph.
struct SideStream {
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[16] = {nullptr};
    int next = 0;
};
SideStream* side_stream() {
    static SideStream per_dev[64];
    static std::mutex mu;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    std::lock_guard<std::mutex> lk(mu);
    SideStream& s = per_dev[dev];
    if (s.stream == nullptr) {
        // PTDB200_SIDE_PRIO=1: lowest priority, so that CTAs of the data-gradient chain are placed first when both
        // streams have work pending
        const char* e = getenv("PTDB200_SIDE_PRIO");
        int lo = 0, hi = 0;
        if (e && atoi(e) == 1 && cudaDeviceGetStreamPriorityRange(&lo, &hi) == cudaSuccess) {
            if (cudaStreamCreateWithPriority(&s.stream, cudaStreamNonBlocking, lo) != cudaSuccess) return nullptr;
        } else if (cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) != cudaSuccess) {
            return nullptr;
        }
        for (auto& e : s.ev)
            if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    }
    return &s;
}
bool dual_bn_enabled() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("PTDB200_DUAL_BN"); v = e ? atoi(e) : 1; }
    return v == 1;
}
// BN-backward reduce pass of conv1's / conv2's BatchNorm fused into the data-gradient kernel that produces their dz
// (ptd_conv_dgrad_fused; the y / mask operands are prefetched one chunk ahead).  Measured per layer at batch 256
// (profiles/fused_bn_reduce_experiment.txt): it saves 14-24 us where the kernel runs 64-wide tiles (layer1: 8 epilogue
// warps, registers to spare for the prefetch) and on the stride-1 3x3 of layer2, and LOSES 2-94 us everywhere else (the
// 128 / 256-wide kernels spill under their 96-register cap, the strided 3x3 runs four short launches).  So:
// In the whole step the per-layer gains do not show (18.66-18.67 ms with the selective fusion, 18.65-18.67 ms without, same
// box, 30 steps each, twice): the reduce kernels it removes were overlapping with the weight-gradient side stream.  So:
//   PTDB200_FUSE_BN_REDUCE = 0 (default) never, 1 only where a layer was measured to win, 2 everywhere (A/B testing).
int fuse_bn_reduce_mode() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("PTDB200_FUSE_BN_REDUCE"); v = e ? atoi(e) : 0; }
    return v;
}
// `ch` = channels of the gradient tensor (the BatchNorm's channels), `k`/`stride` = the data-gradient convolution
bool fuse_bn_reduce_here(int ch, int k, int stride) {
    const int mode = fuse_bn_reduce_mode();
    if (mode <= 0) return false;
    if (mode >= 2) return true;
    return ch == 64 ? (stride == 1) : (ch == 128 && k == 3 && stride == 1);
}
bool overlap_enabled() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("PTDB200_WGRAD_OVERLAP"); v = e ? atoi(e) : 1; }
    return v == 1;
}

// dst[k][i] += src[off_k + i]: adds the per-unit [dbeta | dgamma] sums of one block into up to 8 gradient tensors
// (DDP bucket views) with one launch.
struct ScatterAdd {
    const float* src;
    float* dst[8];
    int off[8];
    int n[8];
};
__global__ void __launch_bounds__(256) scatter_add_kernel(ScatterAdd s) {
    const int k = blockIdx.y;
    float* dst = s.dst[k];
    if (dst == nullptr) return;
    const float* src = s.src + s.off[k];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < s.n[k]; i += gridDim.x * blockDim.x) dst[i] += src[i];
}

}  // namespace

extern "C" int ptd_bottleneck_fwd(const PtdBottleneckFwd* d, void* stream) {
    PTD_REQUIRE(d != nullptr, "bottleneck_fwd: null descriptor");
    UnitGeom g[4];
    geometry(d->N, d->H, d->W, d->inplanes, d->planes, d->stride, g);
    const int nu = d->has_ds ? 4 : 3;
    int off[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < nu; ++i) off[i + 1] = off[i] + 2 * g[i].cout;
    PTD_CUDA_CHECK(cudaMemsetAsync(d->stats, 0, (size_t)off[nu] * sizeof(float), (cudaStream_t)stream));
    // order: conv1, conv2, (downsample), conv3 - conv3's BN consumes the identity / downsample output
    const int order[4] = {0, 1, 3, 2};
    for (int oi = 0; oi < 4; ++oi) {
        const int i = order[oi];
        if (i == 3 && !d->has_ds) continue;
        const void* in = (i == 0 || i == 3) ? d->x : d->z[i - 1];
        float* st = d->stats + off[i];
        float* sv = d->saved + off[i];
        int rc = ptd_conv_fprop(in, d->w[i], d->y[i], st, st + g[i].cout, d->N, g[i].h_in, g[i].w_in, g[i].cin, g[i].cout,
                                g[i].k, g[i].k, g[i].stride, g[i].pad, stream);
        if (rc) return rc;
        if (i == 3 && dual_bn_enabled()) continue;          // the shortcut's BN is applied inside conv3's BN below
        if (i == 2 && d->has_ds && dual_bn_enabled()) {
            float* st3 = d->stats + off[3];
            float* sv3 = d->saved + off[3];
            rc = ptd_bn_apply_dual(d->y[2], d->y[3], d->z[2], d->mask[2], st, st + g[2].cout, d->gamma[2], d->beta[2],
                                   d->running_mean[2], d->running_var[2], sv, sv + g[2].cout, st3, st3 + g[3].cout,
                                   d->gamma[3], d->beta[3], d->running_mean[3], d->running_var[3], sv3, sv3 + g[3].cout,
                                   g[2].m_out, g[2].cout, d->eps, d->momentum, stream);
            if (rc) return rc;
            continue;
        }
        const void* residual = nullptr;
        int relu = 1;
        if (i == 2) residual = d->has_ds ? d->z[3] : d->x;
        if (i == 3) relu = 0;
        rc = ptd_bn_apply_mask(d->y[i], residual, d->z[i], d->mask[i], st, st + g[i].cout, d->gamma[i], d->beta[i],
                               d->running_mean[i], d->running_var[i], sv, sv + g[i].cout, g[i].m_out, g[i].cout, d->eps,
                               d->momentum, relu, stream);
        if (rc) return rc;
    }
    return PTD_OK;
}

extern "C" int ptd_bottleneck_bwd(const PtdBottleneckBwd* d, void* stream) {
    PTD_REQUIRE(d != nullptr, "bottleneck_bwd: null descriptor");
    UnitGeom g[4];
    geometry(d->N, d->H, d->W, d->inplanes, d->planes, d->stride, g);
    const int nu = d->has_ds ? 4 : 3;
    int off[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < nu; ++i) off[i + 1] = off[i] + 2 * g[i].cout;
    cudaStream_t st = (cudaStream_t)stream;
    PTD_CUDA_CHECK(cudaMemsetAsync(d->sums, 0, (size_t)off[nu] * sizeof(float), st));
    if (d->dw_all_bytes > 0) PTD_CUDA_CHECK(cudaMemsetAsync(d->dw_all, 0, (size_t)d->dw_all_bytes, st));
    int rc;
#define BN_BWD(i, dz_, relu_, gout_)                                                                                   \
    do {                                                                                                               \
        float* sm = d->sums + off[i];                                                                                  \
        const float* sv = d->saved + off[i];                                                                           \
        const bool bits = relu_ && d->mask[i] != nullptr;    /* 1 bit per element instead of re-reading z */           \
        const void* zm = bits ? d->mask[i] : d->z[i];                                                                  \
        const int mode = bits ? 2 : relu_;                                                                             \
        rc = ptd_bn_bwd_reduce(dz_, zm, d->y[i], sv, sv + g[i].cout, sm, sm + g[i].cout, g[i].m_out, g[i].cout,        \
                               mode, stream);                                                                          \
        if (rc) return rc;                                                                                             \
        rc = ptd_bn_bwd_apply(dz_, zm, d->y[i], sv, sv + g[i].cout, d->gamma[i], sm, sm + g[i].cout, d->dy[i],        \
                              gout_, g[i].m_out, g[i].cout, mode, stream);                                             \
        if (rc) return rc;                                                                                             \
    } while (0)
    /* BatchNorm backward of unit i with its reduce pass already done by the kernel that produced dz_ (sums are final) */
#define BN_APPLY(i, dz_)                                                                                               \
    do {                                                                                                               \
        float* sm = d->sums + off[i];                                                                                  \
        const float* sv = d->saved + off[i];                                                                           \
        rc = ptd_bn_bwd_apply(dz_, d->mask[i], d->y[i], sv, sv + g[i].cout, d->gamma[i], sm, sm + g[i].cout, d->dy[i], \
                              nullptr, g[i].m_out, g[i].cout, 2, stream);                                              \
        if (rc) return rc;                                                                                             \
    } while (0)
    /* data gradient of unit i into dx_ = the gradient behind unit j's BatchNorm + ReLU, with unit j's reduce pass fused
       into its epilogue; falls back to the separate kernels where the fused variant does not apply */
#define DGRAD_BN(i, dx_, j)                                                                                            \
    do {                                                                                                               \
        bool fused = false;                                                                                            \
        if (d->mask[j] != nullptr && fuse_bn_reduce_here(g[i].cin, g[i].k, g[i].stride)) {                                                                        \
            float* sm = d->sums + off[j];                                                                              \
            const float* sv = d->saved + off[j];                                                                       \
            rc = ptd_conv_dgrad_fused(d->dy[i], d->w[i], dx_, nullptr, nullptr, d->y[j], d->mask[j], sv, sv + g[j].cout, \
                                      sm, sm + g[j].cout, d->N, g[i].h_in, g[i].w_in, g[i].cin, g[i].cout, g[i].k,      \
                                      g[i].k, g[i].stride, g[i].pad, stream);                                          \
            if (rc == PTD_OK) fused = true;                                                                            \
            else if (rc != PTD_ERR_UNSUPPORTED) return rc;                                                             \
        }                                                                                                              \
        if (!fused) DGRAD(i, dx_, nullptr);                                                                            \
        fused_unit[j] = fused;                                                                                         \
    } while (0)
#define DGRAD(i, dx_, addend_)                                                                                         \
    do {                                                                                                               \
        rc = ptd_conv_dgrad(d->dy[i], d->w[i], dx_, addend_, d->N, g[i].h_in, g[i].w_in, g[i].cin, g[i].cout, g[i].k,   \
                            g[i].k, g[i].stride, g[i].pad, stream);                                                    \
        if (rc) return rc;                                                                                             \
    } while (0)
#define WGRAD(i, in_)                                                                                                  \
    do {                                                                                                               \
        void* ws = stream;                                                                                             \
        if (side != nullptr) {      /* fork: the side stream waits for everything issued so far (dy[i] is ready) */     \
            cudaEvent_t e = side->ev[side->next++ & 15];                                                               \
            PTD_CUDA_CHECK(cudaEventRecord(e, st));                                                                    \
            PTD_CUDA_CHECK(cudaStreamWaitEvent(side->stream, e, 0));                                                   \
            ws = (void*)side->stream;                                                                                  \
        }                                                                                                              \
        rc = ptd_conv_wgrad(in_, d->dy[i], d->dw[i], d->N, g[i].h_in, g[i].w_in, g[i].cin, g[i].cout, g[i].k, g[i].k,   \
                            g[i].stride, g[i].pad, ws);                                                                \
        if (rc) return rc;                                                                                             \
    } while (0)

    SideStream* side = overlap_enabled() ? side_stream() : nullptr;
    // The identity-branch gradient is dz behind the block's final ReLU.  With the forward's bit mask it is taken
    // from dz where it is needed (gated addend of conv1's data gradient, or the shortcut BN's backward in mask
    // mode); without it the last BN's backward writes it out as g_id.
    static int gate_on = -1;
    if (gate_on < 0) { const char* e = getenv("PTDB200_GATE_IDENTITY"); gate_on = e ? atoi(e) : 1; }
    const bool gate = d->mask[2] != nullptr && gate_on;
    const bool dual = gate && d->has_ds && dual_bn_enabled();     // last BN and shortcut BN share g = dz * mask
    if (dual) {
        const float* sv2 = d->saved + off[2];
        const float* sv3 = d->saved + off[3];
        rc = ptd_bn_bwd_dual(d->dz, d->mask[2], d->y[2], d->y[3], sv2, sv2 + g[2].cout, d->gamma[2], sv3, sv3 + g[3].cout,
                             d->gamma[3], d->sums + off[2], d->sums + off[3], d->dy[2], d->dy[3], g[2].m_out, g[2].cout,
                             stream);
        if (rc) return rc;
    } else {
        BN_BWD(2, d->dz, 1, gate ? nullptr : d->g_id);
    }
    bool fused_unit[4] = {false, false, false, false};
    DGRAD_BN(2, d->dz2, 1);                 // dz2 = gradient behind bn2 + ReLU: its reduce pass rides on this kernel
    WGRAD(2, d->z[1]);
    if (fused_unit[1]) BN_APPLY(1, d->dz2);
    else BN_BWD(1, d->dz2, 1, nullptr);
    DGRAD_BN(1, d->dz1, 0);                 // likewise bn1's on the 3x3 data gradient
    WGRAD(1, d->z[0]);
    if (fused_unit[0]) BN_APPLY(0, d->dz1);
    else BN_BWD(0, d->dz1, 1, nullptr);
    if (d->has_ds) {
        if (dual) {
            /* the shortcut BN's backward ran together with the last BN's */
        } else if (gate) {
            float* sm = d->sums + off[3];
            const float* sv = d->saved + off[3];
            rc = ptd_bn_bwd_reduce(d->dz, d->mask[2], d->y[3], sv, sv + g[3].cout, sm, sm + g[3].cout, g[3].m_out,
                                   g[3].cout, 2, stream);
            if (rc) return rc;
            rc = ptd_bn_bwd_apply(d->dz, d->mask[2], d->y[3], sv, sv + g[3].cout, d->gamma[3], sm, sm + g[3].cout,
                                  d->dy[3], nullptr, g[3].m_out, g[3].cout, 2, stream);
            if (rc) return rc;
        } else {
            BN_BWD(3, d->g_id, 0, nullptr);
        }
        WGRAD(3, d->x);
        DGRAD(0, d->dx, nullptr);
        // the shortcut convolution adds into dx in place (stride 2 touches only the strided quarter)
        rc = ptd_conv_dgrad_accum(d->dy[3], d->w[3], d->dx, d->N, g[3].h_in, g[3].w_in, g[3].cin, g[3].cout, 1, 1,
                                  g[3].stride, 0, stream);
        if (rc) return rc;
    } else if (gate) {
        rc = ptd_conv_dgrad_masked_addend(d->dy[0], d->w[0], d->dx, d->dz, d->mask[2], d->N, g[0].h_in, g[0].w_in, g[0].cin,
                                          g[0].cout, stream);
        if (rc) return rc;
    } else {
        DGRAD(0, d->dx, d->g_id);
    }
    WGRAD(0, d->x);
    {   // every BN backward of the block has run: hand dbeta / dgamma to their gradient tensors
        ScatterAdd sa;
        bool any = false;
        sa.src = d->sums;
        for (int i = 0; i < 4; ++i)
            for (int j = 0; j < 2; ++j) {
                const int k = 2 * i + j;
                sa.dst[k] = i < nu ? d->gdst[k] : nullptr;
                sa.off[k] = i < nu ? off[i] + j * g[i].cout : 0;
                sa.n[k] = i < nu ? g[i].cout : 0;
                any = any || sa.dst[k] != nullptr;
            }
        if (any) {
            scatter_add_kernel<<<dim3(2, 8), 256, 0, st>>>(sa);
            PTD_CUDA_CHECK(cudaGetLastError());
        }
    }
    if (side != nullptr) {          // join: the caller's stream continues only after the last weight gradient
        cudaEvent_t e = side->ev[side->next++ & 15];
        PTD_CUDA_CHECK(cudaEventRecord(e, side->stream));
        PTD_CUDA_CHECK(cudaStreamWaitEvent(st, e, 0));
    }
#undef BN_BWD
#undef BN_APPLY
#undef DGRAD_BN
#undef DGRAD
#undef WGRAD
    return PTD_OK;
}
