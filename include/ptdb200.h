/* pytorchdistributed_b200 - C ABI of the sm_100a kernel library (libptdb200.so).
 *
 * Every entry point returns 0 on success and a negative code on failure; ptd_last_error() gives
 * the message of the calling thread's last failure.  Pointers are device pointers unless a
 * comment says otherwise; `stream` is a cudaStream_t passed as void*.  Tensors are NHWC
 * ("channels last") bf16 unless stated.  Citations `R/...` name the place in the reference
 * repository (JoeyOL/PytorchDistributed) whose behaviour the entry point serves; the reference
 * itself has no native code and reaches these ops through PyTorch / cuDNN / cuBLAS / NCCL.
 */
#ifndef PTDB200_H
#define PTDB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* ptd_last_error(void);
int ptd_version(void);
int ptd_device_sm_count(void);
/* Programmatic dependent launch of the hot kernels (convolution and BatchNorm kernels start their prologue while the
 * previous kernel of the stream drains; each waits with griddepcontrol.wait before its first global access).
 * 0 = off, 2 = on (default, or PTDB200_PDL), 1 = on + early release of dependents by the GEMM kernels; returns the
 * previous mode. */
int ptd_set_pdl_mode(int mode);

/* ---- convolution as implicit GEMM on tcgen05/TMEM (R/03 模型并行/03_model_parallel.ipynb#cell4-5:
 *      every Conv2d of ResNet(Bottleneck,[3,4,6,3]); cuDNN in the reference) --------------------- */
/* y[N,OH,OW,Cout] = conv(x[N,H,W,Cin], w[Cout,R,S,Cin]);  stat_sum/stat_sq (optional, fp32[Cout],
 * caller-zeroed) receive per-channel sum and sum of squares of the fp32 accumulators. */
int ptd_conv_fprop(const void* x, const void* w, void* y, float* stat_sum, float* stat_sq, int N, int H, int W,
                   int Cin, int Cout, int R, int S, int stride, int pad, void* stream);
/* dx[N,H,W,Cin] = conv_transpose(dy[N,OH,OW,Cout], w) (+ addend[N,H,W,Cin] if not NULL). */
int ptd_conv_dgrad(const void* dy, const void* w, void* dx, const void* addend, int N, int H, int W, int Cin,
                   int Cout, int R, int S, int stride, int pad, void* stream);
/* 1x1 stride-1 data gradient whose addend passes a ReLU first: dx = dy*W + (mask bit ? addend : 0), mask = the
 * u8 bit tensor of ptd_bn_apply_mask over dx's shape (a residual block's identity-branch gradient, taken from dz) */
int ptd_conv_dgrad_masked_addend(const void* dy, const void* w, void* dx, const void* addend, const void* mask, int N,
                                 int H, int W, int Cin, int Cout, void* stream);
/* dx += conv_transpose(dy, w), in place (the 1x1 stride-2 shortcut only touches the strided quarter of dx). */
/* Data gradient with an optional (mask-gated) addend AND the reduction pass of the BatchNorm backward that consumes dx
 * fused into its epilogue: sum_g[c] += sum(g), sum_gx[c] += sum(g * xhat), g = bn_mask ? dx : 0, xhat from bn_y / mean /
 * invstd (all NULL = plain data gradient).  Saves ptd_bn_bwd_reduce's second read of dx, y and the mask (torchvision
 * Bottleneck backward, notebook 03 cell 5).  PTD_ERR_UNSUPPORTED (-3), before any launch, for shapes on the gather fallback. */
int ptd_conv_dgrad_fused(const void* dy, const void* w, void* dx, const void* addend, const void* addend_mask,
                         const void* bn_y, const void* bn_mask, const float* bn_mean, const float* bn_invstd,
                         float* sum_g, float* sum_gx, int N, int H, int W, int Cin, int Cout, int R, int S, int stride,
                         int pad, void* stream);
int ptd_conv_dgrad_accum(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int R, int S,
                         int stride, int pad, void* stream);
/* dw[Cout,R,S,Cin] (fp32) += dy^T * im2col(x).  The caller zeroes dw (or keeps an accumulated value). */
int ptd_conv_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int R, int S,
                   int stride, int pad, void* stream);
/* 7x7/2 stem on a 4-channel NHWC image; w4 is [64][8][8][4] zero-padded, dw4 the same shape in fp32. */
int ptd_stem_fprop(const void* x4, const void* w4, void* y, float* stat_sum, float* stat_sq, int N, int H, int W,
                   void* stream);
int ptd_stem_wgrad(const void* x4, const void* dy, float* dw4, int N, int H, int W, void* stream);
/* The same stem on a space-to-depth input xs = bf16 [N][H/2+3][W/2+3][16] (ptd_nchw3_f32_to_s2d_bf16): a 4x4/1
 * convolution over 16 channels whose im2col rows are contiguous 128-byte windows, fetched by tiled TMA.
 * w16 / dw16 are [64][4][4][16] (ptd_stem_pack_weight_s2d).  Needs H % 16 == 0 and W % 32 == 0. */
int ptd_stem_s2d_fprop(const void* xs, const void* w16, void* y, float* stat_sum, float* stat_sq, int N, int H, int W,
                       void* stream);
int ptd_stem_s2d_wgrad(const void* xs, const void* dy, float* dw16, int N, int H, int W, void* stream);

/* ---- a whole residual bottleneck per call (conv1x1-BN-ReLU, conv3x3-BN-ReLU, conv1x1-BN (+identity | +downsample
 *      conv-BN) -ReLU; torchvision Bottleneck, R/03 模型并行/03_model_parallel.ipynb#cell5).  Index 0..2 = conv1..conv3,
 *      3 = downsample.  The caller owns every buffer; scratch / gradient buffers are zeroed by the call. ------------- */
typedef struct {
    int N, H, W, inplanes, planes, stride, has_ds;
    float eps, momentum;
    const void* x;                 /* block input  [N,H,W,inplanes] bf16 */
    const void* w[4];              /* bf16 weights [Cout][R][S][Cin] */
    const float* gamma[4];
    const float* beta[4];
    float* running_mean[4];
    float* running_var[4];
    void* y[4];                    /* conv outputs (kept for the backward) */
    void* z[4];                    /* BN(+ReLU) outputs; z[2] is the block output */
    float* stats;                  /* scratch, 2*sum(Cout) floats */
    float* saved;                  /* out: per unit [mean | invstd], 2*sum(Cout) floats */
    void* mask[4];                 /* out (optional, NULL = none): ReLU bit masks of units 0..2, u8[M][Cout/8] */
} PtdBottleneckFwd;
typedef struct {
    int N, H, W, inplanes, planes, stride, has_ds;
    const void* x;
    const void* dz;                /* gradient of the block output */
    const void* w[4];
    const float* gamma[4];
    const void* y[4];
    const void* z[4];
    const float* saved;
    float* sums;                   /* out: per unit [dbeta | dgamma], 2*sum(Cout) floats */
    void* dw_all;                  /* one buffer holding dw[0..3]; zeroed here */
    long long dw_all_bytes;
    float* dw[4];                  /* out: fp32 weight gradients [Cout][R][S][Cin] */
    void* dy[4];                   /* scratch: gradients w.r.t. the conv outputs */
    void* g_id;                    /* scratch: masked incoming gradient (identity-branch gradient) */
    void* dz2;                     /* scratch: gradient w.r.t. z[1] */
    void* dz1;                     /* scratch: gradient w.r.t. z[0] */
    void* dx;                      /* out: gradient w.r.t. x */
    const void* mask[4];           /* the forward's ReLU bit masks; a NULL entry recomputes the mask from z[i] */
    float* gdst[8];                /* optional: where unit i's dbeta (gdst[2i]) and dgamma (gdst[2i+1]) are ADDED at the end
                                      of the call (DDP gradient-bucket views); NULL entries are skipped.  With dw[] pointing
                                      into the buckets as well and dw_all_bytes == 0 the weight-gradient kernels accumulate
                                      straight into the buckets: no temporary, no AccumulateGrad pass (ddp_gpus.py:41). */
} PtdBottleneckBwd;
int ptd_bottleneck_fwd(const PtdBottleneckFwd* d, void* stream);
int ptd_bottleneck_bwd(const PtdBottleneckBwd* d, void* stream);

/* ---- dense layers (fc 2048->1000 of ResNet-50; cuBLAS in the reference) ----------------------- */
int ptd_linear_fwd(const void* a, const void* w, const float* bias, float* out, int M, int N, int K, void* stream);
int ptd_linear_dgrad(const void* dout, const void* w, void* da, int M, int N, int K, void* stream);
int ptd_linear_wgrad(const void* a, const void* dout, float* dw, int M, int N, int K, void* stream);

/* ---- BatchNorm (training) fused with ReLU / residual add, NHWC bf16, fp32 statistics
 *      (nn.BatchNorm2d of torchvision Bottleneck, R/03 模型并行/03_model_parallel.ipynb#cell5; cuDNN/ATen there) ---- */
/* per-channel sum / sum of squares of y[M][C] (when the producing convolution did not fuse them) */
int ptd_bn_stats(const void* y, float* sum, float* sq, long long M, int C, void* stream);
/* z = act(bn(y) [+ residual]); updates running stats (may be NULL), writes mean / invstd for the backward */
int ptd_bn_apply(const void* y, const void* residual, void* z, const float* sum, const float* sq, const float* gamma,
                 const float* beta, float* running_mean, float* running_var, float* save_mean, float* save_invstd,
                 long long M, int C, float eps, float momentum, int relu, void* stream);
/* same; with relu and a non-NULL mask also writes one bit per element (u8[M][C/8]; bit i of byte (row, g) is set when
 * element (row, 8g+i) is > 0 after the ReLU) for the backward passes to read in place of z */
int ptd_bn_apply_mask(const void* y, const void* residual, void* z, void* mask, const float* sum, const float* sq,
                      const float* gamma, const float* beta, float* running_mean, float* running_var, float* save_mean,
                      float* save_invstd, long long M, int C, float eps, float momentum, int relu, void* stream);
/* z = relu(bn(y) + bn2(y2)): last BN of a residual block fused with the BN of its conv shortcut (y2 = the shortcut
 * convolution's output); both sets of running statistics and saved mean / invstd are written */
int ptd_bn_apply_dual(const void* y, const void* y2, void* z, void* mask, const float* sum, const float* sq,
                      const float* gamma, const float* beta, float* running_mean, float* running_var, float* save_mean,
                      float* save_invstd, const float* sum2, const float* sq2, const float* gamma2, const float* beta2,
                      float* running_mean2, float* running_var2, float* save_mean2, float* save_invstd2, long long M,
                      int C, float eps, float momentum, void* stream);
/* backward of ptd_bn_apply_dual (both passes, dz and the mask read once for the two BatchNorms) */
int ptd_bn_bwd_dual(const void* dz, const void* mask, const void* y, const void* y2, const float* mean,
                    const float* invstd, const float* gamma, const float* mean2, const float* invstd2,
                    const float* gamma2, float* sums, float* sums2, void* dy, void* dy2, long long M, int C, void* stream);
/* z = act(y * scale[c] + shift[c] [+ residual])  (inference-mode BN) */
int ptd_affine_act(const void* y, const void* residual, void* z, const float* scale, const float* shift, long long M,
                   int C, int relu, void* stream);
/* backward pass 1: sum_g[c] += g, sum_gx[c] += g * xhat (caller-zeroed fp32[C] each), g = dz * (z > 0 if relu).
 * relu = 2 (both passes): `z` is the bit mask ptd_bn_apply_mask wrote, not the activation */
int ptd_bn_bwd_reduce(const void* dz, const void* z, const void* y, const float* mean, const float* invstd,
                      float* sum_g, float* sum_gx, long long M, int C, int relu, void* stream);
/* backward pass 2: dy from the sums; g_out (optional) = masked incoming gradient (identity-branch gradient) */
int ptd_bn_bwd_apply(const void* dz, const void* z, const void* y, const float* mean, const float* invstd,
                     const float* gamma, const float* sum_g, const float* sum_gx, void* dy, void* g_out, long long M,
                     int C, int relu, void* stream);
/* BN + ReLU + 3x3/2 max-pool (pad 1) fused: the ResNet stem tail (bn1, relu, maxpool of torchvision ResNet).  H, W even.
 * idx: u8 per pooled element, the winning tap 0..8 or 15 when the ReLU blocks the gradient.  The backward rebuilds the
 * pooled gradient on the fly in both BN passes; sums = [dbeta | dgamma], 2*C floats zeroed by the caller. */
int ptd_bn_relu_maxpool_fwd(const void* y, void* p, void* idx, const float* sum, const float* sq, const float* gamma,
                            const float* beta, float* running_mean, float* running_var, float* save_mean,
                            float* save_invstd, int N, int H, int W, int C, float eps, float momentum, void* stream);
int ptd_bn_relu_maxpool_bwd(const void* dp, const void* idx, const void* y, const float* mean, const float* invstd,
                            const float* gamma, float* sums, void* dy, int N, int H, int W, int C, void* stream);
/* BatchNorm backward (both passes) for callers that must not allocate: the 2*C-float scratch lives in a fixed array
 * inside the library image instead of being passed in.  C must be a power of two in [64, 2048] (the cap is what the
 * 16 KiB-per-GPU array holds), anything else is rejected; relu is 0 or 1 (z = the activation; no bit-mask mode).  The first
 * call on a device looks up the array's address - make it before any stream capture on that device.  All calls share
 * the array: overlapping calls on one device race, order them.  No counterpart in the reference (it has no native
 * code and never manages device memory itself). */
int ptd_bn_backward_static(const void* dz, const void* z, const void* y, const float* mean, const float* invstd,
                           const float* gamma, void* dy, void* g_out, float* dgamma, float* dbeta, long long M, int C,
                           int relu, void* stream);

/* ---- pooling / layout (nn.MaxPool2d(3,2,1), nn.AdaptiveAvgPool2d(1), `.to(device)` input staging) ---- */
int ptd_maxpool3x3s2_fwd(const void* x, void* y, void* argmax_u8, int N, int H, int W, int C, void* stream);
int ptd_maxpool3x3s2_bwd(const void* dy, const void* argmax_u8, void* dx, int N, int H, int W, int C, void* stream);
int ptd_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream);
int ptd_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream);
int ptd_nchw3_f32_to_nhwc4_bf16(const float* x, void* y, int N, int H, int W, void* stream);
int ptd_nchw_f32_to_nhwc_bf16(const float* x, void* y, int N, int C, int HW, void* stream);
int ptd_nhwc_bf16_to_nchw_f32(const void* x, float* y, int N, int C, int HW, void* stream);
int ptd_cast_f32_to_bf16(const float* x, void* y, long long n, void* stream);
int ptd_nchw3_f32_to_s2d_bf16(const float* x, void* xs, int N, int H, int W, void* stream);
int ptd_stem_pack_weight_s2d(const float* w, void* w16, long long sk, long long sc, long long sr, long long ss,
                             void* stream);
int ptd_stem_unpack_wgrad_s2d(const float* dw16, float* dw, long long sk, long long sc, long long sr, long long ss,
                              int accumulate, void* stream);
int ptd_stem_pack_weight(const float* w, void* w4, long long sk, long long sc, long long sr, long long ss, void* stream);
int ptd_stem_unpack_wgrad(const float* dw4, float* dw, long long sk, long long sc, long long sr, long long ss,
                          int accumulate, void* stream);

/* ---- loss and optimizers (F.cross_entropy with float targets, R/02 DDP基本概念/ddp_gpus.py:40;
 *      torch.optim.SGD(lr=1e-3) ddp_gpus.py:78; torch.optim.Adam(lr=1e-3) notebook 03 cell 7) ---- */
/* loss_sum += scale * sum_rows CE(logits, target); dlogits = scale * dCE/dlogits.  Exactly one target kind. */
int ptd_softmax_ce(const float* logits, const float* soft_targets, const long long* index_targets, float* loss_sum,
                   void* dlogits, int dlogits_is_bf16, int rows, int C, int ldd, float scale, void* stream);
/* table: device array of {param, grad, state1, state2, bf16 shadow, numel} records; blockmap: (tensor, chunk) per block */
int ptd_multi_sgd(const void* table, const void* blockmap, int n_blocks, float lr, float momentum, float weight_decay,
                  float grad_scale, int first_step, void* stream);
int ptd_multi_adam(const void* table, const void* blockmap, int n_blocks, float lr, float beta1, float beta2, float eps,
                   float weight_decay, int step, float grad_scale, void* stream);
/* The same with the learning rate (lr_dev: one float) and Adam's step count (step_dev: one int, incremented by the call)
 * in device memory, so that a training step captured into a CUDA graph keeps following the schedule and the bias
 * corrections keep decaying on replay (torch.optim.Adam(capturable=True); notebook 03 cell 7 builds an Adam per call). */
int ptd_multi_sgd_dev(const void* table, const void* blockmap, int n_blocks, const float* lr_dev, float momentum,
                      float weight_decay, float grad_scale, int first_step, void* stream);
int ptd_multi_adam_dev(const void* table, const void* blockmap, int n_blocks, const float* lr_dev, float beta1,
                       float beta2, float eps, float weight_decay, int* step_dev, float grad_scale, void* stream);
int ptd_multi_cast_bf16(const void* table, const void* blockmap, int n_blocks, void* stream);
int ptd_multi_zero_grad(const void* table, const void* blockmap, int n_blocks, void* stream);
int ptd_opt_chunk_elems(void);

/* ---- tutorial-sized dense layers, fp32 on CUDA cores (Linear(20,1) ddp_gpus.py:77; MLP notebook 01 cell 4) ---- */
int ptd_small_linear_fwd(const float* x, const float* w, const float* b, float* y, int B, int K, int N, int relu,
                         void* stream);
int ptd_small_linear_bwd(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db, int B,
                         int K, int N, void* stream);
/* weights / biases: host arrays of device pointers; dims: host array of n_layers + 1 widths */
int ptd_mlp_fwd(const float* x, float* y, const float* const* weights, const float* const* biases, const int* dims,
                int n_layers, int B, void* stream);

/* ---- peer-to-peer runtime and collectives over NVLink (replaces the NCCL calls behind
 *      DDP(model, device_ids=[gpu_id]) ddp_gpus.py:35,41 and nn.DataParallel notebook 01 cell 5-6) ---- */
int ptd_p2p_malloc(long long bytes, void** out);
int ptd_p2p_free(void* p);
int ptd_p2p_ipc_handle_size(void);
int ptd_p2p_ipc_get(void* p, void* handle_out);
int ptd_p2p_ipc_open(const void* handle, void** out);
int ptd_p2p_ipc_close(void* p);
int ptd_p2p_enable_peer_access(int dev, int peer);
int ptd_p2p_pad_bytes(void);
int ptd_p2p_max_blocks(void);
/* bufs / pads: host arrays of `world` device pointers (this process's mapping of every rank's buffer / signal pad).
 * In-place average (scale = 1/world) or any scaled sum; kernels sharing a pad must be stream-ordered. */
int ptd_allreduce_inplace(const void* const* bufs, const void* const* pads, int rank, int world, long long numel,
                          int is_bf16, float scale, int blocks, void* stream);
int ptd_broadcast_inplace(const void* const* bufs, const void* const* pads, int rank, int world, int root,
                          long long nbytes, int blocks, void* stream);
/* Sharded collectives on a symmetric buffer (shard s = rank s's part): op 0 all-gather, 1 gather to root, 2 scatter
 * from root, 3 reduce-scatter, 4 reduce to root, 5 one-shot all-reduce for messages up to 4 MiB (3-5: fp32, sum * scale).
 * NCCL's all_gather / gather / scatter / reduce_scatter / reduce in the reference's stack. */
int ptd_shard_collective(const void* const* bufs, const void* const* pads, int rank, int world, int root, int op,
                         long long nbytes, float scale, int blocks, void* stream);
int ptd_p2p_barrier(const void* const* pads, int rank, int world, void* stream);
/* single process, several devices: copy / sum over peer pointers from a kernel on the current device */
int ptd_peer_copy(void* dst, const void* src, long long nbytes, void* stream);
/* same through the bulk-copy engine (cp.async.bulk both ways); nbytes a multiple of 16 KiB */
int ptd_peer_copy_bulk(void* dst, const void* src, long long nbytes, int blocks, void* stream);
int ptd_reduce_sum_f32(float* dst, const void* const* srcs, int n_src, long long numel, float scale, int accumulate,
                       void* stream);

/* ---- symmetric memory on the VMM API + NVSwitch multicast (NVLS) mapping (csrc/vmm.cu).  What NCCL sets up for itself
 *      behind init_process_group("nccl") (ddp_gpus.py:22): registered peer buffers and a multicast object. ---- */
int ptd_vmm_caps(void);                                   /* bit 0: VMM + POSIX-fd export, bit 1: multicast */
int ptd_vmm_create(long long bytes, int rank, int world, long long* padded_out, int* fd_out);   /* -> block id > 0 */
int ptd_vmm_import_peer(int id, int peer, int fd);
int ptd_vmm_mc_create(int id, int* fd_out);               /* rank 0 */
int ptd_vmm_mc_import(int id, int fd);                    /* other ranks */
int ptd_vmm_mc_add_device(int id);                        /* all ranks, then a group barrier */
int ptd_vmm_mc_bind(int id);                              /* all ranks: bind local memory, map the multicast range */
int ptd_vmm_mc_drop(int id);
int ptd_vmm_ptrs(int id, void** uc_out, void** mc_out, long long* bytes_out);
int ptd_vmm_destroy(int id);
/* In-place all-reduce through the switch: multimem.ld_reduce + multimem.st on the multicast mapping `mc_buf`; `mc_pad`
 * (multicast mapping of the pads, may be NULL) selects the one-store switch barrier.  DDP's bucket all-reduce,
 * ddp_gpus.py:41 (loss.backward() under DDP). */
int ptd_allreduce_mc(void* mc_buf, const void* const* pads, const void* mc_pad, int rank, int world, long long numel,
                     int is_bf16, float scale, int blocks, void* stream);
/* The same two reductions (mc_buf given: through the switch; NULL: two-shot) with the SGD step of the bucket fused behind
 * them: parameters, momentum buffers and bf16 shadows are flat arrays in the BUCKET'S element order, updated by the block
 * that has just seen the reduction of those elements complete.  `loss.backward(); optimizer.step()` of ddp_gpus.py:41-42
 * with the communication kernel as the last touch of every gradient. */
typedef struct {
    float* param;            /* fp32 masters, bucket layout */
    float* momentum_buf;     /* NULL = no momentum; starts at zero */
    void* shadow_bf16;       /* NULL = none; bf16 copy of param refreshed in the same pass */
    const float* lr_dev;     /* learning rate, one float in device memory */
    float momentum, weight_decay;
} PtdSgdEpilogue;
int ptd_allreduce_fused_sgd(const void* const* bufs, void* mc_buf, const void* const* pads, const void* mc_pad, int rank,
                            int world, long long numel, float scale, int blocks, const PtdSgdEpilogue* sgd, void* stream);
/* Broadcast through the switch: the root stores each vector once (multimem.st), the NVSwitch writes every copy.  DDP's
 * construction-time parameter broadcast and per-forward buffer broadcast (ddp_gpus.py:35). */
int ptd_broadcast_mc(const void* const* bufs, void* mc_buf, const void* const* pads, const void* mc_pad, int rank,
                     int world, int root, long long nbytes, int blocks, void* stream);
/* fp32 gradients, bf16 on the wire: cast + reduce (switch or peer loads, fp32 accumulation) + cast back in one kernel. */
int ptd_allreduce_wire_bf16(float* grad, const void* const* stages, void* mc_stage, const void* const* pads,
                            const void* mc_pad, int rank, int world, long long numel, float scale, int blocks,
                            void* stream);

/* ---- gradient bucket planner / ready-tracking reducer (torch's C++ Reducer behind DDP) ---- */
int ptd_plan_buckets(const long long* sizes_bytes, int n, long long first_cap, long long cap, int* bucket_of);
int ptd_reducer_create(const int* bucket_of, int n_params, int n_buckets);
int ptd_reducer_mark_ready(int id, int param);
int ptd_reducer_all_done(int id);
int ptd_reducer_ready_order(int id, int* out, int n);
int ptd_reducer_incomplete(int id, int* out, int n);
int ptd_reducer_reset(int id);
int ptd_reducer_destroy(int id);
/* What the reduction of one bucket needs, bound once per bucket; ptd_reducer_mark_ready_launch then issues it natively the
 * moment the bucket's last gradient is marked (record an event on the compute stream, make the communication stream wait,
 * launch the all-reduce - with the fused SGD step when fuse_sgd -, record the completion event) and ptd_reducer_finalize
 * makes the compute stream wait for all of them: torch's C++ Reducer::mark_variable_ready / finalize_backward. */
typedef struct {
    int algo;                      /* 1 two-shot peer-to-peer, 2 through the switch (multimem), 3 bf16 wire format */
    int rank, world, blocks;
    long long numel;               /* fp32 elements, padded as the kernels require */
    float scale;
    void* local;                   /* this rank's bucket */
    const void* bufs[8];           /* unicast mappings of every rank's bucket */
    void* mc_buf;                  /* multicast mapping (algo 2) */
    const void* stages[8];         /* bf16 staging buffers (algo 3) */
    void* mc_stage;
    const void* pads[8];
    const void* mc_pad;
    int fuse_sgd;
    PtdSgdEpilogue sgd;
} PtdBucketComm;
int ptd_reducer_bind_bucket(int id, int bucket, const PtdBucketComm* c);
int ptd_reducer_mark_ready_launch(int id, int param, void* compute_stream, void* comm_stream);
int ptd_reducer_finalize(int id, void* compute_stream);

#ifdef __cplusplus
}
#endif
#endif /* PTDB200_H */
