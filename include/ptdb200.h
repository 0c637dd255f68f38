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

/* ---- convolution as implicit GEMM on tcgen05/TMEM (R/03 模型并行/03_model_parallel.ipynb#cell4-5:
 *      every Conv2d of ResNet(Bottleneck,[3,4,6,3]); cuDNN in the reference) --------------------- */
/* y[N,OH,OW,Cout] = conv(x[N,H,W,Cin], w[Cout,R,S,Cin]);  stat_sum/stat_sq (optional, fp32[Cout],
 * caller-zeroed) receive per-channel sum and sum of squares of the fp32 accumulators. */
int ptd_conv_fprop(const void* x, const void* w, void* y, float* stat_sum, float* stat_sq, int N, int H, int W,
                   int Cin, int Cout, int R, int S, int stride, int pad, void* stream);
/* dx[N,H,W,Cin] = conv_transpose(dy[N,OH,OW,Cout], w) (+ addend[N,H,W,Cin] if not NULL). */
int ptd_conv_dgrad(const void* dy, const void* w, void* dx, const void* addend, int N, int H, int W, int Cin,
                   int Cout, int R, int S, int stride, int pad, void* stream);
/* dx += conv_transpose(dy, w), in place (the 1x1 stride-2 shortcut only touches the strided quarter of dx). */
int ptd_conv_dgrad_accum(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int R, int S,
                         int stride, int pad, void* stream);
/* dw[Cout,R,S,Cin] (fp32) += dy^T * im2col(x).  The caller zeroes dw (or keeps an accumulated value). */
int ptd_conv_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int R, int S,
                   int stride, int pad, void* stream);
/* 7x7/2 stem on a 4-channel NHWC image; w4 is [64][8][8][4] zero-padded, dw4 the same shape in fp32. */
int ptd_stem_fprop(const void* x4, const void* w4, void* y, float* stat_sum, float* stat_sq, int N, int H, int W,
                   void* stream);
int ptd_stem_wgrad(const void* x4, const void* dy, float* dw4, int N, int H, int W, void* stream);

/* ---- dense layers (fc 2048->1000 of ResNet-50; cuBLAS in the reference) ----------------------- */
int ptd_linear_fwd(const void* a, const void* w, const float* bias, float* out, int M, int N, int K, void* stream);
int ptd_linear_dgrad(const void* dout, const void* w, void* da, int M, int N, int K, void* stream);
int ptd_linear_wgrad(const void* a, const void* dout, float* dw, int M, int N, int K, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PTDB200_H */
