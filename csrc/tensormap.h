// Host helper: encode a 2-D bf16 TMA descriptor with 128-byte swizzle.
#pragma once
#include <cuda.h>
#include <stdint.h>

// `inner` is the contiguous extent in elements, `rows` the outer extent, `row_pitch_bytes` the
// distance between rows.  The box is `box_inner` x `box_rows` elements (box_inner * 2 <= 128).
// Out-of-bounds elements of a box read as zero.
int ptd_make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t inner, uint64_t rows, uint64_t row_pitch_bytes,
                          uint32_t box_inner, uint32_t box_rows);

// im2col-mode descriptor over an NHWC bf16 activation (see runtime.cu for the walk it describes).
int ptd_make_tmap_im2col_bf16(CUtensorMap* out, const void* base, int C, int W, int H, int N, int lower_w, int lower_h,
                              int upper_w, int upper_h, int pixels, int stride);

// General 4-D tiled descriptor (128-byte swizzle): dims[0] is the contiguous extent in elements, strides_bytes[i] the
// byte stride of dimension i+1.  Strides may be smaller than the extent below them (overlapping windows).
int ptd_make_tmap_4d_bf16(CUtensorMap* out, const void* base, const uint64_t dims[4], const uint64_t strides_bytes[3],
                          const uint32_t box[4]);

// 2-D descriptor for STORING [box_rows][32 bf16] tiles held in shared memory with the 64-byte swizzle (16-byte chunk
// index ^= (row >> 1) & 3).  Rows past `rows` are clipped by the hardware.
int ptd_make_tmap_2d_bf16_store64(CUtensorMap* out, const void* base, uint64_t inner, uint64_t rows,
                                  uint64_t row_pitch_bytes, uint32_t box_rows);
