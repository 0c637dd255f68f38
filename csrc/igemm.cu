// Implicit-GEMM engine on tcgen05 / TMEM / TMA for sm_100a.
//
// Two warp-specialised kernels cover every GEMM-shaped op of the framework:
//
//   igemm_f  ("forward-like", persistent):  D[M,N] = A[M,K] * B[N,K]^T
//       A rows are output pixels.  The pixel-row operand reaches shared memory by TMA in three forms - a plain
//       K-major tile (1x1 stride-1 convolutions, FC), an im2col walk over the NHWC tensor (3x3, strided 1x1, data
//       gradients: cuTensorMapEncodeIm2col, zero fill in the padding) or an overlapping-window patch (the
//       space-to-depth stem) - always as the 128-byte-swizzled image tcgen05 reads.  A cp.async gather by four extra
//       producer warps survives as the fallback for stem inputs the patch form cannot take (and as an A/B switch,
//       PTD_IGEMM_FORCE_GATHER).  B (weights) always arrives by TMA, K-major (fprop) or MN-major (dgrad reads the
//       forward weight tensor in place - no transposed copy exists).  CTA pairs (cta_group::2) share 256-row tiles
//       on the K-heavy 1x1 layers.
//       Epilogue: TMEM -> registers -> bf16 staged in shared memory -> TMA store (or LDS + STG when rows are
//       remapped / an addend is fused), fused per-channel sum / sum-of-squares for BatchNorm, optional (mask-gated)
//       residual addend, or fp32 + bias for the classifier.
//
//   igemm_w  ("weight-gradient", split-K):  dW[Cout,Kw] += dY[P,Cout]^T * Xcol[P,Kw]
//       Both operands are MN-major (the reduction runs over pixels).  dY arrives by TMA, Xcol by TMA (tiled, im2col
//       or patch).  fp32 partial tiles are reduced with vector red.global.add - straight into whatever buffer the
//       caller names (with DDP: the gradient bucket itself).
//
// Roles per CTA: 8 or 16 epilogue warps (TMEM lane quarter = warp % 4), one TMA producer warp, one warp that allocates
// TMEM and issues tcgen05.mma from a single thread, and (gather fallback only) four cp.async producer warps.
// The reference reaches these ops through cuDNN/cuBLAS (call sites
// R/03 模型并行/03_model_parallel.ipynb#cell4-7,12).
#include <stdlib.h>

#include "common.cuh"
#include "ptdb200.h"
#include "tensormap.h"

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;              // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int A_TILE_BYTES = BLOCK_M * BLOCK_K * 2;   // 16 KiB
constexpr int SUB_TILE_BYTES = 64 * 64 * 2;           // 8 KiB : [64 rows][64 elems] swizzled image
#ifndef PTD_EPI_WARPS_WIDE
#define PTD_EPI_WARPS_WIDE 16
#endif
constexpr int NUM_EPI_WARPS = 8;
constexpr int TMA_WARP = 8;
constexpr int MMA_WARP = 9;
constexpr int GATHER_WARP0 = 10;
constexpr int NUM_GATHER_THREADS = 128;

enum { GM_FPROP = 0, GM_DGRAD = 1, GM_STEM = 2 };

struct ConvGeom {
    const bf16* src;     // NHWC tensor the rows are gathered from
    int H, W, C;         // its spatial size and channel count
    int OH, OW;          // row space: row = (n*OH + oh)*OW + ow
    int R, S;            // filter taps
    int stride, pad;
    int mode;            // GM_*
    int cblocks;         // C / 64
};

struct alignas(64) FParams {
    CUtensorMap tmA;
    CUtensorMap tmB;
    CUtensorMap tmO;     // output [M][ldo] bf16, box 32 x 32 with the 64-byte swizzle (valid when tma_store != 0)
    ConvGeom g;
    int tma_store;       // epilogue hands each staged 32 x 32 tile to the TMA store engine instead of LDS + STG
    int M, N, num_kb;
    int n_blocks, m_blocks;
    int b_tap_cols;      // MN-major B: column offset per filter tap (= N of the full weight slice)
    int b_cblocks;       // MN-major B: k-blocks per tap
    bf16* out;
    const bf16* addend;
    const uint8_t* addend_mask;   // optional: bit (element & 7) of byte (element >> 3) gates the addend element
    float* stat_sum;
    float* stat_sq;
    float* out_f32;
    const float* bias;
    int ldo;
    // BatchNorm-backward reduction fused into a data-gradient epilogue (BNR kernels): the tile this kernel writes is the
    // gradient dz behind a BatchNorm + ReLU whose conv output `bn_y` and ReLU bit mask `bn_mask` have the output's shape;
    // the epilogue accumulates sum(g) and sum(g * xhat), g = mask ? dz : 0, per channel - what bn_bwd_reduce would
    // otherwise re-read dz, y and the mask for.
    const bf16* bn_y;
    const uint8_t* bn_mask;
    const float* bn_mean;
    const float* bn_invstd;
    float* bn_sum_g;     // += sum g         (dbeta)
    float* bn_sum_gx;    // += sum g * xhat  (dgamma)
    int dbg;             // experiment switches (PTD_IGEMM_DBG), 0 in production
    int stats_smem;      // BN partial sums: 1 = column sums read back from the bf16 staging tile, 0 = register transpose
    // im2col tap table: k-block kb belongs to table entry kb / cblocks; `wtap` is the filter tap (r*S+s) whose
    // weights it multiplies, (off_w, off_h) the unsigned offsets handed to the im2col load.
    int ntaps;
    signed char tap_wtap[9], tap_offw[9], tap_offh[9];
    // Output row map (strided data gradient): GEMM row (n, a, b) of a [sub_oh x sub_ow] grid lands on
    // pixel (n, a*mul + ph, b*mul + pw) of an [oH x oW] tensor.  mul == 0: identity.
    int map_mul, map_ph, map_pw, map_sub_ow, map_sub_ohw, map_oH, map_oW;
};

struct alignas(64) WParams {
    CUtensorMap tmA;     // dY  [P][Cout], box 64x64
    CUtensorMap tmB;     // X   [P][Cin],  box 64x64 (1x1 stride-1 only)
    ConvGeom g;
    int P, Cout, Kw;
    int m_blocks, n_blocks, splits, kb_per_split, num_kb;
    float* dw;
    int dbg;             // experiment switches (PTD_IGEMM_DBG): 1 = skip B loads, 2 = skip MMA issue
};

// ------------------------------------------------------------------------------------------
// Gather producers
// ------------------------------------------------------------------------------------------
// A gather thread owns NR rows of the tile and the same 16-byte column in each of them: lanes
// 8k..8k+7 of a warp cover one contiguous 128-byte row segment, so one warp-level cp.async touches
// 4 full cache lines.
template <int NR>
struct RowSet {
    int nbase[NR];   // element offset of the row's image inside the source tensor
    int hw[NR];      // (hi0 << 16) | (wi0 & 0xffff): tap-independent part of the source coordinate
};

__device__ __forceinline__ int pack_hw(int hi0, int wi0) { return (hi0 << 16) | (wi0 & 0xffff); }

// Rows first_row + i*ROW_STEP, i < NR.  One division pair for the first row, the rest by stepping.
template <int NR, int ROW_STEP>
__device__ __forceinline__ void rowset_decode(const ConvGeom& g, long long first_row, long long nrows, RowSet<NR>& rs) {
    const int hw = g.OH * g.OW;
    long long fr = first_row < nrows ? first_row : 0;
    int n = (int)(fr / hw);
    int rem = (int)(fr - (long long)n * hw);
    int oh = rem / g.OW;
    int ow = rem - oh * g.OW;
    const int img = g.H * g.W * g.C;
#pragma unroll
    for (int i = 0; i < NR; ++i) {
        const bool valid = first_row + (long long)i * ROW_STEP < nrows;
        int hi0, wi0;
        if (g.mode == GM_FPROP) { hi0 = oh * g.stride - g.pad; wi0 = ow * g.stride - g.pad; }
        else if (g.mode == GM_DGRAD) { hi0 = oh + g.pad; wi0 = ow + g.pad; }
        else { hi0 = oh * 2 - 3; wi0 = ow * 2 - 3; }
        rs.nbase[i] = n * img;
        rs.hw[i] = valid ? pack_hw(hi0, wi0) : pack_hw(-16384, -16384);   // every tap out of range
        ow += ROW_STEP;
        while (ow >= g.OW) {
            ow -= g.OW;
            if (++oh == g.OH) { oh = 0; ++n; }
        }
    }
}

// One k-block (filter tap `tap`, channel block `cb`) of NR rows: 16 bytes per row, zero-filled when the tap
// falls into the padding (or, for the data gradient, between the strides).
template <int NR, int DST_STEP>
__device__ __forceinline__ void gather_conv(const ConvGeom& g, const RowSet<NR>& rs, int tap, int cb, int chunk,
                                            uint32_t dst0) {
    const int r = tap / g.S;
    const int s = tap - r * g.S;
    const bf16* base = g.src + cb * 64 + chunk * 8;
#pragma unroll
    for (int i = 0; i < NR; ++i) {
        const int hi0 = rs.hw[i] >> 16;
        const int wi0 = (int)(short)(rs.hw[i] & 0xffff);
        int hi, wi;
        bool ok;
        if (g.mode == GM_FPROP) {
            hi = hi0 + r;
            wi = wi0 + s;
            ok = (unsigned)hi < (unsigned)g.H && (unsigned)wi < (unsigned)g.W;
        } else {
            hi = hi0 - r;
            wi = wi0 - s;
            ok = hi >= 0 && wi >= 0;
            if (g.stride == 2) {
                ok = ok && !((hi | wi) & 1);
                hi >>= 1;
                wi >>= 1;
            }
            ok = ok && hi < g.H && wi < g.W;
        }
        const bf16* src = ok ? base + rs.nbase[i] + (hi * g.W + wi) * g.C : g.src;
        cp_async_16(dst0 + i * DST_STEP, src, ok ? 16u : 0u);
    }
}

// Stem (7x7 stride 2 pad 3 on an NHWC4 image): k-block kb holds filter rows 2kb, 2kb+1; a filter row is
// 8 taps x 4 channels = 64 bytes (tap 7 and filter row 7 are zero padding).  `piece` (0..15) selects
// (filter row parity, tap): 8 bytes per cp.async, 16 consecutive lanes cover one 128-byte tile row.
template <int NR, int DST_STEP>
__device__ __forceinline__ void gather_stem(const ConvGeom& g, const RowSet<NR>& rs, int kb, int piece, uint32_t dst0) {
    const int r = 2 * kb + (piece >> 3);
    const int s = piece & 7;
    const bool tap_ok = r < 7 && s < 7;
#pragma unroll
    for (int i = 0; i < NR; ++i) {
        const int hi = (rs.hw[i] >> 16) + r;
        const int wi = (int)(short)(rs.hw[i] & 0xffff) + s;
        const bool ok = tap_ok && (unsigned)hi < (unsigned)g.H && (unsigned)wi < (unsigned)g.W;
        const bf16* src = ok ? g.src + rs.nbase[i] + (hi * g.W + wi) * 4 : g.src;
        cp_async_8(dst0 + i * DST_STEP, src, ok ? 8u : 0u);
    }
}

// ------------------------------------------------------------------------------------------
// Shared-memory carve-up
// ------------------------------------------------------------------------------------------
template <int STAGES>
struct Barriers {
    uint64_t full[STAGES];
    uint64_t empty[STAGES];
    uint64_t tmem_full[2];
    uint64_t tmem_empty[2];
    uint32_t tmem_base;
};

// ------------------------------------------------------------------------------------------
// Column sums across the 32 rows held by a warp (one row per lane, 32 columns per lane):
// recursive halving, 31 shuffles; on return v[0] of lane l is the sum of column l.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float transpose_reduce32(float (&v)[32], uint32_t lane) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        const bool up = (lane & o) != 0;
#pragma unroll
        for (int i = 0; i < o; ++i) {
            float send = up ? v[i] : v[i + o];
            float keep = up ? v[i + o] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
    return v[0];
}

// ------------------------------------------------------------------------------------------
// igemm_f
// ------------------------------------------------------------------------------------------
// How the pixel-row operand reaches shared memory.  AM_PATCH: a 4-D tiled descriptor whose box is a [rows x 16]
// patch of output pixels (x 64 contiguous elements); used by the space-to-depth stem, where the row of output pixel
// (oh, ow) for k-block r' is the 128-byte window at pixel (oh + r', ow) of the input.
enum { AM_GATHER = 0, AM_TILED = 1, AM_IM2COL = 2, AM_PATCH = 3 };

// PAIR: two CTAs of a cluster share one 256 x BLOCK_N tile (tcgen05 cta_group::2): each loads its own 128 rows of A
// and HALF of B, the leader issues M = 256 MMAs that read B from both shared memories.  Per SM that is a third fewer
// operand bytes per MMA - what the K-heavy layers are bound by (L2 -> SM ingress, ~40 B/clk/SM measured).
template <int BLOCK_N, int A_MODE, bool B_MN, int EPI /*0 = bf16 (+stats/+addend), 1 = fp32 + bias*/, bool PAIR = false>
struct FCfg {
    static constexpr int B_TILE_BYTES = BLOCK_N * BLOCK_K * 2 / (PAIR ? 2 : 1);      // bytes this CTA loads
    static constexpr int STAGE_BYTES = A_TILE_BYTES + B_TILE_BYTES;
    static constexpr int STAGES = PAIR ? (BLOCK_N == 256 ? 6 : 8) : ((BLOCK_N == 256) ? 4 : (BLOCK_N == 128 ? 6 : 8));
    // Epilogue warps: a warp may only read TMEM lanes 32 * (warp % 4) .. + 31, so warps come in groups of four and
    // each group takes a column range.  Wide tiles use four groups: the epilogue of a memory-bound layer is a chain
    // of long-latency steps per warp (tcgen05.ld, staging, stores) and gains from more warps in flight.
    static constexpr int EPI_WARPS = (BLOCK_N >= 128 && EPI == 0) ? PTD_EPI_WARPS_WIDE : 8;
    static constexpr int EPI_GROUPS = EPI_WARPS / 4;
    static constexpr int TMA_WARP = EPI_WARPS;
    static constexpr int MMA_WARP = EPI_WARPS + 1;
    static constexpr int GATHER_WARP0 = EPI_WARPS + 2;
    static constexpr int THREADS = A_MODE != AM_GATHER ? (MMA_WARP + 1) * 32 : (GATHER_WARP0 + 4) * 32;
    static constexpr int TMEM_COLS = 2 * BLOCK_N;
    static constexpr int STAGING_OFF = STAGES * STAGE_BYTES + 1024;        // barriers live in the 1 KiB before it
    static constexpr size_t SMEM_BYTES = 1024 + (size_t)STAGING_OFF + EPI_WARPS * 2048;
    static constexpr int COLS_PER_GROUP = BLOCK_N / EPI_GROUPS;
    static constexpr int CHUNKS_PER_WARP = COLS_PER_GROUP / 32;   // 32-column chunks per epilogue warp
};

// Gather-producer loop of igemm_f (128 threads).  conv: thread owns the 16-byte column (tid & 7) of rows
// (tid >> 3) + 16 i, i < 8;  stem: the 8-byte piece (tid & 15) of rows (tid >> 4) + 8 i, i < 16.
// Completion of a thread's copies arrives on the stage's full barrier by itself
// (cp.async.mbarrier.arrive.noinc): the producer never blocks on its own loads, so up to STAGES
// k-blocks are in flight.  (A wait_group + fence.proxy.async per k-block serialised the ring to one
// memory latency per k-block - measured with ncu, profiles/.)
template <class Cfg, bool STEM>
__device__ __forceinline__ void f_gather_loop(const FParams& p, Barriers<Cfg::STAGES>* bars, uint32_t smem_base,
                                              int num_tiles) {
    constexpr int STAGES = Cfg::STAGES;
    constexpr int NR = STEM ? 16 : 8;
    constexpr int ROW_STEP = STEM ? 8 : 16;
    const int tid = threadIdx.x - Cfg::GATHER_WARP0 * 32;
    const int chunk = tid & 7;
    const int sub = STEM ? (tid >> 4) : (tid >> 3);
    const uint32_t dst_off = STEM ? (uint32_t)(sub * 128 + ((((tid & 15) >> 1) ^ (sub & 7)) << 4) + (tid & 1) * 8)
                                  : (uint32_t)(sub * 128 + ((chunk ^ (sub & 7)) << 4));
    int stage = 0;
    uint32_t phase = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        const long long m0 = (long long)(t / p.n_blocks) * BLOCK_M;
        RowSet<NR> rs;
        rowset_decode<NR, ROW_STEP>(p.g, m0 + sub, p.M, rs);
        for (int kb = 0; kb < p.num_kb; ++kb) {
            mbar_wait(&bars->empty[stage], phase ^ 1);
            const uint32_t dst0 = smem_base + stage * Cfg::STAGE_BYTES + dst_off;
            if (!(p.dbg & 1)) {
                if (STEM) {
                    gather_stem<NR, ROW_STEP * 128>(p.g, rs, kb, tid & 15, dst0);
                } else {
                    const int tap = (p.dbg & 2) ? 0 : kb / p.g.cblocks;
                    gather_conv<NR, ROW_STEP * 128>(p.g, rs, tap, (p.dbg & 2) ? 0 : kb - tap * p.g.cblocks, chunk, dst0);
                }
            }
            if (p.dbg & 4) {
                __syncwarp();
                if ((tid & 31) == 0) cp_async_mbar_arrive_noinc(&bars->full[stage]);
            } else {
                cp_async_mbar_arrive_noinc(&bars->full[stage]);
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
    }
}

template <int BLOCK_N, int A_MODE, bool B_MN, int EPI, bool PAIR = false, bool BNR = false>
__global__ void __launch_bounds__(FCfg<BLOCK_N, A_MODE, B_MN, EPI, PAIR>::THREADS, 1)
igemm_f_kernel(const __grid_constant__ FParams p) {
    using Cfg = FCfg<BLOCK_N, A_MODE, B_MN, EPI, PAIR>;
    static_assert(!PAIR || (A_MODE == AM_TILED && EPI == 0 && BLOCK_N >= 128), "CTA pairs: tiled 1x1 path only");
    constexpr bool A_TMA = A_MODE != AM_GATHER;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    Barriers<STAGES>* bars = reinterpret_cast<Barriers<STAGES>*>(smem + (size_t)STAGES * Cfg::STAGE_BYTES);
    const uint32_t smem_base = smem_u32(smem);

    const int warp = threadIdx.x >> 5;
    const uint32_t lane = lane_id();
    // PAIR: the two CTAs of a cluster are one worker; m_blocks then counts 256-row tiles
    const uint32_t crank = PAIR ? cluster_ctarank() : 0u;
    const int worker = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int n_workers = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    constexpr int TILE_M = PAIR ? 2 * BLOCK_M : BLOCK_M;
    const int num_tiles = p.m_blocks * p.n_blocks;
    const int n_blk = worker % p.n_blocks;          // workers are a multiple of n_blocks: fixed per CTA
    const int n0 = n_blk * BLOCK_N;

    if (warp == Cfg::TMA_WARP && lane == 0) {
        if (A_TMA) tma_prefetch_desc(&p.tmA);
        tma_prefetch_desc(&p.tmB);
        if (EPI == 0 && p.tma_store != 0) tma_prefetch_desc(&p.tmO);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&bars->full[s], A_TMA ? 1 : 1 + ((p.dbg & 4) ? 4 : NUM_GATHER_THREADS));
            mbar_init(&bars->empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&bars->tmem_full[a], 1);
            mbar_init(&bars->tmem_empty[a], Cfg::EPI_WARPS * 32 * (PAIR ? 2 : 1));   // PAIR: both CTAs' epilogues
        }
        mbar_fence_init();
    }
    if (warp == Cfg::MMA_WARP) {
        if (PAIR) tmem_alloc_2sm<Cfg::TMEM_COLS>(&bars->tmem_base);
        else tmem_alloc<Cfg::TMEM_COLS>(&bars->tmem_base);
    }
    tc_fence_before();
    __syncthreads();
    if (PAIR) cluster_sync_all();         // the peer's barriers exist before anything signals them
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;
    pdl_wait();     // barriers, TMEM and descriptor prefetch above overlap the previous kernel's tail (ptd_launch)

    if (warp == Cfg::TMA_WARP) {
        // ===================== TMA producer (one lane) =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = worker; t < num_tiles; t += n_workers) {
                const int m0 = (t / p.n_blocks) * TILE_M + (int)crank * BLOCK_M;
                int iw0 = 0, ih0 = 0, in0 = 0;          // im2col: input-space position of the tile's first row
                if (A_MODE == AM_IM2COL) {
                    const int hw = p.g.OH * p.g.OW;
                    in0 = m0 / hw;
                    const int rem = m0 - in0 * hw;
                    const int oh = rem / p.g.OW, ow = rem - oh * p.g.OW;
                    // input-space position of the first row's tap-table origin (g.pad holds the origin shift)
                    iw0 = ow * p.g.stride - p.g.pad;
                    ih0 = oh * p.g.stride - p.g.pad;
                }
                int px0 = 0, py0 = 0, pn = 0;           // patch mode: first pixel of the tile's 8 x 16 patch
                if (A_MODE == AM_PATCH) {
                    const int mt = t / p.n_blocks, tiles_x = p.g.OW >> 4, per_img = tiles_x * (p.g.OH >> 3);
                    pn = mt / per_img;
                    const int rem = mt - pn * per_img;
                    const int ty = rem / tiles_x;
                    py0 = ty * 8;
                    px0 = (rem - ty * tiles_x) * 16;
                }
                for (int kb = 0; kb < p.num_kb; ++kb) {
                    mbar_wait(&bars->empty[stage], phase ^ 1);
                    const uint32_t a_dst = smem_base + stage * Cfg::STAGE_BYTES;
                    const uint32_t b_dst = a_dst + A_TILE_BYTES;
                    if (PAIR) {
                        // both CTAs' bytes are counted on the leader's barrier, which the leader arms for the pair
                        if (crank == 0) mbar_arrive_expect_tx(&bars->full[stage], 2 * Cfg::STAGE_BYTES);
                        tma_load_2d_2sm(a_dst, &p.tmA, &bars->full[stage], kb * BLOCK_K, m0);
                        if (!B_MN) {
                            tma_load_2d_2sm(b_dst, &p.tmB, &bars->full[stage], kb * BLOCK_K, n0 + (int)crank * (BLOCK_N / 2));
                        } else {
                            const int te = kb / p.b_cblocks;
                            const int cb = kb - te * p.b_cblocks;
#pragma unroll
                            for (int j = 0; j < BLOCK_N / 128; ++j)
                                tma_load_2d_2sm(b_dst + j * SUB_TILE_BYTES, &p.tmB, &bars->full[stage],
                                                te * p.b_tap_cols + n0 + (int)crank * (BLOCK_N / 2) + j * 64, cb * 64);
                        }
                        if (++stage == STAGES) { stage = 0; phase ^= 1; }
                        continue;
                    }
                    mbar_arrive_expect_tx(&bars->full[stage], Cfg::B_TILE_BYTES + (A_TMA ? A_TILE_BYTES : 0));
                    if (A_MODE == AM_TILED) tma_load_2d(a_dst, &p.tmA, &bars->full[stage], kb * BLOCK_K, m0);
                    if (A_MODE == AM_PATCH) tma_load_4d(a_dst, &p.tmA, &bars->full[stage], 0, px0, py0 + kb, pn);
                    if (A_MODE == AM_IM2COL) {
                        const int te = kb / p.g.cblocks;
                        const int cb = kb - te * p.g.cblocks;
                        tma_load_im2col(a_dst, &p.tmA, &bars->full[stage], cb * 64, iw0, ih0, in0,
                                        (uint32_t)p.tap_offw[te], (uint32_t)p.tap_offh[te]);
                    }
                    if (!B_MN) {
                        tma_load_2d(b_dst, &p.tmB, &bars->full[stage], kb * BLOCK_K, n0);
                    } else {
                        const int te = kb / p.b_cblocks;
                        const int cb = kb - te * p.b_cblocks;
                        const int tap = (A_MODE == AM_IM2COL) ? (int)p.tap_wtap[te] : te;
#pragma unroll
                        for (int j = 0; j < BLOCK_N / 64; ++j)
                            tma_load_2d(b_dst + j * SUB_TILE_BYTES, &p.tmB, &bars->full[stage],
                                        tap * p.b_tap_cols + n0 + j * 64, cb * 64);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
            if (!(p.dbg & 16)) pdl_trigger();      // last operands requested: what is left is the final epilogue
        }
    } else if (warp == Cfg::MMA_WARP) {
        // ===================== MMA issuer (one lane) =====================
        if (lane == 0 && (!PAIR || crank == 0)) {
            constexpr uint32_t idesc = umma_idesc_bf16(TILE_M, BLOCK_N, 0, B_MN ? 1 : 0);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int t = worker; t < num_tiles; t += n_workers, ++it) {
                const int acc = it & 1;
                const uint32_t acc_phase = (it >> 1) & 1;
                mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
                for (int kb = 0; kb < p.num_kb; ++kb) {
                    mbar_wait(&bars->full[stage], phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_base + stage * Cfg::STAGE_BYTES;
                    const uint32_t b_addr = a_addr + A_TILE_BYTES;
#pragma unroll
                    for (int kk = 0; kk < BLOCK_K / UMMA_K; ++kk) {
                        const uint64_t adesc = umma_smem_desc(a_addr + kk * 32, 16, 1024);
                        const uint64_t bdesc = B_MN ? umma_smem_desc(b_addr + kk * 2048, SUB_TILE_BYTES, 1024)
                                                    : umma_smem_desc(b_addr + kk * 32, 16, 1024);
                        if (PAIR) umma_f16_2sm(d_tmem, adesc, bdesc, idesc, (kb | kk) != 0 ? 1u : 0u);
                        else umma_f16(d_tmem, adesc, bdesc, idesc, (kb | kk) != 0 ? 1u : 0u);
                    }
                    if (PAIR) umma_commit_2sm(&bars->empty[stage]);       // frees the stage in both CTAs
                    else umma_commit(&bars->empty[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                if (PAIR) umma_commit_2sm(&bars->tmem_full[acc]);
                else umma_commit(&bars->tmem_full[acc]);
            }
        }
    } else if (warp < Cfg::EPI_WARPS) {
        // ===================== Epilogue =====================
        const int q = warp & 3;
        const int half = warp >> 2;       // column group of this warp
        // per-chunk BN partial sums, kept in four scalar pairs so the chunk loop need not be unrolled (the unrolled
        // butterfly made the epilogue stall on instruction fetch: ncu stall_no_inst 27 %)
        // BN partial sums: lane (cp = lane & 15, par = lane >> 4) accumulates columns 2cp, 2cp+1 of each of the warp's
        // 32-column chunks over the rows of parity `par`, read back from the bf16 staging tile (the statistics then
        // describe exactly the values that were stored).  A cross-lane transpose-reduce of the fp32 registers cost
        // ~250 instructions per chunk; this is ~110 and needs no shuffles.
        float4 acc0 = make_float4(0.f, 0.f, 0.f, 0.f), acc1 = acc0, acc2 = acc0, acc3 = acc0;   // (s_lo, s_hi, q_lo, q_hi)
        // The register transpose-reduce (one column per lane: .x = sum, .z = sum of squares) stays for tiles with a
        // long K loop: there the epilogue hides under the MMAs either way, and the denser instruction stream of the
        // shared-memory variant was measured to slow the main loop down by 7-12 % (profiles/README.md).
        const bool do_stats = (EPI == 0) && !BNR && p.stat_sum != nullptr;      // (a data gradient feeds no BN statistics)
        const bool stats_smem = p.stats_smem != 0;
        // BNR: lane (pc = lane & 3) owns columns pc*8 .. pc*8+7 of each of the warp's (at most two) 32-column chunks and
        // accumulates raw sums over every row it sees: S1 = sum g, S2 = sum g*y; xhat enters at the very end
        // (sum g*xhat = invstd * (S2 - mean * S1)), so the loop needs no per-channel constants.
        float bs1a[8], bs2a[8], bs1b[8], bs2b[8];
        if (BNR) {
#pragma unroll
            for (int e = 0; e < 8; ++e) { bs1a[e] = bs2a[e] = bs1b[e] = bs2b[e] = 0.f; }
        }
        static_assert(!BNR || (EPI == 0 && Cfg::CHUNKS_PER_WARP <= 2), "BNR: at most two 32-column chunks per epilogue warp");
        // BNR: the y pieces and mask bytes a lane needs for a chunk are requested ONE CHUNK AHEAD and carried in
        // registers: with 8-16 epilogue warps per SM and one chunk in flight each, loads issued where they are consumed
        // (or at the top of the same chunk) leave ~16 KB in flight per SM - the kernel then waits on DRAM latency
        // (ncu: long_sb, 42 % of DRAM instead of 75 %).
        uint4 bn_yp[4];
        uint32_t bn_mb[4];
        auto bn_prefetch = [&](int t_, int c_) {
#pragma unroll
            for (int i = 0; i < 4; ++i) { bn_yp[i] = make_uint4(0u, 0u, 0u, 0u); bn_mb[i] = 0u; }
            if (t_ >= num_tiles) return;
            const long long mb0 = (long long)(t_ / p.n_blocks) * TILE_M + crank * BLOCK_M + q * 32;
            const int col_ = half * Cfg::COLS_PER_GROUP + c_ * 32;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int piece = (int)lane + 32 * i;
                const int row = piece >> 2, pc = piece & 3;
                const long long mr = mb0 + row;
                if (mr < p.M) {
                    long long orow = mr;
                    if (p.map_mul != 0) {
                        const int n_ = (int)(mr / p.map_sub_ohw);
                        const int rem_ = (int)(mr - (long long)n_ * p.map_sub_ohw);
                        const int a_ = rem_ / p.map_sub_ow, b_ = rem_ - a_ * p.map_sub_ow;
                        orow = ((long long)n_ * p.map_oH + a_ * p.map_mul + p.map_ph) * p.map_oW + b_ * p.map_mul + p.map_pw;
                    }
                    const size_t off = (size_t)orow * p.ldo + n0 + col_ + pc * 8;
                    bn_yp[i] = *reinterpret_cast<const uint4*>(p.bn_y + off);
                    bn_mb[i] = p.bn_mask[off >> 3];
                }
            }
        };
        if (BNR) bn_prefetch(worker, 0);
        int it = 0;
        for (int t = worker; t < num_tiles; t += n_workers, ++it) {
            const int acc = it & 1;
            const uint32_t acc_phase = (it >> 1) & 1;
            const long long m = (long long)(t / p.n_blocks) * TILE_M + crank * BLOCK_M + q * 32 + lane;
            const bool row_ok = m < p.M;      // fp32 epilogue: one row per lane
            mbar_wait(&bars->tmem_full[acc], acc_phase);
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < Cfg::CHUNKS_PER_WARP; ++c) {
                const int col = half * Cfg::COLS_PER_GROUP + c * 32;
                uint32_t raw[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * BLOCK_N + col, raw);
                tmem_ld_wait();
                float v[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[i]);
                if (EPI == 0) {
                    // Stage the warp's 32 rows x 64 B through shared memory so that global stores are
                    // contiguous: 4 lanes cover one row's 64 B (two full sectors) instead of every lane
                    // writing 16 B of a different row.  XOR on the 16-byte slot keeps both phases conflict-free.
                    const uint32_t stg = smem_base + Cfg::STAGING_OFF + warp * 2048;
                    const bool tma_out = p.tma_store != 0;
                    if (tma_out) {              // the previous tile of this warp has been read out of the buffer
                        if (lane == 0) bulk_wait_group_read<0>();
                        __syncwarp();
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint32_t a = stg + lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4);
                        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a),
                                     "r"(pack_bf16x2(v[j * 8 + 0], v[j * 8 + 1])), "r"(pack_bf16x2(v[j * 8 + 2], v[j * 8 + 3])),
                                     "r"(pack_bf16x2(v[j * 8 + 4], v[j * 8 + 5])), "r"(pack_bf16x2(v[j * 8 + 6], v[j * 8 + 7]))
                                     : "memory");
                    }
                    const long long m_base = (long long)(t / p.n_blocks) * TILE_M + crank * BLOCK_M + q * 32;
                    if (tma_out) {
                        // The staging tile already is the 64-byte-swizzle image of a [32 rows][32 channels] box: one
                        // lane hands it to the TMA store engine (rows past M are clipped), which takes the LDS + STG
                        // wavefronts of the copy-out off the load/store pipe - ncu had that pipe 72 % busy and the
                        // copy-out was 51 of its 83 wavefronts per tile (profiles/README.md).
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0) {
                            tma_store_2d(&p.tmO, stg, n0 + col, (int)m_base);
                            bulk_commit_group();
                        }
                    } else {
                    __syncwarp();
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int piece = (int)lane + 32 * i;
                        const int row = piece >> 2, pc = piece & 3;
                        uint4 o;
                        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                                     : "=r"(o.x), "=r"(o.y), "=r"(o.z), "=r"(o.w)
                                     : "r"(stg + row * 64 + ((pc ^ ((row >> 1) & 3)) << 4)));
                        const long long mr = m_base + row;
                        if (mr < p.M) {
                            long long orow = mr;
                            if (A_MODE == AM_PATCH) {
                                // tile = 8 x 16 patch of output pixels, row r of the tile = pixel (r / 16, r % 16) of it
                                const int mt = t / p.n_blocks, tiles_x = p.g.OW >> 4, per_img = tiles_x * (p.g.OH >> 3);
                                const int n_ = mt / per_img, rem_ = mt - n_ * per_img;
                                const int ty = rem_ / tiles_x, tx = rem_ - ty * tiles_x;
                                const int r_ = q * 32 + row;
                                orow = ((long long)n_ * p.g.OH + ty * 8 + (r_ >> 4)) * p.g.OW + tx * 16 + (r_ & 15);
                            } else if (p.map_mul != 0) {
                                const int n_ = (int)(mr / p.map_sub_ohw);
                                const int rem_ = (int)(mr - (long long)n_ * p.map_sub_ohw);
                                const int a_ = rem_ / p.map_sub_ow, b_ = rem_ - a_ * p.map_sub_ow;
                                orow = ((long long)n_ * p.map_oH + a_ * p.map_mul + p.map_ph) * p.map_oW + b_ * p.map_mul + p.map_pw;
                            }
                            const size_t off = (size_t)orow * p.ldo + n0 + col + pc * 8;
                            if (p.addend != nullptr) {
                                const uint4 a4 = *reinterpret_cast<const uint4*>(p.addend + off);
                                uint32_t aw[4] = {a4.x, a4.y, a4.z, a4.w};
                                if (p.addend_mask != nullptr) {
                                    // the addend is a gradient that still has to pass a ReLU: one mask bit per element
                                    const uint32_t mb = p.addend_mask[off >> 3];
#pragma unroll
                                    for (int e = 0; e < 4; ++e)
                                        aw[e] &= (((mb >> (2 * e)) & 1u) ? 0x0000ffffu : 0u) |
                                                 (((mb >> (2 * e + 1)) & 1u) ? 0xffff0000u : 0u);
                                }
                                uint32_t ow[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
                                for (int e = 0; e < 4; ++e) {
                                    const float2 f = unpack_bf16x2(aw[e]), g = unpack_bf16x2(ow[e]);
                                    ow[e] = pack_bf16x2(f.x + g.x, f.y + g.y);
                                }
                                o = make_uint4(ow[0], ow[1], ow[2], ow[3]);
                                if (BNR)       // the fused BatchNorm reduction below reads the FINAL gradient from the tile
                                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(stg + row * 64 + ((pc ^ ((row >> 1) & 3)) << 4)),
                                                 "r"(o.x), "r"(o.y), "r"(o.z), "r"(o.w) : "memory");
                            }
                            *reinterpret_cast<uint4*>(p.out + off) = o;
                        }
                    }
                    }       // !tma_out
                    if (do_stats && !stats_smem) {
                        float sq[32];
#pragma unroll
                        for (int i = 0; i < 32; ++i) sq[i] = v[i] * v[i];
                        const float cs = transpose_reduce32(v, lane), cq = transpose_reduce32(sq, lane);
                        acc0.x += c == 0 ? cs : 0.f; acc0.z += c == 0 ? cq : 0.f;
                        acc1.x += c == 1 ? cs : 0.f; acc1.z += c == 1 ? cq : 0.f;
                        acc2.x += c == 2 ? cs : 0.f; acc2.z += c == 2 ? cq : 0.f;
                        acc3.x += c == 3 ? cs : 0.f; acc3.z += c == 3 ? cq : 0.f;
                    }
                    if (do_stats && stats_smem) {
                        // Rows past M were zero-filled by the producers, so they add nothing.
                        const uint32_t cp = lane & 15, par = lane >> 4;
                        const uint32_t slot = cp >> 2, word = (cp & 3) << 2;
                        float s_lo = 0.f, s_hi = 0.f, q_lo = 0.f, q_hi = 0.f;
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            // row 2i + par: rows 2i and 2i+1 are 64 B apart -> the two half-warps hit disjoint banks
                            const uint32_t a = stg + (2 * i + par) * 64 + ((slot ^ (uint32_t)(i & 3)) << 4) + word;
                            uint32_t w2;
                            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w2) : "r"(a));
                            const float2 f = unpack_bf16x2(w2);
                            s_lo += f.x; s_hi += f.y;
                            q_lo = fmaf(f.x, f.x, q_lo); q_hi = fmaf(f.y, f.y, q_hi);
                        }
                        switch (c) {
                            case 0: acc0.x += s_lo; acc0.y += s_hi; acc0.z += q_lo; acc0.w += q_hi; break;
                            case 1: acc1.x += s_lo; acc1.y += s_hi; acc1.z += q_lo; acc1.w += q_hi; break;
                            case 2: acc2.x += s_lo; acc2.y += s_hi; acc2.z += q_lo; acc2.w += q_hi; break;
                            default: acc3.x += s_lo; acc3.y += s_hi; acc3.z += q_lo; acc3.w += q_hi; break;
                        }
                    }
                    if (BNR) {
                        // Same 16-byte pieces as the copy-out: 4 lanes cover one row's 64 bytes.  dz comes from the bf16
                        // staging tile (the values that were stored; when an addend is fused the copy-out loop above
                        // wrote the sums back into the tile), y and the mask bits from global memory at the same offset.
                        __syncwarp();
                        float t1[8], t2[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) { t1[e] = 0.f; t2[e] = 0.f; }
#pragma unroll
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int piece = (int)lane + 32 * i;
                            const int row = piece >> 2, pc = piece & 3;
                            uint4 dzp = make_uint4(0u, 0u, 0u, 0u);
                            if (m_base + row < p.M)
                                asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                                             : "=r"(dzp.x), "=r"(dzp.y), "=r"(dzp.z), "=r"(dzp.w)
                                             : "r"(stg + row * 64 + ((pc ^ ((row >> 1) & 3)) << 4)));
                            const uint32_t dw_[4] = {dzp.x, dzp.y, dzp.z, dzp.w};
                            const uint32_t yw_[4] = {bn_yp[i].x, bn_yp[i].y, bn_yp[i].z, bn_yp[i].w};
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                const float2 d2 = unpack_bf16x2(dw_[e]), y2 = unpack_bf16x2(yw_[e]);
                                const float g0 = ((bn_mb[i] >> (2 * e)) & 1u) ? d2.x : 0.f;
                                const float g1 = ((bn_mb[i] >> (2 * e + 1)) & 1u) ? d2.y : 0.f;
                                t1[2 * e] += g0; t1[2 * e + 1] += g1;
                                t2[2 * e] = fmaf(g0, y2.x, t2[2 * e]); t2[2 * e + 1] = fmaf(g1, y2.y, t2[2 * e + 1]);
                            }
                        }
                        // this chunk's y / mask registers are consumed: request the next chunk's (same tile or next tile)
                        if (c + 1 < Cfg::CHUNKS_PER_WARP) bn_prefetch(t, c + 1);
                        else bn_prefetch(t + n_workers, 0);
                        if (c == 0) {
#pragma unroll
                            for (int e = 0; e < 8; ++e) { bs1a[e] += t1[e]; bs2a[e] += t2[e]; }
                        } else {
#pragma unroll
                            for (int e = 0; e < 8; ++e) { bs1b[e] += t1[e]; bs2b[e] += t2[e]; }
                        }
                    }
                    __syncwarp();
                } else {
                    if (row_ok) {
                        float* dst = p.out_f32 + (size_t)m * p.ldo + n0 + col;
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            const int n = n0 + col + i;
                            if (n < p.N) dst[i] = v[i] + (p.bias ? p.bias[n] : 0.f);
                        }
                    }
                }
            }
            tc_fence_before();
            if (PAIR) mbar_arrive_leader(&bars->tmem_empty[acc]);     // the leader's MMA thread waits for both CTAs
            else mbar_arrive(&bars->tmem_empty[acc]);
        }
        if (EPI == 0 && p.tma_store != 0 && lane == 0) bulk_wait_group<0>();      // this warp's stores are complete
        if (BNR) {
            // lanes with equal (lane & 3) hold partial sums of the same 8 columns: fold them, lanes 0..3 publish
#pragma unroll
            for (int c = 0; c < Cfg::CHUNKS_PER_WARP; ++c) {
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    float s1 = c == 0 ? bs1a[e] : bs1b[e], s2 = c == 0 ? bs2a[e] : bs2b[e];
#pragma unroll
                    for (int o = 4; o <= 16; o <<= 1) {
                        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
                        s2 += __shfl_xor_sync(0xffffffffu, s2, o);
                    }
                    if (lane < 4) {
                        const int n = n0 + half * Cfg::COLS_PER_GROUP + c * 32 + (int)lane * 8 + e;
                        atomicAdd(p.bn_sum_g + n, s1);
                        atomicAdd(p.bn_sum_gx + n, p.bn_invstd[n] * (s2 - p.bn_mean[n] * s1));
                    }
                }
            }
        }
        if (do_stats) {
            const float4 fa[4] = {acc0, acc1, acc2, acc3};
#pragma unroll
            for (int c = 0; c < Cfg::CHUNKS_PER_WARP; ++c) {
                const int nb = n0 + half * Cfg::COLS_PER_GROUP + c * 32;
                if (stats_smem) {
                    const int n = nb + 2 * (int)(lane & 15);
                    atomicAdd(p.stat_sum + n, fa[c].x);
                    atomicAdd(p.stat_sum + n + 1, fa[c].y);
                    atomicAdd(p.stat_sq + n, fa[c].z);
                    atomicAdd(p.stat_sq + n + 1, fa[c].w);
                } else {
                    atomicAdd(p.stat_sum + nb + lane, fa[c].x);
                    atomicAdd(p.stat_sq + nb + lane, fa[c].z);
                }
            }
        }
    } else if (!A_TMA) {
        // ===================== Gather producers (128 threads, one tile row each) =====================
        if (p.g.mode == GM_STEM) f_gather_loop<Cfg, true>(p, bars, smem_base, num_tiles);
        else f_gather_loop<Cfg, false>(p, bars, smem_base, num_tiles);
    }

    tc_fence_before();
    __syncthreads();
    if (PAIR) cluster_sync_all();          // neither CTA frees TMEM or exits while its partner may still use it
    if (warp == Cfg::MMA_WARP) {
        if (PAIR) tmem_dealloc_2sm<Cfg::TMEM_COLS>(tmem_base);
        else tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
    }
}

// ------------------------------------------------------------------------------------------
// igemm_w
// ------------------------------------------------------------------------------------------
// M2: the CTA owns TWO 128-row Cout blocks (two TMEM accumulators) that share every im2col / activation tile it
// loads - a third fewer bytes into the SM per MMA (the kernel is bound by loads: ncu + the skip-MMA experiment).
template <int NSUB, int B_MODE, bool M2 = false>
struct WCfg {
    static constexpr int BLOCK_N = 64 * NSUB;
    static constexpr int MSUB = M2 ? 4 : 2;                           // 64-wide Cout sub-tiles of A
    static constexpr int A_BYTES = MSUB * SUB_TILE_BYTES;             // [64 pixels][128 or 256 cout]
    static constexpr int B_BYTES = NSUB * SUB_TILE_BYTES;              // [64 pixels][64*NSUB]
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = M2 ? 3 : (NSUB >= 3 ? 4 : 6);
    static constexpr int THREADS = B_MODE != AM_GATHER ? (MMA_WARP + 1) * 32 : (GATHER_WARP0 + 4) * 32;
    static constexpr int TMEM_COLS = M2 ? 512 : (NSUB == 1 ? 64 : (NSUB == 2 ? 128 : 256));
    static constexpr size_t SMEM_BYTES = 1024 + (size_t)STAGES * STAGE_BYTES + sizeof(Barriers<STAGES>);
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

// Gather-producer loop of igemm_w: 64 pixel rows per k-block, NSUB sub-tiles (tap, channel block) each.
template <class Cfg, int NSUB, bool STEM>
__device__ __forceinline__ void w_gather_loop(const WParams& p, Barriers<Cfg::STAGES>* bars, uint32_t smem_base,
                                              int n_blk, int kb_begin, int kb_end) {
    constexpr int STAGES = Cfg::STAGES;
    constexpr int NR = STEM ? 8 : 4;
    constexpr int ROW_STEP = STEM ? 8 : 16;
    const int tid = threadIdx.x - GATHER_WARP0 * 32;
    const int chunk = tid & 7;
    const int sub = STEM ? (tid >> 4) : (tid >> 3);
    const uint32_t dst_off = STEM ? (uint32_t)(sub * 128 + ((((tid & 15) >> 1) ^ (sub & 7)) << 4) + (tid & 1) * 8)
                                  : (uint32_t)(sub * 128 + ((chunk ^ (sub & 7)) << 4));
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = kb_begin; kb < kb_end; ++kb) {
        RowSet<NR> rs;
        rowset_decode<NR, ROW_STEP>(p.g, (long long)kb * 64 + sub, p.P, rs);
        mbar_wait(&bars->empty[stage], phase ^ 1);
        const uint32_t b_dst = smem_base + stage * Cfg::STAGE_BYTES + Cfg::A_BYTES + dst_off;
#pragma unroll
        for (int j = 0; j < NSUB; ++j) {
            const int sj = n_blk * NSUB + j;
            if (STEM) {
                gather_stem<NR, ROW_STEP * 128>(p.g, rs, sj, tid & 15, b_dst + j * SUB_TILE_BYTES);
            } else {
                const int tap = sj / p.g.cblocks;
                gather_conv<NR, ROW_STEP * 128>(p.g, rs, tap, sj - tap * p.g.cblocks, chunk, b_dst + j * SUB_TILE_BYTES);
            }
        }
        cp_async_mbar_arrive_noinc(&bars->full[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
}

// (A variant that ran CTA pairs on neighbouring Cout blocks and multicast the shared activation tile into both was
// measured no faster - the kernel is bound by per-SM ingress, which multicast does not reduce - and is gone: TMA
// multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung
// the GPU, and a weight-gradient operand handed in through the C ABI may be peer memory.)
template <int NSUB, int B_MODE, bool M2>
__global__ void __launch_bounds__(WCfg<NSUB, B_MODE, M2>::THREADS, 1) igemm_w_kernel(const __grid_constant__ WParams p) {
    using Cfg = WCfg<NSUB, B_MODE, M2>;
    constexpr int MROWS = M2 ? 256 : 128;
    constexpr bool B_TMA = B_MODE != AM_GATHER;
    constexpr int STAGES = Cfg::STAGES;
    constexpr int BLOCK_N = Cfg::BLOCK_N;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    Barriers<STAGES>* bars = reinterpret_cast<Barriers<STAGES>*>(smem + (size_t)STAGES * Cfg::STAGE_BYTES);
    const uint32_t smem_base = smem_u32(smem);
    const int warp = threadIdx.x >> 5;
    const uint32_t lane = lane_id();

    // CTA -> (split, m block, n block)
    int bid = blockIdx.x;
    const int n_blk = bid % p.n_blocks; bid /= p.n_blocks;
    const int m_blk = bid % p.m_blocks; bid /= p.m_blocks;
    const int split = bid;
    const int kb_begin = split * p.kb_per_split;
    const int kb_end = min(p.num_kb, kb_begin + p.kb_per_split);
    const int nkb = kb_end - kb_begin;     // host guarantees >= 1

    if (warp == TMA_WARP && lane == 0) {
        tma_prefetch_desc(&p.tmA);
        if (B_TMA) tma_prefetch_desc(&p.tmB);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&bars->full[s], B_TMA ? 1 : 1 + NUM_GATHER_THREADS);
            mbar_init(&bars->empty[s], 1);
        }
        mbar_init(&bars->tmem_full[0], 1);
        mbar_fence_init();
    }
    if (warp == MMA_WARP) tmem_alloc<Cfg::TMEM_COLS>(&bars->tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;
    pdl_wait();

    if (warp == TMA_WARP) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = kb_begin; kb < kb_end; ++kb) {
                mbar_wait(&bars->empty[stage], phase ^ 1);
                const uint32_t a_dst = smem_base + stage * Cfg::STAGE_BYTES;
                const uint32_t b_dst = a_dst + Cfg::A_BYTES;
                const bool skip_b = (p.dbg & 1) != 0;
                mbar_arrive_expect_tx(&bars->full[stage], Cfg::A_BYTES + ((B_TMA && !skip_b) ? Cfg::B_BYTES : 0));
                int px0 = 0, py0 = 0, pn = 0;           // patch mode: k-block = a 4 x 16 patch of output pixels
                if (B_MODE == AM_PATCH) {
                    const int patches_x = p.g.OW >> 4, per_img = patches_x * (p.g.OH >> 2);
                    pn = kb / per_img;
                    const int rem = kb - pn * per_img;
                    const int py = rem / patches_x;
                    py0 = py * 4;
                    px0 = (rem - py * patches_x) * 16;
                }
#pragma unroll
                for (int j = 0; j < Cfg::MSUB; ++j) {
                    if (B_MODE == AM_PATCH)
                        tma_load_4d(a_dst + j * SUB_TILE_BYTES, &p.tmA, &bars->full[stage], m_blk * MROWS + j * 64, px0, py0, pn);
                    else
                        tma_load_2d(a_dst + j * SUB_TILE_BYTES, &p.tmA, &bars->full[stage], m_blk * MROWS + j * 64, kb * 64);
                }
                if (B_MODE == AM_PATCH && !skip_b) {
#pragma unroll
                    for (int j = 0; j < NSUB; ++j)      // sub-tile j of this n block = filter row r' = n_blk * NSUB + j
                        tma_load_4d(b_dst + j * SUB_TILE_BYTES, &p.tmB, &bars->full[stage], 0, px0, py0 + n_blk * NSUB + j, pn);
                }
                if (B_MODE == AM_TILED && !skip_b) {
#pragma unroll
                    for (int j = 0; j < NSUB; ++j)
                        tma_load_2d(b_dst + j * SUB_TILE_BYTES, &p.tmB, &bars->full[stage], (n_blk * NSUB + j) * 64, kb * 64);
                }
                if (B_MODE == AM_IM2COL && !skip_b) {
                    // 64 output pixels starting at kb*64: one im2col load per (tap, channel block) sub-tile
                    const int hw = p.g.OH * p.g.OW;
                    const int p0 = kb * 64;
                    const int n0 = p0 / hw;
                    const int rem = p0 - n0 * hw;
                    const int oh = rem / p.g.OW, ow = rem - oh * p.g.OW;
                    const int iw0 = ow * p.g.stride - p.g.pad, ih0 = oh * p.g.stride - p.g.pad;
#pragma unroll
                    for (int j = 0; j < NSUB; ++j) {
                        const int sj = n_blk * NSUB + j;
                        const int tap = sj / p.g.cblocks;
                        const int cb = sj - tap * p.g.cblocks;
                        const int r = tap / p.g.S, s_ = tap - r * p.g.S;
                        tma_load_im2col(b_dst + j * SUB_TILE_BYTES, &p.tmB, &bars->full[stage], cb * 64, iw0, ih0, n0, s_, r);
                    }
                }
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
            if (!(p.dbg & 16)) pdl_trigger();
        }
    } else if (warp == MMA_WARP) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(128, BLOCK_N, 1, 1);
            int stage = 0;
            uint32_t phase = 0;
            for (int i = 0; i < nkb; ++i) {
                mbar_wait(&bars->full[stage], phase);
                tc_fence_after();
                const uint32_t a_addr = smem_base + stage * Cfg::STAGE_BYTES;
                const uint32_t b_addr = a_addr + Cfg::A_BYTES;
#pragma unroll
                for (int h = 0; h < (M2 ? 2 : 1); ++h) {
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {
                        if (p.dbg & 2) break;
                        const uint64_t adesc = umma_smem_desc(a_addr + h * 2 * SUB_TILE_BYTES + kk * 2048, SUB_TILE_BYTES, 1024);
                        const uint64_t bdesc = umma_smem_desc(b_addr + kk * 2048, SUB_TILE_BYTES, 1024);
                        umma_f16(tmem_base + h * BLOCK_N, adesc, bdesc, idesc, (i | kk) != 0 ? 1u : 0u);
                    }
                }
                umma_commit(&bars->empty[stage]);
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
            umma_commit(&bars->tmem_full[0]);
        }
    } else if (warp < NUM_EPI_WARPS) {
        const int q = warp & 3;
        const int half = warp >> 2;
        mbar_wait(&bars->tmem_full[0], 0);
        tc_fence_after();
        // 32-column chunks split between the two warps sharing a lane quarter
        constexpr int NCHUNK = BLOCK_N / 32;
        for (int h = 0; h < (M2 ? 2 : 1); ++h) {
            const int cout = m_blk * MROWS + h * 128 + q * 32 + lane;
            for (int c = half; c < NCHUNK; c += 2) {
                uint32_t raw[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + h * BLOCK_N + c * 32, raw);
                tmem_ld_wait();
                if (cout < p.Cout) {
                    float* dst = p.dw + (size_t)cout * p.Kw + (size_t)n_blk * BLOCK_N + c * 32;
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        red_add_v4(dst + j * 4, __uint_as_float(raw[j * 4]), __uint_as_float(raw[j * 4 + 1]),
                                   __uint_as_float(raw[j * 4 + 2]), __uint_as_float(raw[j * 4 + 3]));
                }
            }
        }
        tc_fence_before();
    } else if (!B_TMA) {
        if (p.g.mode == GM_STEM) w_gather_loop<Cfg, NSUB, true>(p, bars, smem_base, n_blk, kb_begin, kb_end);
        else w_gather_loop<Cfg, NSUB, false>(p, bars, smem_base, n_blk, kb_begin, kb_end);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

// ------------------------------------------------------------------------------------------
// Host launchers
// ------------------------------------------------------------------------------------------
template <int BLOCK_N, int A_MODE, bool B_MN, int EPI, bool BNR = false>
int launch_f(FParams& p, cudaStream_t stream) {
    using Cfg = FCfg<BLOCK_N, A_MODE, B_MN, EPI>;
    auto kern = igemm_f_kernel<BLOCK_N, A_MODE, B_MN, EPI, false, BNR>;
    static bool configured[64] = {false};
    int dev = 0;
    PTD_CUDA_CHECK(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        PTD_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM_BYTES));
        configured[dev & 63] = true;
    }
    p.m_blocks = ceil_div_i(p.M, BLOCK_M);
    p.n_blocks = ceil_div_i(p.N, BLOCK_N);
    { const char* e = getenv("PTD_IGEMM_DBG"); p.dbg = (e ? atoi(e) : 0) | (ptd_pdl_mode() == 2 ? 16 : 0); }
    {   // PTD_STATS_SMEM_KB: largest K loop (in k-blocks) that still uses the shared-memory column sums
        static int kb_max = -1;
        if (kb_max < 0) { const char* e = getenv("PTD_STATS_SMEM_KB"); kb_max = e ? atoi(e) : 4; }
        p.stats_smem = p.num_kb <= kb_max ? 1 : 0;
    }
    const int tiles = p.m_blocks * p.n_blocks;
    int grid = ptd_num_sms();
    grid = (grid / p.n_blocks) * p.n_blocks;      // fixed n block per CTA (register-resident BN partial sums)
    if (grid < p.n_blocks) grid = p.n_blocks;
    if (grid > tiles) grid = tiles;               // tiles is a multiple of n_blocks
    PTD_CUDA_CHECK(ptd_launch(kern, dim3(grid), dim3(Cfg::THREADS), Cfg::SMEM_BYTES, stream, 1, p));
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

// CTA-pair launch (cluster of 2, tcgen05 cta_group::2): 256 x BLOCK_N tiles, tiled 1x1 path.
template <int BLOCK_N, bool B_MN, bool BNR = false>
int launch_f_pair(FParams& p, cudaStream_t stream) {
    using Cfg = FCfg<BLOCK_N, AM_TILED, B_MN, 0, true>;
    auto kern = igemm_f_kernel<BLOCK_N, AM_TILED, B_MN, 0, true, BNR>;
    static bool configured[64] = {false};
    int dev = 0;
    PTD_CUDA_CHECK(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        PTD_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM_BYTES));
        configured[dev & 63] = true;
    }
    p.m_blocks = ceil_div_i(p.M, 2 * BLOCK_M);
    p.n_blocks = ceil_div_i(p.N, BLOCK_N);
    { const char* e = getenv("PTD_IGEMM_DBG"); p.dbg = (e ? atoi(e) : 0) | (ptd_pdl_mode() == 2 ? 16 : 0); }
    p.stats_smem = 0;
    const int tiles = p.m_blocks * p.n_blocks;
    int workers = ptd_num_sms() / 2;
    workers = (workers / p.n_blocks) * p.n_blocks;      // fixed n block per worker
    if (workers < p.n_blocks) workers = p.n_blocks;
    if (workers > tiles) workers = tiles;
    PTD_CUDA_CHECK(ptd_launch(kern, dim3(2 * workers), dim3(Cfg::THREADS), Cfg::SMEM_BYTES, stream, 2, p));
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

// K loop length (k-blocks) from which a 1x1 layer runs on CTA pairs; 0 = never.
int pair_min_kb() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("PTD_PAIR_MIN_KB"); v = e ? atoi(e) : 8; }
    return v;
}
bool use_pair(const FParams& p, int block_n) {
    const int mk = pair_min_kb();
    return mk > 0 && block_n >= 128 && p.num_kb >= mk && p.addend == nullptr && p.map_mul == 0;
}
template <bool B_MN>
int dispatch_f_pair(FParams& p, int block_n, cudaStream_t stream) {
    if constexpr (B_MN) {                 // data gradients may carry the fused BatchNorm-backward reduction
        if (p.bn_y != nullptr) {
            if (block_n == 256) return launch_f_pair<256, B_MN, true>(p, stream);
            return launch_f_pair<128, B_MN, true>(p, stream);
        }
    }
    if (block_n == 256) return launch_f_pair<256, B_MN>(p, stream);
    return launch_f_pair<128, B_MN>(p, stream);
}

template <int A_MODE, bool B_MN, int EPI>
int dispatch_f(FParams& p, int block_n, cudaStream_t stream) {
    if constexpr (B_MN && EPI == 0 && A_MODE != AM_GATHER) {
        if (p.bn_y != nullptr) {
            switch (block_n) {
                case 64: return launch_f<64, A_MODE, B_MN, EPI, true>(p, stream);
                case 128: return launch_f<128, A_MODE, B_MN, EPI, true>(p, stream);
                case 256: return launch_f<256, A_MODE, B_MN, EPI, true>(p, stream);
            }
        }
    }
    if (p.bn_y != nullptr) {
        ptd_set_error("igemm_f: this kernel variant cannot carry the fused BatchNorm reduction");
        return PTD_ERR_UNSUPPORTED;
    }
    switch (block_n) {
        case 64: return launch_f<64, A_MODE, B_MN, EPI>(p, stream);
        case 128: return launch_f<128, A_MODE, B_MN, EPI>(p, stream);
        case 256: return launch_f<256, A_MODE, B_MN, EPI>(p, stream);
    }
    ptd_set_error("igemm_f: unsupported BLOCK_N %d", block_n);
    return PTD_ERR_ARG;
}

int pick_block_n(int n) {
    if (n % 256 == 0) return 256;
    if (n % 128 == 0) return 128;
    return 64;
}

template <int NSUB, int B_MODE, bool M2>
int launch_w_impl(WParams& p, int grid, cudaStream_t stream) {
    using Cfg = WCfg<NSUB, B_MODE, M2>;
    auto kern = igemm_w_kernel<NSUB, B_MODE, M2>;
    static bool configured[64] = {false};
    int dev = 0;
    PTD_CUDA_CHECK(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        PTD_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM_BYTES));
        configured[dev & 63] = true;
    }
    PTD_CUDA_CHECK(ptd_launch(kern, dim3(grid), dim3(Cfg::THREADS), Cfg::SMEM_BYTES, stream, 1, p));
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

template <int NSUB, int B_MODE>
int launch_w(WParams& p, cudaStream_t stream) {
    { const char* e = getenv("PTD_IGEMM_DBG"); p.dbg = (e ? atoi(e) : 0) | (ptd_pdl_mode() == 2 ? 16 : 0); }
    static int dual_m = -1;
    if (dual_m < 0) { const char* e = getenv("PTD_WGRAD_M2"); dual_m = e ? atoi(e) : 1; }
    // measured per layer: im2col weight gradients 99 -> 72 us; plain 1x1 ones lose a few us (fewer stages), so only im2col
    constexpr bool M2_OK = B_MODE == AM_IM2COL && NSUB >= 3;          // 2 x (64*NSUB) TMEM columns, 3 stages of <= 64 KiB
    const bool m2 = M2_OK && dual_m && p.Cout >= 256;
    p.m_blocks = ceil_div_i(p.Cout, m2 ? 256 : 128);
    p.n_blocks = p.Kw / (64 * NSUB);
    p.num_kb = ceil_div_i(p.P, 64);
    const int tiles = p.m_blocks * p.n_blocks;
    // One CTA per SM (shared memory) => size the grid to whole waves.  ncu: a 2.07-wave grid (306 CTAs) took three
    // CTA lifetimes, each with ~20 us of fixed cost (TMEM alloc, pipeline fill, 32K reduction atomics) around a
    // ~17 us main loop; one wave with a longer K range per CTA halves the kernel.
    static float waves = -1.f;
    if (waves < 0.f) { const char* e = getenv("PTD_WGRAD_WAVES"); waves = e ? (float)atof(e) : 1.f; }
    int splits = (int)(ptd_num_sms() * waves) / tiles;
    if (splits > p.num_kb) splits = p.num_kb;
    if (splits < 1) splits = 1;
    p.kb_per_split = ceil_div_i(p.num_kb, splits);
    p.splits = ceil_div_i(p.num_kb, p.kb_per_split);            // no empty split
    const int grid = tiles * p.splits;
    if (m2) return launch_w_impl<NSUB, B_MODE, M2_OK>(p, grid, stream);
    return launch_w_impl<NSUB, B_MODE, false>(p, grid, stream);
}

template <int B_MODE>
int dispatch_w(WParams& p, int nsub, cudaStream_t stream) {
    switch (nsub) {
        case 1: return launch_w<1, B_MODE>(p, stream);
        case 2: return launch_w<2, B_MODE>(p, stream);
        case 3: return launch_w<3, B_MODE>(p, stream);
        case 4: return launch_w<4, B_MODE>(p, stream);
    }
    ptd_set_error("igemm_w: unsupported sub-tile count %d", nsub);
    return PTD_ERR_ARG;
}

int pick_nsub(int ksub) {   // ksub = Kw / 64
    { static int f = -1; if (f < 0) { const char* e = getenv("PTD_WGRAD_NSUB"); f = e ? atoi(e) : 0; }
      if (f > 0 && ksub % f == 0) return f; }
    if (ksub % 4 == 0) return 4;
    if (ksub % 3 == 0) return 3;
    if (ksub % 2 == 0) return 2;
    return 1;
}

// Output through the TMA store engine: possible when GEMM row m is output row m (no row map) and nothing is added
// to the tile on its way out.
int enable_tma_store(FParams& p) {
    static int on = -1;
    if (on < 0) { const char* e = getenv("PTD_TMA_STORE"); on = e ? atoi(e) : 1; }
    if (!on || p.addend != nullptr || p.map_mul != 0 || p.out == nullptr) return PTD_OK;
    int rc = ptd_make_tmap_2d_bf16_store64(&p.tmO, p.out, (uint64_t)p.ldo, (uint64_t)p.M, (uint64_t)p.ldo * 2, 32);
    if (rc) return rc;
    p.tma_store = 1;
    return PTD_OK;
}

bool force_gather() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("PTD_IGEMM_FORCE_GATHER"); v = (e && atoi(e)) ? 1 : 0; }
    return v == 1;
}

// Natural tap order (r, s); offsets = the tap itself (fprop / wgrad) or mirrored (stride-1 data gradient).
void fill_taps(FParams& p, int R, int S, bool mirrored) {
    if (R * S > 9) { R = 1; S = 1; }      // callers check; keeps the table in bounds
    p.ntaps = R * S;
    for (int r = 0; r < R; ++r)
        for (int s = 0; s < S; ++s) {
            const int t = r * S + s;
            p.tap_wtap[t] = (signed char)t;
            p.tap_offw[t] = (signed char)(mirrored ? S - 1 - s : s);
            p.tap_offh[t] = (signed char)(mirrored ? R - 1 - r : r);
        }
}

void fill_geom(ConvGeom& g, const bf16* src, int H, int W, int C, int OH, int OW, int R, int S, int stride, int pad,
               int mode) {
    g.src = src; g.H = H; g.W = W; g.C = C; g.OH = OH; g.OW = OW; g.R = R; g.S = S;
    g.stride = stride; g.pad = pad; g.mode = mode; g.cblocks = C / 64;
}

}  // namespace

// ------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------
extern "C" int ptd_conv_fprop(const void* x, const void* w, void* y, float* stat_sum, float* stat_sq, int N, int H,
                              int W, int Cin, int Cout, int R, int S, int stride, int pad, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
    PTD_REQUIRE(Cin % 64 == 0 && Cout % 64 == 0, "conv_fprop: Cin (%d) and Cout (%d) must be multiples of 64", Cin, Cout);
    FParams p;
    memset(&p, 0, sizeof(p));
    p.M = N * OH * OW; p.N = Cout; p.num_kb = R * S * Cin / 64;
    p.out = (bf16*)y; p.ldo = Cout; p.stat_sum = stat_sum; p.stat_sq = stat_sq;
    const int bn = pick_block_n(Cout);
    const bool plain = (R == 1 && S == 1 && stride == 1 && pad == 0);
    const bool pair = plain && use_pair(p, bn);
    // a CTA of a pair loads half of the weight rows of the tile
    int rc = ptd_make_tmap_2d_bf16(&p.tmB, w, (uint64_t)R * S * Cin, Cout, (uint64_t)R * S * Cin * 2, 64, pair ? bn / 2 : bn);
    if (rc) return rc;
    rc = enable_tma_store(p);
    if (rc) return rc;
    if (plain) {
        rc = ptd_make_tmap_2d_bf16(&p.tmA, x, Cin, p.M, (uint64_t)Cin * 2, 64, BLOCK_M);
        if (rc) return rc;
        if (pair) return dispatch_f_pair<false>(p, bn, stream);
        return dispatch_f<AM_TILED, false, 0>(p, bn, stream);
    }
    fill_geom(p.g, (const bf16*)x, H, W, Cin, OH, OW, R, S, stride, pad, GM_FPROP);
    if (!force_gather() && R * S <= 9) {
        // walk output pixels over the padded input with TMA im2col: corners as in CUTLASS' fprop
        rc = ptd_make_tmap_im2col_bf16(&p.tmA, x, Cin, W, H, N, -pad, -pad, pad - (S - 1), pad - (R - 1), BLOCK_M, stride);
        if (rc) return rc;
        fill_taps(p, R, S, false);
        return dispatch_f<AM_IM2COL, false, 0>(p, bn, stream);
    }
    return dispatch_f<AM_GATHER, false, 0>(p, bn, stream);
}

namespace {
// Fused BatchNorm-backward reduction riding on a data gradient (see FParams::bn_*): all six pointers or none.
struct BnReduce {
    const void* y;
    const void* mask;
    const float* mean;
    const float* invstd;
    float* sum_g;
    float* sum_gx;
};
void set_bn(FParams& p, const BnReduce* bn) {
    if (bn == nullptr || bn->y == nullptr) return;
    p.bn_y = (const bf16*)bn->y; p.bn_mask = (const uint8_t*)bn->mask; p.bn_mean = bn->mean; p.bn_invstd = bn->invstd;
    p.bn_sum_g = bn->sum_g; p.bn_sum_gx = bn->sum_gx;
}
int conv_dgrad_impl(const void* dy, const void* w, void* dx, const void* addend, const void* addend_mask, const BnReduce* bnr,
                    int N, int H, int W, int Cin, int Cout, int R, int S, int stride, int pad, void* stream_);
}  // namespace

extern "C" int ptd_conv_dgrad(const void* dy, const void* w, void* dx, const void* addend, int N, int H, int W,
                              int Cin, int Cout, int R, int S, int stride, int pad, void* stream_) {
    return conv_dgrad_impl(dy, w, dx, addend, nullptr, nullptr, N, H, W, Cin, Cout, R, S, stride, pad, stream_);
}

// Data gradient with everything the residual block's backward fuses into it: an addend (optionally gated by a ReLU bit
// mask: the identity-branch gradient), and the reduction pass of the BatchNorm backward that CONSUMES this gradient -
// dx is the gradient behind BatchNorm + ReLU of a layer whose conv output is bn_y (dx's shape) and whose ReLU bits are
// bn_mask; sum_g[c] += sum(g), sum_gx[c] += sum(g * xhat) with g = mask ? dx : 0, exactly what ptd_bn_bwd_reduce
// (relu = 2) computes from a second read of dx, y and the mask.  Returns PTD_ERR_UNSUPPORTED before launching anything
// when the shape takes the gather fallback (the caller then runs the separate reduce kernel).
extern "C" int ptd_conv_dgrad_fused(const void* dy, const void* w, void* dx, const void* addend, const void* addend_mask,
                                    const void* bn_y, const void* bn_mask, const float* bn_mean, const float* bn_invstd,
                                    float* sum_g, float* sum_gx, int N, int H, int W, int Cin, int Cout, int R, int S,
                                    int stride, int pad, void* stream_) {
    BnReduce bn = {bn_y, bn_mask, bn_mean, bn_invstd, sum_g, sum_gx};
    if (bn_y != nullptr)
        PTD_REQUIRE(bn_mask && bn_mean && bn_invstd && sum_g && sum_gx, "conv_dgrad_fused: the BatchNorm reduction needs all six pointers");
    return conv_dgrad_impl(dy, w, dx, addend, addend_mask, bn_y ? &bn : nullptr, N, H, W, Cin, Cout, R, S, stride, pad, stream_);
}

namespace {
int conv_dgrad_impl(const void* dy, const void* w, void* dx, const void* addend, const void* addend_mask, const BnReduce* bnr,
                    int N, int H, int W, int Cin, int Cout, int R, int S, int stride, int pad, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
    PTD_REQUIRE(Cin % 64 == 0 && Cout % 64 == 0, "conv_dgrad: Cin (%d) and Cout (%d) must be multiples of 64", Cin, Cout);
    FParams p;
    memset(&p, 0, sizeof(p));
    p.M = N * H * W; p.N = Cin; p.num_kb = R * S * Cout / 64;
    p.out = (bf16*)dx; p.ldo = Cin; p.addend = (const bf16*)addend; p.addend_mask = (const uint8_t*)addend_mask;
    p.b_tap_cols = Cin; p.b_cblocks = Cout / 64;
    PTD_REQUIRE(addend_mask == nullptr || addend != nullptr, "conv_dgrad: an addend mask needs an addend");
    const bool plain = (R == 1 && S == 1 && stride == 1 && pad == 0);
    const bool tma_path = !force_gather() && R == S && R * S <= 9 &&
                          (plain || stride == 1 ||
                           (stride == 2 && R == 3 && (H % 2 == 0) && (W % 2 == 0) &&
                            (H + 2 * pad - R) / 2 + 1 == H / 2 && (W + 2 * pad - S) / 2 + 1 == W / 2));
    if (bnr != nullptr && !tma_path) {
        ptd_set_error("conv_dgrad_fused: this shape takes the gather fallback, which cannot carry the BatchNorm reduction");
        return PTD_ERR_UNSUPPORTED;
    }
    if (addend_mask != nullptr && !plain) {
        ptd_set_error("conv_dgrad_fused: a gated addend is only supported on the 1x1 stride-1 path");
        return PTD_ERR_UNSUPPORTED;
    }
    set_bn(p, bnr);
    const int bn = pick_block_n(Cin);
    // forward weight tensor viewed as [Cout rows][R*S*Cin cols]; a 64x64 box at (tap*Cin + n, cout) is the
    // MN-major B sub-tile of that tap.
    int rc = ptd_make_tmap_2d_bf16(&p.tmB, w, (uint64_t)R * S * Cin, Cout, (uint64_t)R * S * Cin * 2, 64, 64);
    if (rc) return rc;
    if (plain) {
        rc = ptd_make_tmap_2d_bf16(&p.tmA, dy, Cout, p.M, (uint64_t)Cout * 2, 64, BLOCK_M);
        if (rc) return rc;
        rc = enable_tma_store(p);
        if (rc) return rc;
        if (use_pair(p, bn)) return dispatch_f_pair<true>(p, bn, stream);
        return dispatch_f<AM_TILED, true, 0>(p, bn, stream);
    }
    fill_geom(p.g, (const bf16*)dy, OH, OW, Cout, H, W, R, S, stride, pad, GM_DGRAD);
    if (force_gather() || R != S || R * S > 9) return dispatch_f<AM_GATHER, true, 0>(p, bn, stream);
    if (stride == 1) {
        // dx pixel (h, w) reads dy at (h + pad - r, w + pad - s): im2col over dy with the tap order mirrored
        rc = ptd_make_tmap_im2col_bf16(&p.tmA, dy, Cout, OW, OH, N, pad - (S - 1), pad - (R - 1),
                                       pad - (S - 1) + (W - OW), pad - (R - 1) + (H - OH), BLOCK_M, 1);
        if (rc) return rc;
        fill_taps(p, R, S, true);
        p.g.pad = (S - 1) - pad;          // tile origin = (w - ((S-1) - pad), h - ...)
        rc = enable_tma_store(p);
        if (rc) return rc;
        return dispatch_f<AM_IM2COL, true, 0>(p, bn, stream);
    }
    if (stride == 2 && (H % 2 == 0) && (W % 2 == 0) && OH == H / 2 && OW == W / 2) {
        // Strided data gradient without wasted taps: split dx pixels by parity (ph, pw).  Pixel (2a+ph, 2b+pw)
        // only sees taps with (ph + pad - r) even, reading dy at (a + (ph+pad-r)/2, b + ...): each parity class
        // is a dense stride-1 walk over dy (im2col, zero-filled past the edge) whose rows land on a strided
        // subset of dx.  The four classes together use every tap exactly once.
        if (R == 1) {
            // 1x1: only class (0,0) receives anything; the rest of dx is the addend (or zero)
            if (addend != nullptr && addend != dx)
                PTD_CUDA_CHECK(cudaMemcpyAsync(dx, addend, (size_t)N * H * W * Cin * 2, cudaMemcpyDeviceToDevice, stream));
            else if (addend == nullptr)
                PTD_CUDA_CHECK(cudaMemsetAsync(dx, 0, (size_t)N * H * W * Cin * 2, stream));
            p.M = N * OH * OW;
            p.addend = (const bf16*)dx;       // in place: every mapped element is read and written by one thread
            p.map_mul = 2; p.map_ph = 0; p.map_pw = 0; p.map_sub_ow = OW; p.map_sub_ohw = OH * OW; p.map_oH = H; p.map_oW = W;
            rc = ptd_make_tmap_2d_bf16(&p.tmA, dy, Cout, p.M, (uint64_t)Cout * 2, 64, BLOCK_M);
            if (rc) return rc;
            return dispatch_f<AM_TILED, true, 0>(p, bn, stream);
        }
        rc = ptd_make_tmap_im2col_bf16(&p.tmA, dy, Cout, OW, OH, N, 0, 0, 0, 0, BLOCK_M, 1);
        if (rc) return rc;
        bool ok = true;
        for (int ph = 0; ph < 2 && ok; ++ph)
            for (int pw = 0; pw < 2 && ok; ++pw) {
                FParams q = p;
                q.ntaps = 0;
                for (int r = 0; r < R; ++r) {
                    const int th = ph + pad - r;
                    if (th < 0 || (th & 1)) continue;
                    for (int s2 = 0; s2 < S; ++s2) {
                        const int tw = pw + pad - s2;
                        if (tw < 0 || (tw & 1)) continue;
                        q.tap_wtap[q.ntaps] = (signed char)(r * S + s2);
                        q.tap_offh[q.ntaps] = (signed char)(th / 2);
                        q.tap_offw[q.ntaps] = (signed char)(tw / 2);
                        ++q.ntaps;
                    }
                }
                q.M = N * (H / 2) * (W / 2);
                q.g.OH = H / 2; q.g.OW = W / 2; q.g.stride = 1; q.g.pad = 0;
                q.map_mul = 2; q.map_ph = ph; q.map_pw = pw; q.map_sub_ow = W / 2; q.map_sub_ohw = (H / 2) * (W / 2);
                q.map_oH = H; q.map_oW = W;
                if (q.ntaps == 0) { ok = false; break; }      // a class without taps would need a zero fill
                q.num_kb = q.ntaps * (Cout / 64);
                rc = dispatch_f<AM_IM2COL, true, 0>(q, bn, stream);
                if (rc) return rc;
            }
        if (ok) return PTD_OK;
    }
    return dispatch_f<AM_GATHER, true, 0>(p, bn, stream);
}
}  // namespace

// 1x1 stride-1 data gradient with a gated addend: dx = dy * W + (mask bit ? addend : 0).  `mask` is the ReLU bit
// tensor of ptd_bn_apply_mask over dx's shape: the residual block's identity-branch gradient is the block-output
// gradient behind the final ReLU, so it can be taken from dz directly instead of from a materialised copy.
extern "C" int ptd_conv_dgrad_masked_addend(const void* dy, const void* w, void* dx, const void* addend,
                                            const void* mask, int N, int H, int W, int Cin, int Cout, void* stream_) {
    PTD_REQUIRE(addend != nullptr && mask != nullptr, "conv_dgrad_masked_addend: addend and mask are required");
    return conv_dgrad_impl(dy, w, dx, addend, mask, nullptr, N, H, W, Cin, Cout, 1, 1, 1, 0, stream_);
}

// dx += conv_transpose(dy, w)  (in place).  For the 1x1 stride-2 shortcut convolution this touches only the
// strided quarter of dx - the reason the fused bottleneck backward calls it after the main branch wrote dx.
extern "C" int ptd_conv_dgrad_accum(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout,
                                    int R, int S, int stride, int pad, void* stream_) {
    return ptd_conv_dgrad(dy, w, dx, dx, N, H, W, Cin, Cout, R, S, stride, pad, stream_);
}

extern "C" int ptd_conv_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int R,
                              int S, int stride, int pad, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
    PTD_REQUIRE(Cin % 64 == 0 && Cout % 64 == 0, "conv_wgrad: Cin (%d) and Cout (%d) must be multiples of 64", Cin, Cout);
    WParams p;
    memset(&p, 0, sizeof(p));
    p.P = N * OH * OW; p.Cout = Cout; p.Kw = R * S * Cin; p.dw = dw;
    int rc = ptd_make_tmap_2d_bf16(&p.tmA, dy, Cout, p.P, (uint64_t)Cout * 2, 64, 64);
    if (rc) return rc;
    const int nsub = pick_nsub(p.Kw / 64);
    const bool plain = (R == 1 && S == 1 && stride == 1 && pad == 0);
    if (plain) {
        rc = ptd_make_tmap_2d_bf16(&p.tmB, x, Cin, p.P, (uint64_t)Cin * 2, 64, 64);
        if (rc) return rc;
        return dispatch_w<AM_TILED>(p, nsub, stream);
    }
    fill_geom(p.g, (const bf16*)x, H, W, Cin, OH, OW, R, S, stride, pad, GM_FPROP);
    if (!force_gather()) {
        rc = ptd_make_tmap_im2col_bf16(&p.tmB, x, Cin, W, H, N, -pad, -pad, pad - (S - 1), pad - (R - 1), 64, stride);
        if (rc) return rc;
        return dispatch_w<AM_IM2COL>(p, nsub, stream);
    }
    return dispatch_w<AM_GATHER>(p, nsub, stream);
}

// Stem: x4 is NHWC with 4 channels (3 real + 1 zero), w4 is [64][8][8][4] (filter row 7 / tap 7 / channel 3 zero).
extern "C" int ptd_stem_fprop(const void* x4, const void* w4, void* y, float* stat_sum, float* stat_sq, int N, int H,
                              int W, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    const int OH = (H + 6 - 7) / 2 + 1, OW = (W + 6 - 7) / 2 + 1;
    FParams p;
    memset(&p, 0, sizeof(p));
    p.M = N * OH * OW; p.N = 64; p.num_kb = 4;
    p.out = (bf16*)y; p.ldo = 64; p.stat_sum = stat_sum; p.stat_sq = stat_sq;
    int rc = ptd_make_tmap_2d_bf16(&p.tmB, w4, 256, 64, 512, 64, 64);
    if (rc) return rc;
    fill_geom(p.g, (const bf16*)x4, H, W, 4, OH, OW, 7, 7, 2, 3, GM_STEM);
    return dispatch_f<AM_GATHER, false, 0>(p, 64, stream);
}

extern "C" int ptd_stem_wgrad(const void* x4, const void* dy, float* dw4, int N, int H, int W, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    const int OH = (H + 6 - 7) / 2 + 1, OW = (W + 6 - 7) / 2 + 1;
    WParams p;
    memset(&p, 0, sizeof(p));
    p.P = N * OH * OW; p.Cout = 64; p.Kw = 256; p.dw = dw4;
    int rc = ptd_make_tmap_2d_bf16(&p.tmA, dy, 64, p.P, 128, 64, 64);
    if (rc) return rc;
    fill_geom(p.g, (const bf16*)x4, H, W, 4, OH, OW, 7, 7, 2, 3, GM_STEM);
    return dispatch_w<AM_GATHER>(p, 4, stream);
}

// Stem on the space-to-depth input (pool_layout.cu: nchw3_to_s2d_kernel): xs is [N][OH+3][OW+3][16].
namespace {
bool stem_s2d_shape_ok(int H, int W) { return H % 16 == 0 && W % 32 == 0; }      // OH % 8 == 0, OW % 16 == 0
// rows of output pixels x 128-byte windows: element (e, ow, hs, n) sits at ((n*Hs + hs)*Ws + ow)*16 + e, e < 64
int stem_window_map(CUtensorMap* out, const void* xs, int N, int OH, int OW, int box_rows) {
    const uint64_t Hs = OH + 3, Ws = OW + 3;
    const uint64_t dims[4] = {64, (uint64_t)OW, Hs, (uint64_t)N};
    const uint64_t strides[3] = {32, Ws * 32, Hs * Ws * 32};
    const uint32_t box[4] = {64, 16, (uint32_t)box_rows, 1};
    return ptd_make_tmap_4d_bf16(out, xs, dims, strides, box);
}
}  // namespace

extern "C" int ptd_stem_s2d_fprop(const void* xs, const void* w16, void* y, float* stat_sum, float* stat_sq, int N,
                                  int H, int W, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    PTD_REQUIRE(stem_s2d_shape_ok(H, W), "stem_s2d: needs H %% 16 == 0 and W %% 32 == 0 (got %d x %d)", H, W);
    const int OH = H / 2, OW = W / 2;
    FParams p;
    memset(&p, 0, sizeof(p));
    p.M = N * OH * OW; p.N = 64; p.num_kb = 4;
    p.out = (bf16*)y; p.ldo = 64; p.stat_sum = stat_sum; p.stat_sq = stat_sq;
    int rc = ptd_make_tmap_2d_bf16(&p.tmB, w16, 256, 64, 512, 64, 64);
    if (rc) return rc;
    rc = stem_window_map(&p.tmA, xs, N, OH, OW, 8);
    if (rc) return rc;
    fill_geom(p.g, (const bf16*)xs, OH + 3, OW + 3, 16, OH, OW, 4, 4, 1, 0, GM_FPROP);
    return dispatch_f<AM_PATCH, false, 0>(p, 64, stream);
}

extern "C" int ptd_stem_s2d_wgrad(const void* xs, const void* dy, float* dw16, int N, int H, int W, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    PTD_REQUIRE(stem_s2d_shape_ok(H, W), "stem_s2d: needs H %% 16 == 0 and W %% 32 == 0 (got %d x %d)", H, W);
    const int OH = H / 2, OW = W / 2;
    WParams p;
    memset(&p, 0, sizeof(p));
    p.P = N * OH * OW; p.Cout = 64; p.Kw = 256; p.dw = dw16;
    // dY as [4 x 16 pixel patch][64 channels] so that its rows pair up with the window rows of the same patch
    const uint64_t dims[4] = {64, (uint64_t)OW, (uint64_t)OH, (uint64_t)N};
    const uint64_t strides[3] = {128, (uint64_t)OW * 128, (uint64_t)OH * OW * 128};
    const uint32_t box[4] = {64, 16, 4, 1};
    int rc = ptd_make_tmap_4d_bf16(&p.tmA, dy, dims, strides, box);
    if (rc) return rc;
    rc = stem_window_map(&p.tmB, xs, N, OH, OW, 4);
    if (rc) return rc;
    fill_geom(p.g, (const bf16*)xs, OH + 3, OW + 3, 16, OH, OW, 4, 4, 1, 0, GM_FPROP);
    return dispatch_w<AM_PATCH>(p, 4, stream);
}

// Classifier / generic dense layers.  a: [M][K] bf16, w: [N][K] bf16.
//   forward : out_f32[M][N] = a * w^T + bias           (N need not be a multiple of the tile)
extern "C" int ptd_linear_fwd(const void* a, const void* w, const float* bias, float* out, int M, int N, int K,
                              void* stream_) {
    PTD_REQUIRE(K % 8 == 0, "linear_fwd: K (%d) must be a multiple of 8 (16-byte TMA row pitch)", K);
    FParams p;
    memset(&p, 0, sizeof(p));
    p.M = M; p.N = N; p.num_kb = ceil_div_i(K, 64);
    p.out_f32 = out; p.bias = bias; p.ldo = N;
    const int bn = 128;
    int rc = ptd_make_tmap_2d_bf16(&p.tmA, a, K, M, (uint64_t)K * 2, 64, BLOCK_M);
    if (rc) return rc;
    rc = ptd_make_tmap_2d_bf16(&p.tmB, w, K, N, (uint64_t)K * 2, 64, bn);
    if (rc) return rc;
    return dispatch_f<AM_TILED, false, 1>(p, bn, (cudaStream_t)stream_);
}

//   data gradient : da_bf16[M][K] = dout_bf16[M][N] * w[N][K]        (K % 64 == 0)
extern "C" int ptd_linear_dgrad(const void* dout, const void* w, void* da, int M, int N, int K, void* stream_) {
    PTD_REQUIRE(K % 64 == 0 && N % 8 == 0, "linear_dgrad: K (%d) %% 64 and N (%d) %% 8 must be 0", K, N);
    FParams p;
    memset(&p, 0, sizeof(p));
    p.M = M; p.N = K; p.num_kb = ceil_div_i(N, 64);
    p.out = (bf16*)da; p.ldo = K;
    p.b_tap_cols = 0; p.b_cblocks = p.num_kb;
    const int bn = pick_block_n(K);
    int rc = ptd_make_tmap_2d_bf16(&p.tmA, dout, N, M, (uint64_t)N * 2, 64, BLOCK_M);
    if (rc) return rc;
    rc = ptd_make_tmap_2d_bf16(&p.tmB, w, K, N, (uint64_t)K * 2, 64, 64);
    if (rc) return rc;
    return dispatch_f<AM_TILED, true, 0>(p, bn, (cudaStream_t)stream_);
}

//   weight gradient : dw_f32[N][K] += dout_bf16[M][N]^T * a_bf16[M][K]   (K % 64 == 0)
extern "C" int ptd_linear_wgrad(const void* a, const void* dout, float* dw, int M, int N, int K, void* stream_) {
    PTD_REQUIRE(K % 64 == 0 && N % 8 == 0, "linear_wgrad: K (%d) %% 64 and N (%d) %% 8 must be 0", K, N);
    WParams p;
    memset(&p, 0, sizeof(p));
    p.P = M; p.Cout = N; p.Kw = K; p.dw = dw;
    int rc = ptd_make_tmap_2d_bf16(&p.tmA, dout, N, M, (uint64_t)N * 2, 64, 64);
    if (rc) return rc;
    rc = ptd_make_tmap_2d_bf16(&p.tmB, a, K, M, (uint64_t)K * 2, 64, 64);
    if (rc) return rc;
    return dispatch_w<AM_TILED>(p, pick_nsub(K / 64), (cudaStream_t)stream_);
}
