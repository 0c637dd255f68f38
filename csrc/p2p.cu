// Peer-to-peer runtime and collective kernels over NVLink 5 / NVSwitch.
//
// Every rank owns a "symmetric" buffer (cudaMalloc, exported with CUDA IPC, mapped by every peer) and a
// signal pad.  Collectives are single kernels that read and write peers' buffers with plain
// ld/st.global on the mapped pointers and synchronise ranks with sequence-numbered flags
// (st.release.sys / ld.acquire.sys) - there is no NCCL call anywhere in this file.
//
//   all-reduce (average), in place, "two-shot": rank r owns slice r of the buffer; for each 16-byte
//   element of its slice it loads the element from all N buffers, adds in fp32, applies the 1/N (or any)
//   scale and stores the result into all N buffers.  Every GPU's NVLink port carries 2(N-1)/N of the buffer
//   in each direction (its own reads and writes plus serving its peers') - the minimum for an all-reduce
//   without in-switch reduction - and the kernel needs two flag barriers (inputs ready / outputs complete).  The flatten step of DDP is free (gradients already live in the buffer), the averaging
//   and the optional bf16 storage format are fused into the same pass.
//   all-reduce through the switch (NVLS): when the buffer also has a multicast mapping (csrc/vmm.cu), rank r runs
//   `multimem.ld_reduce` over slice r - the NVSwitch fetches the element from all N copies and returns the fp32 sum -
//   and `multimem.st`s the scaled result to all N copies.  Each port then carries ~(1 + 1/N) of the buffer per
//   direction instead of 2(N-1)/N, and the SMs issue one load and one store per 16 bytes instead of N of each.
//   bf16 wire format: fp32 gradients are cast (pre-scaled by 1/N) into a bf16 staging buffer, reduced there by
//   either path (fp32 accumulation), and cast back - all three phases in ONE kernel, half the NVLink bytes.
//   broadcast: non-root ranks pull the root's buffer.    reduce-to-root / peer copy: used by the
//   single-process DataParallel wrapper (replicate, gather, gradient reduction).
//
// Replaces what the reference reaches through NCCL: DDP's bucket all-reduce and parameter/buffer
// broadcast (R/02 DDP基本概念/ddp_gpus.py:35,41) and nn.DataParallel's
// broadcast_coalesced / gather / reduce_add_coalesced (R/01 数据并行/...ipynb#cell5-6; K6-K9).
#include <stdlib.h>

#include "common.cuh"
#include "ptdb200.h"

namespace {

constexpr int MAX_WORLD = 8;
constexpr int MAX_BLOCKS = 128;
constexpr int AR_THREADS = 512;

// Signal pad layout (uint32 words), one pad per rank and channel:
//   [0 ..)            MAX_BLOCKS x MAX_WORLD arrival flags written by the peers (flag barrier)
//   [PAD_SEQ ..)      MAX_BLOCKS local sequence counters of the flag barrier
//   [PAD_MC_CNT ..)   MAX_BLOCKS arrival counters incremented by every rank with ONE multimem.red (switch barrier)
//   [PAD_MC_EPOCH ..) MAX_BLOCKS local counts of switch barriers passed
constexpr int PAD_SEQ = MAX_BLOCKS * MAX_WORLD;
constexpr int PAD_MC_CNT = PAD_SEQ + MAX_BLOCKS;
constexpr int PAD_MC_EPOCH = PAD_MC_CNT + MAX_BLOCKS;
constexpr int PAD_WORDS = PAD_MC_EPOCH + MAX_BLOCKS;

struct PeerPtrs {
    void* buf[MAX_WORLD];
    uint32_t* pad[MAX_WORLD];
    uint32_t* mc_pad;                 // multicast mapping of the pads (nullptr: flag barrier over the unicast mappings)
    unsigned long long timeout_ns;    // a barrier that waits longer than this reports the missing peer and traps
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Sequence numbers live in device memory next to the signal pad (one counter per block index, local to
// the rank): a kernel reads its block's counter, uses counter+1 / counter+2 for its two barriers and
// stores counter+2 back.  Every rank runs the same collectives with the same grids on a given pad, so
// the counters advance in lock step without any host involvement - which also makes the collectives
// replayable from a CUDA graph (a host-side sequence argument would be frozen at capture time).
__device__ __forceinline__ uint32_t seq_begin(const PeerPtrs& pp, int rank) {
    __shared__ uint32_t s_seq;
    if (threadIdx.x == 0) s_seq = pp.pad[rank][PAD_SEQ + blockIdx.x];
    __syncthreads();
    return s_seq + 1;
}
__device__ __forceinline__ void seq_end(const PeerPtrs& pp, int rank, uint32_t last_used) {
    if (threadIdx.x == 0) pp.pad[rank][PAD_SEQ + blockIdx.x] = last_used;
}

// Block-level barrier across ranks: block b of every rank meets block b of every other rank.
// Flags only grow, so a late waiter still passes.
__device__ __forceinline__ void rank_barrier(const PeerPtrs& pp, int rank, int world, uint32_t seq) {
    __syncthreads();
    if ((int)threadIdx.x < world) {
        const int peer = threadIdx.x;
        __threadfence_system();
        st_release_sys(pp.pad[peer] + blockIdx.x * MAX_WORLD + rank, seq);
        const uint32_t* mine = pp.pad[rank] + blockIdx.x * MAX_WORLD + peer;
        const unsigned long long t0 = global_timer_ns();
        while ((int32_t)(ld_acquire_sys(mine) - seq) < 0) {
            if (global_timer_ns() - t0 > pp.timeout_ns) {      // a peer died or never launched: do not hang the GPU
                printf("ptd: rank %d block %d timed out waiting for rank %d (seq %u)\n", rank, blockIdx.x, peer, seq);
                __trap();
            }
        }
    }
    __syncthreads();
}

// ---- multimem (NVLS) primitives: the address is a multicast mapping, the NVSwitch does the fan-out / reduction ----
__device__ __forceinline__ float4 mm_ld_reduce_f32x4(const void* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mm_st_f32x4(void* mc, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// 8 bf16 per access, summed in fp32 inside the switch, rounded once on the way back
__device__ __forceinline__ uint4 mm_ld_reduce_bf16x8(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mm_st_b32x4(void* mc, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 ::"l"(mc), "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)),
                   "f"(__uint_as_float(v.w)) : "memory");
}

// Switch barrier: every rank adds 1 to the same counter word of ALL pads with one multimem.red (release), then
// waits until its own copy has seen `world` arrivals for this epoch.  One store through the switch instead of N
// remote stores; the epoch count lives next to the counter so a captured graph can replay it.
__device__ __forceinline__ void rank_barrier_mc(const PeerPtrs& pp, int rank, int world) {
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t* local = pp.pad[rank];
        const uint32_t epoch = local[PAD_MC_EPOCH + blockIdx.x] + 1;
        local[PAD_MC_EPOCH + blockIdx.x] = epoch;
        __threadfence_system();
        asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(pp.mc_pad + PAD_MC_CNT + blockIdx.x), "r"(1u)
                     : "memory");
        const uint32_t target = epoch * (uint32_t)world;
        const unsigned long long t0 = global_timer_ns();
        while ((int32_t)(ld_acquire_sys(local + PAD_MC_CNT + blockIdx.x) - target) < 0) {
            if (global_timer_ns() - t0 > pp.timeout_ns) {
                printf("ptd: rank %d block %d timed out in the switch barrier (epoch %u)\n", rank, blockIdx.x, epoch);
                __trap();
            }
        }
    }
    __syncthreads();
}

// Either barrier, chosen per launch.  `seq` is only consumed by the flag barrier.
__device__ __forceinline__ void rank_sync(const PeerPtrs& pp, int rank, int world, uint32_t seq) {
    if (pp.mc_pad != nullptr) rank_barrier_mc(pp, rank, world);
    else rank_barrier(pp, rank, world, seq);
}

__device__ __forceinline__ float4 ld_f4(const float4* p) {
    float4 v;
    asm volatile("ld.global.relaxed.sys.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_f4(float4* p, const float4& v) {
    asm volatile("st.global.relaxed.sys.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// Weak variants: the data itself needs no per-access system-scope ordering - the release/acquire flag barriers
// (plus __threadfence_system) order it.  L1::no_allocate keeps once-read peer lines out of L1.
__device__ __forceinline__ float4 ld_f4_weak(const float4* p) {
    float4 v;
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_f4_weak(float4* p, const float4& v) {
    asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_u4(const uint4* p) {
    uint4 v;
    asm volatile("ld.global.relaxed.sys.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_u4(uint4* p, const uint4& v) {
    asm volatile("st.global.relaxed.sys.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---- optimizer update fused behind the reduction (DDP: `loss.backward(); optimizer.step()`, ddp_gpus.py:41-42) ----------
// The parameters, momentum buffers and bf16 weight shadows of a gradient bucket live in flat arrays with the BUCKET'S
// layout (element i of the bucket belongs to element i of each), so the SGD step is a pure element-wise pass:
//     g = grad[i] + wd * p[i];   buf[i] = mom * buf[i] + g  (buf starts at zero, which makes the first step buf = g);
//     p[i] -= lr * g;   shadow[i] = bf16(p[i])
// It runs at the END of the all-reduce kernel, on exactly the vectors whose reduction this block has seen complete:
// block b of rank r reduces vectors {begin_r + b*T + t + k*G*T}; the closing rank barrier tells block b of every rank
// that block b of every OTHER rank has stored ITS such vectors everywhere - so block b may update the same index
// pattern of all N slices without any grid-wide synchronisation.  The gradient stays in the bucket untouched.
struct SgdEpilogue {
    float* param;            // nullptr = no fused update
    float* momentum_buf;     // nullptr = plain SGD
    __nv_bfloat162* shadow;  // nullptr = no bf16 shadow to refresh
    const float* lr_dev;     // learning rate in device memory (a captured step follows the host's schedule)
    float momentum, weight_decay;
};

__device__ __forceinline__ void sgd_epilogue(const SgdEpilogue& sg, const float4* grad, int world, long long n4) {
    const float lr = *sg.lr_dev;
    const long long per = (n4 + world - 1) / world;
    const long long stride = (long long)gridDim.x * AR_THREADS;
    const long long t0 = (long long)blockIdx.x * AR_THREADS + threadIdx.x;
    float4* p4 = reinterpret_cast<float4*>(sg.param);
    float4* m4 = reinterpret_cast<float4*>(sg.momentum_buf);
    for (int s = 0; s < world; ++s) {
        const long long b = min(n4, per * s), e = min(n4, b + per);
        for (long long i = b + t0; i < e; i += stride) {
            const float4 g0 = ld_f4(grad + i);          // system-scope load: written by a peer (or the switch) just now
            float4 p = p4[i];
            float4 g = make_float4(fmaf(sg.weight_decay, p.x, g0.x), fmaf(sg.weight_decay, p.y, g0.y),
                                   fmaf(sg.weight_decay, p.z, g0.z), fmaf(sg.weight_decay, p.w, g0.w));
            if (m4 != nullptr) {
                const float4 m = m4[i];
                g = make_float4(fmaf(sg.momentum, m.x, g.x), fmaf(sg.momentum, m.y, g.y), fmaf(sg.momentum, m.z, g.z),
                                fmaf(sg.momentum, m.w, g.w));
                m4[i] = g;
            }
            p.x -= lr * g.x; p.y -= lr * g.y; p.z -= lr * g.z; p.w -= lr * g.w;
            p4[i] = p;
            if (sg.shadow != nullptr) {
                sg.shadow[2 * i] = __floats2bfloat162_rn(p.x, p.y);
                sg.shadow[2 * i + 1] = __floats2bfloat162_rn(p.z, p.w);
            }
        }
    }
}

// ---- two-shot all-reduce, fp32 storage ------------------------------------------------------------
// n4 = number of float4 elements (buffer length is padded to a multiple of 4 floats by the caller).
template <int WORLD, bool WEAK>
__global__ void __launch_bounds__(AR_THREADS) allreduce_f32_kernel(PeerPtrs pp, int rank, long long n4, float scale,
                                                                   SgdEpilogue sg) {
    const uint32_t seq = seq_begin(pp, rank);
    rank_barrier(pp, rank, WORLD, seq);
    const long long per = (n4 + WORLD - 1) / WORLD;
    const long long begin = min(n4, per * rank), end = min(n4, begin + per);
    const long long stride = (long long)gridDim.x * AR_THREADS;
    // All WORLD x UNROLL loads are issued before the first add: NVLink round trips are ~2-3 us, so the amount of
    // data in flight per thread is what sets the achieved bandwidth.
    constexpr int UNROLL = WORLD <= 4 ? 4 : 2;
    long long i = begin + (long long)blockIdx.x * AR_THREADS + threadIdx.x;
    for (; i + (UNROLL - 1) * stride < end; i += UNROLL * stride) {
        float4 v[WORLD][UNROLL];
#pragma unroll
        for (int r = 0; r < WORLD; ++r) {
            const int src = (rank + r) % WORLD;        // start at the local copy, spread peers over time
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) v[r][u] = (WEAK ? ld_f4_weak : ld_f4)(reinterpret_cast<const float4*>(pp.buf[src]) + i + u * stride);
        }
        float4 acc[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            acc[u] = v[0][u];
#pragma unroll
            for (int r = 1; r < WORLD; ++r) { acc[u].x += v[r][u].x; acc[u].y += v[r][u].y; acc[u].z += v[r][u].z; acc[u].w += v[r][u].w; }
            acc[u].x *= scale; acc[u].y *= scale; acc[u].z *= scale; acc[u].w *= scale;
        }
#pragma unroll
        for (int r = 0; r < WORLD; ++r) {
            const int dst = (rank + r) % WORLD;
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) (WEAK ? st_f4_weak : st_f4)(reinterpret_cast<float4*>(pp.buf[dst]) + i + u * stride, acc[u]);
        }
    }
    for (; i < end; i += stride) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < WORLD; ++r) {
            const float4 v = (WEAK ? ld_f4_weak : ld_f4)(reinterpret_cast<const float4*>(pp.buf[(rank + r) % WORLD]) + i);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
#pragma unroll
        for (int r = 0; r < WORLD; ++r) (WEAK ? st_f4_weak : st_f4)(reinterpret_cast<float4*>(pp.buf[(rank + r) % WORLD]) + i, acc);
    }
    rank_barrier(pp, rank, WORLD, seq + 1);
    seq_end(pp, rank, seq + 1);
    if (sg.param != nullptr) sgd_epilogue(sg, reinterpret_cast<const float4*>(pp.buf[rank]), WORLD, n4);
}

// ---- two-shot all-reduce, bf16 storage (fp32 accumulate) -----------------------------------------------
template <int WORLD>
__global__ void __launch_bounds__(AR_THREADS) allreduce_bf16_kernel(PeerPtrs pp, int rank, long long n8, float scale) {
    const uint32_t seq = seq_begin(pp, rank);
    rank_barrier(pp, rank, WORLD, seq);
    const long long per = (n8 + WORLD - 1) / WORLD;
    const long long begin = min(n8, per * rank), end = min(n8, begin + per);
    const long long stride = (long long)gridDim.x * AR_THREADS;
    for (long long i = begin + (long long)blockIdx.x * AR_THREADS + threadIdx.x; i < end; i += stride) {
        float acc[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = 0.f;
        uint4 v[WORLD];
#pragma unroll
        for (int r = 0; r < WORLD; ++r) v[r] = ld_u4(reinterpret_cast<const uint4*>(pp.buf[(rank + r) % WORLD]) + i);
#pragma unroll
        for (int r = 0; r < WORLD; ++r) {
            const uint32_t w[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 f = unpack_bf16x2(w[k]);
                acc[2 * k] += f.x;
                acc[2 * k + 1] += f.y;
            }
        }
        uint4 o;
        o.x = pack_bf16x2(acc[0] * scale, acc[1] * scale);
        o.y = pack_bf16x2(acc[2] * scale, acc[3] * scale);
        o.z = pack_bf16x2(acc[4] * scale, acc[5] * scale);
        o.w = pack_bf16x2(acc[6] * scale, acc[7] * scale);
#pragma unroll
        for (int r = 0; r < WORLD; ++r) st_u4(reinterpret_cast<uint4*>(pp.buf[(rank + r) % WORLD]) + i, o);
    }
    rank_barrier(pp, rank, WORLD, seq + 1);
    seq_end(pp, rank, seq + 1);
}

// ---- all-reduce through the switch (NVLS), in place, fp32 or bf16 storage --------------------------------------
// mc = multicast mapping of the buffer; nvec = 16-byte vectors.  Rank r reduces slice r: one multimem.ld_reduce (the
// switch sums the N copies in fp32) and one multimem.st (the switch writes all N copies) per vector.
template <bool BF16>
__global__ void __launch_bounds__(AR_THREADS) allreduce_mc_kernel(PeerPtrs pp, uint4* mc, int rank, int world,
                                                                  long long nvec, float scale, SgdEpilogue sg) {
    const uint32_t seq = seq_begin(pp, rank);
    rank_sync(pp, rank, world, seq);                    // every rank's contribution is final
    const long long per = (nvec + world - 1) / world;
    const long long begin = min(nvec, per * rank), end = min(nvec, begin + per);
    const long long stride = (long long)gridDim.x * AR_THREADS;
    constexpr int UNROLL = 4;
    const bool rescale = scale != 1.f;
    long long i = begin + (long long)blockIdx.x * AR_THREADS + threadIdx.x;
    for (; i + (UNROLL - 1) * stride < end; i += UNROLL * stride) {
        if (BF16) {
            uint4 v[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) v[u] = mm_ld_reduce_bf16x8(mc + i + u * stride);
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                if (rescale) {
                    uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
                    for (int k = 0; k < 4; ++k) { const float2 f = unpack_bf16x2(w[k]); w[k] = pack_bf16x2(f.x * scale, f.y * scale); }
                    v[u] = make_uint4(w[0], w[1], w[2], w[3]);
                }
                mm_st_b32x4(mc + i + u * stride, v[u]);
            }
        } else {
            float4 v[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) v[u] = mm_ld_reduce_f32x4(mc + i + u * stride);
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                v[u].x *= scale; v[u].y *= scale; v[u].z *= scale; v[u].w *= scale;
                mm_st_f32x4(mc + i + u * stride, v[u]);
            }
        }
    }
    for (; i < end; i += stride) {
        if (BF16) {
            uint4 v = mm_ld_reduce_bf16x8(mc + i);
            if (rescale) {
                uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) { const float2 f = unpack_bf16x2(w[k]); w[k] = pack_bf16x2(f.x * scale, f.y * scale); }
                v = make_uint4(w[0], w[1], w[2], w[3]);
            }
            mm_st_b32x4(mc + i, v);
        } else {
            float4 v = mm_ld_reduce_f32x4(mc + i);
            v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
            mm_st_f32x4(mc + i, v);
        }
    }
    rank_sync(pp, rank, world, seq + 1);                // every rank's stores have landed everywhere
    seq_end(pp, rank, seq + 1);
    // pp.buf[rank] = the local (unicast) mapping of the bucket, only needed for the fused update
    if (!BF16 && sg.param != nullptr) sgd_epilogue(sg, reinterpret_cast<const float4*>(pp.buf[rank]), world, nvec);
}

// ---- all-reduce of an fp32 buffer with a bf16 wire format, one kernel, three phases ----------------------------
//   A  every rank casts scale * grad (fp32, local) into its bf16 staging buffer (symmetric);
//   B  rank r reduces slice r of the staging buffers - through the switch (MC) or by loading it from all N copies -
//      with fp32 accumulation, rounds once, and stores the bf16 result into all N staging buffers;
//   C  every rank casts its (now complete) staging buffer back into the fp32 buffer.
// Only block-level rank barriers are needed although A and C run inside the kernel: block b of every rank handles
// the SAME vector indices of every slice in all three phases, so what block b of rank r reads in B was written in A
// by block b of the peers (and likewise for C).  pp.buf = the staging buffers; n8 = vectors of 8 elements.
// Half the NVLink bytes of the fp32 kernels; results are identical on every rank (the slice owner rounds).
template <int WORLD, bool MC>
__global__ void __launch_bounds__(AR_THREADS) allreduce_wire_bf16_kernel(PeerPtrs pp, uint4* mc_stage, float4* grad,
                                                                         int rank, long long n8, float scale) {
    const uint32_t seq = seq_begin(pp, rank);
    const long long per = (n8 + WORLD - 1) / WORLD;
    const long long stride = (long long)gridDim.x * AR_THREADS;
    const long long t0 = (long long)blockIdx.x * AR_THREADS + threadIdx.x;
    uint4* stage = reinterpret_cast<uint4*>(pp.buf[rank]);
    // A: (the previous collective on this pad ended with a barrier, so no peer still reads our staging buffer)
    for (int s = 0; s < WORLD; ++s) {
        const long long b = min(n8, per * s), e = min(n8, b + per);
        for (long long i = b + t0; i < e; i += stride) {
            const float4 lo = grad[2 * i], hi = grad[2 * i + 1];
            uint4 o;
            o.x = pack_bf16x2(lo.x * scale, lo.y * scale); o.y = pack_bf16x2(lo.z * scale, lo.w * scale);
            o.z = pack_bf16x2(hi.x * scale, hi.y * scale); o.w = pack_bf16x2(hi.z * scale, hi.w * scale);
            stage[i] = o;
        }
    }
    rank_sync(pp, rank, WORLD, seq);
    // B
    {
        const long long b = min(n8, per * rank), e = min(n8, b + per);
        if (MC) {
            long long i = b + t0;
            for (; i + 3 * stride < e; i += 4 * stride) {
                uint4 v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) v[u] = mm_ld_reduce_bf16x8(mc_stage + i + u * stride);
#pragma unroll
                for (int u = 0; u < 4; ++u) mm_st_b32x4(mc_stage + i + u * stride, v[u]);
            }
            for (; i < e; i += stride) mm_st_b32x4(mc_stage + i, mm_ld_reduce_bf16x8(mc_stage + i));
        } else {
            for (long long i = b + t0; i < e; i += stride) {
                uint4 v[WORLD];
#pragma unroll
                for (int r = 0; r < WORLD; ++r) v[r] = ld_u4(reinterpret_cast<const uint4*>(pp.buf[(rank + r) % WORLD]) + i);
                float acc[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) acc[k] = 0.f;
#pragma unroll
                for (int r = 0; r < WORLD; ++r) {
                    const uint32_t w[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
#pragma unroll
                    for (int k = 0; k < 4; ++k) { const float2 f = unpack_bf16x2(w[k]); acc[2 * k] += f.x; acc[2 * k + 1] += f.y; }
                }
                const uint4 o = make_uint4(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]),
                                           pack_bf16x2(acc[4], acc[5]), pack_bf16x2(acc[6], acc[7]));
#pragma unroll
                for (int r = 0; r < WORLD; ++r) st_u4(reinterpret_cast<uint4*>(pp.buf[(rank + r) % WORLD]) + i, o);
            }
        }
    }
    rank_sync(pp, rank, WORLD, seq + 1);
    // C: system-scope loads - the lines were written by peers (or the switch) while this SM may hold stale copies
    for (int s = 0; s < WORLD; ++s) {
        const long long b = min(n8, per * s), e = min(n8, b + per);
        for (long long i = b + t0; i < e; i += stride) {
            const uint4 v = ld_u4(stage + i);
            const float2 a = unpack_bf16x2(v.x), bq = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z), d = unpack_bf16x2(v.w);
            grad[2 * i] = make_float4(a.x, a.y, bq.x, bq.y);
            grad[2 * i + 1] = make_float4(c.x, c.y, d.x, d.y);
        }
    }
    // The next collective on this pad starts with phase A writing the staging buffer that peers read in B: it may not
    // begin before every peer is past B.  They are: barrier seq + 1 above is only passed once all ranks finished B.
    seq_end(pp, rank, seq + 1);
}

// ---- broadcast: every non-root rank pulls the root's buffer ----------------------------------------------
__global__ void __launch_bounds__(AR_THREADS) broadcast_pull_kernel(PeerPtrs pp, int rank, int world, int root,
                                                                    long long n16) {
    const uint32_t seq = seq_begin(pp, rank);
    rank_barrier(pp, rank, world, seq);                 // root's data is final
    if (rank != root) {
        const uint4* src = reinterpret_cast<const uint4*>(pp.buf[root]);
        uint4* dst = reinterpret_cast<uint4*>(pp.buf[rank]);
        const long long stride = (long long)gridDim.x * AR_THREADS;
        long long i = (long long)blockIdx.x * AR_THREADS + threadIdx.x;
        for (; i + 3 * stride < n16; i += 4 * stride) {
            uint4 v0 = ld_u4(src + i), v1 = ld_u4(src + i + stride), v2 = ld_u4(src + i + 2 * stride), v3 = ld_u4(src + i + 3 * stride);
            dst[i] = v0; dst[i + stride] = v1; dst[i + 2 * stride] = v2; dst[i + 3 * stride] = v3;
        }
        for (; i < n16; i += stride) dst[i] = ld_u4(src + i);
    }
    rank_barrier(pp, rank, world, seq + 1);             // root may overwrite its buffer again
    seq_end(pp, rank, seq + 1);
}

// ---- broadcast through the switch: the root stores every vector ONCE with multimem.st, the NVSwitch writes all N copies
// (the root's port carries the message once instead of N-1 times as with the pull kernel above) ---------------------
__global__ void __launch_bounds__(AR_THREADS) broadcast_mc_kernel(PeerPtrs pp, uint4* mc, int rank, int world, int root,
                                                                  long long n16) {
    const uint32_t seq = seq_begin(pp, rank);
    rank_sync(pp, rank, world, seq);                    // nobody still reads the previous contents of its copy
    if (rank == root) {
        const uint4* src = reinterpret_cast<const uint4*>(pp.buf[rank]);
        const long long stride = (long long)gridDim.x * AR_THREADS;
        long long i = (long long)blockIdx.x * AR_THREADS + threadIdx.x;
        for (; i + 3 * stride < n16; i += 4 * stride) {
            const uint4 v0 = src[i], v1 = src[i + stride], v2 = src[i + 2 * stride], v3 = src[i + 3 * stride];
            mm_st_b32x4(mc + i, v0); mm_st_b32x4(mc + i + stride, v1);
            mm_st_b32x4(mc + i + 2 * stride, v2); mm_st_b32x4(mc + i + 3 * stride, v3);
        }
        for (; i < n16; i += stride) mm_st_b32x4(mc + i, src[i]);
    }
    rank_sync(pp, rank, world, seq + 1);                // the root's stores have landed everywhere
    seq_end(pp, rank, seq + 1);
}

// ---- sharded collectives on a symmetric buffer of n16 16-byte vectors ------------------------------------
// The buffer is cut into `world` shards of per = ceil(n16 / world) vectors; shard s belongs to rank s.
//   SC_ALL_GATHER     every rank pulls shard s from rank s                   (each shard valid at its owner before)
//   SC_GATHER         only `root` pulls
//   SC_SCATTER        rank r pulls shard r out of root's buffer
//   SC_REDUCE_SCATTER rank r: shard r = scale * sum over ranks of shard r    (fp32; other shards left as they were)
//   SC_REDUCE         root: whole buffer = scale * sum over ranks            (fp32)
//   SC_ALLREDUCE_1    one-shot all-reduce for small messages: every rank loads the whole buffer of every rank,
//                     all ranks meet (nobody may overwrite what a peer has yet to read), then each stores its
//                     own copy.  One NVLink round trip instead of the two of the two-shot kernel; needs the whole
//                     message in the grid's registers (n16 <= gridDim.x * AR_THREADS * 4).
// Every op starts with a barrier (inputs are final everywhere) and ends with one (peers are done reading).
enum { SC_ALL_GATHER = 0, SC_GATHER = 1, SC_SCATTER = 2, SC_REDUCE_SCATTER = 3, SC_REDUCE = 4, SC_ALLREDUCE_1 = 5 };

__device__ __forceinline__ void pull_range(uint4* dst, const uint4* src, long long begin, long long end) {
    const long long stride = (long long)gridDim.x * AR_THREADS;
    long long i = begin + (long long)blockIdx.x * AR_THREADS + threadIdx.x;
    for (; i + 3 * stride < end; i += 4 * stride) {
        const uint4 v0 = ld_u4(src + i), v1 = ld_u4(src + i + stride), v2 = ld_u4(src + i + 2 * stride),
                    v3 = ld_u4(src + i + 3 * stride);
        dst[i] = v0; dst[i + stride] = v1; dst[i + 2 * stride] = v2; dst[i + 3 * stride] = v3;
    }
    for (; i < end; i += stride) dst[i] = ld_u4(src + i);
}

__device__ __forceinline__ void reduce_range(const PeerPtrs& pp, int rank, int world, long long begin, long long end,
                                             float scale) {
    const long long stride = (long long)gridDim.x * AR_THREADS;
    float4* dst = reinterpret_cast<float4*>(pp.buf[rank]);
    for (long long i = begin + (long long)blockIdx.x * AR_THREADS + threadIdx.x; i < end; i += stride) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int r = 0; r < world; ++r) {
            const float4 v = ld_f4(reinterpret_cast<const float4*>(pp.buf[(rank + r) % world]) + i);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
        dst[i] = acc;
    }
}

__global__ void __launch_bounds__(AR_THREADS) shard_collective_kernel(PeerPtrs pp, int rank, int world, int root, int op,
                                                                      long long n16, float scale) {
    const uint32_t seq = seq_begin(pp, rank);
    rank_barrier(pp, rank, world, seq);
    const long long per = (n16 + world - 1) / world;
    uint4* mine = reinterpret_cast<uint4*>(pp.buf[rank]);
    if (op == SC_ALL_GATHER || (op == SC_GATHER && rank == root)) {
        for (int k = 1; k < world; ++k) {
            const int s = (rank + k) % world;
            pull_range(mine, reinterpret_cast<const uint4*>(pp.buf[s]), min(n16, per * s), min(n16, per * (s + 1)));
        }
    } else if (op == SC_SCATTER) {
        if (rank != root)
            pull_range(mine, reinterpret_cast<const uint4*>(pp.buf[root]), min(n16, per * rank), min(n16, per * (rank + 1)));
    } else if (op == SC_REDUCE_SCATTER) {
        reduce_range(pp, rank, world, min(n16, per * rank), min(n16, per * (rank + 1)), scale);
    } else if (op == SC_REDUCE) {
        // the root's own buffer is both an input and the output: element i is read and written by the same thread
        if (rank == root) reduce_range(pp, rank, world, 0, n16, scale);
    } else if (op == SC_ALLREDUCE_1) {
        const long long stride = (long long)gridDim.x * AR_THREADS;
        const long long i0 = (long long)blockIdx.x * AR_THREADS + threadIdx.x;
        float4 acc[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            const long long i = i0 + u * stride;
            if (i < n16) {
                for (int r = 0; r < world; ++r) {
                    const float4 v = ld_f4(reinterpret_cast<const float4*>(pp.buf[(rank + r) % world]) + i);
                    acc[u].x += v.x; acc[u].y += v.y; acc[u].z += v.z; acc[u].w += v.w;
                }
            }
        }
        rank_barrier(pp, rank, world, seq + 1);         // every rank holds its sums: the inputs may be overwritten
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const long long i = i0 + u * stride;
            if (i < n16) {
                acc[u].x *= scale; acc[u].y *= scale; acc[u].z *= scale; acc[u].w *= scale;
                reinterpret_cast<float4*>(mine)[i] = acc[u];
            }
        }
        seq_end(pp, rank, seq + 1);                     // outputs are local: no closing barrier needed
        return;
    }
    rank_barrier(pp, rank, world, seq + 1);
    seq_end(pp, rank, seq + 1);
}

__global__ void __launch_bounds__(AR_THREADS) barrier_only_kernel(PeerPtrs pp, int rank, int world) {
    const uint32_t seq = seq_begin(pp, rank);
    rank_barrier(pp, rank, world, seq);
    seq_end(pp, rank, seq);
}

// ---- single-process helpers (DataParallel): device-issued copies / reductions over peer pointers ----------
__global__ void __launch_bounds__(256) peer_copy_kernel(uint4* __restrict__ dst, const uint4* __restrict__ src,
                                                        long long n16, const uint8_t* src_tail, uint8_t* dst_tail,
                                                        int tail) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n16; i += 4 * stride) {
        uint4 v0 = src[i], v1 = src[i + stride], v2 = src[i + 2 * stride], v3 = src[i + 3 * stride];
        dst[i] = v0; dst[i + stride] = v1; dst[i + 2 * stride] = v2; dst[i + 3 * stride] = v3;
    }
    for (; i < n16; i += stride) dst[i] = src[i];
    if (blockIdx.x == 0 && (int)threadIdx.x < tail) dst_tail[threadIdx.x] = src_tail[threadIdx.x];
}

// Copy through shared memory with the bulk-copy engine on both sides: cp.async.bulk global -> smem (mbarrier
// completion), then cp.async.bulk smem -> global (bulk-group completion).  One elected thread drives a CTA;
// used to move data to a peer without tying up load/store slots (and to measure NVLink write bandwidth).
constexpr int BULK_CHUNK = 16384;
constexpr int BULK_STAGES = 4;
__global__ void __launch_bounds__(32) bulk_copy_kernel(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src,
                                                       long long nchunks) {
    extern __shared__ __align__(128) uint8_t bsm[];
    __shared__ uint64_t bar[BULK_STAGES];
    if (threadIdx.x == 0) {
        for (int s = 0; s < BULK_STAGES; ++s) mbar_init(&bar[s], 1);
        mbar_fence_init();
        uint32_t phase = 0;
        int stage = 0;
        long long issued = 0;
        for (long long c = blockIdx.x; c < nchunks; c += gridDim.x, ++issued) {
            // the store that last used this stage must have finished READING smem before we overwrite it
            if (issued >= BULK_STAGES) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(BULK_STAGES - 1) : "memory");
            const uint32_t sm = smem_u32(bsm + stage * BULK_CHUNK);
            mbar_arrive_expect_tx(&bar[stage], BULK_CHUNK);
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(sm), "l"(src + c * BULK_CHUNK), "n"(BULK_CHUNK), "r"(smem_u32(&bar[stage])) : "memory");
            mbar_wait(&bar[stage], phase);
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                         ::"l"(dst + c * BULK_CHUNK), "r"(sm), "n"(BULK_CHUNK) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            if (++stage == BULK_STAGES) { stage = 0; phase ^= 1; }
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

struct SrcList {
    const float* p[MAX_WORLD];
};
// dst[i] = scale * (accumulate * dst[i] + sum_k src_k[i]).  16-byte accesses, every source's load of a vector issued
// before the first add (the sources are peer GPUs: what limits the kernel is bytes in flight over NVLink), two
// vectors per thread and iteration.  Scalar tail for n % 4 and for unaligned pointers.
template <int NSRC>
__global__ void __launch_bounds__(256) reduce_sum_kernel(float* __restrict__ dst, SrcList srcs, long long n,
                                                         float scale, int accumulate, int vec_ok) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long n4 = vec_ok ? n / 4 : 0;
    float4* d4 = reinterpret_cast<float4*>(dst);
    long long i = tid;
    for (; i + stride < n4; i += 2 * stride) {
        float4 v[NSRC > 0 ? NSRC : 1][2];
#pragma unroll
        for (int k = 0; k < NSRC; ++k) {
            v[k][0] = ld_f4_weak(reinterpret_cast<const float4*>(srcs.p[k]) + i);
            v[k][1] = ld_f4_weak(reinterpret_cast<const float4*>(srcs.p[k]) + i + stride);
        }
        float4 a0 = accumulate ? d4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        float4 a1 = accumulate ? d4[i + stride] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < NSRC; ++k) {
            a0.x += v[k][0].x; a0.y += v[k][0].y; a0.z += v[k][0].z; a0.w += v[k][0].w;
            a1.x += v[k][1].x; a1.y += v[k][1].y; a1.z += v[k][1].z; a1.w += v[k][1].w;
        }
        d4[i] = make_float4(a0.x * scale, a0.y * scale, a0.z * scale, a0.w * scale);
        d4[i + stride] = make_float4(a1.x * scale, a1.y * scale, a1.z * scale, a1.w * scale);
    }
    for (; i < n4; i += stride) {
        float4 a = accumulate ? d4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < NSRC; ++k) {
            const float4 v = ld_f4_weak(reinterpret_cast<const float4*>(srcs.p[k]) + i);
            a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
        }
        d4[i] = make_float4(a.x * scale, a.y * scale, a.z * scale, a.w * scale);
    }
    for (long long j = n4 * 4 + tid; j < n; j += stride) {
        float acc = accumulate ? dst[j] : 0.f;
#pragma unroll
        for (int k = 0; k < NSRC; ++k) acc += srcs.p[k][j];
        dst[j] = acc * scale;
    }
}

int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}

int pick_blocks(long long work_items, int requested) {
    if (requested > 0) return requested > MAX_BLOCKS ? MAX_BLOCKS : requested;
    long long b = (work_items + AR_THREADS * 4 - 1) / (AR_THREADS * 4);
    if (b < 1) b = 1;
    // enough loads in flight to fill NVLink (measured: 32 CTAs 455 GB/s, 64 CTAs 700, 128 CTAs 715 on a pull)
    static int cap = -1;
    if (cap < 0) cap = env_int("PTDB200_AR_MAX_BLOCKS", 128);
    if (b > cap) b = cap;
    return (int)b;
}

// Through the switch the SMs issue one load and one store per vector (not N of each): fewer CTAs saturate the
// port, and every CTA not taken is an SM left to the backward kernels the all-reduce overlaps with.
int pick_blocks_mc(long long work_items, int requested) {
    if (requested > 0) return requested > MAX_BLOCKS ? MAX_BLOCKS : requested;
    long long b = (work_items + AR_THREADS * 4 - 1) / (AR_THREADS * 4);
    if (b < 1) b = 1;
    static int cap = -1;
    if (cap < 0) cap = env_int("PTDB200_MC_MAX_BLOCKS", 16);     // measured at N = 4 and 8: 8-16 CTAs already give the full rate (profiles/allreduce_sweep_n8.json)
    if (b > cap) b = cap;
    return (int)b;
}

unsigned long long barrier_timeout_ns() {
    static unsigned long long v = 0;
    if (v == 0) {
        // Long enough for a peer that is still capturing a graph, loading data or writing a checkpoint; a dead peer
        // is reported after this time instead of hanging the GPU forever.
        const char* e = getenv("PTDB200_P2P_TIMEOUT_S");
        double sec = e ? atof(e) : 600.0;
        if (sec < 0.001) sec = 0.001;
        v = (unsigned long long)(sec * 1e9);
    }
    return v;
}

int fill_peers(PeerPtrs& pp, const void* const* bufs, const void* const* pads, int world, const void* mc_pad = nullptr) {
    if (world < 1 || world > MAX_WORLD) {
        ptd_set_error("p2p: world size %d outside [1,%d]", world, MAX_WORLD);
        return PTD_ERR_ARG;
    }
    for (int r = 0; r < MAX_WORLD; ++r) {
        pp.buf[r] = (r < world && bufs != nullptr) ? const_cast<void*>(bufs[r]) : nullptr;
        pp.pad[r] = r < world ? (uint32_t*)const_cast<void*>(pads[r]) : nullptr;
    }
    pp.mc_pad = (uint32_t*)const_cast<void*>(mc_pad);
    pp.timeout_ns = barrier_timeout_ns();
    return PTD_OK;
}

}  // namespace

// ---- memory / IPC -------------------------------------------------------------------------------------------
extern "C" int ptd_p2p_malloc(long long bytes, void** out) {
    PTD_CUDA_CHECK(cudaMalloc(out, (size_t)bytes));
    PTD_CUDA_CHECK(cudaMemset(*out, 0, (size_t)bytes));
    return PTD_OK;
}
extern "C" int ptd_p2p_free(void* p) {
    PTD_CUDA_CHECK(cudaFree(p));
    return PTD_OK;
}
extern "C" int ptd_p2p_ipc_handle_size() { return (int)sizeof(cudaIpcMemHandle_t); }
extern "C" int ptd_p2p_ipc_get(void* p, void* handle_out) {
    cudaIpcMemHandle_t h;
    PTD_CUDA_CHECK(cudaIpcGetMemHandle(&h, p));
    memcpy(handle_out, &h, sizeof(h));
    return PTD_OK;
}
extern "C" int ptd_p2p_ipc_open(const void* handle, void** out) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    PTD_CUDA_CHECK(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
    return PTD_OK;
}
extern "C" int ptd_p2p_ipc_close(void* p) {
    PTD_CUDA_CHECK(cudaIpcCloseMemHandle(p));
    return PTD_OK;
}
// Single process, several devices: make `peer` memory addressable from kernels running on `dev`.
extern "C" int ptd_p2p_enable_peer_access(int dev, int peer) {
    int can = 0;
    PTD_CUDA_CHECK(cudaDeviceCanAccessPeer(&can, dev, peer));
    PTD_REQUIRE(can, "device %d cannot access device %d", dev, peer);
    int cur = 0;
    PTD_CUDA_CHECK(cudaGetDevice(&cur));
    PTD_CUDA_CHECK(cudaSetDevice(dev));
    cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
    cudaSetDevice(cur);
    PTD_CUDA_CHECK(e);
    return PTD_OK;
}
extern "C" int ptd_p2p_pad_bytes() { return PAD_WORDS * (int)sizeof(uint32_t); }
extern "C" int ptd_p2p_max_blocks() { return MAX_BLOCKS; }

// ---- collectives --------------------------------------------------------------------------------------------
// bufs / pads: HOST arrays of `world` device pointers (this rank's mapping of every rank's buffer / signal pad).
// Kernels sharing a signal pad must be stream-ordered with respect to each other.
static int launch_two_shot(const void* const* bufs, const void* const* pads, int rank, int world, long long numel,
                           int is_bf16, float scale, int blocks, void* stream, const SgdEpilogue& sg) {
    PeerPtrs pp;
    int rc = fill_peers(pp, bufs, pads, world);
    if (rc) return rc;
    const int per_vec = is_bf16 ? 8 : 4;
    PTD_REQUIRE(numel % per_vec == 0, "allreduce: numel (%lld) must be a multiple of %d", numel, per_vec);
    const long long nvec = numel / per_vec;
    const int grid = pick_blocks((nvec + world - 1) / world, blocks);
    cudaStream_t st = (cudaStream_t)stream;
    static int weak = -1;
    if (weak < 0) { const char* e = getenv("PTDB200_AR_WEAK"); weak = e ? atoi(e) : 1; }
#define PTD_AR(W)                                                                                   \
    case W:                                                                                         \
        if (is_bf16) allreduce_bf16_kernel<W><<<grid, AR_THREADS, 0, st>>>(pp, rank, nvec, scale); \
        else if (weak) allreduce_f32_kernel<W, true><<<grid, AR_THREADS, 0, st>>>(pp, rank, nvec, scale, sg); \
        else allreduce_f32_kernel<W, false><<<grid, AR_THREADS, 0, st>>>(pp, rank, nvec, scale, sg);      \
        break;
    switch (world) {
        PTD_AR(1) PTD_AR(2) PTD_AR(3) PTD_AR(4) PTD_AR(5) PTD_AR(6) PTD_AR(7) PTD_AR(8)
    }
#undef PTD_AR
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

extern "C" int ptd_allreduce_inplace(const void* const* bufs, const void* const* pads, int rank, int world,
                                     long long numel, int is_bf16, float scale, int blocks, void* stream) {
    SgdEpilogue none;
    memset(&none, 0, sizeof(none));
    return launch_two_shot(bufs, pads, rank, world, numel, is_bf16, scale, blocks, stream, none);
}

extern "C" int ptd_broadcast_inplace(const void* const* bufs, const void* const* pads, int rank, int world, int root,
                                     long long nbytes, int blocks, void* stream) {
    PeerPtrs pp;
    int rc = fill_peers(pp, bufs, pads, world);
    if (rc) return rc;
    PTD_REQUIRE(nbytes % 16 == 0, "broadcast: nbytes (%lld) must be a multiple of 16", nbytes);
    const int grid = pick_blocks(nbytes / 16, blocks);
    broadcast_pull_kernel<<<grid, AR_THREADS, 0, (cudaStream_t)stream>>>(pp, rank, world, root, nbytes / 16);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

// Broadcast through the switch; bufs = unicast mappings (only the root's own is read), mc_buf = multicast mapping.
extern "C" int ptd_broadcast_mc(const void* const* bufs, void* mc_buf, const void* const* pads, const void* mc_pad,
                                int rank, int world, int root, long long nbytes, int blocks, void* stream) {
    PeerPtrs pp;
    int rc = fill_peers(pp, bufs, pads, world, mc_pad);
    if (rc) return rc;
    PTD_REQUIRE(nbytes % 16 == 0 && mc_buf != nullptr, "broadcast_mc: nbytes (%lld) must be a multiple of 16, mc_buf set", nbytes);
    PTD_REQUIRE(root >= 0 && root < world, "broadcast_mc: root %d outside the group", root);
    const int grid = pick_blocks_mc(nbytes / 16, blocks);
    broadcast_mc_kernel<<<grid, AR_THREADS, 0, (cudaStream_t)stream>>>(pp, (uint4*)mc_buf, rank, world, root, nbytes / 16);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

// op: 0 all-gather, 1 gather (to root), 2 scatter (from root), 3 reduce-scatter, 4 reduce (to root), 5 one-shot
// all-reduce (ops 3-5: fp32 data, `scale` applied to the sum).  nbytes a multiple of 16; shard s of the buffer
// (ceil(nbytes/16/world) vectors) belongs to rank s.  See shard_collective_kernel.
extern "C" int ptd_shard_collective(const void* const* bufs, const void* const* pads, int rank, int world, int root,
                                    int op, long long nbytes, float scale, int blocks, void* stream) {
    PeerPtrs pp;
    int rc = fill_peers(pp, bufs, pads, world);
    if (rc) return rc;
    PTD_REQUIRE(nbytes % 16 == 0, "shard_collective: nbytes (%lld) must be a multiple of 16", nbytes);
    PTD_REQUIRE(op >= SC_ALL_GATHER && op <= SC_ALLREDUCE_1, "shard_collective: unknown op %d", op);
    PTD_REQUIRE(root >= 0 && root < world, "shard_collective: root %d outside the group", root);
    const long long n16 = nbytes / 16;
    int grid = pick_blocks(op == SC_REDUCE || op == SC_ALLREDUCE_1 ? n16 : (n16 + world - 1) / world, blocks);
    if (op == SC_ALLREDUCE_1) {
        const long long need = (n16 + AR_THREADS * 4 - 1) / (AR_THREADS * 4);
        PTD_REQUIRE(need <= MAX_BLOCKS, "one-shot all-reduce: %lld bytes exceed the %d-byte limit (use the two-shot kernel)",
                    nbytes, MAX_BLOCKS * AR_THREADS * 4 * 16);
        if (grid < need) grid = (int)need;
    }
    shard_collective_kernel<<<grid, AR_THREADS, 0, (cudaStream_t)stream>>>(pp, rank, world, root, op, n16, scale);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

extern "C" int ptd_p2p_barrier(const void* const* pads, int rank, int world, void* stream) {
    PeerPtrs pp;
    int rc = fill_peers(pp, nullptr, pads, world);
    if (rc) return rc;
    barrier_only_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pp, rank, world);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

extern "C" int ptd_peer_copy(void* dst, const void* src, long long nbytes, void* stream) {
    PTD_REQUIRE(((uintptr_t)dst & 15) == 0 && ((uintptr_t)src & 15) == 0, "peer_copy: pointers must be 16-byte aligned");
    const long long n16 = nbytes / 16;
    const int tail = (int)(nbytes - n16 * 16);
    long long blocks = (n16 + 256 * 4 - 1) / (256 * 4);
    if (blocks < 1) blocks = 1;
    if (blocks > 4LL * ptd_num_sms()) blocks = 4LL * ptd_num_sms();
    peer_copy_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>((uint4*)dst, (const uint4*)src, n16,
                                                                    (const uint8_t*)src + n16 * 16,
                                                                    (uint8_t*)dst + n16 * 16, tail);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

// nbytes must be a multiple of 16 KiB; runs on the current device, either pointer may be peer memory.
extern "C" int ptd_peer_copy_bulk(void* dst, const void* src, long long nbytes, int blocks, void* stream) {
    PTD_REQUIRE(nbytes % BULK_CHUNK == 0 && ((uintptr_t)dst & 127) == 0 && ((uintptr_t)src & 127) == 0,
                "peer_copy_bulk: size must be a multiple of %d and pointers 128-byte aligned", BULK_CHUNK);
    static bool configured[64] = {false};
    int dev = 0;
    PTD_CUDA_CHECK(cudaGetDevice(&dev));
    const int smem = BULK_CHUNK * BULK_STAGES;
    if (!configured[dev & 63]) {
        PTD_CUDA_CHECK(cudaFuncSetAttribute(bulk_copy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        configured[dev & 63] = true;
    }
    if (blocks <= 0) blocks = 2 * ptd_num_sms();
    bulk_copy_kernel<<<blocks, 32, smem, (cudaStream_t)stream>>>((uint8_t*)dst, (const uint8_t*)src, nbytes / BULK_CHUNK);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

extern "C" int ptd_reduce_sum_f32(float* dst, const void* const* srcs, int n_src, long long numel, float scale,
                                  int accumulate, void* stream) {
    PTD_REQUIRE(n_src >= 0 && n_src <= MAX_WORLD, "reduce_sum: %d sources (max %d)", n_src, MAX_WORLD);
    SrcList sl;
    int vec_ok = ((uintptr_t)dst & 15) == 0;
    for (int k = 0; k < MAX_WORLD; ++k) {
        sl.p[k] = k < n_src ? (const float*)srcs[k] : nullptr;
        if (k < n_src && ((uintptr_t)srcs[k] & 15) != 0) vec_ok = 0;
    }
    long long blocks = (numel / 4 + 2 * 256 - 1) / (2 * 256);
    if (blocks < 1) blocks = 1;
    if (blocks > 2LL * ptd_num_sms()) blocks = 2LL * ptd_num_sms();
    cudaStream_t st = (cudaStream_t)stream;
#define PTD_RS(K) case K: reduce_sum_kernel<K><<<(int)blocks, 256, 0, st>>>(dst, sl, numel, scale, accumulate, vec_ok); break;
    switch (n_src) {
        PTD_RS(0) PTD_RS(1) PTD_RS(2) PTD_RS(3) PTD_RS(4) PTD_RS(5) PTD_RS(6) PTD_RS(7) PTD_RS(8)
    }
#undef PTD_RS
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

// ---- NVLS / bf16-wire entry points --------------------------------------------------------------------------
// In-place all-reduce through the switch.  mc_buf: multicast mapping of the symmetric buffer (csrc/vmm.cu); pads:
// unicast pads; mc_pad: multicast mapping of the pads (NULL = flag barrier).  scale as in ptd_allreduce_inplace.
static int launch_mc(void* mc_buf, const void* const* bufs, const void* const* pads, const void* mc_pad, int rank,
                     int world, long long numel, int is_bf16, float scale, int blocks, void* stream,
                     const SgdEpilogue& sg) {
    PeerPtrs pp;
    int rc = fill_peers(pp, bufs, pads, world, mc_pad);
    if (rc) return rc;
    PTD_REQUIRE(mc_buf != nullptr && ((uintptr_t)mc_buf & 15) == 0, "allreduce_mc: needs a 16-byte aligned multicast address");
    const int per_vec = is_bf16 ? 8 : 4;
    PTD_REQUIRE(numel % per_vec == 0, "allreduce_mc: numel (%lld) must be a multiple of %d", numel, per_vec);
    const long long nvec = numel / per_vec;
    const int grid = pick_blocks_mc((nvec + world - 1) / world, blocks);
    cudaStream_t st = (cudaStream_t)stream;
    if (is_bf16) allreduce_mc_kernel<true><<<grid, AR_THREADS, 0, st>>>(pp, (uint4*)mc_buf, rank, world, nvec, scale, sg);
    else allreduce_mc_kernel<false><<<grid, AR_THREADS, 0, st>>>(pp, (uint4*)mc_buf, rank, world, nvec, scale, sg);
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}

extern "C" int ptd_allreduce_mc(void* mc_buf, const void* const* pads, const void* mc_pad, int rank, int world,
                                long long numel, int is_bf16, float scale, int blocks, void* stream) {
    SgdEpilogue none;
    memset(&none, 0, sizeof(none));
    return launch_mc(mc_buf, nullptr, pads, mc_pad, rank, world, numel, is_bf16, scale, blocks, stream, none);
}

// fp32 all-reduce (through the switch when mc_buf is given, two-shot otherwise) with the SGD step of the bucket's
// parameters fused behind it - see SgdEpilogue.  bufs[rank] must be the local mapping of the bucket in both cases.
extern "C" int ptd_allreduce_fused_sgd(const void* const* bufs, void* mc_buf, const void* const* pads, const void* mc_pad,
                                       int rank, int world, long long numel, float scale, int blocks,
                                       const PtdSgdEpilogue* sgd, void* stream) {
    PTD_REQUIRE(sgd != nullptr && sgd->param != nullptr && sgd->lr_dev != nullptr,
                "allreduce_fused_sgd: param and lr_dev are required");
    PTD_REQUIRE(bufs != nullptr && bufs[rank] != nullptr, "allreduce_fused_sgd: the local mapping of the bucket is required");
    PTD_REQUIRE((((uintptr_t)sgd->param | (uintptr_t)sgd->momentum_buf | (uintptr_t)sgd->shadow_bf16) & 15) == 0,
                "allreduce_fused_sgd: param / momentum / shadow arrays must be 16-byte aligned");
    SgdEpilogue sg;
    sg.param = sgd->param; sg.momentum_buf = sgd->momentum_buf; sg.shadow = (__nv_bfloat162*)sgd->shadow_bf16;
    sg.lr_dev = sgd->lr_dev; sg.momentum = sgd->momentum; sg.weight_decay = sgd->weight_decay;
    if (mc_buf != nullptr) return launch_mc(mc_buf, bufs, pads, mc_pad, rank, world, numel, 0, scale, blocks, stream, sg);
    return launch_two_shot(bufs, pads, rank, world, numel, 0, scale, blocks, stream, sg);
}

// All-reduce of the LOCAL fp32 buffer `grad` (numel a multiple of 8) with bf16 on the wire: `stages` are the unicast
// mappings of a symmetric bf16 staging buffer of at least numel elements, `mc_stage` its multicast mapping (NULL =
// peers' staging buffers are loaded one by one).  grad <- scale * sum over ranks, rounded to bf16 once per phase.
extern "C" int ptd_allreduce_wire_bf16(float* grad, const void* const* stages, void* mc_stage, const void* const* pads,
                                       const void* mc_pad, int rank, int world, long long numel, float scale, int blocks,
                                       void* stream) {
    PeerPtrs pp;
    int rc = fill_peers(pp, stages, pads, world, mc_pad);
    if (rc) return rc;
    PTD_REQUIRE(numel % 8 == 0 && ((uintptr_t)grad & 15) == 0, "allreduce_wire_bf16: numel (%lld) %% 8 and a 16-byte aligned buffer", numel);
    const long long n8 = numel / 8;
    // phases A and C are local HBM passes over the whole buffer: they want a full-size grid whatever phase B uses
    const int grid = pick_blocks(n8, blocks);
    cudaStream_t st = (cudaStream_t)stream;
#define PTD_WIRE(W)                                                                                                      \
    case W:                                                                                                              \
        if (mc_stage) allreduce_wire_bf16_kernel<W, true><<<grid, AR_THREADS, 0, st>>>(pp, (uint4*)mc_stage, (float4*)grad, rank, n8, scale); \
        else allreduce_wire_bf16_kernel<W, false><<<grid, AR_THREADS, 0, st>>>(pp, nullptr, (float4*)grad, rank, n8, scale);         \
        break;
    switch (world) {
        PTD_WIRE(1) PTD_WIRE(2) PTD_WIRE(3) PTD_WIRE(4) PTD_WIRE(5) PTD_WIRE(6) PTD_WIRE(7) PTD_WIRE(8)
    }
#undef PTD_WIRE
    PTD_CUDA_CHECK(cudaGetLastError());
    return PTD_OK;
}
