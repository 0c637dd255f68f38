// Host-side gradient bucket planner and reducer: ready tracking AND, for buckets bound to a communication recipe, the
// whole launch sequence of a finished bucket (event on the compute stream, wait on the communication stream, the
// all-reduce kernel - optionally with the fused SGD step -, completion event) in ONE native call per parameter.
//
// Serves the role of torch's C++ Reducer + _compute_bucket_assignment_by_size that the reference
// reaches through DDP(model, device_ids=[gpu_id]) (R/02 DDP基本概念/ddp_gpus.py:35):
//   * parameters are assigned to buckets in the order given (the wrapper passes reverse
//     registration order, then the observed ready order); the first bucket closes at `first_cap`
//     bytes, later ones at `cap` bytes (1 MiB / 25 MiB by default) - a bucket closes as soon as its
//     size reaches the limit;
//   * during backward each parameter is marked ready once; when the last parameter of a bucket
//     arrives the bucket index is returned so the caller can launch that bucket's all-reduce;
//   * the order in which parameters became ready is recorded for the one-time bucket rebuild;
//   * ptd_reducer_bind_bucket attaches what a bucket's reduction needs (mappings, pads, scale, algorithm);
//     ptd_reducer_mark_ready_launch then issues it the moment the bucket completes and ptd_reducer_finalize makes the
//     compute stream wait for everything issued in this backward - the Python wrapper only counts.
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "common.cuh"
#include "ptdb200.h"

namespace {

struct BucketComm {
    bool bound = false;
    bool launched = false;
    PtdBucketComm c;
    cudaEvent_t ready = nullptr, done = nullptr;
};

struct Reducer {
    std::vector<BucketComm> comm;     // per bucket
    std::vector<int> bucket_of;
    std::vector<int> bucket_size;     // parameters per bucket
    std::vector<int> pending;         // parameters not yet ready, per bucket
    std::vector<uint8_t> ready;
    std::vector<int> ready_order;
    int buckets_done = 0;
};

std::mutex g_mu;
std::unordered_map<int, Reducer> g_reducers;
int g_next_id = 1;

}  // namespace

extern "C" int ptd_plan_buckets(const long long* sizes_bytes, int n, long long first_cap, long long cap,
                                int* bucket_of) {
    if (n < 0 || (n > 0 && (!sizes_bytes || !bucket_of))) {
        ptd_set_error("plan_buckets: bad arguments");
        return PTD_ERR_ARG;
    }
    long long limit = first_cap > 0 ? first_cap : cap;
    long long cur = 0;
    int b = 0;
    bool open = false;
    for (int i = 0; i < n; ++i) {
        bucket_of[i] = b;
        cur += sizes_bytes[i];
        open = true;
        if (cur >= limit) {
            ++b;
            cur = 0;
            open = false;
            limit = cap;
        }
    }
    return open ? b + 1 : b;
}

extern "C" int ptd_reducer_create(const int* bucket_of, int n_params, int n_buckets) {
    if (n_params <= 0 || n_buckets <= 0 || !bucket_of) {
        ptd_set_error("reducer_create: bad arguments");
        return PTD_ERR_ARG;
    }
    Reducer r;
    r.bucket_of.assign(bucket_of, bucket_of + n_params);
    r.bucket_size.assign(n_buckets, 0);
    for (int i = 0; i < n_params; ++i) {
        if (bucket_of[i] < 0 || bucket_of[i] >= n_buckets) {
            ptd_set_error("reducer_create: parameter %d has bucket %d outside [0,%d)", i, bucket_of[i], n_buckets);
            return PTD_ERR_ARG;
        }
        r.bucket_size[bucket_of[i]]++;
    }
    r.pending = r.bucket_size;
    r.comm.resize(n_buckets);
    r.ready.assign(n_params, 0);
    r.ready_order.reserve(n_params);
    std::lock_guard<std::mutex> lk(g_mu);
    int id = g_next_id++;
    g_reducers.emplace(id, std::move(r));
    return id;
}

// Returns the bucket index that just became complete, -1 if its bucket still waits for other
// parameters, -2 if the parameter was already marked in this iteration, < -2 on error.
extern "C" int ptd_reducer_mark_ready(int id, int param) {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_reducers.find(id);
    if (it == g_reducers.end()) return -3;
    Reducer& r = it->second;
    if (param < 0 || param >= (int)r.ready.size()) return -4;
    if (r.ready[param]) return -2;
    r.ready[param] = 1;
    r.ready_order.push_back(param);
    int b = r.bucket_of[param];
    if (--r.pending[b] == 0) {
        r.buckets_done++;
        return b;
    }
    return -1;
}

// 1 when every bucket completed in this iteration.
extern "C" int ptd_reducer_all_done(int id) {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_reducers.find(id);
    if (it == g_reducers.end()) return -3;
    return it->second.buckets_done == (int)it->second.bucket_size.size() ? 1 : 0;
}

// Copies the ready order of the finished iteration into `out` (n entries) and returns how many were recorded.
extern "C" int ptd_reducer_ready_order(int id, int* out, int n) {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_reducers.find(id);
    if (it == g_reducers.end()) return -3;
    const auto& ro = it->second.ready_order;
    int m = (int)ro.size() < n ? (int)ro.size() : n;
    for (int i = 0; i < m; ++i) out[i] = ro[i];
    return (int)ro.size();
}

// Buckets that did not complete (unused parameters): writes their indices, returns the count.
extern "C" int ptd_reducer_incomplete(int id, int* out, int n) {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_reducers.find(id);
    if (it == g_reducers.end()) return -3;
    int c = 0;
    for (int b = 0; b < (int)it->second.pending.size(); ++b)
        if (it->second.pending[b] != 0) {
            if (c < n) out[c] = b;
            ++c;
        }
    return c;
}

extern "C" int ptd_reducer_reset(int id) {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_reducers.find(id);
    if (it == g_reducers.end()) return -3;
    Reducer& r = it->second;
    r.pending = r.bucket_size;
    std::fill(r.ready.begin(), r.ready.end(), 0);
    r.ready_order.clear();
    r.buckets_done = 0;
    return 0;
}

extern "C" int ptd_reducer_destroy(int id) {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_reducers.find(id);
    if (it == g_reducers.end()) return -3;
    for (auto& bc : it->second.comm) {
        if (bc.ready) cudaEventDestroy(bc.ready);
        if (bc.done) cudaEventDestroy(bc.done);
    }
    g_reducers.erase(it);
    return 0;
}

// Attach the communication recipe of `bucket` (copied).  algo: 1 = two-shot peer-to-peer, 2 = through the switch
// (multimem), 3 = bf16 wire format; fuse_sgd != 0 adds the fused SGD step (algo 1 and 2).  Creates the bucket's two
// events on the current device.
extern "C" int ptd_reducer_bind_bucket(int id, int bucket, const PtdBucketComm* c) {
    PTD_REQUIRE(c != nullptr && c->algo >= 1 && c->algo <= 3, "reducer_bind_bucket: algo must be 1, 2 or 3");
    PTD_REQUIRE(c->world >= 2 && c->world <= 8 && c->rank >= 0 && c->rank < c->world, "reducer_bind_bucket: bad rank / world");
    PTD_REQUIRE(!(c->fuse_sgd && c->algo == 3), "reducer_bind_bucket: the fused SGD step needs fp32 on the wire");
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_reducers.find(id);
    PTD_REQUIRE(it != g_reducers.end(), "reducer_bind_bucket: unknown reducer %d", id);
    Reducer& r = it->second;
    PTD_REQUIRE(bucket >= 0 && bucket < (int)r.comm.size(), "reducer_bind_bucket: bucket %d outside [0,%d)", bucket, (int)r.comm.size());
    BucketComm& bc = r.comm[bucket];
    bc.c = *c;
    if (!bc.ready) PTD_CUDA_CHECK(cudaEventCreateWithFlags(&bc.ready, cudaEventDisableTiming));
    if (!bc.done) PTD_CUDA_CHECK(cudaEventCreateWithFlags(&bc.done, cudaEventDisableTiming));
    bc.bound = true;
    bc.launched = false;
    return PTD_OK;
}

static int launch_bucket(BucketComm& bc, cudaStream_t compute, cudaStream_t comm) {
    const PtdBucketComm& c = bc.c;
    PTD_CUDA_CHECK(cudaEventRecord(bc.ready, compute));          // the bucket's gradients are final on the compute stream
    PTD_CUDA_CHECK(cudaStreamWaitEvent(comm, bc.ready, 0));
    int rc;
    if (c.algo == 3) {
        rc = ptd_allreduce_wire_bf16((float*)c.local, c.stages, c.mc_stage, c.pads, c.mc_pad, c.rank, c.world, c.numel,
                                     c.scale, c.blocks, comm);
    } else if (c.fuse_sgd) {
        rc = ptd_allreduce_fused_sgd(c.bufs, c.algo == 2 ? c.mc_buf : nullptr, c.pads, c.mc_pad, c.rank, c.world, c.numel,
                                     c.scale, c.blocks, &c.sgd, comm);
    } else if (c.algo == 2) {
        rc = ptd_allreduce_mc(c.mc_buf, c.pads, c.mc_pad, c.rank, c.world, c.numel, 0, c.scale, c.blocks, comm);
    } else {
        rc = ptd_allreduce_inplace(c.bufs, c.pads, c.rank, c.world, c.numel, 0, c.scale, c.blocks, comm);
    }
    if (rc) return rc;
    PTD_CUDA_CHECK(cudaEventRecord(bc.done, comm));
    bc.launched = true;
    return PTD_OK;
}

// ptd_reducer_mark_ready + launch: when `param` completes a bound bucket, its reduction is issued on `comm_stream`
// behind everything recorded so far on `compute_stream`.  Same return values as ptd_reducer_mark_ready; -5 if the
// launch itself failed (ptd_last_error tells why).
extern "C" int ptd_reducer_mark_ready_launch(int id, int param, void* compute_stream, void* comm_stream) {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_reducers.find(id);
    if (it == g_reducers.end()) return -3;
    Reducer& r = it->second;
    if (param < 0 || param >= (int)r.ready.size()) return -4;
    if (r.ready[param]) return -2;
    r.ready[param] = 1;
    r.ready_order.push_back(param);
    const int b = r.bucket_of[param];
    if (--r.pending[b] != 0) return -1;
    r.buckets_done++;
    if (r.comm[b].bound && launch_bucket(r.comm[b], (cudaStream_t)compute_stream, (cudaStream_t)comm_stream) != PTD_OK)
        return -5;
    return b;
}

// End of a backward pass: `compute_stream` waits for every reduction issued in it.  Returns 1 when every bucket
// completed, 0 when some did not (ptd_reducer_incomplete lists them), negative on error.  Does not reset the ready state.
extern "C" int ptd_reducer_finalize(int id, void* compute_stream) {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_reducers.find(id);
    if (it == g_reducers.end()) return -3;
    Reducer& r = it->second;
    for (auto& bc : r.comm) {
        if (!bc.launched) continue;
        if (cudaStreamWaitEvent((cudaStream_t)compute_stream, bc.done, 0) != cudaSuccess) {
            ptd_set_error("reducer_finalize: cudaStreamWaitEvent failed: %s", cudaGetErrorString(cudaGetLastError()));
            return -5;
        }
        bc.launched = false;
    }
    return r.buckets_done == (int)r.bucket_size.size() ? 1 : 0;
}
