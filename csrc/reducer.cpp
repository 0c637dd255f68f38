// Host-side gradient bucket planner and ready-tracking reducer (no CUDA in this file).
//
// Serves the role of torch's C++ Reducer + _compute_bucket_assignment_by_size that the reference
// reaches through DDP(model, device_ids=[gpu_id]) (R/02 DDP基本概念/ddp_gpus.py:35):
//   * parameters are assigned to buckets in the order given (the wrapper passes reverse
//     registration order, then the observed ready order); the first bucket closes at `first_cap`
//     bytes, later ones at `cap` bytes (1 MiB / 25 MiB by default) - a bucket closes as soon as its
//     size reaches the limit;
//   * during backward each parameter is marked ready once; when the last parameter of a bucket
//     arrives the bucket index is returned so the caller can launch that bucket's all-reduce;
//   * the order in which parameters became ready is recorded for the one-time bucket rebuild.
#include <stdint.h>

#include <algorithm>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "common.cuh"
#include "ptdb200.h"

namespace {

struct Reducer {
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
    return g_reducers.erase(id) ? 0 : -3;
}
