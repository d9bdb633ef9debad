// Fusion-bucket manager (reference scripts/train.py:114 -> Horovod C++ core).
//
// Horovod negotiates ready tensors between ranks, copies them into a fusion buffer and all-reduces
// that.  In this framework every rank records the same static plan, the gradient arena already is
// the fusion buffer (parameters sit in backward-readiness order), so the "negotiation" reduces to
// two deterministic pieces of bookkeeping, kept native next to the kernels that consume them:
//   * hstd_bucket_plan    — cut the arena into contiguous buckets at parameter boundaries;
//   * hstd_bucket_tracker — given "gradients of these parameters now exist", release the buckets
//                           that are complete, strictly front to back (a fused all-reduce kernel
//                           is launched per released bucket on the communication stream).
#include <cstdint>
#include <string>
#include <vector>

#include "common.cuh"
#include "hstd_b200.h"

extern "C" HSTD_API int hstd_bucket_plan(const int64_t* offsets, const int64_t* ends, int n, int64_t bucket_elems,
                                         int64_t* starts, int64_t* bucket_ends, int32_t* param_bucket, int max_buckets) {
    if (n <= 0 || !offsets || !ends) return hstd::set_error(-1, "hstd_bucket_plan: no parameters");
    if (bucket_elems <= 0) return hstd::set_error(-1, "hstd_bucket_plan: bucket_elems must be positive");
    for (int i = 0; i < n; ++i) {
        if (ends[i] < offsets[i]) return hstd::set_error(-1, "hstd_bucket_plan: end < offset");
        if (i && offsets[i] < ends[i - 1]) return hstd::set_error(-1, "hstd_bucket_plan: parameters must be in arena order");
    }
    int nb = 0;
    int64_t cur_start = 0;
    bool open = false;
    for (int i = 0; i < n; ++i) {
        if (open && ends[i] - cur_start > bucket_elems) {
            if (nb >= max_buckets) return hstd::set_error(-3, "hstd_bucket_plan: max_buckets too small");
            starts[nb] = cur_start;
            bucket_ends[nb] = offsets[i];
            ++nb;
            cur_start = offsets[i];
            open = false;
        }
        param_bucket[i] = nb;
        open = true;
    }
    if (nb >= max_buckets) return hstd::set_error(-3, "hstd_bucket_plan: max_buckets too small");
    starts[nb] = cur_start;
    bucket_ends[nb] = ends[n - 1];
    return nb + 1;
}

struct hstd_bucket_tracker {
    std::vector<int32_t> param_bucket;
    std::vector<uint8_t> exempt;
    std::vector<int32_t> need;  // per bucket: non-exempt parameters
    std::vector<int32_t> left;  // per bucket: still missing
    std::vector<uint8_t> seen;  // per parameter
    int next = 0;               // first bucket not yet released
};

extern "C" HSTD_API hstd_bucket_tracker* hstd_bucket_tracker_create(const int32_t* param_bucket, const uint8_t* exempt,
                                                                    int n, int num_buckets) {
    if (n <= 0 || num_buckets <= 0 || !param_bucket) {
        hstd::set_error(-1, "hstd_bucket_tracker_create: empty plan");
        return nullptr;
    }
    auto* t = new hstd_bucket_tracker();
    t->param_bucket.assign(param_bucket, param_bucket + n);
    t->exempt.assign(n, 0);
    if (exempt) t->exempt.assign(exempt, exempt + n);
    t->need.assign(num_buckets, 0);
    for (int i = 0; i < n; ++i) {
        int b = param_bucket[i];
        if (b < 0 || b >= num_buckets) {
            delete t;
            hstd::set_error(-1, "hstd_bucket_tracker_create: bucket index out of range");
            return nullptr;
        }
        if (!t->exempt[i]) t->need[b] += 1;
    }
    t->left = t->need;
    t->seen.assign(n, 0);
    return t;
}

extern "C" HSTD_API void hstd_bucket_tracker_destroy(hstd_bucket_tracker* t) { delete t; }

extern "C" HSTD_API void hstd_bucket_tracker_reset(hstd_bucket_tracker* t) {
    if (!t) return;
    t->left = t->need;
    t->seen.assign(t->seen.size(), 0);
    t->next = 0;
}

extern "C" HSTD_API int hstd_bucket_tracker_mark(hstd_bucket_tracker* t, const int32_t* params, int k, int32_t* ready_out) {
    if (!t) return hstd::set_error(-1, "hstd_bucket_tracker_mark: null tracker");
    const int n = static_cast<int>(t->param_bucket.size());
    for (int j = 0; j < k; ++j) {
        int p = params[j];
        if (p < 0 || p >= n) return hstd::set_error(-1, "hstd_bucket_tracker_mark: parameter index out of range");
        if (t->seen[p]) continue;
        t->seen[p] = 1;
        if (!t->exempt[p]) t->left[t->param_bucket[p]] -= 1;
    }
    int out = 0;
    const int nb = static_cast<int>(t->need.size());
    while (t->next < nb && t->left[t->next] == 0) ready_out[out++] = t->next++;
    return out;
}

extern "C" HSTD_API int hstd_bucket_tracker_done(const hstd_bucket_tracker* t) { return t ? t->next : 0; }
