// Prefetching batch loader (reference scripts/train.py:84-86,98-100).
//
// tf.data.Dataset.from_tensor_slices(...).batch(bs) is executed by TensorFlow's C++ runtime; this
// is the equivalent piece of the native host runtime.  The tokenised set lives in host memory as
// dense int32 tables.  One worker thread walks the (already sharded) index list batch by batch and
// gathers rows into a ring of caller-owned slots — pinned tensors allocated by the Python side —
// so that the H2D copy of step k+1 never waits for host-side batch assembly.  The consumer calls
// next() (blocks until a slot is staged), issues its async copies, and release()s the slot once
// those copies have been enqueued and synchronised by the training loop.
#include <algorithm>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "common.cuh"
#include "hstd_b200.h"

struct hstd_loader {
    const int32_t* ids = nullptr;
    const int32_t* mask = nullptr;
    const int32_t* labels = nullptr;
    int64_t n_rows = 0;
    int seq_len = 0;
    std::vector<int64_t> indices;
    int batch = 0, num_batches = 0, num_slots = 0;
    std::vector<int32_t*> slot_ids, slot_mask, slot_labels;
    int32_t pad_id = 0, ignore_label = -1;

    std::mutex mu;
    std::condition_variable cv_free, cv_ready;
    std::vector<int> state;  // 0 = free, 1 = staged, 2 = held by the consumer
    std::vector<int> ready;  // staged slots in batch order (FIFO)
    size_t ready_head = 0;
    int produced = 0, consumed = 0;
    bool stop = false, running = false;
    std::thread worker;

    void fill(int slot, int b) {
        // batches past the end of this rank's shard are all padding (ranks whose shard is one sample
        // shorter still take the common number of steps)
        const int64_t n_idx = static_cast<int64_t>(indices.size());
        const int64_t lo = std::min<int64_t>(static_cast<int64_t>(b) * batch, n_idx);
        const int64_t hi = std::min<int64_t>(lo + batch, n_idx);
        const size_t row_bytes = sizeof(int32_t) * static_cast<size_t>(seq_len);
        int32_t* oi = slot_ids[slot];
        int32_t* om = slot_mask[slot];
        int32_t* ol = slot_labels[slot];
        for (int64_t r = lo; r < hi; ++r) {
            const int64_t src = indices[r];
            std::memcpy(oi + (r - lo) * seq_len, ids + src * seq_len, row_bytes);
            std::memcpy(om + (r - lo) * seq_len, mask + src * seq_len, row_bytes);
            ol[r - lo] = labels[src];
        }
        for (int64_t r = hi - lo; r < batch; ++r) {  // short final batch: well-formed ignored rows
            for (int s = 0; s < seq_len; ++s) {
                oi[r * seq_len + s] = pad_id;
                om[r * seq_len + s] = 0;
            }
            om[r * seq_len] = 1;
            ol[r] = ignore_label;
        }
    }

    void run() {
        for (int b = 0; b < num_batches; ++b) {
            int slot = -1;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv_free.wait(lk, [&] {
                    if (stop) return true;
                    for (int s = 0; s < num_slots; ++s)
                        if (state[s] == 0) return true;
                    return false;
                });
                if (stop) return;
                for (int s = 0; s < num_slots; ++s)
                    if (state[s] == 0) {
                        slot = s;
                        break;
                    }
                state[slot] = 3;  // being filled
            }
            fill(slot, b);
            {
                std::lock_guard<std::mutex> lk(mu);
                state[slot] = 1;
                ready.push_back(slot);
                ++produced;
            }
            cv_ready.notify_all();
        }
    }

    void join() {
        {
            std::lock_guard<std::mutex> lk(mu);
            stop = true;
        }
        cv_free.notify_all();
        cv_ready.notify_all();
        if (worker.joinable()) worker.join();
        running = false;
    }
};

extern "C" HSTD_API hstd_loader* hstd_loader_create(const int32_t* ids, const int32_t* mask, const int32_t* labels,
                                                    int64_t n_rows, int seq_len, const int64_t* indices, int64_t n_indices,
                                                    int batch, int num_batches, int num_slots, int32_t* const* slot_ids,
                                                    int32_t* const* slot_mask, int32_t* const* slot_labels, int32_t pad_id,
                                                    int32_t ignore_label) {
    if (!ids || !mask || !labels || n_rows <= 0 || seq_len <= 0 || batch <= 0 || num_slots < 2 || n_indices < 0) {
        hstd::set_error(-1, "hstd_loader_create: bad arguments (need tables, batch > 0, at least 2 slots)");
        return nullptr;
    }
    const int64_t max_batches = (n_indices + batch - 1) / batch;
    if (num_batches < 0) num_batches = static_cast<int>(max_batches);
    for (int64_t i = 0; i < n_indices; ++i)
        if (indices[i] < 0 || indices[i] >= n_rows) {
            hstd::set_error(-1, "hstd_loader_create: index out of range");
            return nullptr;
        }
    auto* l = new hstd_loader();
    l->ids = ids;
    l->mask = mask;
    l->labels = labels;
    l->n_rows = n_rows;
    l->seq_len = seq_len;
    l->indices.assign(indices, indices + n_indices);
    l->batch = batch;
    l->num_batches = num_batches;
    l->num_slots = num_slots;
    l->slot_ids.assign(slot_ids, slot_ids + num_slots);
    l->slot_mask.assign(slot_mask, slot_mask + num_slots);
    l->slot_labels.assign(slot_labels, slot_labels + num_slots);
    l->pad_id = pad_id;
    l->ignore_label = ignore_label;
    l->state.assign(num_slots, 0);
    return l;
}

extern "C" HSTD_API void hstd_loader_destroy(hstd_loader* l) {
    if (!l) return;
    l->join();
    delete l;
}

extern "C" HSTD_API int hstd_loader_start_epoch(hstd_loader* l) {
    if (!l) return hstd::set_error(-1, "hstd_loader_start_epoch: null loader");
    l->join();  // abandon whatever was left of the previous epoch
    l->state.assign(l->num_slots, 0);
    l->ready.clear();
    l->ready_head = 0;
    l->produced = l->consumed = 0;
    l->stop = false;
    l->running = true;
    l->worker = std::thread([l] { l->run(); });
    return 0;
}

extern "C" HSTD_API int hstd_loader_next(hstd_loader* l) {
    if (!l) return hstd::set_error(-2, "hstd_loader_next: null loader");
    if (!l->running) return hstd::set_error(-2, "hstd_loader_next: start_epoch() not called");
    std::unique_lock<std::mutex> lk(l->mu);
    if (l->consumed >= l->num_batches) return -1;
    l->cv_ready.wait(lk, [&] { return l->stop || l->ready_head < l->ready.size(); });
    if (l->ready_head >= l->ready.size()) return -1;
    int slot = l->ready[l->ready_head++];
    l->state[slot] = 2;
    ++l->consumed;
    return slot;
}

extern "C" HSTD_API int hstd_loader_release(hstd_loader* l, int slot) {
    if (!l) return hstd::set_error(-1, "hstd_loader_release: null loader");
    if (slot < 0 || slot >= l->num_slots) return hstd::set_error(-1, "hstd_loader_release: slot out of range");
    {
        std::lock_guard<std::mutex> lk(l->mu);
        if (l->state[slot] != 2) return hstd::set_error(-1, "hstd_loader_release: slot is not held by the consumer");
        l->state[slot] = 0;
    }
    l->cv_free.notify_all();
    return 0;
}
