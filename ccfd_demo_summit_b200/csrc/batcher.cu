// Micro-batcher: Kafka-style transaction records -> pinned session slots -> CUDA-graph launches.
//
// In the demo every transaction is one Kafka message on topic odh-demo (reference
// deploy/kafka/ProducerDeployment.yaml:88-97, README.md:547-548) and the router turns each into one
// model call.  BASELINE configs[1] asks for the B200 shape of that: records are collected into
// 1,024-row micro-batches in pinned host memory and each batch is one pre-captured graph launch.
// This file is that ring.  The slots of a ccfd_session ARE the ring: the producer writes records
// straight into the pinned buffer the GPU copies from (no intermediate queue, no per-record
// allocation), a full slot is submitted at once, and the next slot fills while it is in flight.
// A collector thread waits for submitted slots in submission order and stamps "result visible";
// per-batch latency = that stamp minus the arrival of the batch's last record (SURVEY.md §8d,
// config 2).  Back-pressure: when every slot is filling, in flight or waiting to be polled, push
// blocks.
//
// One producer thread (push / flush / replay) and one consumer thread (poll) at a time.
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <climits>
#include <cstdint>
#include <cstring>
#include <deque>
#include <thread>

#include "ccfd_internal.h"

using Clock = std::chrono::steady_clock;

struct ccfd_batcher {
    ccfd_session* s = nullptr;
    int64_t batch = 0;
    int32_t F = 0, n_slots = 0;
    enum State : int { FREE, FILLING, IN_FLIGHT, DONE };
    struct Slot {
        State st = FREE;
        float* x = nullptr;
        double* p = nullptr;
        int64_t first_row = 0, n = 0;
        Clock::time_point t_last;
        double latency_us = 0.0;
        int rc = CCFD_OK;
    };
    std::vector<Slot> slots;
    std::deque<int> in_flight, done;  // submission order
    int cur = -1;                     // slot being filled (producer side)
    std::atomic<int64_t> fill{0}, next_row{0};  // written by the producer thread only; stats may read them
    std::mutex mu;
    std::condition_variable cv_free, cv_flight, cv_done;
    std::thread collector;
    bool stop = false;
    int64_t batches_done = 0;
};

namespace {

void collector_main(ccfd_batcher* b) {
    for (;;) {
        int k;
        {
            std::unique_lock<std::mutex> lk(b->mu);
            b->cv_flight.wait(lk, [&] { return b->stop || !b->in_flight.empty(); });
            if (b->in_flight.empty()) return;  // stop requested and nothing left to collect
            k = b->in_flight.front();
        }
        const int rc = ccfd_session_wait(b->s, k);  // results are in the slot's pinned p buffer from here on
        const Clock::time_point t = Clock::now();
        {
            std::lock_guard<std::mutex> lk(b->mu);
            ccfd_batcher::Slot& sl = b->slots[(size_t)k];
            sl.rc = rc;
            sl.latency_us = std::chrono::duration<double, std::micro>(t - sl.t_last).count();
            sl.st = ccfd_batcher::DONE;
            b->in_flight.pop_front();
            b->done.push_back(k);
            ++b->batches_done;
        }
        b->cv_done.notify_all();
    }
}

// producer side: make sure a slot is being filled; blocks while none is free
int acquire_slot(ccfd_batcher* b) {
    if (b->cur >= 0) return CCFD_OK;
    std::unique_lock<std::mutex> lk(b->mu);
    int k = -1;
    b->cv_free.wait(lk, [&] {
        for (int i = 0; i < b->n_slots; ++i)
            if (b->slots[(size_t)i].st == ccfd_batcher::FREE) {
                k = i;
                return true;
            }
        return b->stop;
    });
    if (k < 0) return ccfd_fail(CCFD_ESTATE, "batcher is shutting down");
    b->slots[(size_t)k].st = ccfd_batcher::FILLING;
    b->cur = k;
    b->fill = 0;
    return CCFD_OK;
}

int submit_current(ccfd_batcher* b) {
    if (b->cur < 0 || b->fill == 0) return CCFD_OK;
    const int k = b->cur;
    ccfd_batcher::Slot& sl = b->slots[(size_t)k];
    const int64_t fill = b->fill.load();
    sl.n = fill;
    sl.first_row = b->next_row.load();
    const int rc = ccfd_session_submit(b->s, k, fill);
    if (rc != CCFD_OK) return rc;
    {
        std::lock_guard<std::mutex> lk(b->mu);
        sl.st = ccfd_batcher::IN_FLIGHT;
        b->in_flight.push_back(k);
    }
    b->cv_flight.notify_one();
    b->next_row += fill;
    b->cur = -1;
    b->fill = 0;
    return CCFD_OK;
}

int push_impl(ccfd_batcher* b, const float* records, int64_t n, int64_t ld) {
    int64_t i = 0;
    while (i < n) {
        int rc = acquire_slot(b);
        if (rc != CCFD_OK) return rc;
        ccfd_batcher::Slot& sl = b->slots[(size_t)b->cur];
        const int64_t take = std::min<int64_t>(n - i, b->batch - b->fill);
        // Records go into the pinned slot with streaming stores (the copy engine reads them next; out of a CPU
        // cache it would pull them at a quarter of its speed), checked for +-inf on the way: scikit-learn refuses
        // +-inf before scoring (SK/ensemble/_forest.py:606-625); a record stream has nobody to raise to, so the
        // records are refused here.
        int32_t bad = 0;
        const int rc_copy = ccfd_cast_span(records + i * ld, 0, take, b->F, ld, sl.x + b->fill * b->F, &bad, INT64_MAX, take * b->F >= 1024);
        if (rc_copy != CCFD_OK) return rc_copy;
        if (bad) return ccfd_fail(CCFD_EINVAL, "a record in [%lld, %lld) holds +-infinity", (long long)(b->next_row + b->fill), (long long)(b->next_row + b->fill + take));
        b->fill += take;
        i += take;
        sl.t_last = Clock::now();  // arrival of the (so far) last record of this batch
        if (b->fill == b->batch) {
            rc = submit_current(b);
            if (rc != CCFD_OK) return rc;
        }
    }
    return CCFD_OK;
}

}  // namespace

extern "C" int ccfd_batcher_create(ccfd_session* session, int64_t batch_rows, ccfd_batcher** out) {
    return ccfd_guarded("ccfd_batcher_create", [&]() -> int {
        if (!session || !out) return ccfd_fail(CCFD_EINVAL, "null argument");
        *out = nullptr;
        int64_t max_batch = 0;
        int32_t n_slots = 0, F = 0;
        int rc = ccfd_session_geometry(session, &max_batch, &n_slots, &F);
        if (rc != CCFD_OK) return rc;
        if (batch_rows < 1 || batch_rows > max_batch) return ccfd_fail(CCFD_EINVAL, "batch_rows must be in 1..%lld (got %lld)", (long long)max_batch, (long long)batch_rows);
        if (n_slots < 2) return ccfd_fail(CCFD_EINVAL, "a batcher needs a session with at least 2 slots (one fills while one is in flight)");
        ccfd_batcher* b = new ccfd_batcher();
        b->s = session;
        b->batch = batch_rows;
        b->F = F;
        b->n_slots = n_slots;
        b->slots.resize((size_t)n_slots);
        for (int i = 0; i < n_slots; ++i) {
            rc = ccfd_session_buffers(session, i, &b->slots[(size_t)i].x, &b->slots[(size_t)i].p);
            if (rc != CCFD_OK) {
                delete b;
                return rc;
            }
        }
        try {
            b->collector = std::thread(collector_main, b);
        } catch (...) {
            delete b;
            throw;
        }
        *out = b;
        return CCFD_OK;
    });
}

extern "C" int ccfd_batcher_destroy(ccfd_batcher* b) {
    return ccfd_guarded("ccfd_batcher_destroy", [&]() -> int {
        if (!b) return CCFD_OK;
        {
            std::lock_guard<std::mutex> lk(b->mu);
            b->stop = true;
        }
        b->cv_flight.notify_all();
        b->cv_free.notify_all();
        b->cv_done.notify_all();
        if (b->collector.joinable()) b->collector.join();  // drains what is in flight first
        delete b;
        return CCFD_OK;
    });
}

extern "C" int ccfd_batcher_push(ccfd_batcher* b, const float* records, int64_t n, int64_t ld) {
    return ccfd_guarded("ccfd_batcher_push", [&]() -> int {
        if (!b || (!records && n > 0)) return ccfd_fail(CCFD_EINVAL, "null argument");
        if (n < 0 || ld < b->F) return ccfd_fail(CCFD_EINVAL, "bad sizes (n %lld, record stride %lld floats, n_features %d)", (long long)n, (long long)ld, b->F);
        return push_impl(b, records, n, ld);
    });
}

extern "C" int ccfd_batcher_flush(ccfd_batcher* b) {
    return ccfd_guarded("ccfd_batcher_flush", [&]() -> int {
        if (!b) return ccfd_fail(CCFD_EINVAL, "null argument");
        return submit_current(b);
    });
}

extern "C" int ccfd_batcher_poll(ccfd_batcher* b, int32_t timeout_ms, int64_t* first_row, int64_t* n_rows, double* proba, int64_t cap_rows,
                                 double* latency_us) {
    return ccfd_guarded("ccfd_batcher_poll", [&]() -> int {
        if (!b || !n_rows) return ccfd_fail(CCFD_EINVAL, "null argument");
        *n_rows = 0;
        int k = -1;
        {
            std::unique_lock<std::mutex> lk(b->mu);
            if (timeout_ms > 0) b->cv_done.wait_for(lk, std::chrono::milliseconds(timeout_ms), [&] { return !b->done.empty() || b->stop; });
            if (b->done.empty()) return 0;
            k = b->done.front();
        }
        ccfd_batcher::Slot& sl = b->slots[(size_t)k];
        if (sl.rc != CCFD_OK) {
            std::lock_guard<std::mutex> lk(b->mu);
            b->done.pop_front();
            sl.st = ccfd_batcher::FREE;
            b->cv_free.notify_all();
            return ccfd_fail(sl.rc, "scoring of the batch starting at record %lld failed", (long long)sl.first_row);
        }
        if (proba) {
            if (cap_rows < sl.n) return ccfd_fail(CCFD_EINVAL, "proba holds %lld rows, the batch has %lld", (long long)cap_rows, (long long)sl.n);
            memcpy(proba, sl.p, sizeof(double) * 2 * (size_t)sl.n);
        }
        if (first_row) *first_row = sl.first_row;
        if (latency_us) *latency_us = sl.latency_us;
        *n_rows = sl.n;
        {
            std::lock_guard<std::mutex> lk(b->mu);
            b->done.pop_front();
            sl.st = ccfd_batcher::FREE;
        }
        b->cv_free.notify_all();
        return 1;
    });
}

// Replay tool: what the demo's producer does with creditcard.csv, at a chosen rate.  Record i is pushed
// when its due time i / records_per_s has come (busy-wait pacing on the steady clock; records_per_s <= 0:
// as fast as the ring takes them) in bursts of `burst` records, the way a consumer poll delivers them.
extern "C" int ccfd_batcher_replay(ccfd_batcher* b, const float* records, int64_t n, int64_t ld, double records_per_s, int32_t burst) {
    return ccfd_guarded("ccfd_batcher_replay", [&]() -> int {
        if (!b || (!records && n > 0)) return ccfd_fail(CCFD_EINVAL, "null argument");
        if (n < 0 || ld < b->F || burst < 1) return ccfd_fail(CCFD_EINVAL, "bad sizes");
        const Clock::time_point t0 = Clock::now();
        for (int64_t i = 0; i < n; i += burst) {
            const int64_t take = std::min<int64_t>(burst, n - i);
            if (records_per_s > 0) {
                const auto due = t0 + std::chrono::duration_cast<Clock::duration>(std::chrono::duration<double>((double)(i + take - 1) / records_per_s));
                while (Clock::now() < due) {
                }
            }
            const int rc = push_impl(b, records + i * ld, take, ld);
            if (rc != CCFD_OK) return rc;
        }
        return submit_current(b);
    });
}

extern "C" int ccfd_batcher_stats(ccfd_batcher* b, int64_t* batches_done, int64_t* records_in) {
    return ccfd_guarded("ccfd_batcher_stats", [&]() -> int {
        if (!b) return ccfd_fail(CCFD_EINVAL, "null argument");
        std::lock_guard<std::mutex> lk(b->mu);
        if (batches_done) *batches_done = b->batches_done;
        if (records_in) *records_in = b->next_row + b->fill;
        return CCFD_OK;
    });
}
