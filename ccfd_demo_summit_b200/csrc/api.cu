// Scoring entry points and the latency session (host code above the kernels).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "ccfd_internal.h"

namespace {

constexpr int64_t kHostChunkRows = 1 << 20;  // 120 MB of fp32 rows per staged chunk
constexpr int64_t kDirectOutRows = 4096;     // session buckets up to this size write results straight to pinned host memory

int ensure_staging(ccfd_forest* f, int64_t rows) {
    if (f->stage_rows >= rows) return CCFD_OK;
    for (int i = 0; i < 2; ++i) {
        if (f->stage_stream[i]) CCFD_CUDA(cudaStreamSynchronize(f->stage_stream[i]));
        cudaFree(f->stage_x[i]);
        cudaFree(f->stage_p[i]);
        cudaFree(f->stage_d[i]);
        f->stage_x[i] = nullptr;
        f->stage_p[i] = nullptr;
        f->stage_d[i] = nullptr;
    }
    f->stage_rows = 0;
    for (int i = 0; i < 2; ++i) {
        if (!f->stage_stream[i]) CCFD_CUDA(cudaStreamCreateWithFlags(&f->stage_stream[i], cudaStreamNonBlocking));
        CCFD_CUDA(cudaMalloc(&f->stage_x[i], sizeof(float) * (size_t)rows * (size_t)f->dev.n_features));
        CCFD_CUDA(cudaMalloc(&f->stage_p[i], sizeof(double) * 2 * (size_t)rows));
        CCFD_CUDA(cudaMalloc(&f->stage_d[i], (size_t)rows));
    }
    f->stage_rows = rows;
    return CCFD_OK;
}

// Enqueue one scoring of device-resident rows on `stream`; SPLIT takes its scratch from the
// stream-ordered allocator so concurrent callers never share it.
int enqueue_device(const ccfd_forest* f, const ScoreArgs& a, uint32_t kernel, cudaStream_t stream) {
    if (kernel == CCFD_KERNEL_AUTO) kernel = ccfd_pick_kernel(f, a.n);
    void* scratch = nullptr;
    ScoreArgs args = a;
    int rc = CCFD_OK;
    if (kernel == CCFD_KERNEL_SPLIT) {
        CCFD_CUDA(cudaMallocAsync(&scratch, (size_t)ccfd_split_scratch_bytes(f, a.n), stream));
    } else if (ccfd_row_sort_wanted(f, kernel, a.n)) {
        // bulk GLOBAL launch: order the rows first so that a warp's rows share tree paths (sort.cu)
        CCFD_CUDA(cudaMallocAsync(&scratch, (size_t)ccfd_sort_scratch_bytes(f, a.n), stream));
        rc = ccfd_launch_row_sort(f, a.X, a.n, a.ld, scratch, &args.perm, stream);
    }
    if (rc == CCFD_OK) rc = ccfd_launch_score(f, args, kernel, scratch, stream);
    if (scratch) {
        cudaError_t e = cudaFreeAsync(scratch, stream);
        if (rc == CCFD_OK && e != cudaSuccess) return ccfd_cuda_fail(e, "cudaFreeAsync");
    }
    return rc;
}

// Rows per pipeline step of the pageable-host path.  Measured on the B200 box with the 284,807-row batch:
// steps of 65,536 rows cost 0.23 ms of enqueueing each (copy-in, launch, copy-out) and 4.0-4.4 ms in all,
// one step for the whole batch 2.5 ms: the steps only pay once a request is far beyond a million rows.
constexpr int64_t kPageableChunkRows = 1 << 20;

int ensure_pinned(ccfd_forest* f, int64_t rows) {
    if (f->pin_rows >= rows) return CCFD_OK;
    for (int i = 0; i < 2; ++i) {
        if (f->stage_stream[i]) CCFD_CUDA(cudaStreamSynchronize(f->stage_stream[i]));
        cudaFreeHost(f->pin_x[i]);
        cudaFreeHost(f->pin_p[i]);
        cudaFreeHost(f->pin_d[i]);
        f->pin_x[i] = nullptr;
        f->pin_p[i] = nullptr;
        f->pin_d[i] = nullptr;
    }
    f->pin_rows = 0;
    for (int i = 0; i < 2; ++i) {
        CCFD_CUDA(cudaHostAlloc(&f->pin_x[i], sizeof(float) * (size_t)rows * (size_t)f->dev.n_features, cudaHostAllocDefault));
        CCFD_CUDA(cudaHostAlloc(&f->pin_p[i], sizeof(double) * 2 * (size_t)rows, cudaHostAllocDefault));
        CCFD_CUDA(cudaHostAlloc(&f->pin_d[i], (size_t)rows, cudaHostAllocDefault));
    }
    f->pin_rows = rows;
    return CCFD_OK;
}

// Rows in PAGEABLE host memory (what a Seldon wrapper hands predict(): a numpy matrix, float64 after JSON
// decoding or float32).  A cudaMemcpyAsync from pageable memory is staged by the driver on one thread; here
// the library does the staging itself, fused with the float64 -> float32 narrowing and the infinity scan of
// ForestClassifier._validate_X_predict (SK/ensemble/_forest.py:606-625), on its host thread pool, two
// pinned buffers deep: while the GPU copies in, scores and copies out chunk k, the CPU casts chunk k+1 and
// hands chunk k-1's results to the caller's array.  *has_inf (may be null) reports +-inf / float32 overflow
// in the input; scoring stops at the chunk that holds it.
int score_pageable(ccfd_forest* f, const void* X, int32_t is_f64, int64_t n, int64_t ld, double* out, uint8_t* decision, double threshold,
                   int ge, uint32_t kernel, int32_t* has_inf) {
    std::lock_guard<std::mutex> lock(f->host_mu);
    const int F = f->dev.n_features;
    const int64_t chunk = std::min<int64_t>(n, kPageableChunkRows);
    int rc = ensure_staging(f, chunk);
    if (rc == CCFD_OK) rc = ensure_pinned(f, chunk);
    if (rc != CCFD_OK) return rc;
    if (has_inf) *has_inf = 0;
    struct InFlight {
        int64_t r0 = 0, rows = 0;
        bool live = false;
    } fl[2];
    auto finish = [&](int s) -> int {  // wait for slot s and hand its results over
        if (!fl[s].live) return CCFD_OK;
        fl[s].live = false;
        CCFD_CUDA(cudaStreamSynchronize(f->stage_stream[s]));
        // the caller's result array is usually fresh memory: every page faults on first touch, so the
        // copy is spread over the host pool (4.5 MB of results: 1.5 ms on one thread, a fifth of that on the pool)
        auto spread = [](void* dst, const void* src, size_t bytes) {
            constexpr size_t kPart = 256 * 1024;
            const int parts = (int)std::min<size_t>((bytes + kPart - 1) / kPart, (size_t)ccfd_host_threads());
            if (parts <= 1) {
                memcpy(dst, src, bytes);
                return;
            }
            ccfd_parallel_for(parts, [&](int k) {
                const size_t lo = bytes * (size_t)k / (size_t)parts / 64 * 64, hi = k + 1 == parts ? bytes : bytes * (size_t)(k + 1) / (size_t)parts / 64 * 64;
                memcpy(static_cast<char*>(dst) + lo, static_cast<const char*>(src) + lo, hi - lo);
            });
        };
        if (out) spread(out + 2 * fl[s].r0, f->pin_p[s], sizeof(double) * 2 * (size_t)fl[s].rows);
        if (decision) spread(decision + fl[s].r0, f->pin_d[s], (size_t)fl[s].rows);
        return CCFD_OK;
    };
    const bool trace = getenv("CCFD_TRACE_PAGEABLE") != nullptr;  // development aid: where a call's time goes
    double t_finish = 0, t_cast = 0, t_enq = 0;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_begin = now();
    int k = 0;
    for (int64_t r0 = 0; r0 < n && rc == CCFD_OK; r0 += chunk, ++k) {
        const int s = k & 1;
        const int64_t rows = std::min<int64_t>(chunk, n - r0);
        double t0 = now();
        rc = finish(s);  // chunk k-2 used this slot
        t_finish += now() - t0;
        if (rc != CCFD_OK) break;
        int32_t bad = 0;
        t0 = now();
        const void* src = is_f64 ? static_cast<const void*>(static_cast<const double*>(X) + r0 * ld)
                                 : static_cast<const void*>(static_cast<const float*>(X) + r0 * ld);
        rc = ccfd_cast_span(src, is_f64, rows, F, ld, f->pin_x[s], &bad, 65536, getenv("CCFD_NO_STREAM_STORES") == nullptr);  // every pool thread takes part
        if (rc != CCFD_OK) break;
        t_cast += now() - t0;
        if (bad && has_inf) {
            *has_inf = 1;
            break;
        }
        t0 = now();
        cudaStream_t st = f->stage_stream[s];
        cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
        if (trace) {
            for (auto& e : ev) cudaEventCreate(&e);
            cudaEventRecord(ev[0], st);
        }
        CCFD_CUDA(cudaMemcpyAsync(f->stage_x[s], f->pin_x[s], sizeof(float) * (size_t)rows * F, cudaMemcpyHostToDevice, st));
        if (trace) cudaEventRecord(ev[1], st);
        ScoreArgs a{f->stage_x[s], rows, F, out ? f->stage_p[s] : nullptr, decision ? f->stage_d[s] : nullptr, threshold, ge};
        rc = enqueue_device(f, a, kernel, st);
        if (rc != CCFD_OK) break;
        if (trace) cudaEventRecord(ev[2], st);
        if (out) CCFD_CUDA(cudaMemcpyAsync(f->pin_p[s], f->stage_p[s], sizeof(double) * 2 * (size_t)rows, cudaMemcpyDeviceToHost, st));
        if (trace) {
            cudaEventRecord(ev[3], st);
            cudaEventSynchronize(ev[3]);
            float a01 = 0, a12 = 0, a23 = 0;
            cudaEventElapsedTime(&a01, ev[0], ev[1]);
            cudaEventElapsedTime(&a12, ev[1], ev[2]);
            cudaEventElapsedTime(&a23, ev[2], ev[3]);
            fprintf(stderr, "[ccfd pageable] step %d: h2d %.3f ms, kernel %.3f ms, d2h %.3f ms\n", k, a01, a12, a23);
            for (auto& e : ev) cudaEventDestroy(e);
        }
        if (decision) CCFD_CUDA(cudaMemcpyAsync(f->pin_d[s], f->stage_d[s], (size_t)rows, cudaMemcpyDeviceToHost, st));
        fl[s].r0 = r0;
        fl[s].rows = rows;
        fl[s].live = true;
        t_enq += now() - t0;
    }
    const double t_tail = now();
    for (int j = 0; j < 2; ++j) {  // oldest first
        const int rc2 = finish((k + j) & 1);
        if (rc == CCFD_OK) rc = rc2;
    }
    if (trace)
        fprintf(stderr, "[ccfd pageable] %lld rows, %d chunks: total %.3f ms = cast %.3f + enqueue %.3f + waits %.3f + tail %.3f\n", (long long)n, k,
                now() - t_begin, t_cast, t_enq, t_finish, now() - t_tail);
    return rc;
}

bool is_pageable(const void* p) {
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
        cudaGetLastError();
        return true;
    }
    return attr.type == cudaMemoryTypeUnregistered;
}

int score_impl(ccfd_forest* f, const float* X, int64_t n, int64_t ld, double* out, uint8_t* decision, double threshold, int ge,
               uint32_t flags, void* stream_) {
    if (!f) return ccfd_fail(CCFD_EINVAL, "null forest");
    if (n < 0) return ccfd_fail(CCFD_EINVAL, "n must be >= 0 (got %lld)", (long long)n);
    if (n == 0) return CCFD_OK;
    if (!X) return ccfd_fail(CCFD_EINVAL, "X is null");
    if (!out && !decision) return ccfd_fail(CCFD_EINVAL, "nothing to write: out and decision are both null");
    if (ld < f->dev.n_features) return ccfd_fail(CCFD_EINVAL, "row stride %lld is smaller than n_features %d", (long long)ld, f->dev.n_features);
    if (flags & ~(CCFD_DEVICE_PTRS | CCFD_ASYNC | CCFD_KERNEL_MASK)) return ccfd_fail(CCFD_EINVAL, "unknown flag bits 0x%x", flags);
    if (threshold != threshold) return ccfd_fail(CCFD_EINVAL, "threshold is NaN");
    const uint32_t kernel = flags & CCFD_KERNEL_MASK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    DeviceGuard guard(f->device);

    if (flags & CCFD_DEVICE_PTRS) {
        ScoreArgs a{X, n, ld, out, decision, threshold, ge};
        int rc = enqueue_device(f, a, kernel, stream);
        if (rc != CCFD_OK) return rc;
        if (!(flags & CCFD_ASYNC)) CCFD_CUDA(cudaStreamSynchronize(stream));
        return CCFD_OK;
    }

    if (flags & CCFD_ASYNC) return ccfd_fail(CCFD_EINVAL, "CCFD_ASYNC needs CCFD_DEVICE_PTRS (host buffers are copied in chunks and the call must wait)");
    // Host rows: chunks alternate between two streams, so the copy-in of chunk k+1 and the
    // copy-out of chunk k-1 overlap the scoring of chunk k.  Every kernel produces the same bits,
    // so results do not depend on how the request is chunked.
    if (n >= 8192 && is_pageable(X)) return score_pageable(f, X, 0, n, ld, out, decision, threshold, ge, kernel, nullptr);
    std::lock_guard<std::mutex> lock(f->host_mu);
    const int64_t chunk = std::min<int64_t>(n, kHostChunkRows);
    int rc = ensure_staging(f, chunk);
    if (rc != CCFD_OK) return rc;
    const int F = f->dev.n_features;
    int k = 0;
    for (int64_t r0 = 0; r0 < n; r0 += chunk, ++k) {
        const int s = k & 1;
        const int64_t rows = std::min<int64_t>(chunk, n - r0);
        cudaStream_t st = f->stage_stream[s];
        if (ld == F)
            CCFD_CUDA(cudaMemcpyAsync(f->stage_x[s], X + r0 * ld, sizeof(float) * (size_t)rows * F, cudaMemcpyHostToDevice, st));
        else
            CCFD_CUDA(cudaMemcpy2DAsync(f->stage_x[s], sizeof(float) * F, X + r0 * ld, sizeof(float) * (size_t)ld, sizeof(float) * F, (size_t)rows, cudaMemcpyHostToDevice, st));
        ScoreArgs a{f->stage_x[s], rows, F, out ? f->stage_p[s] : nullptr, decision ? f->stage_d[s] : nullptr, threshold, ge};
        rc = enqueue_device(f, a, kernel, st);
        if (rc != CCFD_OK) break;
        if (out) CCFD_CUDA(cudaMemcpyAsync(out + 2 * r0, f->stage_p[s], sizeof(double) * 2 * (size_t)rows, cudaMemcpyDeviceToHost, st));
        if (decision) CCFD_CUDA(cudaMemcpyAsync(decision + r0, f->stage_d[s], (size_t)rows, cudaMemcpyDeviceToHost, st));
    }
    for (int s = 0; s < 2; ++s) {
        cudaError_t e = cudaStreamSynchronize(f->stage_stream[s]);
        if (rc == CCFD_OK && e != cudaSuccess) rc = ccfd_cuda_fail(e, "cudaStreamSynchronize");
    }
    return rc;
}

}  // namespace

static int ccfd_predict_proba_impl(ccfd_forest* f, const float* X, int64_t n, int64_t ld, double* out, uint32_t flags, void* stream) {
    if (!out && n > 0) return ccfd_fail(CCFD_EINVAL, "out is null");
    return score_impl(f, X, n, ld, out, nullptr, 0.5, 0, flags, stream);
}

static int ccfd_predict_decision_impl(ccfd_forest* f, const float* X, int64_t n, int64_t ld, double* out, uint8_t* decision, double threshold,
                                     int ge, uint32_t flags, void* stream) {
    if (!decision && n > 0) return ccfd_fail(CCFD_EINVAL, "decision is null");
    return score_impl(f, X, n, ld, out, decision, threshold, ge, flags, stream);
}

static int ccfd_predict_proba_rows_impl(ccfd_forest* f, const void* X, int32_t x_is_f64, int64_t n, int64_t ld, double* out, uint32_t flags,
                                        int32_t* has_inf) {
    if (!f) return ccfd_fail(CCFD_EINVAL, "null forest");
    if (n < 0) return ccfd_fail(CCFD_EINVAL, "n must be >= 0 (got %lld)", (long long)n);
    if (has_inf) *has_inf = 0;
    if (n == 0) return CCFD_OK;
    if (!X || !out) return ccfd_fail(CCFD_EINVAL, "X and out must not be null");
    if (ld < f->dev.n_features) return ccfd_fail(CCFD_EINVAL, "row stride %lld is smaller than n_features %d", (long long)ld, f->dev.n_features);
    if (flags & ~CCFD_KERNEL_MASK) return ccfd_fail(CCFD_EINVAL, "only a kernel selector may be given (host rows, synchronous)");
    DeviceGuard guard(f->device);
    int32_t bad = 0;
    const int rc = score_pageable(f, X, x_is_f64 ? 1 : 0, n, ld, out, nullptr, 0.5, 0, flags & CCFD_KERNEL_MASK, &bad);
    if (has_inf) *has_inf = bad;
    if (rc == CCFD_OK && bad) return ccfd_fail(CCFD_EINVAL, "Input X contains infinity or a value too large for dtype('float32').");
    return rc;
}

// ===================================================================================== session
struct ccfd_slot {
    float* hx = nullptr;   // pinned, max_batch * F
    double* hp = nullptr;  // pinned, max_batch * 2
    float* dx = nullptr;
    double* dp = nullptr;
    void* scratch = nullptr;
    cudaStream_t stream = nullptr;
    std::vector<cudaGraphExec_t> exec;  // one per bucket
    bool busy = false;
};

struct ccfd_session {
    ccfd_forest* f = nullptr;
    int64_t max_batch = 0;
    std::vector<int64_t> buckets;  // ascending capacities; a request runs the smallest that holds it
    std::vector<ccfd_slot> slots;
};

namespace {

void free_slot(ccfd_slot& s) {
    if (s.stream) cudaStreamSynchronize(s.stream);
    for (cudaGraphExec_t e : s.exec)
        if (e) cudaGraphExecDestroy(e);
    s.exec.clear();
    cudaFreeHost(s.hx);
    cudaFreeHost(s.hp);
    cudaFree(s.dx);
    cudaFree(s.dp);
    cudaFree(s.scratch);
    if (s.stream) cudaStreamDestroy(s.stream);
    s = ccfd_slot();
}

int build_slot(ccfd_session* ss, ccfd_slot& s) {
    const ccfd_forest* f = ss->f;
    const int F = f->dev.n_features;
    const size_t xb = sizeof(float) * (size_t)ss->max_batch * F, pb = sizeof(double) * 2 * (size_t)ss->max_batch;
    // requests are short: give them the highest stream priority so they overtake bulk scoring
    int prio_lo = 0, prio_hi = 0;
    CCFD_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    CCFD_CUDA(cudaStreamCreateWithPriority(&s.stream, cudaStreamNonBlocking, prio_hi));
    CCFD_CUDA(cudaHostAlloc(&s.hx, xb, cudaHostAllocDefault));
    CCFD_CUDA(cudaHostAlloc(&s.hp, pb, cudaHostAllocMapped));
    double* hp_dev = nullptr;  // the same pinned result buffer as the GPU sees it
    CCFD_CUDA(cudaHostGetDevicePointer(&hp_dev, s.hp, 0));
    memset(s.hx, 0, xb);
    CCFD_CUDA(cudaMalloc(&s.dx, xb));
    CCFD_CUDA(cudaMalloc(&s.dp, pb));
    int64_t scratch_bytes = 0;
    for (int64_t cap : ss->buckets)
        if (ccfd_pick_kernel(f, cap) == CCFD_KERNEL_SPLIT) scratch_bytes = std::max(scratch_bytes, ccfd_split_scratch_bytes(f, cap));
    if (scratch_bytes) CCFD_CUDA(cudaMalloc(&s.scratch, (size_t)scratch_bytes));

    // One graph per bucket: copy in -> score -> copy out, all addresses fixed at capture.  Small
    // buckets skip the copy-out: their last kernel writes the 16 bytes per row straight into the
    // pinned host buffer (one graph node and one DMA set-up less on the critical path).
    for (int64_t cap : ss->buckets) {
        const bool direct_out = cap <= kDirectOutRows;
        cudaGraph_t graph = nullptr;
        CCFD_CUDA(cudaStreamBeginCapture(s.stream, cudaStreamCaptureModeThreadLocal));
        cudaError_t e = cudaMemcpyAsync(s.dx, s.hx, sizeof(float) * (size_t)cap * F, cudaMemcpyHostToDevice, s.stream);
        int rc = CCFD_OK;
        if (e == cudaSuccess) {
            ScoreArgs a{s.dx, cap, F, direct_out ? hp_dev : s.dp, nullptr, 0.5, 0};
            rc = ccfd_launch_score(f, a, ccfd_pick_kernel(f, cap), s.scratch, s.stream);
        }
        if (e == cudaSuccess && rc == CCFD_OK && !direct_out)
            e = cudaMemcpyAsync(s.hp, s.dp, sizeof(double) * 2 * (size_t)cap, cudaMemcpyDeviceToHost, s.stream);
        cudaError_t e2 = cudaStreamEndCapture(s.stream, &graph);
        if (rc != CCFD_OK) {
            if (graph) cudaGraphDestroy(graph);
            return rc;
        }
        if (e != cudaSuccess) {
            if (graph) cudaGraphDestroy(graph);
            return ccfd_cuda_fail(e, "graph capture");
        }
        if (e2 != cudaSuccess) return ccfd_cuda_fail(e2, "cudaStreamEndCapture");
        cudaGraphExec_t exec = nullptr;
        e = cudaGraphInstantiate(&exec, graph, 0);
        cudaGraphDestroy(graph);
        if (e != cudaSuccess) return ccfd_cuda_fail(e, "cudaGraphInstantiate");
        s.exec.push_back(exec);
        // first launch uploads the graph; do it now so no request pays for it
        CCFD_CUDA(cudaGraphLaunch(exec, s.stream));
    }
    CCFD_CUDA(cudaStreamSynchronize(s.stream));
    return CCFD_OK;
}

}  // namespace

static int ccfd_session_create_impl(ccfd_forest* f, int64_t max_batch, int32_t n_slots, ccfd_session** out) {
    if (!f || !out) return ccfd_fail(CCFD_EINVAL, "null argument");
    *out = nullptr;
    if (max_batch < 1 || max_batch > (1 << 22)) return ccfd_fail(CCFD_EINVAL, "max_batch must be in 1..4194304 (got %lld)", (long long)max_batch);
    if (n_slots < 1 || n_slots > 64) return ccfd_fail(CCFD_EINVAL, "n_slots must be in 1..64 (got %d)", n_slots);
    ccfd_session* ss = new (std::nothrow) ccfd_session();
    if (!ss) return ccfd_fail(CCFD_ENOMEM, "out of host memory");
    ss->f = f;
    ss->max_batch = max_batch;
    for (int64_t cap = 32; cap < max_batch; cap *= 2) ss->buckets.push_back(cap);
    ss->buckets.push_back(max_batch);
    DeviceGuard guard(f->device);
    ss->slots.resize((size_t)n_slots);
    for (auto& s : ss->slots) {
        int rc = build_slot(ss, s);
        if (rc != CCFD_OK) {
            ccfd_session_destroy(ss);
            return rc;
        }
    }
    *out = ss;
    return CCFD_OK;
}

static int ccfd_session_destroy_impl(ccfd_session* ss) {
    if (!ss) return CCFD_OK;
    DeviceGuard guard(ss->f->device);
    for (auto& s : ss->slots) free_slot(s);
    delete ss;
    return CCFD_OK;
}

static int ccfd_session_buffers_impl(ccfd_session* ss, int32_t slot, float** x, double** p) {
    if (!ss) return ccfd_fail(CCFD_EINVAL, "null session");
    if (slot < 0 || slot >= (int32_t)ss->slots.size()) return ccfd_fail(CCFD_EINVAL, "slot %d out of range", slot);
    if (x) *x = ss->slots[(size_t)slot].hx;
    if (p) *p = ss->slots[(size_t)slot].hp;
    return CCFD_OK;
}

static int ccfd_session_submit_impl(ccfd_session* ss, int32_t slot, int64_t n_rows) {
    if (!ss) return ccfd_fail(CCFD_EINVAL, "null session");
    if (slot < 0 || slot >= (int32_t)ss->slots.size()) return ccfd_fail(CCFD_EINVAL, "slot %d out of range", slot);
    if (n_rows < 1 || n_rows > ss->max_batch) return ccfd_fail(CCFD_EINVAL, "n_rows must be in 1..%lld (got %lld)", (long long)ss->max_batch, (long long)n_rows);
    ccfd_slot& s = ss->slots[(size_t)slot];
    if (s.busy) return ccfd_fail(CCFD_ESTATE, "slot %d is in flight: wait before submitting again", slot);
    size_t b = 0;
    while (ss->buckets[b] < n_rows) ++b;
    DeviceGuard guard(ss->f->device);
    CCFD_CUDA(cudaGraphLaunch(s.exec[b], s.stream));
    s.busy = true;
    return CCFD_OK;
}

static int ccfd_session_wait_impl(ccfd_session* ss, int32_t slot) {
    if (!ss) return ccfd_fail(CCFD_EINVAL, "null session");
    if (slot < 0 || slot >= (int32_t)ss->slots.size()) return ccfd_fail(CCFD_EINVAL, "slot %d out of range", slot);
    ccfd_slot& s = ss->slots[(size_t)slot];
    if (!s.busy) return ccfd_fail(CCFD_ESTATE, "slot %d has nothing in flight", slot);
    s.busy = false;
    CCFD_CUDA(cudaStreamSynchronize(s.stream));
    return CCFD_OK;
}

// One request, start to finish, in one foreign-function call: narrow / copy the caller's rows into the
// slot's pinned buffer (float64 or float32, any row stride; scikit-learn's infinity check on the way;
// streaming stores for all but tiny requests, so the copy engine finds the rows in memory, not in a cache),
// launch the slot's graph, wait, copy the probabilities out.
static int ccfd_session_score_impl(ccfd_session* ss, int32_t slot, const void* X, int32_t x_is_f64, int64_t n, int64_t ld, double* out,
                                   int32_t* has_inf) {
    if (!ss) return ccfd_fail(CCFD_EINVAL, "null session");
    if (slot < 0 || slot >= (int32_t)ss->slots.size()) return ccfd_fail(CCFD_EINVAL, "slot %d out of range", slot);
    if (!X || !out) return ccfd_fail(CCFD_EINVAL, "X and out must not be null");
    if (n < 1 || n > ss->max_batch) return ccfd_fail(CCFD_EINVAL, "n must be in 1..%lld (got %lld)", (long long)ss->max_batch, (long long)n);
    const int F = ss->f->dev.n_features;
    if (ld < F) return ccfd_fail(CCFD_EINVAL, "row stride %lld is smaller than n_features %d", (long long)ld, F);
    ccfd_slot& s = ss->slots[(size_t)slot];
    if (s.busy) return ccfd_fail(CCFD_ESTATE, "slot %d is in flight: wait before submitting again", slot);
    int32_t bad = 0;
    int rc = ccfd_cast_span(X, x_is_f64, n, F, ld, s.hx, &bad, (int64_t)1 << 19, n * F >= 4096);
    if (has_inf) *has_inf = bad;
    if (rc != CCFD_OK) return rc;
    if (bad) return ccfd_fail(CCFD_EINVAL, "Input X contains infinity or a value too large for dtype('float32').");
    rc = ccfd_session_run(ss, slot, n);
    if (rc != CCFD_OK) return rc;
    memcpy(out, s.hp, sizeof(double) * 2 * (size_t)n);
    return CCFD_OK;
}

int ccfd_session_geometry(const ccfd_session* ss, int64_t* max_batch, int32_t* n_slots, int32_t* n_features) {
    if (!ss) return ccfd_fail(CCFD_EINVAL, "null session");
    if (max_batch) *max_batch = ss->max_batch;
    if (n_slots) *n_slots = (int32_t)ss->slots.size();
    if (n_features) *n_features = ss->f->dev.n_features;
    return CCFD_OK;
}

static int ccfd_session_run_impl(ccfd_session* ss, int32_t slot, int64_t n_rows) {
    int rc = ccfd_session_submit(ss, slot, n_rows);
    if (rc != CCFD_OK) return rc;
    return ccfd_session_wait(ss, slot);
}

// ---- the C ABI: every body above runs inside ccfd_guarded (no C++ exception crosses the boundary)
extern "C" int ccfd_predict_proba(ccfd_forest* f, const float* X, int64_t n, int64_t ld, double* out, uint32_t flags, void* stream) {
    return ccfd_guarded("ccfd_predict_proba", [&]() -> int { return ccfd_predict_proba_impl(f, X, n, ld, out, flags, stream); });
}
extern "C" int ccfd_predict_proba_rows(ccfd_forest* f, const void* X, int32_t x_is_f64, int64_t n, int64_t ld, double* out, uint32_t flags,
                                       int32_t* has_inf) {
    return ccfd_guarded("ccfd_predict_proba_rows", [&]() -> int { return ccfd_predict_proba_rows_impl(f, X, x_is_f64, n, ld, out, flags, has_inf); });
}
extern "C" int ccfd_predict_decision(ccfd_forest* f, const float* X, int64_t n, int64_t ld, double* out, uint8_t* decision, double threshold, int ge, uint32_t flags, void* stream) {
    return ccfd_guarded("ccfd_predict_decision", [&]() -> int { return ccfd_predict_decision_impl(f, X, n, ld, out, decision, threshold, ge, flags, stream); });
}
extern "C" int ccfd_session_create(ccfd_forest* f, int64_t max_batch, int32_t n_slots, ccfd_session** out) {
    return ccfd_guarded("ccfd_session_create", [&]() -> int { return ccfd_session_create_impl(f, max_batch, n_slots, out); });
}
extern "C" int ccfd_session_destroy(ccfd_session* ss) {
    return ccfd_guarded("ccfd_session_destroy", [&]() -> int { return ccfd_session_destroy_impl(ss); });
}
extern "C" int ccfd_session_buffers(ccfd_session* ss, int32_t slot, float** x, double** p) {
    return ccfd_guarded("ccfd_session_buffers", [&]() -> int { return ccfd_session_buffers_impl(ss, slot, x, p); });
}
extern "C" int ccfd_session_submit(ccfd_session* ss, int32_t slot, int64_t n_rows) {
    return ccfd_guarded("ccfd_session_submit", [&]() -> int { return ccfd_session_submit_impl(ss, slot, n_rows); });
}
extern "C" int ccfd_session_wait(ccfd_session* ss, int32_t slot) {
    return ccfd_guarded("ccfd_session_wait", [&]() -> int { return ccfd_session_wait_impl(ss, slot); });
}
extern "C" int ccfd_session_score(ccfd_session* ss, int32_t slot, const void* X, int32_t x_is_f64, int64_t n, int64_t ld, double* out,
                                  int32_t* has_inf) {
    return ccfd_guarded("ccfd_session_score", [&]() -> int { return ccfd_session_score_impl(ss, slot, X, x_is_f64, n, ld, out, has_inf); });
}
extern "C" int ccfd_session_run(ccfd_session* ss, int32_t slot, int64_t n_rows) {
    return ccfd_guarded("ccfd_session_run", [&]() -> int { return ccfd_session_run_impl(ss, slot, n_rows); });
}

// Which kernel CCFD_KERNEL_AUTO resolves to for a batch of n rows (one of CCFD_KERNEL_STREAM / SPLIT / GLOBAL).
extern "C" int ccfd_kernel_for(const ccfd_forest* f, int64_t n, uint32_t* kernel, int32_t* rows_presorted) {
    return ccfd_guarded("ccfd_kernel_for", [&]() -> int {
        if (!f || !kernel) return ccfd_fail(CCFD_EINVAL, "null argument");
        *kernel = ccfd_pick_kernel(f, n);
        if (rows_presorted) *rows_presorted = ccfd_row_sort_wanted(f, *kernel, n) ? 1 : 0;
        return CCFD_OK;
    });
}
