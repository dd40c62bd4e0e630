// Text -> float32 rows, written straight into a caller buffer (typically a session's pinned slot),
// the response encoder, the input cast, and the host worker pool they share.
//
// Two producers feed the model in the demo: the Camel router POSTs Seldon JSON
// ({"data":{"names":[...],"ndarray":[[...],...]}} or {"data":{"tensor":{"shape":[n,30],"values":[...]}}})
// to SELDON_URL/SELDON_ENDPOINT (reference deploy/router.yaml:65-68), and the Kafka producer replays
// creditcard.csv one row per message (reference deploy/kafka/ProducerDeployment.yaml:88-97; 31
// columns, the last being Class).  Once scoring takes microseconds, turning that text into floats
// is the request's main cost, so it is done here in one pass with no intermediate objects.
//
// The decoder is STRUCTURE-AWARE: a row is what the payload says is a row (an innermost [...] or
// {...}, a line of CSV, shape[1] values of a tensor), and a row with any other number of columns
// than the model expects, a non-numeric cell, a 1-D ndarray or a tensor whose shape disagrees with
// its values is an error -- the Seldon wrapper + scikit-learn answer those with a ValueError (HTTP
// 400), and so must this path, instead of re-cutting the numbers every n_features and scoring
// misaligned columns.
//
// Numbers are parsed with std::from_chars (correctly rounded to float64, several times faster than
// strtod) and then narrowed to float32: the same two roundings json.loads +
// numpy.asarray(dtype=float32) perform, so the rows are bit-identical to what the Python wrapper
// would have produced.
#include <atomic>
#include <cctype>
#include <cerrno>
#include <charconv>
#include <cmath>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <thread>

#include "ccfd_internal.h"

// ============================================================================ host worker pool
namespace {

class HostPool {
  public:
    static HostPool& get() {
        static HostPool* p = new HostPool();  // never destroyed: workers may outlive static destructors
        return *p;
    }
    int threads() const { return (int)workers_ + 1; }

    void run(int n_parts, const std::function<void(int)>& fn) {
        if (n_parts <= 1 || workers_ == 0 || !busy_.try_lock()) {  // small job, no pool, or pool taken: do it here
            for (int i = 0; i < n_parts; ++i) fn(i);
            return;
        }
        std::exception_ptr err;
        {
            std::unique_lock<std::mutex> lk(mu_);
            fn_ = &fn;
            n_parts_ = n_parts;
            pending_ = n_parts;
            err_ = nullptr;
            // release: a worker still leaving the previous job's loop may pick a part of this one as soon as it
            // sees the counter reset, and must then see this job's function and size too
            next_.store(0, std::memory_order_release);
            ++generation_;
        }
        cv_work_.notify_all();
        work_loop();  // the caller takes parts too
        {
            std::unique_lock<std::mutex> lk(mu_);
            cv_done_.wait(lk, [&] { return pending_ == 0; });
            fn_ = nullptr;
            err = err_;
        }
        busy_.unlock();
        if (err) std::rethrow_exception(err);
    }

  private:
    HostPool() {
        unsigned hw = std::thread::hardware_concurrency();
        unsigned want = hw > 1 ? std::min(hw, 32u) - 1 : 0;
        if (const char* e = getenv("CCFD_HOST_THREADS")) want = (unsigned)std::max(0, atoi(e) - 1);
        for (unsigned i = 0; i < want; ++i) {
            try {
                std::thread([this] { worker(); }).detach();
                ++workers_;
            } catch (...) {
                break;  // fewer threads (or none): the pool still works
            }
        }
    }
    void work_loop() {
        for (;;) {
            const int i = next_.fetch_add(1, std::memory_order_acq_rel);
            if (i >= n_parts_) return;
            try {
                (*fn_)(i);
            } catch (...) {
                std::lock_guard<std::mutex> lk(mu_);
                if (!err_) err_ = std::current_exception();
            }
            std::lock_guard<std::mutex> lk(mu_);
            if (--pending_ == 0) cv_done_.notify_all();
        }
    }
    void worker() {
        uint64_t seen = 0;
        for (;;) {
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_work_.wait(lk, [&] { return generation_ != seen; });
                seen = generation_;
            }
            work_loop();
        }
    }
    std::mutex mu_, busy_;
    std::condition_variable cv_work_, cv_done_;
    const std::function<void(int)>* fn_ = nullptr;
    int n_parts_ = 0, pending_ = 0;
    std::atomic<int> next_{0};
    uint64_t generation_ = 0;
    std::exception_ptr err_;
    unsigned workers_ = 0;
};

}  // namespace

void ccfd_parallel_for(int n_parts, const std::function<void(int)>& fn) { HostPool::get().run(n_parts, fn); }
int ccfd_host_threads() { return HostPool::get().threads(); }

// ================================================================================== the decoder
namespace {

inline bool ieq(const char* b, const char* e, const char* w) {  // case-insensitive whole-token compare (w lower case)
    size_t n = (size_t)(e - b);
    if (strlen(w) != n) return false;
    for (size_t i = 0; i < n; ++i)
        if (std::tolower((unsigned char)b[i]) != w[i]) return false;
    return true;
}

// One cell, WHOLE token: 1 = number (value in *v), 0 = a non-numeric word, -1 = a cell that is neither
// (true / false).  "Info" or "Nanos" are words, not inf / nan followed by junk.
inline int classify(const char* b, const char* e, double* v) {
    if (b == e) return 0;
    const char* s = b + (*b == '+' ? 1 : 0);
    if (s == e) return 0;
    const auto res = std::from_chars(s, e, *v);  // also takes inf / infinity / nan, any case, with a sign
    if (res.ptr == e && res.ec == std::errc()) return 1;
    if (res.ptr == e && res.ec == std::errc::result_out_of_range) {
        // beyond float64: strtod's +-inf / +-0 is what json.loads would have produced
        char buf[64];
        const size_t n = (size_t)(e - s);
        if (n >= sizeof(buf)) return 0;
        memcpy(buf, s, n);
        buf[n] = 0;
        char* end = nullptr;
        *v = strtod(buf, &end);
        return end == buf + n ? 1 : 0;
    }
    if (ieq(b, e, "null")) {
        *v = std::numeric_limits<double>::quiet_NaN();
        return 1;
    }
    if (ieq(b, e, "true") || ieq(b, e, "false")) return -1;
    return 0;
}

inline bool is_space(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n'; }
inline bool is_delim(char c) { return c == ',' || c == '[' || c == ']' || c == '{' || c == '}' || c == ':' || c == '"' || c == ';' || is_space(c); }

struct RowSink {
    float* dst;
    int64_t cap_rows;
    int32_t cols_in, cols_keep;
    int64_t rows = 0;
    int col = 0;    // numbers seen in the open row
    int words = 0;  // non-numeric cells seen in the open row
    const char* text;
    int err = CCFD_OK;

    int fail_at(const char* p, const char* what) {
        err = ccfd_fail(CCFD_EINVAL, "%s at byte %lld (row %lld)", what, (long long)(p - text), (long long)rows);
        return err;
    }
    int number(const char* p, double v) {
        if (rows >= cap_rows) {
            err = ccfd_fail(CCFD_EINVAL, "more than %lld rows in the payload", (long long)cap_rows);
            return err;
        }
        if (col >= cols_in) {
            err = ccfd_fail(CCFD_EINVAL, "row %lld has more than %d columns (byte %lld)", (long long)rows, cols_in, (long long)(p - text));
            return err;
        }
        if (col < cols_keep) dst[rows * cols_keep + col] = (float)v;  // double -> float32, round to nearest even
        ++col;
        return CCFD_OK;
    }
    // A row boundary.  header_ok: a line of words only (a CSV header) is skipped rather than refused.
    int close_row(const char* p, bool header_ok) {
        if (col == 0) {
            if (words && !header_ok) return fail_at(p, "non-numeric cell in the payload");
            words = 0;
            return CCFD_OK;
        }
        if (words) return fail_at(p, "non-numeric cell in a row of numbers");
        if (col != cols_in) {
            err = ccfd_fail(CCFD_EINVAL, "row %lld has %d columns, the model expects %d (byte %lld): payload is ragged or ends in the middle of a row",
                            (long long)rows, col, cols_in, (long long)(p - text));
            return err;
        }
        col = 0;
        ++rows;
        return CCFD_OK;
    }
};

// Bracket-aware scan of [b, e).  Rows are closed by the `]` / `}` of an innermost container, or, outside
// every container, by a newline (CSV, JSON lines).  flat_cut > 0: numbers of ONE flat array are cut into
// rows of cols_in (the "values" of a tensor payload).  stop_at_close: return at the bracket that closes
// the first container opened (so what follows a payload, e.g. "meta", is never read).  min_depth: rows
// must sit at least this deep (2 for "ndarray": a 1-D ndarray is not a matrix).
int scan(RowSink& s, const char* b, const char* e, bool flat_cut, bool stop_at_close, int min_depth, const char** stopped, int* end_depth) {
    const char* p = b;
    int depth = 0;
    bool opened = false;
    while (p < e) {
        const char c = *p;
        if (c == '"') {
            const char* q = p + 1;
            while (q < e && *q != '"') q += (*q == '\\' && q + 1 < e) ? 2 : 1;
            if (q >= e) return s.fail_at(p, "unterminated string");
            const char* after = q + 1;
            while (after < e && (*after == ' ' || *after == '\t')) ++after;
            if (after < e && *after == ':') {  // an object key: not a cell
                p = after + 1;
                continue;
            }
            double v;
            const int k = classify(p + 1, q, &v);
            if (k == 1) {
                if (s.number(p, v) != CCFD_OK) return s.err;
            } else if (k == 0) {
                ++s.words;
            } else {
                return s.fail_at(p, "boolean cell");
            }
            p = q + 1;
        } else if (c == '[' || c == '{') {
            if (s.col || s.words) {
                if (depth == 0) {  // "1,2,[3" on a CSV line
                    return s.fail_at(p, "array inside a row");
                }
                return s.fail_at(p, "nested array inside a row");
            }
            ++depth;
            opened = true;
            ++p;
        } else if (c == ']' || c == '}') {
            if (depth == 0) return s.fail_at(p, "unbalanced bracket");
            if (!flat_cut) {
                if (s.col && depth < min_depth) return s.fail_at(p, "expected a matrix (rows of numbers), got a flat array");
                if (s.close_row(p, false) != CCFD_OK) return s.err;
            }
            --depth;
            ++p;
            if (stop_at_close && opened && depth == 0) break;
        } else if (c == '\n') {
            if (depth == 0 && s.close_row(p, true) != CCFD_OK) return s.err;
            ++p;
        } else if (is_delim(c)) {
            ++p;
        } else {
            const char* q = p;
            while (q < e && !is_delim(*q)) ++q;
            double v;
            const int k = classify(p, q, &v);
            if (k == 1) {
                if (flat_cut && s.col == s.cols_in && s.close_row(p, false) != CCFD_OK) return s.err;
                if (s.number(p, v) != CCFD_OK) return s.err;
            } else if (k == 0) {
                ++s.words;
            } else {
                return s.fail_at(p, "boolean cell");
            }
            p = q;
        }
    }
    if (flat_cut) {
        if (s.words) return s.fail_at(p, "non-numeric cell in the payload");
        if (s.col && s.close_row(p, false) != CCFD_OK) return s.err;
    } else if (depth == 0 && p >= e) {
        if (s.close_row(p, true) != CCFD_OK) return s.err;  // last line without a newline
    }
    if (stopped) *stopped = p;
    if (end_depth) *end_depth = depth;
    return CCFD_OK;
}

// `"key"` followed by a colon, searched in [b, e); returns the first byte after the colon or null.
const char* find_key(const char* b, const char* e, const char* key) {
    const size_t n = strlen(key);
    for (const char* q = b; q + n + 2 <= e; ++q) {
        if (*q != '"' || memcmp(q + 1, key, n) != 0 || q[n + 1] != '"') continue;
        const char* a = q + n + 2;
        while (a < e && is_space(*a)) ++a;
        if (a < e && *a == ':') return a + 1;
    }
    return nullptr;
}

}  // namespace

extern "C" int ccfd_parse_rows(const char* text, int64_t len, int32_t cols_in, int32_t cols_keep, float* dst, int64_t cap_rows,
                               int64_t* n_rows) {
    return ccfd_guarded("ccfd_parse_rows", [&]() -> int {
        if (!text || !dst || !n_rows) return ccfd_fail(CCFD_EINVAL, "null argument");
        if (len < 0 || cols_in < 1 || cols_keep < 1 || cols_keep > cols_in || cap_rows < 0)
            return ccfd_fail(CCFD_EINVAL, "bad sizes (len %lld, cols_in %d, cols_keep %d, cap_rows %lld)", (long long)len, cols_in, cols_keep, (long long)cap_rows);
        *n_rows = 0;
        const char* end = text + len;
        RowSink sink{dst, cap_rows, cols_in, cols_keep};
        sink.text = text;

        // ---- Seldon "ndarray": a matrix, rows are its inner arrays
        if (const char* v = find_key(text, end, "ndarray")) {
            int rc = scan(sink, v, end, false, true, 2, nullptr, nullptr);
            if (rc != CCFD_OK) return rc;
            *n_rows = sink.rows;
            return CCFD_OK;
        }
        // ---- Seldon "tensor": {"shape":[n, F], "values":[flat]} in either order
        if (const char* t = find_key(text, end, "tensor")) {
            const char* sh = find_key(t, end, "shape");
            const char* va = find_key(t, end, "values");
            if (!sh || !va) return ccfd_fail(CCFD_EINVAL, "tensor payload needs both \"shape\" and \"values\"");
            int64_t shape[4];
            int nd = 0;
            {
                const char* p = sh;
                while (p < end && is_space(*p)) ++p;
                if (p >= end || *p != '[') return ccfd_fail(CCFD_EINVAL, "tensor shape is not an array");
                ++p;
                while (p < end && *p != ']') {
                    if (is_space(*p) || *p == ',') {
                        ++p;
                        continue;
                    }
                    int64_t d = 0;
                    const auto r = std::from_chars(p, end, d);
                    if (r.ec != std::errc() || d < 0 || nd >= 4) return ccfd_fail(CCFD_EINVAL, "bad tensor shape");
                    shape[nd++] = d;
                    p = r.ptr;
                }
            }
            if (nd != 2) return ccfd_fail(CCFD_EINVAL, "tensor shape has %d dimensions, expected 2", nd);
            if (shape[1] != cols_in) return ccfd_fail(CCFD_EINVAL, "tensor shape says %lld columns, the model expects %d", (long long)shape[1], cols_in);
            int rc = scan(sink, va, end, true, true, 1, nullptr, nullptr);
            if (rc != CCFD_OK) return rc;
            if (sink.rows != shape[0]) return ccfd_fail(CCFD_EINVAL, "tensor shape says %lld rows, values hold %lld", (long long)shape[0], (long long)sink.rows);
            *n_rows = sink.rows;
            return CCFD_OK;
        }

        // ---- records: CSV lines, JSON lines, or bare nested arrays.  Large newline-delimited inputs (a
        // consumer poll of many messages) are cut at line boundaries and parsed by the host pool; anything
        // that turns out not to be line-structured falls back to the serial scan below.
        constexpr int64_t kParallelBytes = 256 * 1024;
        const int threads = ccfd_host_threads();
        if (len >= kParallelBytes && threads > 1) {
            const int parts = (int)std::min<int64_t>(threads * 2, len / (kParallelBytes / 4));
            std::vector<const char*> cut((size_t)parts + 1);
            cut[0] = text;
            for (int k = 1; k < parts; ++k) {
                const char* q = text + len * k / parts;
                q = static_cast<const char*>(memchr(q, '\n', (size_t)(end - q)));
                cut[(size_t)k] = q ? q + 1 : end;
                if (cut[(size_t)k] < cut[(size_t)k - 1]) cut[(size_t)k] = cut[(size_t)k - 1];
            }
            cut[(size_t)parts] = end;
            std::vector<int64_t> lines((size_t)parts), base((size_t)parts + 1, 0), got((size_t)parts, 0);
            std::vector<int> ok((size_t)parts, 0);
            ccfd_parallel_for(parts, [&](int k) {
                const char* const hi = cut[(size_t)k + 1];
                int64_t n = (cut[(size_t)k] < hi && hi[-1] != '\n') ? 1 : 0;  // a last line without a newline
                for (const char* q = cut[(size_t)k]; q < hi;) {
                    q = static_cast<const char*>(memchr(q, '\n', (size_t)(hi - q)));
                    if (!q) break;
                    ++n;
                    ++q;
                }
                lines[(size_t)k] = n;
            });
            for (int k = 0; k < parts; ++k) base[(size_t)k + 1] = base[(size_t)k] + lines[(size_t)k];
            if (base[(size_t)parts] <= cap_rows) {
                ccfd_parallel_for(parts, [&](int k) {
                    RowSink s{dst + base[(size_t)k] * cols_keep, lines[(size_t)k], cols_in, cols_keep};
                    s.text = text;
                    int depth = 0;
                    const int rc = scan(s, cut[(size_t)k], cut[(size_t)k + 1], false, false, 1, nullptr, &depth);
                    ok[(size_t)k] = (rc == CCFD_OK && depth == 0) ? 1 : 0;
                    got[(size_t)k] = s.rows;
                });
                bool all = true;
                for (int k = 0; k < parts; ++k) all = all && ok[(size_t)k];
                if (all) {
                    int64_t rows = 0;
                    for (int k = 0; k < parts; ++k) {  // close the gaps blank / header lines left
                        if (rows != base[(size_t)k] && got[(size_t)k])
                            memmove(dst + rows * cols_keep, dst + base[(size_t)k] * cols_keep, sizeof(float) * (size_t)got[(size_t)k] * (size_t)cols_keep);
                        rows += got[(size_t)k];
                    }
                    *n_rows = rows;
                    return CCFD_OK;
                }
            }
        }
        int depth = 0;
        int rc = scan(sink, text, end, false, false, 1, nullptr, &depth);
        if (rc != CCFD_OK) return rc;
        if (depth != 0) return ccfd_fail(CCFD_EINVAL, "payload ends inside an array (unbalanced brackets)");
        *n_rows = sink.rows;
        return CCFD_OK;
    });
}

// Probabilities -> the "ndarray" value of a Seldon response, "[[p0,p1],[p0,p1],...]", with the
// shortest decimal strings that read back to exactly the same float64 (std::to_chars), which is the
// property json.dumps(repr) has.  Returns the length written (no terminating zero is counted).
extern "C" int ccfd_format_proba(const double* p, int64_t n_rows, char* dst, int64_t cap, int64_t* len) {
    return ccfd_guarded("ccfd_format_proba", [&]() -> int {
        if (!p || !dst || !len || n_rows < 0) return ccfd_fail(CCFD_EINVAL, "null argument");
        char* q = dst;
        char* const end = dst + cap;
        auto need = [&](int64_t k) { return q + k <= end; };
        if (!need(2)) return ccfd_fail(CCFD_EINVAL, "output buffer too small");
        *q++ = '[';
        for (int64_t i = 0; i < n_rows; ++i) {
            if (!need(2 * 32 + 4)) return ccfd_fail(CCFD_EINVAL, "output buffer too small (need up to 68 bytes per row)");
            if (i) *q++ = ',';
            *q++ = '[';
            for (int c = 0; c < 2; ++c) {
                const double v = p[2 * i + c];
                if (!(v == v) || v - v != 0.0) return ccfd_fail(CCFD_EINVAL, "row %lld holds a non-finite probability", (long long)i);
                if (c) *q++ = ',';
                const auto res = std::to_chars(q, end, v);
                if (res.ec != std::errc()) return ccfd_fail(CCFD_EINVAL, "output buffer too small");
                // JSON wants a digit on both sides of the exponent-less form too; to_chars already gives "0.25", "1", "1e-05"
                q = res.ptr;
            }
            *q++ = ']';
        }
        if (!need(1)) return ccfd_fail(CCFD_EINVAL, "output buffer too small");
        *q++ = ']';
        *len = q - dst;
        return CCFD_OK;
    });
}

// float64 (or float32) request matrix -> dense float32 rows + the "contains infinity" verdict, in
// one pass.  Replaces the cast and the finiteness scan of
// ForestClassifier._validate_X_predict (SK/ensemble/_forest.py:606-625) for the common case of a
// C-contiguous matrix as the Seldon wrapper hands it over (float64 after JSON decoding).  The
// narrowing is the C cast (round to nearest even, overflow -> inf), bit-identical to numpy's astype.
// Requests below ~1 M values (every latency-path request) are done on the calling thread: waking other
// threads costs more than the pass; larger ones go to the persistent host pool (no thread is created
// per call).
#if defined(__SSE2__)
#include <emmintrin.h>
namespace {
// Streaming (non-temporal) variants for destinations a DMA engine reads next (the pinned staging of the
// pageable-host path): ordinary stores leave the lines dirty in the CPU caches and the copy engine then pulls
// them out of there at a quarter of its speed (measured on the B200 box: 34 MB host->device in 2.5 ms after
// ordinary stores).  dst must be 16-byte aligned and m a multiple of 4; the tail is the caller's.
inline int stream_copy_f32(const float* sp, float* d, int64_t m) {
    const __m128i absmask = _mm_set1_epi32(0x7FFFFFFF), infbits = _mm_set1_epi32(0x7F800000);
    __m128i bad = _mm_setzero_si128();
    for (int64_t i = 0; i < m; i += 4) {
        const __m128 v = _mm_loadu_ps(sp + i);
        bad = _mm_or_si128(bad, _mm_cmpeq_epi32(_mm_and_si128(_mm_castps_si128(v), absmask), infbits));
        _mm_stream_ps(d + i, v);
    }
    return _mm_movemask_epi8(bad) != 0;
}
inline int stream_cast_f64(const double* sp, float* d, int64_t m) {
    const __m128i absmask = _mm_set1_epi32(0x7FFFFFFF), infbits = _mm_set1_epi32(0x7F800000);
    __m128i bad = _mm_setzero_si128();
    for (int64_t i = 0; i < m; i += 4) {
        const __m128 lo = _mm_cvtpd_ps(_mm_loadu_pd(sp + i)), hi = _mm_cvtpd_ps(_mm_loadu_pd(sp + i + 2));  // round to nearest even
        const __m128 v = _mm_movelh_ps(lo, hi);
        bad = _mm_or_si128(bad, _mm_cmpeq_epi32(_mm_and_si128(_mm_castps_si128(v), absmask), infbits));
        _mm_stream_ps(d + i, v);
    }
    return _mm_movemask_epi8(bad) != 0;
}
}  // namespace
#define CCFD_HAVE_STREAM_STORES 1
#else
#define CCFD_HAVE_STREAM_STORES 0
#endif

int ccfd_cast_span(const void* src, int32_t src_is_f64, int64_t n, int64_t n_cols, int64_t src_ld, float* dst, int32_t* has_inf,
                   int64_t values_per_part, bool streaming_stores) {
    const int nt = (int)std::max<int64_t>(1, std::min<int64_t>(ccfd_host_threads(), n * n_cols / std::max<int64_t>(values_per_part, 1)));
    std::vector<int> inf_seen((size_t)nt, 0);
    // spans of contiguous values; +-inf is "exponent all ones, mantissa zero" (NaN is allowed): an integer
    // test on the bits, which the host compiler vectorises
    auto is_inf = [](float v) {
        uint32_t u;
        memcpy(&u, &v, 4);
        return (int)((u & 0x7FFFFFFFu) == 0x7F800000u);
    };
    auto scan_f32 = [&](const float* sp, int64_t m) {
        int bad = 0;
        for (int64_t i = 0; i < m; ++i) bad |= is_inf(sp[i]);
        return bad;
    };
    auto copy_f32 = [&](const float* sp, float* d, int64_t m) {
        int bad = 0;
        for (int64_t i = 0; i < m; ++i) {
            d[i] = sp[i];
            bad |= is_inf(sp[i]);
        }
        return bad;
    };
    auto cast_f64 = [&](const double* sp, float* d, int64_t m) {
        int bad = 0;
        for (int64_t i = 0; i < m; ++i) {
            const float v = (float)sp[i];
            d[i] = v;
            bad |= is_inf(v);
        }
        return bad;
    };
    const bool stream = CCFD_HAVE_STREAM_STORES && streaming_stores && dst && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0;
    auto span = [&](int64_t r, int64_t m) {  // m contiguous source values starting at row r
        if (!dst) return scan_f32(static_cast<const float*>(src) + r * src_ld, m);
        float* d = dst + r * n_cols;
#if CCFD_HAVE_STREAM_STORES
        if (stream) {
            // head up to a 16-byte boundary of dst and tail below four values with ordinary stores
            int64_t head = (int64_t)((16 - (reinterpret_cast<uintptr_t>(d) & 15u)) & 15u) / 4;
            if (head > m) head = m;
            const int64_t body = (m - head) / 4 * 4;
            int bad = 0;
            if (src_is_f64) {
                const double* sp = static_cast<const double*>(src) + r * src_ld;
                bad |= cast_f64(sp, d, head) | stream_cast_f64(sp + head, d + head, body) | cast_f64(sp + head + body, d + head + body, m - head - body);
            } else {
                const float* sp = static_cast<const float*>(src) + r * src_ld;
                bad |= copy_f32(sp, d, head) | stream_copy_f32(sp + head, d + head, body) | copy_f32(sp + head + body, d + head + body, m - head - body);
            }
            return bad;
        }
#endif
        if (src_is_f64) return cast_f64(static_cast<const double*>(src) + r * src_ld, d, m);
        return copy_f32(static_cast<const float*>(src) + r * src_ld, d, m);
    };
    auto work = [&](int k) {
        const int64_t r0 = n * k / nt, r1 = n * (k + 1) / nt;
        int bad = 0;
        if (src_ld == n_cols) {
            bad = span(r0, (r1 - r0) * n_cols);  // dense rows: one span
        } else {
            for (int64_t r = r0; r < r1; ++r) bad |= span(r, n_cols);
        }
#if CCFD_HAVE_STREAM_STORES
        if (stream) _mm_sfence();  // streaming stores are globally visible before the part is reported done
#endif
        inf_seen[(size_t)k] = bad;
    };
    if (nt == 1)
        work(0);
    else
        ccfd_parallel_for(nt, work);
    *has_inf = 0;
    for (int b : inf_seen) *has_inf |= b;
    return CCFD_OK;
}

extern "C" int ccfd_cast_rows(const void* src, int32_t src_is_f64, int64_t n, int64_t n_cols, int64_t src_ld, float* dst,
                              int32_t* has_inf) {
    return ccfd_guarded("ccfd_cast_rows", [&]() -> int {
        if (!src || !has_inf || n < 0 || n_cols < 1 || src_ld < n_cols) return ccfd_fail(CCFD_EINVAL, "bad argument");
        if (!dst && src_is_f64) return ccfd_fail(CCFD_EINVAL, "dst may be null (scan only) for float32 input only");
        // a request below ~1 M values is done on the calling thread (waking the pool costs more than the pass)
        return ccfd_cast_span(src, src_is_f64, n, n_cols, src_ld, dst, has_inf, (int64_t)1 << 19, getenv("CCFD_CAST_STREAM") != nullptr /* test hook: the DMA-staging variant */);
    });
}
