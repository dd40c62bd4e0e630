// Text -> float32 rows, written straight into a caller buffer (typically a session's pinned slot).
//
// Two producers feed the model in the demo: the Camel router POSTs Seldon JSON
// ({"data":{"names":[...],"ndarray":[[...],...]}} or {"data":{"tensor":{"shape":[n,30],"values":[...]}}})
// to SELDON_URL/SELDON_ENDPOINT (reference deploy/router.yaml:65-68), and the Kafka producer replays
// creditcard.csv one row per message (reference deploy/kafka/ProducerDeployment.yaml:88-97; 31
// columns, the last being Class).  Once scoring takes microseconds, turning that text into floats
// is the request's main cost, so it is done here in one pass with no intermediate objects.
//
// Numbers are parsed with std::from_chars (correctly rounded to float64, several times faster than
// strtod) and then narrowed to float32: the same two roundings json.loads +
// numpy.asarray(dtype=float32) perform, so the rows are bit-identical to what the Python wrapper
// would have produced.
#include <cctype>
#include <cerrno>
#include <charconv>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>

#include "ccfd_internal.h"

namespace {

inline bool number_start(const char* p, const char* end) {
    const char c = *p;
    if (c >= '0' && c <= '9') return true;
    if ((c == '-' || c == '+' || c == '.') && p + 1 < end) {
        const char d = p[1];
        return (d >= '0' && d <= '9') || d == '.' || ((d == 'I' || d == 'i' || d == 'N' || d == 'n') && c != '.');
    }
    return false;
}

// Case-insensitive match of a bare word (NaN, null, Infinity) at p.
inline int word_at(const char* p, const char* end, const char* w) {
    int n = 0;
    while (w[n]) {
        if (p + n >= end || std::tolower((unsigned char)p[n]) != w[n]) return 0;
        ++n;
    }
    return n;
}

}  // namespace

extern "C" int ccfd_parse_rows(const char* text, int64_t len, int32_t cols_in, int32_t cols_keep, float* dst, int64_t cap_rows,
                               int64_t* n_rows) {
    if (!text || !dst || !n_rows) return ccfd_fail(CCFD_EINVAL, "null argument");
    if (len < 0 || cols_in < 1 || cols_keep < 1 || cols_keep > cols_in || cap_rows < 0)
        return ccfd_fail(CCFD_EINVAL, "bad sizes (len %lld, cols_in %d, cols_keep %d, cap_rows %lld)", (long long)len, cols_in, cols_keep, (long long)cap_rows);
    *n_rows = 0;
    const char* p = text;
    const char* end = text + len;
    // Seldon JSON: numbers before the payload ("shape":[n,30]) are not rows.  Start at "ndarray"
    // or "values" when present; plain CSV / JSON-lines text has neither and starts at 0.
    {
        for (const char* q = text; q + 8 <= end; ++q) {
            if (*q != '"') continue;
            if (q + 9 <= end && !memcmp(q, "\"ndarray\"", 9)) {
                p = q + 9;
                break;
            }
            if (!memcmp(q, "\"values\"", 8)) {
                p = q + 8;
                break;
            }
        }
    }
    int64_t rows = 0;
    int col = 0;
    char buf[64];
    while (p < end) {
        const char c = *p;
        double v;
        int w;
        if (c == '"') {
            // a quoted token: either a quoted number ("0" as the CSV's Class column) or a name to skip
            const char* q = p + 1;
            while (q < end && *q != '"') ++q;
            if (q >= end) return ccfd_fail(CCFD_EINVAL, "unterminated string at byte %lld", (long long)(p - text));
            const int64_t n = q - (p + 1);
            bool numeric = n > 0 && n < (int64_t)sizeof(buf) && number_start(p + 1, q + 1);
            if (numeric) {
                const char* b = p + 1 + (p[1] == '+' ? 1 : 0);
                const auto res = std::from_chars(b, q, v);
                numeric = res.ec == std::errc() && res.ptr == q;
            }
            p = q + 1;
            if (!numeric) continue;
        } else if (number_start(p, end)) {
            const char* b = p + (*p == '+' ? 1 : 0);
            const auto res = std::from_chars(b, end, v);
            if (res.ec == std::errc::result_out_of_range) {
                // beyond float64: strtod's +-inf / +-0 is what json.loads would have produced
                int64_t n = 0;
                while (b + n < end && n < (int64_t)sizeof(buf) - 1 && (std::isalnum((unsigned char)b[n]) || b[n] == '+' || b[n] == '-' || b[n] == '.')) ++n;
                memcpy(buf, b, (size_t)n);
                buf[n] = 0;
                char* e = nullptr;
                v = strtod(buf, &e);
                if (e == buf) return ccfd_fail(CCFD_EINVAL, "bad number at byte %lld", (long long)(p - text));
                p = b + (e - buf);
            } else if (res.ec != std::errc()) {
                return ccfd_fail(CCFD_EINVAL, "bad number at byte %lld", (long long)(p - text));
            } else {
                p = res.ptr;
            }
        } else if ((w = word_at(p, end, "nan")) || (w = word_at(p, end, "null"))) {
            v = std::numeric_limits<double>::quiet_NaN();
            p += w;
        } else if ((w = word_at(p, end, "infinity")) || (w = word_at(p, end, "inf"))) {
            v = std::numeric_limits<double>::infinity();
            p += w;
        } else if (std::isalpha((unsigned char)c)) {
            // a bare word that is not a number (CSV header such as Time,V1,...): skip the identifier
            while (p < end && (std::isalnum((unsigned char)*p) || *p == '_')) ++p;
            continue;
        } else {
            ++p;
            continue;
        }
        if (rows >= cap_rows) return ccfd_fail(CCFD_EINVAL, "more than %lld rows in the payload", (long long)cap_rows);
        if (col < cols_keep) dst[rows * cols_keep + col] = (float)v;  // double -> float32, round to nearest even
        if (++col == cols_in) {
            col = 0;
            ++rows;
        }
    }
    if (col != 0) return ccfd_fail(CCFD_EINVAL, "payload ends in the middle of a row (%d of %d columns)", col, cols_in);
    *n_rows = rows;
    return CCFD_OK;
}

// Probabilities -> the "ndarray" value of a Seldon response, "[[p0,p1],[p0,p1],...]", with the
// shortest decimal strings that read back to exactly the same float64 (std::to_chars), which is the
// property json.dumps(repr) has.  Returns the length written (no terminating zero is counted).
extern "C" int ccfd_format_proba(const double* p, int64_t n_rows, char* dst, int64_t cap, int64_t* len) {
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
}

// float64 (or float32) request matrix -> dense float32 rows + the "contains infinity" verdict, in
// one multi-threaded pass.  Replaces the cast and the finiteness scan of
// ForestClassifier._validate_X_predict (SK/ensemble/_forest.py:606-625) for the common case of a
// C-contiguous matrix as the Seldon wrapper hands it over (float64 after JSON decoding).  The
// narrowing is the C cast (round to nearest even, overflow -> inf), bit-identical to numpy's astype.
#include <thread>

extern "C" int ccfd_cast_rows(const void* src, int32_t src_is_f64, int64_t n, int64_t n_cols, int64_t src_ld, float* dst,
                              int32_t* has_inf) {
    if (!src || !has_inf || n < 0 || n_cols < 1 || src_ld < n_cols) return ccfd_fail(CCFD_EINVAL, "bad argument");
    if (!dst && src_is_f64) return ccfd_fail(CCFD_EINVAL, "dst may be null (scan only) for float32 input only");
    unsigned hw = std::thread::hardware_concurrency();
    int nt = (int)std::min<int64_t>(std::max(1u, std::min(hw, 16u)), std::max<int64_t>(1, n * n_cols / 65536));
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
    auto span = [&](int64_t r, int64_t m) {  // m contiguous source values starting at row r
        if (!dst) return scan_f32(static_cast<const float*>(src) + r * src_ld, m);
        if (src_is_f64) return cast_f64(static_cast<const double*>(src) + r * src_ld, dst + r * n_cols, m);
        return copy_f32(static_cast<const float*>(src) + r * src_ld, dst + r * n_cols, m);
    };
    auto work = [&](int k) {
        const int64_t r0 = n * k / nt, r1 = n * (k + 1) / nt;
        int bad = 0;
        if (src_ld == n_cols) {
            bad = span(r0, (r1 - r0) * n_cols);  // dense rows: one span
        } else {
            for (int64_t r = r0; r < r1; ++r) bad |= span(r, n_cols);
        }
        inf_seen[(size_t)k] = bad;
    };
    if (nt == 1) {
        work(0);
    } else {
        std::vector<std::thread> th;
        for (int k = 0; k < nt; ++k) th.emplace_back(work, k);
        for (auto& t : th) t.join();
    }
    *has_inf = 0;
    for (int b : inf_seen) *has_inf |= b;
    return CCFD_OK;
}
