// Text -> float32 rows, written straight into a caller buffer (typically a session's pinned slot).
//
// Two producers feed the model in the demo: the Camel router POSTs Seldon JSON
// ({"data":{"names":[...],"ndarray":[[...],...]}} or {"data":{"tensor":{"shape":[n,30],"values":[...]}}})
// to SELDON_URL/SELDON_ENDPOINT (reference deploy/router.yaml:65-68), and the Kafka producer replays
// creditcard.csv one row per message (reference deploy/kafka/ProducerDeployment.yaml:88-97; 31
// columns, the last being Class).  Once scoring takes microseconds, turning that text into floats
// is the request's main cost, so it is done here in one pass with no intermediate objects.
//
// Numbers are parsed with strtod (correctly rounded) and then narrowed to float32: the same two
// roundings json.loads + numpy.asarray(dtype=float32) perform, so the rows are bit-identical to what
// the Python wrapper would have produced.
#include <cctype>
#include <cerrno>
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
                memcpy(buf, p + 1, (size_t)n);
                buf[n] = 0;
                char* e = nullptr;
                v = strtod(buf, &e);
                numeric = (e == buf + n);
            }
            p = q + 1;
            if (!numeric) continue;
        } else if (number_start(p, end)) {
            int64_t n = 0;
            while (p + n < end && n < (int64_t)sizeof(buf) - 1) {
                const char d = p[n];
                if ((d >= '0' && d <= '9') || d == '+' || d == '-' || d == '.' || d == 'e' || d == 'E' || std::isalpha((unsigned char)d))
                    ++n;
                else
                    break;
            }
            memcpy(buf, p, (size_t)n);
            buf[n] = 0;
            char* e = nullptr;
            v = strtod(buf, &e);
            if (e == buf) return ccfd_fail(CCFD_EINVAL, "bad number at byte %lld", (long long)(p - text));
            p += (e - buf);
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
