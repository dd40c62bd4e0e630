/* ORACLE (test infrastructure, never the product path).
 *
 * Plain-C restatement of scikit-learn 1.9.0's random-forest predict_proba, the
 * arithmetic behind the demo's opaque model image (reference
 * deploy/model/modelfull.json:24).  See rf_oracle.py for the full provenance
 * and the parity-pin statement (unpinned by the reference; pinned against
 * scikit-learn itself through tests/golden).
 *
 *   descent        SK/tree/_tree.pyx:977-994   (float32 x promoted to double for `<=`)
 *   leaf gather    SK/tree/_tree.pyx:939-945
 *   accumulate     SK/ensemble/_forest.py:704-717  (n_jobs=1 order: tree 0..T-1)
 *   zero / divide  SK/ensemble/_forest.py:951-962  (true divide by n_estimators)
 *
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC (see oracle/Makefile).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

int rf_oracle_predict_proba(int64_t n_trees, const int64_t* tree_offsets, const int64_t* left,
                            const int64_t* right, const int64_t* feature, const double* threshold,
                            const uint8_t* missing_go_to_left, const double* value,
                            int64_t n_features, int64_t n_classes, const float* X, int64_t n_rows,
                            double* out) {
    if (n_trees <= 0 || n_features <= 0 || n_classes <= 0 || n_rows < 0) return -1;
    memset(out, 0, sizeof(double) * (size_t)n_rows * (size_t)n_classes);
    /* tree-major, exactly like one joblib task after another with n_jobs=1 */
    for (int64_t t = 0; t < n_trees; ++t) {
        const int64_t base = tree_offsets[t];
        for (int64_t i = 0; i < n_rows; ++i) {
            const float* x = X + i * n_features;
            int64_t node = 0;
            while (left[base + node] != -1) {
                const float xv = x[feature[base + node]];
                if (isnan(xv)) {
                    node = missing_go_to_left[base + node] ? left[base + node] : right[base + node];
                } else if ((double)xv <= threshold[base + node]) {
                    node = left[base + node];
                } else {
                    node = right[base + node];
                }
            }
            const double* v = value + (base + node) * n_classes;
            double* o = out + i * n_classes;
            for (int64_t c = 0; c < n_classes; ++c) o[c] += v[c];
        }
    }
    const double T = (double)n_trees;
    for (int64_t k = 0; k < n_rows * n_classes; ++k) out[k] /= T;
    return 0;
}
