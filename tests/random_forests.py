"""Test helper: random forests built directly as node arrays (no fitting): unbalanced shapes, single-node
trees, duplicate and +inf thresholds, thresholds that are not float32-representable, arbitrary
missing_go_to_left bits, leaf values that are count ratios or arbitrary doubles."""
import numpy as np

from oracle.rf_oracle import RawForest


def random_tree(rng, n_internal, F, skew, exotic):
    left, right, feat, thr, mgl, val = [-1], [-1], [-2], [-2.0], [0], [[0.5, 0.5]]
    leaves = [0]
    for _ in range(n_internal):
        # skew -> 1 keeps splitting the newest leaf (a vine), skew -> 0 picks uniformly (bushy)
        k = len(leaves) - 1 if rng.random() < skew else int(rng.integers(len(leaves)))
        u = leaves.pop(k)
        a, b = len(left), len(left) + 1
        for _c in range(2):
            left.append(-1); right.append(-1); feat.append(-2); thr.append(-2.0); mgl.append(0); val.append([0.5, 0.5])
        left[u], right[u] = (a, b) if rng.random() < 0.5 else (b, a)  # children are not always left-first
        feat[u] = int(rng.integers(F))
        r = rng.random()
        if exotic and r < 0.03:
            thr[u] = np.inf
        elif exotic and r < 0.10:
            thr[u] = float(rng.integers(-2, 3))  # heavy ties, exactly representable
        else:
            lo = np.float32(rng.standard_normal() * 2)
            thr[u] = (float(lo) + float(np.nextafter(lo, np.float32(np.inf)))) / 2  # midpoint: not a float32
        mgl[u] = int(rng.integers(2))
        leaves += [a, b]
    for u in leaves:
        if rng.random() < 0.7:
            n = int(rng.integers(1, 3000)); c = int(rng.integers(0, n + 1))
            val[u] = [(n - c) / n, c / n]
        else:
            p = rng.random()
            val[u] = [1.0 - p, p]
    return left, right, feat, thr, mgl, val


def random_forest(seed, n_trees=7, max_internal=200, F=30, skew=0.3, exotic=True):
    rng = np.random.default_rng(seed)
    offs, L, R, Fe, Th, M, V = [0], [], [], [], [], [], []
    for t in range(n_trees):
        n_int = 0 if (exotic and t % 5 == 4) else int(rng.integers(1, max_internal + 1))
        l, r, f, th, m, v = random_tree(rng, n_int, F, skew, exotic)
        L += l; R += r; Fe += f; Th += th; M += m; V += v
        offs.append(len(L))
    return RawForest(np.array(offs, np.int64), np.array(L, np.int64), np.array(R, np.int64), np.array(Fe, np.int64),
                     np.array(Th, np.float64), np.array(M, np.uint8), np.array(V, np.float64), F, 2)


def random_rows(seed, raw, n):
    rng = np.random.default_rng(seed + 1000)
    X = (rng.standard_normal((n, raw.n_features)) * 2).astype(np.float32)
    # plant rows exactly on thresholds and one float32 step either side
    internal = np.nonzero(raw.left != -1)[0]
    if internal.size:
        pick = rng.choice(internal, size=min(n // 2, internal.size * 3))
        t = raw.threshold[pick]
        finite = np.isfinite(t)
        with np.errstate(over="ignore"):
            t32 = t.astype(np.float32)
        t32 = np.where(t32.astype(np.float64) > t, np.nextafter(t32, np.float32(-np.inf)), t32)
        step = rng.integers(-1, 2, size=len(pick))
        v = np.where(step == 0, t32, np.where(step > 0, np.nextafter(t32, np.float32(np.inf)), np.nextafter(t32, np.float32(-np.inf))))
        rows = rng.integers(n, size=len(pick))
        ok = finite & np.isfinite(v)
        X[rows[ok], raw.feature[pick][ok]] = v[ok]
    X[rng.random(X.shape) < 0.02] = np.nan
    X[rng.random(X.shape) < 0.01] = 0.0
    X[rng.random(X.shape) < 0.005] = -0.0
    return X
