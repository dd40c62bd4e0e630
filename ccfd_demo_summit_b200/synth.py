"""Seeded synthetic Kaggle-shaped credit-card transactions.

The reference replays ``creditcard.csv`` from S3 into Kafka
(reference deploy/kafka/ProducerDeployment.yaml:88-97); the dataset itself is
not in the reference and cannot be downloaded here, so every benchmark and
parity test uses this generator.  Column order follows the Kaggle file:
``Time, V1..V28, Amount`` (30 fp32 features).  The draw order below is the
recipe's and must not change: fixtures under tests/golden depend on it.
"""
from __future__ import annotations

import numpy as np

FEATURE_NAMES = ["Time"] + [f"V{i}" for i in range(1, 29)] + ["Amount"]
N_FEATURES = 30
KAGGLE_ROWS = 284_807


def make_features(n: int = KAGGLE_ROWS, seed: int = 1, rng=None) -> np.ndarray:
    """(n, 30) float32 transactions; ``Time`` sorted like the replayed CSV."""
    rng = np.random.default_rng(seed) if rng is None else rng
    X = np.empty((n, N_FEATURES), dtype=np.float64)
    X[:, 0] = np.sort(rng.uniform(0.0, 172792.0, size=n))
    X[:, 1:29] = rng.standard_normal((n, 28)) * np.linspace(1.96, 0.33, 28)
    X[:, 29] = np.round(rng.lognormal(mean=3.15, sigma=1.65, size=n), 2)
    return X.astype(np.float32)


def make_dataset(n: int = KAGGLE_ROWS, seed: int = 1, labels: str = "hard"):
    """Features plus labels.

    ``hard`` (primary): noisy linear score with one interaction, top decile
    positive -> depth-10 trees of ~1.2-1.8k nodes.  ``easy``: Kaggle-like
    0.172 % positives shifted on a few PCA columns -> trees of tens of nodes.
    """
    rng = np.random.default_rng(seed)
    X = make_features(n, rng=rng)
    if labels == "hard":
        Z = (X - X.mean(axis=0)) / X.std(axis=0)
        w = rng.standard_normal(N_FEATURES)
        s = Z @ w + 0.8 * Z[:, 14] * Z[:, 17] + 1.5 * rng.standard_normal(n)
        y = (s > np.quantile(s, 0.9)).astype(np.int64)
    elif labels == "easy":
        y = (rng.uniform(size=n) < 0.00172).astype(np.int64)
        pos = y == 1
        for col, shift in ((14, -6.0), (17, -6.0), (10, -5.0), (12, -6.0), (4, 4.0), (11, 3.0)):
            X[pos, col] += np.float32(shift)
    else:
        raise ValueError(f"unknown labels={labels!r}")
    return X, y


def fit_forest(X, y, n_estimators=100, max_depth=10, random_state=0, n_jobs=-1, **kw):
    """The demo's model: sklearn RandomForestClassifier (BASELINE.json configs[0])."""
    from sklearn.ensemble import RandomForestClassifier

    rf = RandomForestClassifier(
        n_estimators=n_estimators, max_depth=max_depth, random_state=random_state, n_jobs=n_jobs, **kw
    )
    rf.fit(X, y)
    return rf


def tile_rows(X: np.ndarray, n: int, seed: int = 7) -> np.ndarray:
    """Grow a block to ``n`` distinct rows: tiled repeats with a small fp32 jitter."""
    reps = -(-n // X.shape[0])
    rng = np.random.default_rng(seed)
    out = np.tile(X, (reps, 1))[:n].copy()
    jitter = rng.standard_normal(out.shape, dtype=np.float32) * np.float32(1e-3)
    out += jitter * np.abs(out)
    return out
