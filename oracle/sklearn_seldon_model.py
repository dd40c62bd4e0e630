"""ORACLE / CPU comparator (test infrastructure and the timed CPU baseline only).

A Seldon-shaped model class around the *unmodified* scikit-learn forest: what
the opaque image ``nakfour/modelfull`` (reference deploy/model/modelfull.json:24)
is understood to run behind ``predict(X, feature_names)``.  bench.py times this
class on the host cores (``--impl reference`` and the ``cpu_baseline`` leg); the
tests compare the CUDA model class against it.  The product never imports it.
"""
from __future__ import annotations

import os

import numpy as np


class SklearnSeldonModel:
    class_names = ["proba_0", "proba_1"]

    def __init__(self, rf, n_jobs=None):
        self.rf = rf
        self.rf.n_jobs = os.cpu_count() if n_jobs is None else n_jobs

    @property
    def n_jobs(self):
        return self.rf.n_jobs

    def predict(self, X, feature_names=None, meta=None):
        # SK/ensemble/_forest.py:921-967
        return self.rf.predict_proba(np.asarray(X))

    def decide(self, X, threshold=0.5):
        """Router rule (reference deploy/router.yaml:69-70): fraud iff p_fraud > threshold."""
        return self.predict(X)[:, 1] > threshold
