"""Development aid: a few session requests (1 and 1,024 rows) so ncu can list the latency path's kernels."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccfd_demo_summit_b200 import converter, model, synth  # noqa: E402

X, y = synth.make_dataset(seed=1, labels="hard")
rf = synth.fit_forest(X, y)
forest = model.Forest(converter.from_sklearn(rf))
s = model.Session(forest, 1024, 2)
s.x[0][:] = X[:1024]
for n in (1024, 1, 256):
    lat = []
    for i in range(300):
        t = time.perf_counter()
        s.submit(0, n)
        s.wait(0)
        lat.append((time.perf_counter() - t) * 1e6)
    lat = np.sort(lat[50:])
    print(n, "rows: submit->wait p50 %.1f us p99 %.1f us" % (lat[len(lat) // 2], lat[int(len(lat) * .99)]))

m = model.CcfdForestModel(converter.from_sklearn(rf), session_max_batch=1024)
for n, xx in ((1, X[:1]), (1024, X[:1024]), (1024, X[:1024].astype(np.float64))):
    lat = []
    for i in range(600):
        t = time.perf_counter()
        m.predict(xx, None)
        lat.append((time.perf_counter() - t) * 1e6)
    lat = np.sort(lat[100:])
    print(n, "rows", xx.dtype, ": CcfdForestModel.predict p50 %.1f us p99 %.1f us" % (lat[len(lat) // 2], lat[int(len(lat) * .99)]))
rf.n_jobs = 1
for n in (1, 1024):
    lat = []
    for i in range(60):
        t = time.perf_counter()
        rf.predict_proba(X[:n])
        lat.append((time.perf_counter() - t) * 1e6)
    lat = np.sort(lat[10:])
    print(n, "rows: sklearn predict_proba n_jobs=1 p50 %.1f us p99 %.1f us" % (lat[len(lat) // 2], lat[int(len(lat) * .99)]))
