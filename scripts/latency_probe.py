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

import json
from ccfd_demo_summit_b200 import ingest, seldon_harness as sh
app = sh.SeldonApp(m)
for n in (1, 1024):
    body = json.dumps({"data": {"names": synth.FEATURE_NAMES, "ndarray": X[:n].astype(np.float64).tolist()}}).encode()
    lat = []
    for i in range(400):
        t = time.perf_counter()
        app.handle(body)
        lat.append((time.perf_counter() - t) * 1e6)
    lat = np.sort(lat[50:])
    print(n, "rows: Seldon JSON body -> JSON response (native decode) p50 %.1f us p99 %.1f us, body %d bytes" % (lat[len(lat) // 2], lat[int(len(lat) * .99)], len(body)))
    app2 = sh.SeldonApp(m, fast_decode=False)
    lat = []
    for i in range(200):
        t = time.perf_counter()
        app2.handle(body)
        lat.append((time.perf_counter() - t) * 1e6)
    lat = np.sort(lat[20:])
    print(n, "rows: same through json.loads + numpy p50 %.1f us" % lat[len(lat) // 2])
sess = model.Session(m.forest, 1024, 4)
recs = list(ingest.KafkaShapedProducer(X[:200_000], fmt="csv"))
mb = ingest.MicroBatcher(sess, batch=1024)
t = time.perf_counter()
for i in range(0, len(recs), 500):
    mb.push_many(recs[i:i + 500])
mb.close()
dt = time.perf_counter() - t
print("Kafka-shaped CSV replay, polls of 500 records -> 1,024-row micro-batches: %.0f records/s (%d batches)" % (len(recs) / dt, mb.batches))
