"""Request-path metrics without a lock on the request path.

The demo scrapes the model service's ``/prometheus`` endpoint (reference README.md:292-301): the
ModelPrediction board plots the gauges ``proba_1``, ``Amount``, ``V17``, ``V10`` (reference
deploy/grafana/ModelPrediction.json:96,104,211,322) and the SeldonCore board plots request rates and
latency percentiles from ``seldon_api_engine_server_requests_seconds_*`` and
``seldon_api_engine_client_requests_seconds_bucket`` (reference deploy/grafana/SeldonCore.json:119-531).
Those two families are produced by the Seldon engine in the real deployment; the stand-in harness
emits them under the same names and labels so the shipped panels have something to plot, next to the
model's own counters (requests, rows, batch-size histogram, scoring-time histogram).

Every thread adds into its own cells (a ``threading.local``); a scrape sums the cells of all threads.
Nothing is locked while a request is served; the registry lock is taken once per thread, when that
thread's cells are created.
"""
from __future__ import annotations

import bisect
import threading

# seconds; the latency path lives in the tens of microseconds, bulk requests in milliseconds
LATENCY_BUCKETS = (25e-6, 50e-6, 100e-6, 250e-6, 500e-6, 1e-3, 2.5e-3, 5e-3, 10e-3, 25e-3, 50e-3, 100e-3, 250e-3, 1.0, 5.0)
BATCH_BUCKETS = (1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 4096, 16384, 65536, 1 << 20)


def _labels(labels: dict) -> str:
    if not labels:
        return ""
    return "{" + ",".join(f'{k}="{v}"' for k, v in sorted(labels.items())) + "}"


class Registry:
    def __init__(self):
        self._tls = threading.local()
        self._cells = []  # one dict per thread that has ever observed something
        self._lock = threading.Lock()
        self._meta = {}   # name -> ("counter" | "histogram", buckets)

    def _mine(self) -> dict:
        d = getattr(self._tls, "d", None)
        if d is None:
            d = self._tls.d = {}
            with self._lock:
                self._cells.append(d)
        return d

    def declare(self, name: str, kind: str, buckets=None):
        self._meta[name] = (kind, tuple(buckets) if buckets else None)

    @staticmethod
    def key(name: str, **labels):
        """The cell key of (name, labels): build it once, pass it to inc_key / observe_key on the hot path."""
        return (name, tuple(sorted(labels.items())))

    def inc(self, name: str, value=1, **labels):
        self.inc_key(self.key(name, **labels), value)

    def observe(self, name: str, value: float, **labels):
        self.observe_key(self.key(name, **labels), value)

    def cells(self) -> dict:
        """The calling thread's cells: fetch once per request, then update with plain dictionary operations
        (counters) and observe_in (histograms)."""
        return self._mine()

    def observe_in(self, d: dict, key, value: float):
        buckets = self._meta[key[0]][1]
        cell = d.get(key)
        if cell is None:
            cell = d[key] = [0] * (len(buckets) + 1) + [0.0]
        cell[bisect.bisect_left(buckets, value)] += 1
        cell[-1] += value

    def inc_key(self, key, value=1):
        d = self._mine()
        d[key] = d.get(key, 0) + value

    def observe_key(self, key, value: float):
        buckets = self._meta[key[0]][1]
        d = self._mine()
        cell = d.get(key)
        if cell is None:
            cell = d[key] = [0] * (len(buckets) + 1) + [0.0]  # per-bucket counts (+Inf last), then the sum
        cell[bisect.bisect_left(buckets, value)] += 1
        cell[-1] += value

    def snapshot(self) -> dict:
        out = {}
        with self._lock:
            cells = list(self._cells)
        for d in cells:
            for key, v in list(d.items()):
                if isinstance(v, list):
                    acc = out.setdefault(key, [0] * len(v))
                    for i, x in enumerate(list(v)):
                        acc[i] += x
                else:
                    out[key] = out.get(key, 0) + v
        return out

    def exposition(self) -> str:
        """Prometheus text format 0.0.4."""
        snap = self.snapshot()
        lines, typed = [], set()
        for (name, labels), v in sorted(snap.items()):
            kind, buckets = self._meta.get(name, ("counter", None))
            if name not in typed:
                lines.append(f"# TYPE {name} {kind}")
                typed.add(name)
            lab = dict(labels)
            if kind == "histogram":
                cum = 0
                for ub, c in zip(list(buckets) + ["+Inf"], v[:-1]):
                    cum += c
                    lines.append(f"{name}_bucket{_labels({**lab, 'le': ub})} {cum}")
                lines.append(f"{name}_sum{_labels(lab)} {v[-1]!r}")
                lines.append(f"{name}_count{_labels(lab)} {cum}")
            else:
                lines.append(f"{name}{_labels(lab)} {v}")
        return "\n".join(lines) + ("\n" if lines else "")
