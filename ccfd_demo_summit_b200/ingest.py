"""Kafka-shaped record ingest: one transaction per message into pinned micro-batches.

The demo's producer replays ``creditcard.csv`` from S3 into topic ``odh-demo`` one row per
message (reference deploy/kafka/ProducerDeployment.yaml:88-97, README.md:547-548); the router
extracts the features and calls the model.  The producer's source is not in the reference, so the
wire encoding is unpinned: both plausible encodings are supported, a CSV line of the Kaggle file's
31 columns (``Time,V1..V28,Amount,Class``) and a JSON object/array per line.  There is no broker
on the GPU box; ``KafkaShapedProducer`` replays the synthetic set from memory at a target rate and
``MicroBatcher`` cuts fixed-size batches, parsing each record straight into a session slot's
pinned buffer and overlapping the fill of one slot with the flight of the previous one.
"""
from __future__ import annotations

import ctypes
import time

import numpy as np

from . import _lib
from .seldon_harness import parse_rows_into
from .synth import FEATURE_NAMES

CSV_COLUMNS = FEATURE_NAMES + ["Class"]


def _fmt(v: float) -> str:
    return repr(float(v))  # shortest string that round-trips the float32 value as a double


class KafkaShapedProducer:
    """Iterates over encoded records (bytes).  ``rate`` = records/s (None: as fast as possible)."""

    def __init__(self, X: np.ndarray, y: np.ndarray | None = None, fmt: str = "csv", rate: float | None = None):
        if fmt not in ("csv", "json"):
            raise ValueError("fmt must be 'csv' or 'json'")
        self.X, self.y, self.fmt, self.rate = X, y, fmt, rate

    def encode(self, i: int) -> bytes:
        row = self.X[i]
        cls = int(self.y[i]) if self.y is not None else 0
        if self.fmt == "csv":
            return (",".join(_fmt(v) for v in row) + f',"{cls}"').encode()
        return ("[" + ",".join(_fmt(v) for v in row) + f",{cls}]").encode()

    def __iter__(self):
        t0 = time.perf_counter()
        for i in range(len(self.X)):
            if self.rate:
                due = t0 + i / self.rate
                while time.perf_counter() < due:
                    pass
            yield self.encode(i)


class MicroBatcher:
    """Cuts ``batch``-row micro-batches out of a record stream and scores them through a Session.

    A thin front of the native ring (``ccfd_batcher_*``, csrc/batcher.cu): records are copied into the
    session's pinned slots, a full slot is one CUDA-graph launch, the next slot fills while it is in
    flight, and a native collector thread stamps each batch's completion.
    ``on_result(first_row_index, probabilities, latency_seconds)`` is called per batch, in stream order;
    latency runs from the arrival of the batch's last record to the result being visible on the host
    (SURVEY.md §8d config 2).  Records are text (one CSV line or JSON line per message: ``push`` /
    ``push_many``) or binary (``push_rows``: float32 rows, 120 bytes per transaction).  A record
    holding +-inf raises ValueError, as scikit-learn would."""

    def __init__(self, session, batch: int = 1024, cols_in: int = 31, on_result=None):
        if batch > session.max_batch:
            raise ValueError("batch exceeds the session's max_batch")
        self.s, self.batch, self.cols_in, self.on_result = session, batch, cols_in, on_result
        self._h = ctypes.c_void_p()
        _lib.check(_lib.lib().ccfd_batcher_create(session._h, batch, ctypes.byref(self._h)))
        self._scratch = np.empty((max(batch, 1024), session.forest.n_features), dtype=np.float32)
        self._probs = np.empty((batch, 2), dtype=np.float64)
        self.batches = 0
        self.pushed = 0
        self.returned = 0

    # -- producer side ------------------------------------------------------------------------
    def push_rows(self, rows: np.ndarray):
        """Binary records: (n, n_features) float32, unit column stride."""
        if rows.dtype != np.float32 or rows.ndim != 2 or rows.strides[1] != 4:
            rows = np.ascontiguousarray(rows, dtype=np.float32).reshape(-1, self._scratch.shape[1])
        _lib.check(_lib.lib().ccfd_batcher_push(self._h, rows.ctypes.data_as(ctypes.c_void_p), rows.shape[0], rows.strides[0] // 4))
        self.pushed += rows.shape[0]
        self._drain(0)

    def push(self, record: bytes):
        n = parse_rows_into(record, self._scratch[:1], self.cols_in)
        if n != 1:
            raise ValueError(f"record holds {n} rows, expected 1")
        self.push_rows(self._scratch[:1])

    def push_many(self, records) -> None:
        """A consumer poll usually returns many messages: parse them in one native pass
        (`records`: iterable of bytes, one transaction each)."""
        records = list(records)
        for i in range(0, len(records), len(self._scratch)):
            part = records[i:i + len(self._scratch)]
            n = parse_rows_into(b"\n".join(part), self._scratch[:len(part)], self.cols_in)
            if n != len(part):
                raise ValueError(f"{len(part)} records held {n} rows")
            self.push_rows(self._scratch[:n])

    def replay(self, rows: np.ndarray, records_per_s: float = 0.0, burst: int = 1):
        """Native replay of binary records at a fixed rate (ccfd_batcher_replay); blocks until the last
        record is pushed and flushed.  Results are collected by a concurrent ``drain`` or by ``close``."""
        rows = np.ascontiguousarray(rows, dtype=np.float32)
        _lib.check(_lib.lib().ccfd_batcher_replay(self._h, rows.ctypes.data_as(ctypes.c_void_p), rows.shape[0], rows.shape[1],
                                                  float(records_per_s), int(burst)))
        self.pushed += rows.shape[0]

    def flush(self):
        """Submit the current slot if it holds rows."""
        _lib.check(_lib.lib().ccfd_batcher_flush(self._h))
        self._drain(0)

    # -- consumer side ------------------------------------------------------------------------
    def _drain(self, timeout_ms: int) -> int:
        got = 0
        first, n, lat = ctypes.c_int64(), ctypes.c_int64(), ctypes.c_double()
        while True:
            rc = _lib.check(_lib.lib().ccfd_batcher_poll(self._h, timeout_ms, ctypes.byref(first), ctypes.byref(n),
                                                        self._probs.ctypes.data_as(ctypes.c_void_p), self.batch, ctypes.byref(lat)))
            if rc == 0:
                return got
            got += 1
            self.batches += 1
            self.returned += n.value
            if self.on_result:
                self.on_result(first.value, self._probs[:n.value].copy(), lat.value * 1e-6)
            if timeout_ms:
                return got

    def drain(self, timeout_ms: int = 100) -> int:
        """Collect at most one completed batch, waiting up to timeout_ms for it (for a consumer thread)."""
        return self._drain(timeout_ms)

    def close(self):
        if not self._h:
            return
        try:
            self.flush()
            while self.returned < self.pushed:
                if self._drain(1000) == 0 and self.returned < self.pushed:
                    done, seen = ctypes.c_int64(), ctypes.c_int64()
                    _lib.check(_lib.lib().ccfd_batcher_stats(self._h, ctypes.byref(done), ctypes.byref(seen)))
                    if seen.value <= self.returned:
                        break  # a push failed half-way: nothing more is coming
        finally:
            _lib.lib().ccfd_batcher_destroy(self._h)
            self._h = ctypes.c_void_p()
