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

import time

import numpy as np

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

    ``on_result(first_row_index, probabilities, latency_seconds)`` is called per batch; latency runs
    from the arrival of the batch's last record to the result being visible on the host
    (SURVEY.md §8d config 2)."""

    def __init__(self, session, batch: int = 1024, cols_in: int = 31, on_result=None):
        if batch > session.max_batch:
            raise ValueError("batch exceeds the session's max_batch")
        self.s, self.batch, self.cols_in, self.on_result = session, batch, cols_in, on_result
        self.slot, self.fill = 0, 0
        self.in_flight = None  # (slot, first_row, n, t_last_record)
        self.first_row = 0
        self.batches = 0

    def push(self, record: bytes):
        s = self.s
        n = parse_rows_into(record, s.x[self.slot][self.fill:self.fill + 1], self.cols_in)
        if n != 1:
            raise ValueError(f"record holds {n} rows, expected 1")
        self.fill += 1
        if self.fill == self.batch:
            self.flush()

    def push_many(self, records) -> None:
        """A consumer poll usually returns many messages: parse them in one native pass per slot fill
        instead of one call per record (`records`: iterable of bytes, one transaction each)."""
        records = list(records)
        i = 0
        while i < len(records):
            take = min(len(records) - i, self.batch - self.fill)
            blob = b"\n".join(records[i:i + take])
            n = parse_rows_into(blob, self.s.x[self.slot][self.fill:self.fill + take], self.cols_in)
            if n != take:
                raise ValueError(f"{take} records held {n} rows")
            self.fill += take
            i += take
            if self.fill == self.batch:
                self.flush()

    def flush(self):
        """Submit the current slot (if it holds rows) and collect the previous one."""
        if self.fill == 0:
            return self._collect()
        t_last = time.perf_counter()
        self.s.submit(self.slot, self.fill)
        prev = self.in_flight
        self.in_flight = (self.slot, self.first_row, self.fill, t_last)
        self.first_row += self.fill
        self.slot, self.fill = (self.slot + 1) % self.s.n_slots, 0
        if prev is not None:
            self._finish(prev)

    def _finish(self, rec):
        slot, first, n, t_last = rec
        self.s.wait(slot)
        lat = time.perf_counter() - t_last
        self.batches += 1
        if self.on_result:
            self.on_result(first, self.s.p[slot][:n], lat)

    def _collect(self):
        if self.in_flight is not None:
            rec, self.in_flight = self.in_flight, None
            self._finish(rec)

    def close(self):
        self.flush()
        self._collect()
