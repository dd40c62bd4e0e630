"""Seldon REST wire format around the model class, plus a request coalescer.

``seldon_core`` is not installed in this image, so this module stands in for the Seldon Python
wrapper at the two places the demo calls the model: the engine path
``POST /api/v0.1/predictions`` (reference deploy/router.yaml:65-68) and the raw microservice path
``POST /predict`` (reference deploy/ccd-service.yaml:61-62, README.md:379).  Request bodies are
``{"data": {"names": [...], "ndarray": [[...], ...]}}`` or
``{"data": {"tensor": {"shape": [n, 30], "values": [...]}}}``; the response carries
``names = ["proba_0", "proba_1"]`` and one probability pair per row.

Small requests never become Python lists: ``ccfd_parse_rows`` writes the floats straight into a
session slot's pinned buffer.  ``RequestCoalescer`` merges concurrent single-transaction requests
(the jBPM/Camel pattern of BASELINE configs[4]) into one CUDA-graph launch.
"""
from __future__ import annotations

import ctypes
import json
import os
import threading
import time
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

import numpy as np

from . import _lib, metrics

PREDICT_PATHS = ("/api/v0.1/predictions", "/predict", "/api/v1.0/predictions")


def decode_request(body) -> tuple[np.ndarray, list]:
    """Reference decoding (what the Python wrapper does): JSON -> (float64 matrix, names)."""
    msg = body if isinstance(body, dict) else json.loads(body)
    data = msg.get("data", {})
    names = list(data.get("names", []) or [])
    if "ndarray" in data:
        X = np.asarray(data["ndarray"], dtype=np.float64)
    elif "tensor" in data:
        t = data["tensor"]
        X = np.asarray(t["values"], dtype=np.float64).reshape(t["shape"])
    else:
        raise ValueError("request has neither data.ndarray nor data.tensor")
    return X, names


def parse_rows_into(body: bytes, dst: np.ndarray, cols_in: int | None = None) -> int:
    """Native one-pass decode of a request/record body into ``dst`` (float32, C order, (cap, F)).
    Returns the number of rows written."""
    if dst.dtype != np.float32 or not dst.flags.c_contiguous or dst.ndim != 2:
        raise ValueError("dst must be a C-contiguous float32 matrix")
    if isinstance(body, str):
        body = body.encode()
    n = ctypes.c_int64()
    F = dst.shape[1]
    rc = _lib.lib().ccfd_parse_rows(body, len(body), int(cols_in or F), F, dst.ctypes.data_as(ctypes.c_void_p),
                                    dst.shape[0], ctypes.byref(n))
    _lib.check(rc)
    return int(n.value)


def format_proba(P: np.ndarray) -> bytes:
    """Native ``[[p0,p1],...]`` with shortest round-trip decimals (ccfd_format_proba)."""
    P = np.ascontiguousarray(P, dtype=np.float64).reshape(-1, 2)
    buf = ctypes.create_string_buffer(68 * P.shape[0] + 2)
    n = ctypes.c_int64()
    _lib.check(_lib.lib().ccfd_format_proba(P.ctypes.data_as(ctypes.c_void_p), P.shape[0], buf, len(buf), ctypes.byref(n)))
    return buf.raw[: n.value]


def encode_response(P: np.ndarray, metrics=None, tags=None) -> bytes:
    meta = {}
    if metrics:
        meta["metrics"] = metrics
    if tags:
        meta["tags"] = tags
    return (b'{"data": {"names": ["proba_0", "proba_1"], "ndarray": ' + format_proba(P) + b'}, "meta": '
            + json.dumps(meta).encode() + b"}")


class SeldonApp:
    """Body in, body out: what the wrapper does between the HTTP layer and ``predict()``.

    ``coalesce`` (or ``CCFD_COALESCE=1``): single-transaction requests (the jBPM/Camel pattern, reference
    deploy/ccd-service.yaml:61-66) go through a RequestCoalescer, so requests that arrive while a launch
    is in flight share the next one.  The app also keeps the request histograms the SeldonCore board
    plots (reference deploy/grafana/SeldonCore.json:119-531) under the engine's metric names."""

    LABELS = {"deployment_name": "modelfull", "predictor_name": "modelfull", "predictor_version": "v1",
              "model_name": "modelfull", "model_image": "libccfd_rf", "model_version": "sm_100a"}

    def __init__(self, model, fast_decode: bool = True, coalesce: bool | None = None):
        self.model = model
        self.fast_decode = fast_decode and getattr(model, "_session", None) is not None
        self._scratch = threading.local()
        self.registry = reg = getattr(model, "metrics_registry", None) or metrics.Registry()
        reg.declare("seldon_api_engine_server_requests_seconds", "histogram", metrics.LATENCY_BUCKETS)
        reg.declare("seldon_api_engine_client_requests_seconds", "histogram", metrics.LATENCY_BUCKETS)
        if coalesce is None:
            coalesce = os.environ.get("CCFD_COALESCE", "0") not in ("", "0")
        self.coalescer = None
        if coalesce and self.fast_decode:
            self.coalescer = RequestCoalescer(lambda M: model.predict(M, None), model.n_features_in_,
                                              max_batch=model._session.max_batch, max_delay_s=0.0)

    def _predict(self, X, names):
        t = time.perf_counter()
        try:
            if self.coalescer is not None and len(X) == 1 and isinstance(X, np.ndarray) and X.dtype == np.float32:
                return self.coalescer.score(X)
            return self.model.predict(X, names)
        finally:
            self.registry.observe("seldon_api_engine_client_requests_seconds", time.perf_counter() - t, status="200", **self.LABELS)

    def handle(self, body: bytes) -> bytes:
        m = self.model
        if self.fast_decode and len(body) < (1 << 22):
            buf = getattr(self._scratch, "buf", None)
            if buf is None:
                buf = self._scratch.buf = np.empty((m._session.max_batch, m.n_features_in_), dtype=np.float32)
            try:
                n = parse_rows_into(body, buf)
            except ValueError:
                n = -1  # oversize or irregular payload: let the reference decoder produce the error or the matrix
            if n > 0:
                P = self._predict(buf[:n], None)
                return encode_response(P, m.metrics(), m.tags())
        X, names = decode_request(body)
        return encode_response(self._predict(X, names), m.metrics(), m.tags())

    def serve(self, body: bytes):
        """handle() with the status mapping of the HTTP layer: (status code, response body)."""
        t = time.perf_counter()
        try:
            out, code = self.handle(body), 200
        except ValueError as e:
            out, code = json.dumps({"status": {"code": 400, "info": str(e), "status": "FAILURE"}}).encode(), 400
        except Exception as e:  # noqa: BLE001  (CUDA failure etc.: 5xx, process stays up)
            out, code = json.dumps({"status": {"code": 500, "info": str(e), "status": "FAILURE"}}).encode(), 500
        self.registry.observe("seldon_api_engine_server_requests_seconds", time.perf_counter() - t, status=str(code), **self.LABELS)
        return code, out

    def close(self):
        if self.coalescer is not None:
            self.coalescer.close()
            self.coalescer = None


def make_server(model, host="127.0.0.1", port=8000, coalesce: bool | None = None) -> ThreadingHTTPServer:
    app = SeldonApp(model, coalesce=coalesce)

    class Handler(BaseHTTPRequestHandler):
        def log_message(self, *a):  # quiet
            pass

        def do_POST(self):
            if self.path.split("?")[0] not in PREDICT_PATHS:
                self.send_error(404)
                return
            body = self.rfile.read(int(self.headers.get("Content-Length", "0")))
            code, out = app.serve(body)
            self.send_response(code)
            self.send_header("Content-Type", "application/json")
            self.send_header("Content-Length", str(len(out)))
            self.end_headers()
            self.wfile.write(out)

        def do_GET(self):
            if self.path.startswith("/prometheus") or self.path.startswith("/metrics"):
                out = prometheus_text(model, app.registry).encode()
                self.send_response(200)
                self.send_header("Content-Type", "text/plain; version=0.0.4")
                self.send_header("Content-Length", str(len(out)))
                self.end_headers()
                self.wfile.write(out)
            else:
                self.send_error(404)

    srv = ThreadingHTTPServer((host, port), Handler)
    srv.app = app
    return srv


def prometheus_text(model, registry=None) -> str:
    """The model-level gauges of reference deploy/grafana/ModelPrediction.json:96,104,211,322 in
    Prometheus exposition format (the model service is scraped at /prometheus, README.md:292-301),
    followed by the counters and histograms of the model (requests, rows, request sizes, scoring time)
    and of the request layer (the SeldonCore board's seldon_api_engine_* families)."""
    lines = []
    for g in model.metrics():
        lines.append(f"# TYPE {g['key']} gauge")
        lines.append(f"{g['key']} {g['value']!r}")
    text = "\n".join(lines) + "\n"
    registry = registry or getattr(model, "metrics_registry", None)
    if registry is not None:
        text += registry.exposition()
    return text


class RequestCoalescer:
    """Merge concurrent small requests into one launch.

    Callers block in ``score(rows)``; a dispatcher thread drains whatever has queued (up to
    ``max_batch`` rows), scores it with one ``score_fn(matrix)`` call and hands each caller its
    slice.  ``max_delay_s`` bounds how long the first request of a batch waits for company."""

    def __init__(self, score_fn, n_features: int, max_batch: int = 1024, max_delay_s: float = 50e-6):
        self.score_fn, self.F, self.max_batch, self.max_delay_s = score_fn, n_features, max_batch, max_delay_s
        self._cv = threading.Condition()
        self._pending = []  # (rows, holder)
        self._stop = False
        self.launches = 0
        self.rows_scored = 0
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()

    def score(self, rows: np.ndarray) -> np.ndarray:
        rows = np.ascontiguousarray(rows, dtype=np.float32).reshape(-1, self.F)
        if rows.shape[0] > self.max_batch:
            raise ValueError(f"request of {rows.shape[0]} rows exceeds max_batch {self.max_batch}")
        holder = {"ev": threading.Event(), "out": None, "err": None}
        with self._cv:
            self._pending.append((rows, holder))
            self._cv.notify()
        holder["ev"].wait()
        if holder["err"] is not None:
            raise holder["err"]
        return holder["out"]

    def _run(self):
        while True:
            with self._cv:
                while not self._pending and not self._stop:
                    self._cv.wait()
                if self._stop and not self._pending:
                    return
                if self.max_delay_s > 0:
                    # wait for company until the first request's deadline or a full batch, whichever is first
                    deadline = time.perf_counter() + self.max_delay_s
                    while sum(r.shape[0] for r, _ in self._pending) < self.max_batch:
                        left = deadline - time.perf_counter()
                        if left <= 0 or self._stop:
                            break
                        self._cv.wait(left)
                # max_delay_s == 0: no artificial wait.  Whatever queued while the previous launch was in
                # flight forms this batch, so an idle server adds nothing to a lone request's latency.
                take, total = [], 0
                while self._pending and total + self._pending[0][0].shape[0] <= self.max_batch:
                    r, h = self._pending.pop(0)
                    take.append((r, h))
                    total += r.shape[0]
            if not take:
                continue
            try:
                P = self.score_fn(np.concatenate([r for r, _ in take], axis=0))
                self.launches += 1
                self.rows_scored += total
                k = 0
                for r, h in take:
                    h["out"] = P[k:k + r.shape[0]].copy()
                    k += r.shape[0]
            except Exception as e:  # noqa: BLE001
                for _, h in take:
                    h["err"] = e
            for _, h in take:
                h["ev"].set()

    def close(self):
        with self._cv:
            self._stop = True
            self._cv.notify()
        self._thread.join(timeout=5)
