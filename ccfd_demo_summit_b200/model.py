"""Host side of the scoring path: the Seldon model class and thin handle wrappers.

``CcfdForestModel`` is what goes into the demo's Seldon deployment in place of
the class inside ``nakfour/modelfull`` (reference deploy/model/modelfull.json:20-44):
the Seldon Python wrapper calls ``predict(X, feature_names)`` with the decoded
request matrix and expects class probabilities back, column order
``classes_ = [0, 1]`` (non-fraud, fraud).  Everything numeric happens in
libccfd_rf on the GPU; there is no CPU fallback.
"""
from __future__ import annotations

import ctypes
import os
import pickle
import queue
import threading
import time

import numpy as np

from . import _lib, metrics
from .converter import ForestArrays, from_sklearn
from .synth import FEATURE_NAMES


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


def coerce_X(X, n_features: int) -> np.ndarray:
    """Shape and type checks of ForestClassifier._validate_X_predict (SK/ensemble/_forest.py:606-625),
    before any cast: a 2-D numeric array with at least one row and exactly n_features columns, or the
    ValueError scikit-learn raises."""
    X = np.asarray(X)
    if X.dtype == object or X.dtype.kind not in "fiub":
        try:
            X = X.astype(np.float64)
        except (TypeError, ValueError) as e:
            raise ValueError(f"could not convert input to float: {e}") from None
    if X.ndim != 2:
        raise ValueError(f"Expected 2D array, got {X.ndim}D array instead. Reshape your data using "
                         "array.reshape(1, -1) if it contains a single sample.")
    if X.shape[0] < 1:
        raise ValueError(f"Found array with 0 sample(s) (shape={X.shape}) while a minimum of 1 is required.")
    if X.shape[1] != n_features:
        raise ValueError(f"X has {X.shape[1]} features, but the forest is expecting {n_features} features as input.")
    return X


def _row_major_float(X: np.ndarray) -> bool:
    """float32 / float64 rows with unit column stride and a row stride that is a whole number of elements:
    what the native cast (ccfd_cast_rows, ccfd_predict_proba_rows, ccfd_session_score) reads in place.
    (numpy gives a length-1 axis an arbitrary stride, e.g. 0 for x[None, :]: a single row always qualifies.)"""
    if X.dtype not in (np.float64, np.float32) or (X.shape[1] > 1 and X.strides[1] != X.itemsize):
        return False
    return X.shape[0] == 1 or (X.strides[0] % X.itemsize == 0 and X.strides[0] >= X.shape[1] * X.itemsize)


def _ld(X: np.ndarray) -> int:
    """Row stride in elements of a matrix that passed _row_major_float."""
    return X.shape[1] if X.shape[0] == 1 else X.strides[0] // X.itemsize


def validate_X(X, n_features: int, scratch=None) -> np.ndarray:
    """Mirror ForestClassifier._validate_X_predict (SK/ensemble/_forest.py:606-625).

    Any real dtype / memory order / nested list is accepted and cast to a C-ordered
    float32 matrix; NaN is allowed (trees route it by missing_go_to_left); 1-D
    input, zero rows, a wrong feature count and +-inf (or a float64 that overflows
    float32) raise ValueError, as scikit-learn does.  ``scratch`` (optional, a dict the caller keeps)
    lets repeated calls reuse one float32 buffer instead of faulting in fresh pages every request.
    """
    X = coerce_X(X, n_features)
    if _row_major_float(X) and X.size >= 4096:
        # the usual case (a C-ordered float matrix): cast and finiteness scan in one native pass
        dense32 = X.dtype == np.float32 and X.flags.c_contiguous
        if dense32:
            X32 = X  # dense float32 rows are used in place
        elif scratch is not None:
            buf = scratch.get("buf")
            if buf is None or buf.size < X.size:
                buf = scratch["buf"] = np.empty(X.size, dtype=np.float32)
            X32 = buf[: X.size].reshape(X.shape)
        else:
            X32 = np.empty(X.shape, dtype=np.float32)
        bad = ctypes.c_int32()
        _lib.check(_lib.lib().ccfd_cast_rows(_ptr(X), int(X.dtype == np.float64), X.shape[0], X.shape[1],
                                             _ld(X), None if dense32 else _ptr(X32), ctypes.byref(bad)))
        has_inf = bool(bad.value)
    else:
        with np.errstate(over="ignore"):
            X32 = np.ascontiguousarray(X, dtype=np.float32)
        has_inf = bool(np.isinf(X32).any())
    if has_inf:
        raise ValueError("Input X contains infinity or a value too large for dtype('float32').")
    return X32


def packed_model_bytes(arrays: ForestArrays) -> int:
    """Bytes of the packed device image of a forest (what one launch reads of the model), computed on the
    host by ccfd_forest_pack: the same figure Forest.info["model_bytes"] reports, without a GPU."""
    a = arrays
    keep = [np.ascontiguousarray(x) for x in
            (a.tree_offsets, a.feature, a.threshold, a.left, a.right, a.missing_go_to_left, a.value)]
    d = _lib.ForestDesc(a.n_trees, a.n_features, a.n_classes, 0, *[_ptr(x) for x in keep])
    nbytes = ctypes.c_int64()
    _lib.check(_lib.lib().ccfd_forest_pack(ctypes.byref(d), None, 0, None, None, ctypes.byref(nbytes)))
    return int(nbytes.value) + 32 * a.n_trees


class Forest:
    """A forest resident on one GPU (opaque ``ccfd_forest`` handle)."""

    def __init__(self, arrays: ForestArrays, device: int = 0):
        self._h = ctypes.c_void_p()
        L = _lib.lib()
        a = arrays
        keep = [np.ascontiguousarray(x) for x in
                (a.tree_offsets, a.feature, a.threshold, a.left, a.right, a.missing_go_to_left, a.value)]
        d = _lib.ForestDesc(a.n_trees, a.n_features, a.n_classes, 0, *[_ptr(x) for x in keep])
        _lib.check(L.ccfd_forest_create(ctypes.byref(d), int(device), ctypes.byref(self._h)))
        del keep  # the library copied everything
        info = _lib.ForestInfo()
        _lib.check(L.ccfd_forest_get_info(self._h, ctypes.byref(info)))
        self.info = {k: getattr(info, k) for k, _ in _lib.ForestInfo._fields_}
        self.n_features = a.n_features
        self.n_trees = a.n_trees
        self.device = int(device)

    @property
    def handle(self):
        if not self._h:
            raise RuntimeError("forest is closed")
        return self._h

    def close(self):
        if getattr(self, "_h", None) and _lib is not None:  # _lib is None while the interpreter shuts down
            _lib.lib().ccfd_forest_destroy(self._h)
            self._h = ctypes.c_void_p()

    __del__ = close

    def kernel_for(self, n: int):
        """(kernel name, rows pre-sorted?) that ``kernel="auto"`` means for a batch of n rows."""
        k, srt = ctypes.c_uint32(), ctypes.c_int32()
        _lib.check(_lib.lib().ccfd_kernel_for(self.handle, int(n), ctypes.byref(k), ctypes.byref(srt)))
        return {v: name for name, v in _lib.KERNELS.items()}[k.value], bool(srt.value)

    # -- host buffers ---------------------------------------------------------------------
    def predict_proba(self, X32: np.ndarray, kernel: str = "auto", out: np.ndarray | None = None) -> np.ndarray:
        """Validated float32 rows in host memory (unit column stride) -> (n, 2) float64."""
        n, ld = X32.shape[0], _ld(X32)
        if out is None:
            out = np.empty((n, 2), dtype=np.float64)
        _lib.check(_lib.lib().ccfd_predict_proba(self.handle, _ptr(X32), n, ld, _ptr(out), _lib.KERNELS[kernel], None))
        return out

    def predict_rows(self, X: np.ndarray, kernel: str = "auto", out: np.ndarray | None = None) -> np.ndarray:
        """A float64 / float32 request matrix in ordinary host memory -> (n, 2) float64, validation
        included: cast, infinity check and scoring pipelined in the library (ccfd_predict_proba_rows).
        Raises ValueError for +-inf / float32 overflow, as scikit-learn does."""
        if not _row_major_float(X):
            raise ValueError("predict_rows needs float32 / float64 rows with unit column stride")
        n = X.shape[0]
        if out is None:
            out = np.empty((n, 2), dtype=np.float64)
        _lib.check(_lib.lib().ccfd_predict_proba_rows(self.handle, _ptr(X), int(X.dtype == np.float64), n,
                                                      _ld(X), _ptr(out), _lib.KERNELS[kernel], None))
        return out

    def decide(self, X32: np.ndarray, threshold: float = 0.5, ge: bool = False, want_proba: bool = False,
               kernel: str = "auto"):
        n, ld = X32.shape[0], _ld(X32)
        dec = np.empty(n, dtype=np.uint8)
        out = np.empty((n, 2), dtype=np.float64) if want_proba else None
        _lib.check(_lib.lib().ccfd_predict_decision(self.handle, _ptr(X32), n, ld, _ptr(out) if want_proba else None,
                                                    _ptr(dec), float(threshold), int(ge), _lib.KERNELS[kernel], None))
        return (dec.astype(bool), out) if want_proba else dec.astype(bool)

    # -- device buffers (raw pointers, e.g. torch tensors' data_ptr()) -----------------------
    def predict_proba_device(self, x_ptr: int, n: int, ld: int, out_ptr: int, stream: int = 0, kernel: str = "auto",
                             sync: bool = True):
        flags = _lib.DEVICE_PTRS | _lib.KERNELS[kernel] | (0 if sync else _lib.ASYNC)
        _lib.check(_lib.lib().ccfd_predict_proba(self.handle, ctypes.c_void_p(x_ptr), n, ld, ctypes.c_void_p(out_ptr),
                                                 flags, ctypes.c_void_p(stream)))

    def decide_device(self, x_ptr: int, n: int, ld: int, dec_ptr: int, out_ptr: int = 0, threshold: float = 0.5,
                      ge: bool = False, stream: int = 0, kernel: str = "auto", sync: bool = True):
        flags = _lib.DEVICE_PTRS | _lib.KERNELS[kernel] | (0 if sync else _lib.ASYNC)
        _lib.check(_lib.lib().ccfd_predict_decision(
            self.handle, ctypes.c_void_p(x_ptr), n, ld, ctypes.c_void_p(out_ptr) if out_ptr else None,
            ctypes.c_void_p(dec_ptr), float(threshold), int(ge), flags, ctypes.c_void_p(stream)))


class Session:
    """Pinned ring of request slots with pre-captured CUDA graphs (latency path)."""

    def __init__(self, forest: Forest, max_batch: int = 1024, n_slots: int = 4):
        self.forest, self.max_batch, self.n_slots = forest, int(max_batch), int(n_slots)
        self._h = ctypes.c_void_p()
        L = _lib.lib()
        _lib.check(L.ccfd_session_create(forest.handle, self.max_batch, self.n_slots, ctypes.byref(self._h)))
        self.x, self.p = [], []
        for s in range(self.n_slots):
            xp, pp = ctypes.POINTER(ctypes.c_float)(), ctypes.POINTER(ctypes.c_double)()
            _lib.check(L.ccfd_session_buffers(self._h, s, ctypes.byref(xp), ctypes.byref(pp)))
            self.x.append(np.ctypeslib.as_array(xp, shape=(self.max_batch, forest.n_features)))
            self.p.append(np.ctypeslib.as_array(pp, shape=(self.max_batch, 2)))

    def submit(self, slot: int, n_rows: int):
        _lib.check(_lib.lib().ccfd_session_submit(self._h, slot, n_rows))

    def wait(self, slot: int):
        _lib.check(_lib.lib().ccfd_session_wait(self._h, slot))

    def score(self, slot: int, rows: np.ndarray) -> np.ndarray:
        """One request in one native call (ccfd_session_score): float64 / float32 rows are narrowed into the
        slot's pinned buffer (ValueError on +-inf, as scikit-learn), the slot's graph runs, the probabilities
        come back as a new array."""
        if not _row_major_float(rows):
            rows = np.ascontiguousarray(rows, dtype=np.float32)
        n = rows.shape[0]
        out = np.empty((n, 2), dtype=np.float64)
        _lib.check(_lib.lib().ccfd_session_score(self._h, slot, _ptr(rows), int(rows.dtype == np.float64), n,
                                                 _ld(rows), _ptr(out), None))
        return out

    def close(self):
        if getattr(self, "_h", None) and _lib is not None:
            self.x, self.p = [], []
            _lib.lib().ccfd_session_destroy(self._h)
            self._h = ctypes.c_void_p()

    __del__ = close


class MultiGpuForest:
    """One forest replica per GPU inside one process; a bulk batch is cut into contiguous row ranges
    (sharding.shard_bounds) and every GPU scores its range on its own streams.  Nothing is exchanged
    between GPUs: rows are independent.  ctypes releases the GIL, so the per-GPU
    calls overlap."""

    def __init__(self, arrays: ForestArrays, devices=None, reuse=None):
        """``reuse``: forests that already exist ({device: Forest}); they are used as the replica of their
        device instead of uploading a second copy there, and are not closed by ``close()``."""
        n = _lib.check(_lib.lib().ccfd_device_count())
        self.devices = list(range(n)) if devices is None else [int(d) for d in devices]
        if not self.devices:
            raise RuntimeError("no CUDA device")
        if len(set(self.devices)) != len(self.devices):
            raise ValueError(f"device listed twice: {self.devices}")
        reuse = reuse or {}
        self._owned = [d not in reuse for d in self.devices]
        self.forests = [reuse[d] if d in reuse else Forest(arrays, d) for d in self.devices]
        self.n_features, self.n_trees = arrays.n_features, arrays.n_trees

    def predict_proba(self, X32: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        from .sharding import shard_bounds

        n = X32.shape[0]
        if out is None:
            out = np.empty((n, 2), dtype=np.float64)
        b = shard_bounds(n, len(self.forests))
        errs = []

        def work(g):
            lo, hi = int(b[g]), int(b[g + 1])
            try:
                if hi > lo:
                    self.forests[g].predict_proba(X32[lo:hi], out=out[lo:hi])
            except Exception as e:  # noqa: BLE001
                errs.append(e)

        th = [threading.Thread(target=work, args=(g,)) for g in range(len(self.forests))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        if errs:
            raise errs[0]
        return out

    def close(self):
        for f, mine in zip(self.forests, self._owned):
            if mine:
                f.close()


def load_model_file(path: str) -> ForestArrays:
    """``.npz`` written by ForestArrays.save, or a pickle/joblib of a fitted RandomForestClassifier."""
    if path.endswith(".npz"):
        return ForestArrays.load(path)
    try:
        import joblib

        obj = joblib.load(path)
    except Exception:
        with open(path, "rb") as fh:
            obj = pickle.load(fh)
    return from_sklearn(obj)


class CcfdForestModel:
    """Seldon MODEL class: ``predict(X, feature_names) -> (n, 2) float64 [p_nonfraud, p_fraud]``.

    Configuration keeps the demo's environment-variable idiom (reference
    deploy/router.yaml:54-70): ``CCFD_MODEL_PATH`` (forest file), ``CCFD_DEVICE``
    (GPU index), ``CCFD_DEVICES`` (comma-separated GPU indices: bulk requests are row-sharded over
    them, one replica each), ``CCFD_SESSION_MAX_BATCH`` (requests up to this many rows take the
    CUDA-graph path; 0 disables it), ``CCFD_SESSION_SLOTS`` (how many such requests may be in flight at
    once, default 4), ``FRAUD_THRESHOLD`` (same name and default as the
    router's).  A fitted forest or a ForestArrays can also be passed directly.
    """

    class_names = ["proba_0", "proba_1"]

    def __init__(self, model=None, device: int | None = None, session_max_batch: int | None = None,
                 session_slots: int | None = None):
        if model is None:
            path = os.environ.get("CCFD_MODEL_PATH")
            if not path:
                raise ValueError("no model given and CCFD_MODEL_PATH is not set")
            model = load_model_file(path)
        arrays = model if isinstance(model, ForestArrays) else from_sklearn(model)
        device = int(os.environ.get("CCFD_DEVICE", "0")) if device is None else device
        self.forest = Forest(arrays, device)
        devs = [int(d) for d in os.environ.get("CCFD_DEVICES", "").split(",") if d.strip() != ""]
        # the forest above doubles as the replica of its own device: no second upload there
        self._multi = MultiGpuForest(arrays, devs, reuse={self.forest.device: self.forest}) if len(devs) > 1 else None
        self.n_features_in_ = arrays.n_features
        self.classes_ = np.arange(arrays.n_classes)
        self.threshold = float(os.environ.get("FRAUD_THRESHOLD", "0.5"))
        smb = int(os.environ.get("CCFD_SESSION_MAX_BATCH", "1024")) if session_max_batch is None else session_max_batch
        n_slots = max(1, int(os.environ.get("CCFD_SESSION_SLOTS", "4")) if session_slots is None else int(session_slots))
        self._session = Session(self.forest, smb, n_slots=n_slots) if smb > 0 else None
        # Requests take a free slot and give it back: up to n_slots requests are in flight at once (copy-in,
        # graph and wait of different slots overlap on their own streams); nothing else is shared between
        # them, so no lock is held while a request is on the GPU.
        self._free_slots = queue.SimpleQueue()
        for k in range(n_slots if self._session is not None else 0):
            self._free_slots.put(k)
        self._tls = threading.local()  # per-thread cast buffer (Seldon may run threaded workers)
        self._last = None              # (row, probabilities) of the latest request: one tuple, swapped atomically
        self._in_flight = 0
        self.max_in_flight = 0         # high-water mark of concurrent session requests (observability, tests)
        self._stat_lock = threading.Lock()
        self.metrics_registry = reg = metrics.Registry()
        reg.declare("ccfd_requests_total", "counter")
        reg.declare("ccfd_rows_total", "counter")
        reg.declare("ccfd_request_rows", "histogram", metrics.BATCH_BUCKETS)
        reg.declare("ccfd_score_seconds", "histogram", metrics.LATENCY_BUCKETS)
        reg.declare("ccfd_predict_seconds", "histogram", metrics.LATENCY_BUCKETS)
        self._metric_keys = {p: (reg.key("ccfd_requests_total", path=p), reg.key("ccfd_rows_total", path=p), reg.key("ccfd_request_rows"),
                                 reg.key("ccfd_score_seconds", path=p), reg.key("ccfd_predict_seconds", path=p))
                             for p in ("session", "bulk", "multi_gpu")}

    def predict(self, X, feature_names=None, meta=None):
        """The Seldon wrapper passes the names positionally; like scikit-learn on a forest fitted
        from an ndarray, they are not used to reorder columns."""
        t0 = time.perf_counter()
        tl = self._tls
        if not hasattr(tl, "scratch"):
            tl.scratch = {}
        Xa = np.asarray(X)
        small = self._session is not None and (Xa.ndim != 2 or Xa.shape[0] <= self._session.max_batch)
        if small:  # the latency path (and every malformed input)
            Xa = coerce_X(Xa, self.n_features_in_)
            # float rows go to the slot as they are: the native call narrows them and checks for infinity
            X32, fused = (Xa if _row_major_float(Xa) else validate_X(Xa, self.n_features_in_, tl.scratch)), False
        else:
            Xa = coerce_X(Xa, self.n_features_in_)
            fused = not (self._multi is not None and Xa.shape[0] >= 65536) and _row_major_float(Xa)
            X32 = None if fused else validate_X(Xa, self.n_features_in_, tl.scratch)
        n = Xa.shape[0] if fused else X32.shape[0]
        t1 = time.perf_counter()
        if small:
            slot = self._free_slots.get()  # blocks only when every slot is in flight
            with self._stat_lock:
                self._in_flight += 1
                self.max_in_flight = max(self.max_in_flight, self._in_flight)
            try:
                P = self._session.score(slot, X32)
            finally:
                with self._stat_lock:
                    self._in_flight -= 1
                self._free_slots.put(slot)
            path = "session"
        elif X32 is not None and self._multi is not None and n >= 65536:
            P = self._multi.predict_proba(X32)
            path = "multi_gpu"
        elif fused:
            # bulk request straight from the wrapper's (pageable, float64 or float32) matrix: the cast, the
            # infinity check, the copies and the scoring are one pipelined native call
            P = self.forest.predict_rows(Xa)
            path = "bulk"
        else:
            P = self.forest.predict_proba(X32)
            path = "bulk"
        t2 = time.perf_counter()
        last_row = (Xa[-1] if X32 is None else X32[-1]).astype(np.float32)
        self._last = (last_row, P[-1].copy())  # the pair is built first, then published in one store
        reg, k = self.metrics_registry, self._metric_keys[path]  # keys built once; this thread's cells fetched once
        d = reg.cells()
        d[k[0]] = d.get(k[0], 0) + 1
        d[k[1]] = d.get(k[1], 0) + n
        reg.observe_in(d, k[2], n)
        reg.observe_in(d, k[3], t2 - t1)  # copy-in + kernels + result visible on the host
        reg.observe_in(d, k[4], t2 - t0)  # with validation / cast
        return P

    def decide(self, X, threshold: float | None = None, ge: bool = False):
        """Router rule (reference deploy/router.yaml:69-70): fraud iff p_fraud > FRAUD_THRESHOLD."""
        X32 = validate_X(X, self.n_features_in_)
        return self.forest.decide(X32, self.threshold if threshold is None else threshold, ge)

    def metrics(self):
        """Gauges the shipped ModelPrediction board plots (reference deploy/grafana/ModelPrediction.json:96,104,211,322)."""
        if self._last is None:
            return []
        x, p = self._last
        g = lambda k, v: {"type": "GAUGE", "key": k, "value": float(v)}  # noqa: E731
        out = [g("proba_1", p[1])]
        if self.n_features_in_ == len(FEATURE_NAMES):
            out += [g("Amount", x[29]), g("V17", x[17]), g("V10", x[10])]
        return out

    def tags(self):
        return {"backend": "libccfd_rf/sm_100a", "n_trees": self.forest.n_trees, "device": self.forest.device}

    def close(self):
        if self._session is not None:
            self._session.close()
            self._session = None
        if self._multi is not None:
            self._multi.close()
            self._multi = None
        self.forest.close()
