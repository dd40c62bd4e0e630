"""Seldon wire format, native row parser, request coalescer and Kafka-shaped ingest.
CPU tests cover decoding and host logic; the gpu-marked ones run the full request path."""
import json
import threading

import numpy as np
import pytest

from conftest import bits, load_golden
from ccfd_demo_summit_b200 import ingest, seldon_harness as sh, synth


@pytest.fixture(scope="module")
def rows():
    X = synth.make_features(64, seed=21)
    X[3, 5] = np.float32(1e-42)  # denormal
    X[4, 6] = np.float32(-0.0)
    X[5, 7] = np.float32(3.4e38)
    return X


def _ref(body, cols=30):
    X, _ = sh.decode_request(body)
    return np.ascontiguousarray(X, dtype=np.float32)


def test_native_parser_matches_json_numpy_bitwise(rows):
    for payload in (
        {"data": {"names": synth.FEATURE_NAMES, "ndarray": rows.astype(np.float64).tolist()}},
        {"data": {"ndarray": rows[:1].astype(np.float64).tolist()}},
        {"data": {"tensor": {"shape": [64, 30], "values": rows.astype(np.float64).ravel().tolist()}}},
        {"meta": {"puid": "abc123"}, "data": {"names": ["V1", "V2"], "tensor": {"shape": [2, 30], "values": rows[:2].astype(np.float64).ravel().tolist()}}},
    ):
        body = json.dumps(payload).encode()
        dst = np.zeros((64, 30), dtype=np.float32)
        n = sh.parse_rows_into(body, dst)
        want = _ref(body)
        assert n == len(want)
        assert np.array_equal(dst[:n].view(np.uint32), want.view(np.uint32))


def test_native_parser_double_rounding_equals_numpy():
    # decimal strings that sit close to float32 rounding boundaries
    vals = ["0.1", "16777217.0", "1.00000005960464477539", "3.4028235677973366e38", "1e-45", "-7.006492321624085e-46",
            "123456.7890123456789", "2.9802322387695312e-08"]
    body = ("[" + ",".join(vals) + "]").encode()
    dst = np.zeros((1, len(vals)), dtype=np.float32)
    assert sh.parse_rows_into(body, dst) == 1
    with np.errstate(over="ignore"):
        want = np.asarray([float(v) for v in vals], dtype=np.float64).astype(np.float32)
    assert np.array_equal(dst[0].view(np.uint32), want.view(np.uint32))


def test_csv_and_json_records_drop_the_class_column(rows):
    y = (np.arange(len(rows)) % 3 == 0).astype(np.int64)
    for fmt in ("csv", "json"):
        prod = ingest.KafkaShapedProducer(rows, y, fmt=fmt)
        dst = np.zeros((1, 30), dtype=np.float32)
        for i, rec in enumerate(prod):
            assert sh.parse_rows_into(rec, dst, 31) == 1
            assert np.array_equal(dst[0].view(np.uint32), rows[i].view(np.uint32)), (fmt, i)
    # a whole CSV file with header
    text = (",".join(ingest.CSV_COLUMNS) + "\n" + "\n".join(
        ingest.KafkaShapedProducer(rows, y).encode(i).decode() for i in range(len(rows)))).encode()
    dst = np.zeros((64, 30), dtype=np.float32)
    assert sh.parse_rows_into(text, dst, 31) == 64
    assert np.array_equal(dst.view(np.uint32), rows.view(np.uint32))


def test_parser_special_values_and_errors():
    dst = np.zeros((2, 3), dtype=np.float32)
    assert sh.parse_rows_into(b'{"data":{"ndarray":[[1, NaN, null],[Infinity,-Infinity,-1e3]]}}', dst) == 2
    assert np.isnan(dst[0, 1]) and np.isnan(dst[0, 2]) and dst[1, 0] == np.inf and dst[1, 1] == -np.inf and dst[1, 2] == -1000
    with pytest.raises(ValueError, match="middle of a row"):
        sh.parse_rows_into(b"[[1,2,3],[4,5]]", dst)
    with pytest.raises(ValueError, match="more than 2 rows"):
        sh.parse_rows_into(b"[[1,2,3],[4,5,6],[7,8,9]]", dst)
    with pytest.raises(ValueError, match="unterminated"):
        sh.parse_rows_into(b'[[1,2,"3]]', dst)
    assert sh.parse_rows_into(b"", dst) == 0


def test_parser_is_structure_aware(rows):
    """A row is what the payload says is a row.  Payloads the Seldon wrapper + scikit-learn answer with a
    ValueError must not be re-cut every 30 numbers and scored (they used to be)."""
    dst = np.zeros((64, 30), dtype=np.float32)
    r64 = rows.astype(np.float64)
    wide = np.concatenate([r64[:30], r64[:30, :1]], axis=1)                     # 30 rows x 31 columns = 31 x 30 numbers
    ragged = [r64[0, :29].tolist() + [True], r64[1].tolist() + [1.0]]           # 29 numbers + true, then 31 numbers
    bad = [
        {"data": {"ndarray": wide.tolist()}},
        {"data": {"ndarray": [np.concatenate([r64[0], r64[1]]).tolist()]}},     # one row of 60
        {"data": {"ndarray": ragged}},
        {"data": {"ndarray": r64[0].tolist()}},                                  # 1-D: not a matrix
        {"data": {"ndarray": [r64[0].tolist(), r64[1, :29].tolist() + ["abc"]]}},
        {"data": {"tensor": {"shape": [2, 31], "values": wide[:2].ravel().tolist()}}},
        {"data": {"tensor": {"shape": [3, 30], "values": r64[:2].ravel().tolist()}}},
        {"data": {"tensor": {"shape": [60], "values": r64[:2].ravel().tolist()}}},
        {"data": {"tensor": {"values": r64[:2].ravel().tolist()}}},
    ]
    for payload in bad:
        with pytest.raises(ValueError):
            sh.parse_rows_into(json.dumps(payload).encode(), dst)
    # what follows the payload is not read: numbers in "meta" are not cells
    body = json.dumps({"data": {"names": ["a"], "ndarray": r64[:3].tolist()}, "meta": {"tags": {"k": [1, 2, 3]}, "puid": "7"}}).encode()
    assert sh.parse_rows_into(body, dst) == 3 and np.array_equal(dst[:3].view(np.uint32), rows[:3].view(np.uint32))
    body = json.dumps({"data": {"tensor": {"values": r64[:2].ravel().tolist(), "shape": [2, 30]}}, "meta": {"n": 5}}).encode()
    assert sh.parse_rows_into(body, dst) == 2  # "values" before "shape"
    # pretty-printed JSON: newlines inside an array do not end a row
    body = json.dumps({"data": {"ndarray": r64[:2].tolist()}}, indent=2).encode()
    assert sh.parse_rows_into(body, dst) == 2 and np.array_equal(dst[:2].view(np.uint32), rows[:2].view(np.uint32))
    # CSV: header words are whole tokens ("Info", "Nanos" are not inf / nan), a header line is skipped,
    # a line that mixes words and numbers or has the wrong width is refused
    d3 = np.zeros((4, 3), dtype=np.float32)
    assert sh.parse_rows_into(b"Info,Nanos,Time\n1,2,3\n\n4,nan,-inf\n", d3) == 2
    assert d3[0].tolist() == [1.0, 2.0, 3.0] and d3[1, 0] == 4 and np.isnan(d3[1, 1]) and d3[1, 2] == -np.inf
    for text in (b"1,2,x\n", b"1,2\n", b"1,2,3,4\n", b"1,2,3\n4,5\n", b"1,true,3\n", b"[[1,2,3],[4,[5,6]]]"):
        with pytest.raises(ValueError):
            sh.parse_rows_into(text, d3)
    # JSON object per line: keys are not cells
    assert sh.parse_rows_into(b'{"Time": 1, "V1": 2.5, "Amount": "3"}\n{"Time": 4, "V1": 5, "Amount": 6}', d3) == 2
    assert d3[0].tolist() == [1.0, 2.5, 3.0]


def test_large_record_blobs_are_parsed_on_the_host_pool_with_the_same_result(rows):
    """A consumer poll of many messages (> 256 KB of text) is cut at line boundaries and parsed by the
    persistent host pool; header and blank lines in the middle leave no holes."""
    big = np.tile(rows, (80, 1))  # 5,120 records, ~1.4 MB of CSV
    prod = ingest.KafkaShapedProducer(big, None)
    lines = [prod.encode(i).decode() for i in range(len(big))]
    lines.insert(0, ",".join(ingest.CSV_COLUMNS))
    lines.insert(2500, "")
    lines.insert(4000, ",".join(ingest.CSV_COLUMNS))
    text = "\n".join(lines).encode()
    assert len(text) > (1 << 20)
    dst = np.full((len(big) + 3, 30), -1.0, dtype=np.float32)
    assert sh.parse_rows_into(text, dst, 31) == len(big)
    assert np.array_equal(dst[: len(big)].view(np.uint32), big.view(np.uint32))
    with pytest.raises(ValueError):  # an error deep inside a chunk still surfaces
        sh.parse_rows_into(text.replace(lines[3333].encode(), lines[3333].encode() + b",9"), dst, 31)
    with pytest.raises(ValueError, match="more than"):
        sh.parse_rows_into(text, dst[:100], 31)


class _StubModel:
    """Stands in for CcfdForestModel on a box without a GPU: same validation, constant output."""
    n_features_in_ = 30

    class _session:  # noqa: N801
        max_batch = 64

    def predict(self, X, names=None):
        from ccfd_demo_summit_b200.model import validate_X

        X32 = validate_X(X, 30)
        return np.tile(np.array([[0.75, 0.25]]), (len(X32), 1))

    def metrics(self):
        return []

    def tags(self):
        return {}


def test_seldon_app_refuses_misshapen_payloads_like_the_wrapper_would(rows):
    app = sh.SeldonApp(_StubModel())
    r64 = rows.astype(np.float64)
    ok = json.loads(app.handle(json.dumps({"data": {"ndarray": r64[:3].tolist()}}).encode()))
    assert len(ok["data"]["ndarray"]) == 3
    wide = np.concatenate([r64[:30], r64[:30, :1]], axis=1)
    for payload in ({"data": {"ndarray": wide.tolist()}}, {"data": {"ndarray": [np.concatenate([r64[0], r64[1]]).tolist()]}},
                    {"data": {"ndarray": [r64[0, :29].tolist() + [True], r64[1].tolist() + [1.0]]}}):
        with pytest.raises(ValueError):
            app.handle(json.dumps(payload).encode())


def test_encode_response_shape():
    P = np.array([[0.75, 0.25], [0.1, 0.9]])
    msg = json.loads(sh.encode_response(P, [{"type": "GAUGE", "key": "proba_1", "value": 0.9}], {"k": 1}))
    assert msg["data"]["names"] == ["proba_0", "proba_1"]
    assert msg["data"]["ndarray"] == P.tolist()
    assert msg["meta"]["metrics"][0]["key"] == "proba_1"


def test_native_response_encoding_round_trips_every_float64():
    rng = np.random.default_rng(3)
    P = np.concatenate([rng.random((4000, 2)), np.array([[1.0, 0.0], [0.0, 1.0], [0.5, 0.5], [1e-300, 1 - 1e-16], [5e-324, 0.1],
                                                         [1e-5, 123456789.125], [1e16, 1e22]])])
    txt = sh.format_proba(P)
    back = np.asarray(json.loads(txt), dtype=np.float64)
    assert np.array_equal(back.view(np.uint64), P.view(np.uint64))
    assert sh.format_proba(np.zeros((0, 2))) == b"[]"
    with pytest.raises(ValueError):
        sh.format_proba(np.array([[np.nan, 0.5]]))


def test_coalescer_merges_concurrent_single_row_requests(rows):
    calls = []

    def score_fn(M):
        calls.append(len(M))
        return np.stack([M[:, 0].astype(np.float64), M[:, 1].astype(np.float64)], axis=1)

    co = sh.RequestCoalescer(score_fn, 30, max_batch=16, max_delay_s=0.02)
    out = [None] * 40
    def worker(i):
        out[i] = co.score(rows[i % len(rows)][None, :])
    th = [threading.Thread(target=worker, args=(i,)) for i in range(40)]
    for t in th:
        t.start()
    for t in th:
        t.join(timeout=30)
    co.close()
    for i in range(40):
        r = rows[i % len(rows)]
        assert out[i].shape == (1, 2) and out[i][0, 0] == np.float64(r[0]) and out[i][0, 1] == np.float64(r[1])
    assert sum(calls) == 40 and max(calls) <= 16 and len(calls) < 40  # fewer launches than requests
    with pytest.raises(ValueError):
        sh.RequestCoalescer(score_fn, 30, max_batch=4).score(np.zeros((5, 30), np.float32))


# ----------------------------------------------------------------------------------------- GPU
def _model():
    from ccfd_demo_summit_b200 import converter, model

    raw, X, P, _ = load_golden("small_hard")
    fa = converter.from_arrays(raw.tree_offsets, raw.left, raw.right, raw.feature, raw.threshold,
                               raw.missing_go_to_left, raw.value, raw.n_features, raw.n_classes)
    return model, model.CcfdForestModel(fa, session_max_batch=512), np.ascontiguousarray(X, dtype=np.float32), P


@pytest.mark.gpu
def test_seldon_request_response_matches_golden_over_http():
    import urllib.request

    _, m, X, P = _model()
    srv = sh.make_server(m, port=0)
    port = srv.server_address[1]
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    try:
        for path, payload, sl in (
            ("/api/v0.1/predictions", {"data": {"names": synth.FEATURE_NAMES, "ndarray": X[:7].astype(np.float64).tolist()}}, slice(0, 7)),
            ("/predict", {"data": {"tensor": {"shape": [1, 30], "values": X[9].astype(np.float64).tolist()}}}, slice(9, 10)),
            ("/api/v0.1/predictions", {"data": {"ndarray": X[:900].astype(np.float64).tolist()}}, slice(0, 900)),  # > session batch
        ):
            req = urllib.request.Request(f"http://127.0.0.1:{port}{path}", data=json.dumps(payload).encode(),
                                         headers={"Content-Type": "application/json"})
            msg = json.loads(urllib.request.urlopen(req, timeout=30).read())
            got = np.asarray(msg["data"]["ndarray"], dtype=np.float64)
            assert msg["data"]["names"] == ["proba_0", "proba_1"]
            assert np.array_equal(bits(got), bits(P[sl]))  # repr round-trips float64 exactly
            assert {g["key"] for g in msg["meta"]["metrics"]} == {"proba_1", "Amount", "V17", "V10"}
        bad = urllib.request.Request(f"http://127.0.0.1:{port}/predict", data=b'{"data":{"ndarray":[[1,2,3]]}}')
        with pytest.raises(urllib.error.HTTPError) as ei:
            urllib.request.urlopen(bad, timeout=30)
        assert ei.value.code == 400
        txt = urllib.request.urlopen(f"http://127.0.0.1:{port}/prometheus", timeout=30).read().decode()
        assert "proba_1 " in txt and "# TYPE Amount gauge" in txt
    finally:
        srv.shutdown()
        m.close()


@pytest.mark.gpu
def test_kafka_shaped_replay_through_micro_batcher_matches_golden():
    model, m, X, P = _model()
    sess = model.Session(m.forest, max_batch=256, n_slots=3)
    got = np.full((1000, 2), np.nan)
    lats = []

    def on_result(first, probs, lat):
        got[first:first + len(probs)] = probs
        lats.append(lat)

    mb = ingest.MicroBatcher(sess, batch=256, on_result=on_result)
    for rec in ingest.KafkaShapedProducer(X[:1000], fmt="csv"):
        mb.push(rec)
    mb.close()  # ragged tail of 232 rows
    assert mb.batches == 4 and len(lats) == 4
    assert np.array_equal(bits(got), bits(P[:1000]))
    # the same stream delivered as consumer polls of 37 messages
    got[:] = np.nan
    mb = ingest.MicroBatcher(sess, batch=256, on_result=on_result)
    recs = list(ingest.KafkaShapedProducer(X[:1000], fmt="json"))
    for i in range(0, len(recs), 37):
        mb.push_many(recs[i:i + 37])
    mb.close()
    assert mb.batches == 4 and np.array_equal(bits(got), bits(P[:1000]))
    sess.close()
    m.close()


@pytest.mark.gpu
def test_coalesced_single_transaction_requests_on_the_gpu():
    _, m, X, P = _model()
    co = sh.RequestCoalescer(lambda M: m.predict(M), 30, max_batch=512, max_delay_s=0.005)
    out = [None] * 64
    th = [threading.Thread(target=lambda i=i: out.__setitem__(i, co.score(X[i][None, :]))) for i in range(64)]
    for t in th:
        t.start()
    for t in th:
        t.join(timeout=60)
    co.close()
    assert np.array_equal(bits(np.concatenate(out)), bits(P[:64]))
    assert co.launches < 64
    m.close()


def test_metrics_registry_exposition_is_prometheus_text():
    from ccfd_demo_summit_b200 import metrics

    reg = metrics.Registry()
    reg.declare("ccfd_requests_total", "counter")
    reg.declare("ccfd_score_seconds", "histogram", metrics.LATENCY_BUCKETS)

    def work():
        for _ in range(100):
            reg.inc("ccfd_requests_total", 1, path="session")
            reg.observe("ccfd_score_seconds", 30e-6, path="session")
            reg.observe("ccfd_score_seconds", 0.5, path="session")

    th = [threading.Thread(target=work) for _ in range(4)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    txt = reg.exposition()
    assert '# TYPE ccfd_requests_total counter' in txt and 'ccfd_requests_total{path="session"} 400' in txt
    assert 'ccfd_score_seconds_bucket{le="5e-05",path="session"} 400' in txt   # cumulative
    assert 'ccfd_score_seconds_bucket{le="+Inf",path="session"} 800' in txt
    assert 'ccfd_score_seconds_count{path="session"} 800' in txt


@pytest.mark.gpu
def test_prometheus_endpoint_carries_model_counters_and_seldon_request_histograms():
    """besides the four gauges, batch-size / scoring-time histograms and the request
    families the SeldonCore board plots (reference deploy/grafana/SeldonCore.json:119,499-531)."""
    _, m, X, P = _model()
    app = sh.SeldonApp(m)
    for n in (1, 7, 300, 900):
        code, _ = app.serve(json.dumps({"data": {"ndarray": X[:n].astype(np.float64).tolist()}}).encode())
        assert code == 200
    code, _ = app.serve(b'{"data":{"ndarray":[[1,2,3]]}}')
    assert code == 400
    txt = sh.prometheus_text(m, app.registry)
    for name in ("# TYPE proba_1 gauge", "# TYPE ccfd_requests_total counter", 'ccfd_requests_total{path="session"} 3',
                 'ccfd_requests_total{path="bulk"} 1', 'ccfd_rows_total{path="bulk"} 900', "# TYPE ccfd_request_rows histogram",
                 'ccfd_request_rows_bucket{le="8"} 2', "# TYPE ccfd_score_seconds histogram", "ccfd_predict_seconds_count",
                 "# TYPE seldon_api_engine_server_requests_seconds histogram", "seldon_api_engine_client_requests_seconds_bucket{"):
        assert name in txt, name
    assert 'seldon_api_engine_server_requests_seconds_count{deployment_name="modelfull"' in txt
    assert 'status="400"} 1' in txt and 'status="200"} 4' in txt
    app.close()
    m.close()


@pytest.mark.gpu
def test_concurrent_predict_calls_overlap_on_different_slots():
    """n_slots requests are in flight at once: no lock is held around copy-in + graph + wait."""
    from ccfd_demo_summit_b200 import converter, model

    raw, X, P, _ = load_golden("small_hard")
    fa = converter.from_arrays(raw.tree_offsets, raw.left, raw.right, raw.feature, raw.threshold,
                               raw.missing_go_to_left, raw.value, raw.n_features, raw.n_classes)
    m = model.CcfdForestModel(fa, session_max_batch=1024, session_slots=4)
    X = np.ascontiguousarray(X, dtype=np.float32)
    errs = []

    def worker(k):
        try:
            for i in range(300):
                lo = (k * 131 + i * 17) % 400
                got = m.predict(X[lo:lo + 1024], None)
                if not np.array_equal(bits(got), bits(P[lo:lo + 1024])):
                    raise AssertionError((k, i))
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    th = [threading.Thread(target=worker, args=(k,)) for k in range(8)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs[:1]
    assert m.max_in_flight > 1, "requests never overlapped"
    x_last, p_last = m._last
    assert np.array_equal(bits(m.predict(x_last[None, :])[0]), bits(p_last))  # the gauge's row and probability belong together
    m.close()


@pytest.mark.gpu
def test_native_batcher_binary_records_replay_and_inf_refusal():
    from ccfd_demo_summit_b200 import _lib

    model, m, X, P = _model()
    sess = model.Session(m.forest, max_batch=256, n_slots=3)
    got = np.full((1500, 2), np.nan)
    lats = []

    def on_result(first, probs, lat):
        got[first:first + len(probs)] = probs
        lats.append(lat)

    mb = ingest.MicroBatcher(sess, batch=256, on_result=on_result)
    for i in range(0, 1500, 97):       # binary records in polls of 97
        mb.push_rows(X[i:i + 97])
    mb.close()
    assert mb.batches == 6 and np.array_equal(bits(got), bits(P[:1500]))
    assert all(0 < v < 1.0 for v in lats)
    # native replay at a fixed rate with a consumer thread collecting results
    got[:] = np.nan
    lats.clear()
    mb = ingest.MicroBatcher(sess, batch=256, on_result=on_result)
    t = threading.Thread(target=lambda: mb.replay(X[:1500], records_per_s=2e6, burst=16))
    t.start()
    while t.is_alive() or mb.returned < 1500:
        if mb.drain(200) == 0 and not t.is_alive() and mb.returned >= 1500:
            break
    t.join()
    mb.close()
    assert np.array_equal(bits(got), bits(P[:1500])) and len(lats) == 6
    # a record with +-inf is refused, as scikit-learn refuses it
    mb = ingest.MicroBatcher(sess, batch=256)
    bad = X[:3].copy()
    bad[1, 4] = -np.inf
    with pytest.raises(ValueError, match="infinity"):
        mb.push_rows(bad)
    mb.close()
    with pytest.raises(ValueError, match="2 slots"):
        ingest.MicroBatcher(model.Session(m.forest, max_batch=64, n_slots=1), batch=64)
    sess.close()
    m.close()


@pytest.mark.gpu
def test_http_server_with_coalescing_merges_concurrent_single_transaction_posts():
    """CCFD_COALESCE / make_server(coalesce=True): single-transaction POSTs (the KIE server's pool of
    clients, reference README.md:386-393, deploy/ccd-service.yaml:61-66) that arrive while a launch is in
    flight share the next launch; every caller still gets its own row's probabilities, bit for bit."""
    import urllib.request

    _, m, X, P = _model()
    srv = sh.make_server(m, port=0, coalesce=True)
    port = srv.server_address[1]
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    results, errs = {}, []

    def client(k):
        try:
            for i in range(25):
                row = (k * 25 + i) % 600
                body = json.dumps({"data": {"ndarray": [X[row].astype(np.float64).tolist()]}}).encode()
                req = urllib.request.Request(f"http://127.0.0.1:{port}/predict", data=body, headers={"Content-Type": "application/json"})
                msg = json.loads(urllib.request.urlopen(req, timeout=30).read())
                results[(k, i)] = (row, np.asarray(msg["data"]["ndarray"], dtype=np.float64))
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    try:
        th = [threading.Thread(target=client, args=(k,)) for k in range(16)]
        for t in th:
            t.start()
        for t in th:
            t.join(timeout=120)
        assert not errs, errs[:1]
        assert len(results) == 400
        for row, got in results.values():
            assert got.shape == (1, 2) and np.array_equal(bits(got), bits(P[row:row + 1]))
        co = srv.app.coalescer
        assert co is not None and co.rows_scored == 400 and co.launches <= 400
    finally:
        srv.shutdown()
        srv.app.close()
        m.close()
