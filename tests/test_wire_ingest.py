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
