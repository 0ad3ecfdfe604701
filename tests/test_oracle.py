"""CPU tests of the oracle: the reference-defined pieces against independent closed forms and
the committed golden vectors; the numpy restatement against the C restatement."""
import os

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from oracle import clib, dcn_client as odc, model as om

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


# ---------------------------------------------------------------- DC:46-55 partitionList
@given(st.integers(0, 200000), st.integers(1, 64))
@settings(max_examples=200, deadline=None)
def test_partition_rule_properties(n, p):
    b = odc.partition_bounds(n, p)
    assert len(b) == p and b[0][0] == 0 and b[-1][1] == n
    assert all(b[i][1] == b[i + 1][0] for i in range(p - 1)), "parts are contiguous"
    assert all(e - s == n // p for s, e in b[:-1]), "first P-1 parts have floor(N/P) elements"
    assert b[-1][1] - b[-1][0] == n - (p - 1) * (n // p), "the last part takes the remainder"
    assert b == clib.partition_bounds(n, p), "numpy and C restatements agree"


def test_partition_golden_and_reference_defaults():
    g = np.load(os.path.join(GOLD, "partition_bounds.npz"))
    for k in g.files:
        n, p = map(int, k.split("_"))
        assert np.array_equal(np.array(odc.partition_bounds(n, p), np.int64), g[k])
    # the reference's own run: 1500 candidates x 43 fields over 3 hosts (DC:25, 29, 38)
    assert odc.partition_bounds(1500 * 43, 3) == [(0, 21500), (21500, 43000), (43000, 64500)]
    # element split is candidate-aligned only when P divides C
    assert all(s % 43 == 0 for s, _ in odc.partition_bounds(1500 * 43, 3))
    assert any(s % 43 for s, _ in odc.partition_bounds(1000 * 43, 3))
    with pytest.raises(ZeroDivisionError):
        odc.partition_bounds(10, 0)


def test_demo_inputs_and_malformed_shard():
    ids, wts = odc.get_feature_ids(), odc.get_feature_wts()          # DC:57-74 defaults
    assert [len(x) for x in ids] == [21500] * 3 and [len(x) for x in wts] == [21500] * 3
    flat = np.concatenate(ids)
    assert np.array_equal(flat.reshape(1500, 43), np.tile(np.arange(1, 44), (1500, 1)))
    assert np.all(np.concatenate(wts) == 1.0)
    assert odc.get_tensor_shape(21500) == (500, 43)                   # DC:97
    assert odc.get_tensor_shape(14334) == (333, 43)                   # truncation: declared shape < value count
    model = lambda i, w: np.zeros(i.shape[0], np.float32)
    bad = odc.get_feature_ids(1000, 43, 3)[0]                          # 14333 elements: not a multiple of 43
    with pytest.raises(ValueError):
        odc.send_request(model, bad, np.ones(len(bad), np.float32))


# ---------------------------------------------------------------- DC:195 sort order
def test_java_float_sort_order():
    g = np.load(os.path.join(GOLD, "java_sort.npz"))
    s = odc.java_float_sort(g["x"])
    assert np.array_equal(s.view(np.uint32), g["sorted"].view(np.uint32))
    assert np.array_equal(odc.java_float_sort_key(g["x"]), clib.java_sort_keys(g["x"]))
    # Float.compareTo: -inf < negatives < -0.0 < +0.0 < positives < +inf < NaN
    assert s[0] == -np.inf and np.isnan(s[-1]) and s[-2] == np.inf
    i = int(np.where(s == 0)[0][0])
    assert np.signbit(s[i]) and not np.signbit(s[i + 1])
    rng = np.random.default_rng(0)
    x = rng.normal(size=1000).astype(np.float32)
    assert np.array_equal(odc.java_float_sort(x), np.sort(x))


# ---------------------------------------------------------------- integer hash / procedural weights
def test_mix64_and_procedural_values():
    g = np.load(os.path.join(GOLD, "mix64.npz"))
    assert np.array_equal(om.mix64(g["z"]), g["h"])
    assert np.array_equal(clib.mix64(g["z"]), g["h"])
    # splitmix64 known answer: first output of the generator seeded with 0
    assert int(om.mix64(np.uint64(0))) == 0xE220A8397B1DCDAF
    u = om.proc_unit(1234, 1, g["z"])
    assert np.array_equal(u.view(np.uint32), g["unit"].view(np.uint32))
    assert np.array_equal(clib.proc_unit(1234, 1, g["z"]).view(np.uint32), u.view(np.uint32))
    big = om.proc_unit(7, 2, np.arange(200000, dtype=np.uint64))
    assert big.min() >= -1 and big.max() < 1 and abs(big.mean()) < 0.01 and abs(big.std() - 3 ** -0.5) < 0.01
    emb, w1 = clib.fill_table(1234, 0.25, 0.5, 1000, 16)
    spec = om.ModelSpec(om.LR, 39, 16, 1000, table_mode=om.TABLE_PROCEDURAL, seed=1234, scale_e=0.25, scale_w1=0.5)
    oe, ow = om.Weights(spec).rows(np.arange(1000))
    assert np.array_equal(emb, oe) and np.array_equal(w1, ow)


def test_hashed_rows():
    ids = np.array([0, 1, -1, 2 ** 40, -2 ** 62], np.int64)
    rows = om.id_to_row(ids, 1 << 20, om.ID_HASHED)
    assert rows.max() < (1 << 20)
    out = np.empty(5, np.uint64)
    clib.lib().oracle_id_to_row(ids.ctypes.data_as(clib.C.c_void_p), clib.C.c_int64(5), clib.C.c_int64(1 << 20),
                                clib.C.c_int32(1), out.ctypes.data_as(clib.C.c_void_p))
    assert np.array_equal(rows, out)
    with pytest.raises(IndexError):
        om.id_to_row(np.array([5]), 5, om.ID_DIRECT)


# ---------------------------------------------------------------- wire format (tensor.proto:48,64; predict.proto)
def _pb_varint(v):
    from google.protobuf.internal import encoder
    return encoder._VarintBytes(v & 0xFFFFFFFFFFFFFFFF)


def test_wire_codec_against_protobuf_runtime():
    ids, wts = om.criteo_inputs(7, n_sparse=5, rows_per_field=1 << 20)
    ids[0, 0] = -3                                                      # negative int64 -> 10-byte varint
    buf = clib.encode_predict_request(ids, wts)
    i2, w2 = clib.decode_predict_request(buf, ids.size)
    assert np.array_equal(i2, ids) and np.array_equal(w2, wts)
    # the packed int64_val payload is exactly the protobuf runtime's varints, in order
    payload = b"".join(_pb_varint(int(v)) for v in ids.ravel())
    assert payload in buf
    assert wts.astype("<f4").tobytes() in buf                           # float_val: packed fixed32
    assert b"feat_ids" in buf and b"feat_wts" in buf and b"DCN" in buf and b"serving_default" in buf
    # a generic walk of the top-level message finds field 1 (model_spec) and two field-2 entries (inputs)
    from google.protobuf.internal import decoder
    pos, fields, ends = 0, [], []
    while pos < len(buf):
        tag, pos = decoder._DecodeVarint(buf, pos)
        ln, pos = decoder._DecodeVarint(buf, pos)
        fields.append(tag >> 3)
        pos += ln
        ends.append(pos)
    assert fields == [1, 2, 2] and pos == len(buf)
    s = np.array([0.25, 0.5, 1.0], np.float32)
    assert np.array_equal(clib.decode_predict_response(clib.encode_predict_response(s), 3), s)
    with pytest.raises(ValueError):                                     # feat_wts entry missing
        clib.decode_predict_request(buf[: ends[1]], ids.size)


def test_reference_wire_payload_size():
    """one reference shard (500 x 43, ids <= 43) is ~21.5 KB of varints + 86 KB of floats."""
    ids = np.tile(np.arange(1, 44, dtype=np.int64), (500, 1))
    buf = clib.encode_predict_request(ids, np.ones((500, 43), np.float32))
    assert 21500 + 86000 < len(buf) < 21500 + 86000 + 200


# ---------------------------------------------------------------- model math
@pytest.mark.parametrize("name", ["lr", "wide_deep", "deepfm", "dcn_v2"])
def test_model_golden(name):
    w = om.load_weights(os.path.join(GOLD, name + ".ctrw"))
    io = np.load(os.path.join(GOLD, name + "_io.npz"))
    z = om.forward_logits(w, io["ids"], io["wts"], "fp64")
    assert np.allclose(z, io["logits_fp64"], rtol=0, atol=1e-12)
    assert np.abs(om.forward(w, io["ids"], io["wts"], "fp32") - io["p_fp64"]).max() <= 1e-5
    assert np.abs(om.forward(w, io["ids"], io["wts"], "f16") - io["p_fp64"]).max() <= 2e-3
    if name != "lr":
        assert np.std(io["logits_fp64"]) > 0.5


def test_lr_closed_form_and_identical_candidates():
    """Known answers: LR has a closed form; the reference's demo inputs make every
    candidate identical, so every score is."""
    spec = om.ModelSpec(om.LR, 43, 16, 44)
    w = om.make_weights(spec)
    ids = np.tile(np.arange(1, 44, dtype=np.int64), (1500, 1))
    wts = np.ones((1500, 43), np.float32)
    p = om.forward(w, ids, wts, "fp64")
    z = float(w.b0) + float(w.w1[1:44].astype(np.float64).sum())
    assert np.allclose(p, 1 / (1 + np.exp(-z)), atol=1e-15) and np.all(p == p[0])
    assert np.abs(clib.lr_scores(ids, wts, w.w1, float(w.b0)) - p).max() <= 1e-6
    models = [om.as_model(w, "fp32")] * 3
    merged = odc.process_a_request(models, odc.get_feature_ids(), odc.get_feature_wts(), sort=False)
    assert merged.shape == (1500,) and np.all(merged == merged[0])


def test_fm_identity_and_c_gather():
    spec = om.ModelSpec(om.DEEPFM, 39, 16, om.criteo_num_rows(rows_per_field=64), hidden=(32,), scale_e=0.25)
    w = om.make_weights(spec)
    ids, wts = om.criteo_inputs(20, rows_per_field=64)
    e, w1, wt = om._gather(w, ids, wts)
    x = e.astype(np.float64) * wt[..., None]
    pair = sum((x[:, i] * x[:, j]).sum(1) for i in range(39) for j in range(i + 1, 39))       # sum_{i<j} <x_i, x_j>
    lin = float(w.b0) + (w1.astype(np.float64) * wt).sum(1)
    assert np.allclose(om._wide_part(w, ids, wts), lin + pair, atol=1e-9)
    x0, wide = clib.gather_fp32(ids, wts, spec.num_rows, False, w.emb, w.w1, float(w.b0), True)
    assert np.allclose(x0, x.reshape(20, -1), atol=1e-6) and np.allclose(wide, lin + pair, atol=1e-4)


def test_weight_file_roundtrip_and_product_writer(tmp_path):
    import dtfs_b200  # noqa: F401
    from dtfs_b200 import weights as pw
    spec = om.ModelSpec(om.DCN_V2, 18, 16, om.criteo_num_rows(13, 5, 32), hidden=(64, 32), n_cross=2)
    w = om.make_weights(spec)
    a, b = str(tmp_path / "a.ctrw"), str(tmp_path / "b.ctrw")
    om.save_weights(w, a)
    pw.write_weight_file(b, model_kind=pw.MODEL_DCN_V2, num_fields=18, num_rows=spec.num_rows, b0=float(w.b0),
                         w1=w.w1, emb=w.emb, seed=spec.seed, scale_e=spec.scale_e, scale_w1=spec.scale_w1,
                         mlp=w.mlp, cross=w.cross, w_out=w.w_out, b_out=float(w.b_out))
    assert open(a, "rb").read() == open(b, "rb").read(), "oracle and product writers must be byte-identical"
    w2 = om.load_weights(a)
    ids, wts = om.criteo_inputs(9, n_sparse=5, rows_per_field=32)
    assert np.array_equal(om.forward_logits(w, ids, wts), om.forward_logits(w2, ids, wts))
    h = pw.read_header(a)
    assert h["model_kind"] == 3 and h["hidden"] == [64, 32] and h["n_cross"] == 2 and h["num_fields"] == 18


def test_cpu_baseline_matches_oracle():
    from oracle import cpu_baseline as cb
    spec = om.ModelSpec(om.WIDE_DEEP, 39, 16, om.criteo_num_rows(rows_per_field=256), hidden=(64, 32),
                        table_mode=om.TABLE_PROCEDURAL, scale_e=0.25)
    w = om.make_weights(spec)
    ids, wts = om.criteo_inputs(501, rows_per_field=256)
    c = cb.CpuClient(w, n_shards=2, with_wire=True)
    got = c.process_a_request(ids, wts, sort=False)
    got_sorted = c.process_a_request(ids, wts, sort=True)
    c.close()
    ref = om.forward(w, ids, wts, "fp64")
    assert np.abs(got - ref).max() <= 1e-5
    assert np.array_equal(got_sorted, odc.java_float_sort(got))
