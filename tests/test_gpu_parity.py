"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle on the same
seeded inputs.  Tolerances (BASELINE.json north star): split/merge indices bit-exact;
fp32 arithmetic (LR, first-order, FM) |dp| <= 1e-5; 16-bit tower |dp| <= 2e-3 against the
fp32/fp64 oracle.  A tighter check against the oracle's emulation of the kernel's own
rounding points ("f16" mode) catches layout bugs a loose tolerance would hide."""
import threading

import numpy as np
import pytest

from conftest import make_case

pytestmark = pytest.mark.gpu

TOL_FP32 = 1e-5     # probabilities, fp32 arithmetic path
TOL_F16 = 2e-3      # probabilities, 16-bit tensor-core tower vs the fp64 master
TOL_F16_EMU = 3e-4  # probabilities vs the oracle emulating the kernel's rounding points


@pytest.fixture(scope="module")
def om():
    from oracle import model
    return model


@pytest.fixture(scope="module")
def binding(lib_built):
    from dtfs_b200 import binding
    return binding


def _engine(binding, path, **kw):
    e = binding.Engine(**kw)
    e.load_weights(path)
    return e


# ---------------------------------------------------------------- kernel-level
@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (128, 256, 640), (300, 1024, 640), (1000, 512, 1024),
                                   (77, 128, 128), (257, 640, 624), (129, 384, 200)])
def test_gemm_kernel_alone(binding, om, weight_file, M, N, K):
    """One tower layer through the tcgen05 kernel vs numpy on the same half-rounded operands."""
    w = make_case(om.LR, F=43)
    e = _engine(binding, weight_file(w))
    rng = np.random.default_rng(M * 7 + N)
    A = rng.uniform(-1, 1, (M, K)).astype(np.float32)
    W = rng.uniform(-0.2, 0.2, (N, K)).astype(np.float32)
    b = rng.uniform(-0.5, 0.5, N).astype(np.float32)
    got = e.debug_gemm(A, W, b)
    ref = np.maximum(om.round_f16(A).astype(np.float64) @ om.round_f16(W).astype(np.float64).T + b, 0)
    e.close()
    err = np.abs(got - ref)
    assert err.max() <= 2e-3 * max(1.0, np.abs(ref).max()), "max err %g at %s" % (err.max(), np.unravel_index(err.argmax(), err.shape))


def test_table_rows_bit_exact(binding, om, weight_file):
    """Procedural table on the device == the oracle's formula, bit for bit (integer hash + exact fp32)."""
    for tdt in (om.TABLE_FP32, om.TABLE_F16):
        w = make_case(om.LR, F=39, rows_per_field=1 << 12, table_dtype=tdt)
        e = _engine(binding, weight_file(w))
        rows = np.random.default_rng(5).integers(0, w.spec.num_rows, 4096)
        emb, w1 = e.read_rows(rows)
        oe, ow = w.rows(rows)
        e.close()
        assert np.array_equal(emb.view(np.uint32), oe.view(np.uint32))
        assert np.array_equal(w1.view(np.uint32), ow.view(np.uint32))


def test_raw_table_rows_bit_exact(binding, om, weight_file):
    w = make_case(om.LR, F=43, procedural=False)
    e = _engine(binding, weight_file(w))
    rows = np.arange(w.spec.num_rows)
    emb, w1 = e.read_rows(rows)
    e.close()
    assert np.array_equal(emb, w.emb) and np.array_equal(w1, w.w1)


@pytest.mark.parametrize("kind_name,n", [("wide_deep", 1000), ("deepfm", 333), ("dcn_v2", 64)])
def test_gather_kernel(binding, om, weight_file, kind_name, n):
    kind = {"wide_deep": om.WIDE_DEEP, "deepfm": om.DEEPFM, "dcn_v2": om.DCN_V2}[kind_name]
    w = make_case(kind, hidden=(256, 128), n_cross=2 if kind == om.DCN_V2 else 0)
    e = _engine(binding, weight_file(w))
    ids, wts = om.criteo_inputs(n, rows_per_field=1 << 10)
    x, wide = e.debug_gather(ids, wts)
    e.close()
    emb, w1, wt = om._gather(w, ids, wts)
    x_ref = om.round_f16((emb * wt[..., None]).reshape(n, -1))
    assert np.array_equal(x.view(np.uint32), x_ref.view(np.uint32)), "X differs from half(E*wt)"
    if kind != om.DCN_V2:
        wide_ref = om._wide_part(w, ids, wts)
        assert np.abs(wide - wide_ref).max() <= 2e-5 * max(1.0, np.abs(wide_ref).max())


# ---------------------------------------------------------------- config 1: the reference's own case
def test_lr_reference_demo_inputs(binding, om, weight_file):
    """BASELINE config 1 with the reference's inputs (DCNClient.java:57-74): ids 1..43, wts 1.0.
    Known answers (-ii): every score equal, and equal to the closed form."""
    from dtfs_b200 import dcn_client as dc
    from oracle import dcn_client as odc
    w = make_case(om.LR, F=43, procedural=False)
    path = weight_file(w)
    chans = dc.getChannels([0, 0], path, max_candidates=2000)          # 2 "servers" (config 1)
    ids_l, wts_l = dc.getFeatureIds(2, 1000, 43), dc.getFeatureWts(2, 1000, 43)
    got = dc.processARequest(chans, ids_l, wts_l, sort=False)
    got_sorted = dc.processARequest(chans, ids_l, wts_l, sort=True)
    dc.shutDownChannels(chans)
    z = np.float64(w.b0) + w.w1[1:44].astype(np.float64).sum()
    p = 1.0 / (1.0 + np.exp(-z))
    assert got.shape == (1000,)
    assert np.all(got == got[0]), "identical candidates must get identical scores"
    assert abs(float(got[0]) - p) <= TOL_FP32
    ref = odc.process_a_request([om.as_model(w, "fp32")] * 2, odc.get_feature_ids(1000, 43, 2),
                                odc.get_feature_wts(1000, 43, 2), 43, sort=True)
    assert np.abs(got_sorted - ref).max() <= TOL_FP32


@pytest.mark.parametrize("C,P", [(1000, 2), (1500, 3), (7, 3), (1, 1), (5, 8)])
def test_lr_randomised_split_merge(binding, om, weight_file, C, P):
    """Randomised config 1 (ids uniform in [1,43], wts ~ U(0,1), seed 4321): output index j is
    candidate j whatever the split; ragged splits send the remainder to the last shard."""
    from dtfs_b200 import dcn_client as dc
    w = make_case(om.LR, F=43, procedural=False)
    rng = np.random.default_rng(4321)
    ids = rng.integers(1, 44, (C, 43)).astype(np.int64)
    wts = rng.uniform(0, 1, (C, 43)).astype(np.float32)
    ref = om.forward(w, ids, wts, "fp64")
    path = weight_file(w)
    chans = dc.getChannels([0] * P, path, max_candidates=2048)
    bounds = binding.partition_bounds(C, P)
    got = dc.processARequest(chans, [ids[a:b] for a, b in bounds], [wts[a:b] for a, b in bounds], sort=False)
    dc.shutDownChannels(chans)
    assert got.shape == (C,)
    assert np.abs(got - ref).max() <= TOL_FP32


# ---------------------------------------------------------------- tower models end to end
CASES = {
    "wide_deep": dict(hidden=(1024, 512, 256)),
    "wide_deep_small": dict(hidden=(256, 128)),
    "deepfm": dict(hidden=(512, 512, 512)),
    "dcn_v2": dict(hidden=(1024, 512, 256), n_cross=3),
    "wide_deep_f16_table": dict(hidden=(1024, 512, 256), table_dtype=1),
    "dcn_v2_hashed": dict(hidden=(512, 256), n_cross=2, id_mode=1),
}


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("n", [1, 127, 1000])
def test_tower_models(binding, om, weight_file, name, n):
    kw = dict(CASES[name])
    kind = om.DEEPFM if name.startswith("deepfm") else om.DCN_V2 if name.startswith("dcn") else om.WIDE_DEEP
    w = make_case(kind, **kw)
    e = _engine(binding, weight_file(w), max_candidates=4096)
    ids, wts = om.criteo_inputs(n, seed=4321 + n, rows_per_field=1 << 10)
    got = e.predict(ids, wts)
    e.close()
    p64 = om.forward(w, ids, wts, "fp64")
    pemu = om.forward(w, ids, wts, "f16")
    z64 = om.forward_logits(w, ids, wts, "fp64")
    assert got.shape == (n,)
    assert np.all(np.isfinite(got))
    if n >= 100:
        assert np.std(z64) > 0.5, "logits too flat for an absolute tolerance on p to mean anything"
    assert np.abs(got - pemu).max() <= TOL_F16_EMU, "vs kernel-contract oracle: %g" % np.abs(got - pemu).max()
    assert np.abs(got - p64).max() <= TOL_F16, "vs fp64 master: %g" % np.abs(got - p64).max()


def test_empty_request(binding, om, weight_file):
    w = make_case(om.WIDE_DEEP, hidden=(256, 128))
    e = _engine(binding, weight_file(w))
    out = e.predict(np.zeros((0, 39), np.int64), np.zeros((0, 39), np.float32))
    e.close()
    assert out.shape == (0,)


def test_multi_chunk_and_pinned(binding, om, weight_file, monkeypatch):
    """A request larger than the activation chunk is scored chunk by chunk; pinned caller
    buffers are used in place.  Results must equal the one-shot ones bit for bit."""
    w = make_case(om.WIDE_DEEP, hidden=(256, 128))
    path = weight_file(w)
    n = 3000
    ids, wts = om.criteo_inputs(n, rows_per_field=1 << 10)
    e = _engine(binding, path, max_candidates=4096)
    base = e.predict(ids, wts).copy()
    pi, pw, po = e.pinned((n, 39), np.int64), e.pinned((n, 39), np.float32), e.pinned((n,), np.float32)
    pi.array[:] = ids
    pw.array[:] = wts
    got = e.predict(pi.array, pw.array, out=po.array).copy()
    for b in (pi, pw, po):
        b.free()
    e.close()
    assert np.array_equal(got, base)
    monkeypatch.setenv("CTR_CHUNK_ROWS", "512")
    e = _engine(binding, path, max_candidates=4096)
    chunked = e.predict(ids, wts)
    e.close()
    assert np.array_equal(chunked, base), "chunking changed results"


def test_concurrent_requests(binding, om, weight_file):
    """6 driver threads sharing read-only inputs (DCNClient.java:213-224) get identical answers."""
    w = make_case(om.DEEPFM, hidden=(256, 128))
    e = _engine(binding, weight_file(w), max_candidates=2048, workers_per_device=3)
    ids, wts = om.criteo_inputs(1500, rows_per_field=1 << 10)
    base = e.predict(ids, wts).copy()
    results, errors = [], []

    def drive():
        try:
            for _ in range(10):
                results.append(e.predict(ids, wts).copy())
        except Exception as ex:      # noqa: BLE001
            errors.append(ex)
    ts = [threading.Thread(target=drive) for _ in range(6)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    e.close()
    assert not errors, errors
    assert len(results) == 60 and all(np.array_equal(r, base) for r in results)


def test_errors_are_loud(binding, om, weight_file, tmp_path):
    w = make_case(om.LR, F=43, procedural=False)
    e = _engine(binding, weight_file(w), max_candidates=100)
    with pytest.raises(binding.CtrError) as ei:
        e.predict(np.ones((101, 43), np.int64), np.ones((101, 43), np.float32))
    assert ei.value.code == -6
    with pytest.raises(binding.CtrError) as ei:                        # id == num_rows: out of range
        e.predict(np.full((3, 43), w.spec.num_rows, np.int64), np.ones((3, 43), np.float32))
    assert ei.value.code == -1
    ok = e.predict(np.ones((3, 43), np.int64), np.ones((3, 43), np.float32))   # engine still usable
    assert np.all(np.isfinite(ok))
    bad = tmp_path / "bad.ctrw"
    bad.write_bytes(b"nonsense")
    with pytest.raises(binding.CtrError) as ei:
        e.load_weights(str(bad))
    assert ei.value.code == -3
    e.close()


def test_device_resident_matches_host_path(binding, om, weight_file):
    import torch
    w = make_case(om.WIDE_DEEP, hidden=(1024, 512, 256))
    e = _engine(binding, weight_file(w), max_candidates=20000)
    n = 20000
    ids, wts = om.criteo_inputs(n, rows_per_field=1 << 10)
    host = e.predict(ids, wts).copy()
    d_ids = torch.from_numpy(ids).cuda()
    d_wts = torch.from_numpy(wts).cuda()
    d_out = torch.empty(n, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    e.score_device(d_ids.data_ptr(), d_wts.data_ptr(), n, d_out.data_ptr())
    a = d_out.cpu().numpy()
    st = torch.cuda.Stream()
    d_out2 = torch.empty_like(d_out)
    with torch.cuda.stream(st):
        e.score_device(d_ids.data_ptr(), d_wts.data_ptr(), n, d_out2.data_ptr(), stream=st.cuda_stream)
    st.synchronize()
    e.close()
    assert np.array_equal(a, host)
    assert np.array_equal(d_out2.cpu().numpy(), host)
    assert e is not None


# ---------------------------------------------------------------- A6: the sort of processARequest, on the device
@pytest.mark.parametrize("n", [1, 257, 10000, 70000])
def test_predict_sorted_matches_java_order(binding, om, weight_file, n):
    from oracle import dcn_client as odc
    w = make_case(om.DEEPFM, hidden=(256, 128))
    e = _engine(binding, weight_file(w), max_candidates=70000)
    ids, wts = om.criteo_inputs(n, seed=n, rows_per_field=1 << 10)
    ids[n // 2:] = ids[:n - n // 2]                       # duplicates -> ties, to exercise stability
    wts[n // 2:] = wts[:n - n // 2]
    plain = e.predict(ids, wts).copy()
    srt, order = e.predict_sorted(ids, wts)
    e.close()
    ref_order = np.argsort(odc.java_float_sort_key(plain), kind="stable")
    assert np.array_equal(order, ref_order.astype(np.int32)), "candidate indices of the sorted scores (stable)"
    assert np.array_equal(srt, plain[ref_order]) and np.array_equal(srt, odc.java_float_sort(plain))


def test_predict_sorted_reference_inputs_and_special_values(binding, om, weight_file):
    """With the reference's demo inputs every score is equal: the sort is the identity permutation."""
    w = make_case(om.LR, F=43, procedural=False)
    e = _engine(binding, weight_file(w), max_candidates=2048)
    ids = np.tile(np.arange(1, 44, dtype=np.int64), (1500, 1))
    srt, order = e.predict_sorted(ids, np.ones((1500, 43), np.float32))
    assert np.all(srt == srt[0]) and np.array_equal(order, np.arange(1500, dtype=np.int32))
    # weights chosen to produce +-0.0-like extremes: sigmoid saturates to exactly 0.0 and 1.0
    wts = np.zeros((4, 43), np.float32)
    wts[0, :] = 1e30; wts[1, :] = -1e30; wts[2, :] = 0.0; wts[3, :] = 1.0
    srt, order = e.predict_sorted(np.tile(np.arange(1, 44, dtype=np.int64), (4, 1)), wts)
    e.close()
    assert np.all(np.diff(srt[~np.isnan(srt)]) >= 0) and sorted(order.tolist()) == [0, 1, 2, 3]


# ---------------------------------------------------------------- f2: the RPC itself, wire format in and out
def test_predict_wire_is_a_drop_in_for_the_rpc(binding, om, weight_file):
    from oracle import clib
    w = make_case(om.WIDE_DEEP, hidden=(256, 128))
    e = _engine(binding, weight_file(w), max_candidates=4096)
    ids, wts = om.criteo_inputs(1000, rows_per_field=1 << 10)
    want = e.predict(ids, wts).copy()
    resp = e.predict_wire(clib.encode_predict_request(ids, wts))           # built like DCNClient.java:97-109
    got = clib.decode_predict_response(resp, 1000)                         # read like DCNClient.java:162
    assert np.array_equal(got, want)
    with pytest.raises(binding.CtrError) as ei:                             # shape [n, 38] for a 39-field model
        e.predict_wire(clib.encode_predict_request(ids[:, :38].copy(), wts[:, :38].copy()))
    assert ei.value.code == -1
    e.close()


# ---------------------------------------------------------------- f1: shared-context requests
@pytest.mark.parametrize("n_ctx", [1, 13, 26, 38])
def test_shared_context_equals_replicated_fields(binding, om, weight_file, n_ctx):
    """The context fields given once must score exactly like the reference layout that repeats them
    for every candidate (DCNClient.java:61-63)."""
    w = make_case(om.DEEPFM, hidden=(256, 128))
    e = _engine(binding, weight_file(w), max_candidates=4096)
    ids, wts = om.criteo_inputs(2001, seed=n_ctx, rows_per_field=1 << 10)
    ids[:, :n_ctx] = ids[0, :n_ctx]                         # a real request: same user / context for all candidates
    wts[:, :n_ctx] = wts[0, :n_ctx]
    want = e.predict(ids, wts).copy()
    t, out = e.predict_shared_async(ids[0, :n_ctx], wts[0, :n_ctx], np.ascontiguousarray(ids[:, n_ctx:]), np.ascontiguousarray(wts[:, n_ctx:]))
    got = e.wait(t)
    e.close()
    assert np.array_equal(got, want)
    assert np.abs(want - om.forward(w, ids, wts, "fp64")).max() <= TOL_F16


def test_callback_and_timeout(binding, om, weight_file):
    """Completion callback (what the Java shim hangs its CompletableFuture on) and ctr_wait's timeout."""
    import threading
    w = make_case(om.WIDE_DEEP, hidden=(1024, 512, 256))
    e = _engine(binding, weight_file(w), max_candidates=16384, workers_per_device=1)
    ids, wts = om.criteo_inputs(16000, rows_per_field=1 << 10)
    done = threading.Event()
    seen = {}

    def cb(ticket, status):
        seen["ticket"], seen["status"] = ticket, status
        done.set()
    tickets = [e.predict_async(ids, wts)[0] for _ in range(40)]            # keep the single worker busy for tens of ms
    t_cb, out = e.predict_async(ids, wts, callback=cb)
    with pytest.raises(binding.CtrError) as ei:                             # not ready within 0 ms behind 40 requests
        e.wait(t_cb, timeout_ms=0)
    assert ei.value.code == -5
    assert done.wait(30) and seen == {"ticket": t_cb, "status": 0}
    got = e.wait(t_cb)                                                      # the ticket survived the timeout
    for t in tickets:
        e.wait(t)
    with pytest.raises(binding.CtrError) as ei:                             # released tickets are unknown
        e.wait(t_cb)
    assert ei.value.code == -4
    assert np.array_equal(got, e.predict(ids, wts))
    e.close()


def test_called_back_tickets_are_released(binding, om, weight_file):
    """A caller that completes its future in the callback never waits (the open-loop driver, the Java
    shim): the engine releases those tickets itself, keeping the 4096 most recent for a late ctr_wait."""
    import threading
    w = make_case(om.LR)
    e = _engine(binding, weight_file(w), max_candidates=64)
    ids, wts = om.criteo_inputs(8, rows_per_field=1 << 10)
    total = 4096 + 300
    left = threading.Semaphore(0)
    bad = []

    def cb(ticket, status):
        if status != 0:
            bad.append((ticket, status))
        left.release()
    outs = [np.empty(8, np.float32) for _ in range(64)]
    tickets = []
    for i in range(total):
        if i >= 64:
            assert left.acquire(timeout=30)                                 # at most 64 in flight: outs[] slots are reused
        tickets.append(e.predict_async(ids, wts, out=outs[i % 64], callback=cb)[0])
    for _ in range(64):
        assert left.acquire(timeout=30)
    assert not bad
    e.wait(tickets[-1])                                                     # a recent one is still there
    with pytest.raises(binding.CtrError) as ei:                             # the oldest were released by the engine
        e.wait(tickets[0])
    assert ei.value.code == -4
    assert np.array_equal(outs[(total - 1) % 64], e.predict(ids, wts))
    e.close()


# ---------------------------------------------------------------- odd shapes: padding paths
@pytest.mark.parametrize("kind_name,F,hidden,n_cross", [
    ("wide_deep", 48, (200,), 0),             # widest field count, one hidden layer padded 200 -> 256
    ("wide_deep", 14, (72, 40), 0),           # 13 dense + 1 sparse, tiny layers padded to 128
    ("deepfm", 39, (640, 384, 128, 64), 0),   # four hidden layers, BN = 128 tiles
    ("dcn_v2", 20, (136, 8), 1),              # d = 320 (pad 384), single cross layer, last hidden layer of 8
    ("dcn_v2", 43, (256, 256), 4),            # reference FIELD_NUM, four cross layers (d = 688, pad 768)
])
def test_odd_shapes(binding, om, weight_file, kind_name, F, hidden, n_cross):
    kind = {"wide_deep": om.WIDE_DEEP, "deepfm": om.DEEPFM, "dcn_v2": om.DCN_V2}[kind_name]
    rpf = 1 << 9
    spec = om.ModelSpec(kind, F, 16, om.criteo_num_rows(13, F - 13, rpf), hidden=tuple(hidden), n_cross=n_cross,
                        table_mode=om.TABLE_PROCEDURAL, scale_e=0.25)
    w = om.make_weights(spec, calib_inputs=om.criteo_inputs(256, seed=99, n_sparse=F - 13, rows_per_field=rpf))
    e = _engine(binding, weight_file(w), max_candidates=2048)
    for n in (1, 300, 1025):
        ids, wts = om.criteo_inputs(n, seed=n, n_sparse=F - 13, rows_per_field=rpf)
        got = e.predict(ids, wts)
        p64 = om.forward(w, ids, wts, "fp64")
        pemu = om.forward(w, ids, wts, "f16")
        # 2x the usual emulation tolerance: with four layers (and DeepFM's larger logits) more activations
        # sit on a half-precision rounding boundary, where the accumulation order decides the rounding
        assert np.abs(got - pemu).max() <= 2 * TOL_F16_EMU, "n=%d vs kernel-contract oracle: %g" % (n, np.abs(got - pemu).max())
        assert np.abs(got - p64).max() <= TOL_F16, "n=%d vs fp64 master: %g" % (n, np.abs(got - p64).max())
    e.close()
