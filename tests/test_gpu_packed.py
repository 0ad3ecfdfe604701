"""ctr_predict_packed_async / ctr_pack_candidates: the narrow marshaller (uint32 table rows, weights only
where a field has one, optional shared context) must score bit for bit what the reference layout of
sendRequest (DCNClient.java:97-108: an int64 id and a float weight per field slot) scores."""
import numpy as np
import pytest

from conftest import make_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def om():
    from oracle import model
    return model


@pytest.fixture(scope="module")
def binding(lib_built):
    from dtfs_b200 import binding
    return binding


@pytest.mark.parametrize("kind_name,table_dtype,id_mode", [("wide_deep", 0, 0), ("deepfm", 1, 0), ("dcn_v2", 0, 1), ("lr", 0, 0),
                                                             ("wide_deep", 1, 1)])
@pytest.mark.parametrize("n_ctx", [0, 13, 30])
def test_packed_equals_reference_layout(binding, om, weight_file, kind_name, table_dtype, id_mode, n_ctx):
    kind = {"lr": om.LR, "wide_deep": om.WIDE_DEEP, "deepfm": om.DEEPFM, "dcn_v2": om.DCN_V2}[kind_name]
    hidden = () if kind == om.LR else (256, 128)
    w = make_case(kind, hidden=hidden, n_cross=2 if kind == om.DCN_V2 else 0, table_dtype=table_dtype, id_mode=id_mode)
    e = binding.Engine(device_ids=(0,), max_candidates=4096)
    e.load_weights(weight_file(w))
    try:
        for n in (1, 3, 4, 2001):                                   # ragged and full warp tiles
            ids, wts = om.criteo_inputs(n, seed=100 + n + n_ctx, rows_per_field=1 << 10)
            if id_mode == om.ID_HASHED:
                ids[:, 13:] = ((ids[:, 13:].astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)) >> np.uint64(2)).astype(np.int64)   # 62-bit ids: only their hash fits 32 bits
            ids[:, :n_ctx] = ids[0, :n_ctx]
            wts[:, :n_ctx] = wts[0, :n_ctx]
            want = e.predict(ids, wts).copy()
            lay = e.packed_layout(n_ctx, weighted_fields=range(13))  # dense fields carry weights, sparse ones weigh 1.0
            ctx_i, ctx_w, rows, pw = e.pack_candidates(lay, ids, wts)
            assert rows.dtype == np.uint32 and rows.shape == (n, 39 - n_ctx) and pw.shape == (n, max(0, 13 - n_ctx))
            assert np.array_equal(ctx_i, ids[0, :n_ctx]) and np.array_equal(ctx_w, wts[0, :n_ctx])
            assert np.array_equal(rows.astype(np.uint64), om.id_to_row(ids[:, n_ctx:], w.spec.num_rows, id_mode).astype(np.uint64))
            t, out = e.predict_packed_async(lay, ctx_i, ctx_w, rows, pw)
            got = e.wait(t)
            assert np.array_equal(got, want), "n=%d: max |d| = %g" % (n, np.abs(got - want).max())
            # the dense fields' ids are the same for every candidate: flagged fixed-id they travel once, 4 bytes less per field
            lay2 = e.packed_layout(n_ctx, weighted_fields=range(13), fixed_id_fields=range(13))
            fk2, wk2, nfx = e.packed_shape(lay2)
            assert (fk2, wk2, nfx) == (39 - max(n_ctx, 13), max(0, 13 - n_ctx), max(0, 13 - n_ctx))
            ci2, cw2, rows2, pw2 = e.pack_candidates(lay2, ids, wts)
            assert rows2.shape == (n, fk2) and ci2.size == n_ctx + nfx
            assert np.array_equal(rows2, rows[:, rows.shape[1] - fk2:])
            t, out = e.predict_packed_async(lay2, ci2, cw2, rows2, pw2)
            assert np.array_equal(e.wait(t), want), "fixed-id layout, n=%d" % n
        tol = 1e-5 if kind == om.LR else 2e-3
        assert np.abs(want - om.forward(w, ids, wts, "fp64")).max() <= tol
    finally:
        e.close()


def test_packed_all_fields_weighted_and_pinned_buffers(binding, om, weight_file):
    """Every field weighted (the layout degenerates to uint32 ids + all weights) and buffers from ctr_alloc_pinned."""
    w = make_case(om.DEEPFM, hidden=(128,))
    e = binding.Engine(device_ids=(0,), max_candidates=4096)
    e.load_weights(weight_file(w))
    try:
        n = 1500
        ids, wts = om.criteo_inputs(n, seed=5, rows_per_field=1 << 10)
        wts[:, 13:] = np.random.default_rng(1).uniform(0.5, 1.5, (n, 26)).astype(np.float32)
        want = e.predict(ids, wts).copy()
        lay = e.packed_layout(0, weighted_fields=range(39))
        pr, pw, po = e.pinned((n, 39), np.uint32), e.pinned((n, 39), np.float32), e.pinned((n,), np.float32)
        e.pack_candidates(lay, ids, wts, rows_out=pr.array, wts_out=pw.array)
        t, out = e.predict_packed_async(lay, [], [], pr.array, pw.array, out=po.array)
        assert np.array_equal(e.wait(t), want)
        for b in (pr, pw, po):
            b.free()
    finally:
        e.close()


def test_pack_candidates_refuses_what_it_cannot_represent(binding, om, weight_file):
    w = make_case(om.WIDE_DEEP, hidden=(128,))
    e = binding.Engine(device_ids=(0,), max_candidates=1024)
    e.load_weights(weight_file(w))
    try:
        ids, wts = om.criteo_inputs(16, seed=6, rows_per_field=1 << 10)
        lay = e.packed_layout(0, weighted_fields=range(13))
        bad = wts.copy()
        bad[3, 20] = 0.5                                             # a sparse field with a weight the layout cannot carry
        with pytest.raises(binding.CtrError) as ei:
            e.pack_candidates(lay, ids, bad)
        assert ei.value.code == -1 and "unweighted" in str(ei.value)
        lay2 = e.packed_layout(5, weighted_fields=range(13))
        ids2 = ids.copy()
        ids2[7, 2] += 1                                              # context differs between candidates
        with pytest.raises(binding.CtrError) as ei:
            e.pack_candidates(lay2, ids2, wts)
        assert ei.value.code == -1 and "context" in str(ei.value)
        # an id the direct table cannot hold survives packing as an out-of-range row and fails the REQUEST
        lay3 = e.packed_layout(0, weighted_fields=range(13), fixed_id_fields=(0, 1, 20))
        with pytest.raises(binding.CtrError) as ei:                  # field 20 is sparse: its id changes from candidate to candidate
            e.pack_candidates(lay3, ids, wts)
        assert ei.value.code == -1 and "fixed-id" in str(ei.value)
        ids3 = ids.copy()
        ids3[2, 30] = 1 << 40
        ci, cw, rows, pw = e.pack_candidates(lay, ids3, wts)
        t, _ = e.predict_packed_async(lay, ci, cw, rows, pw)
        with pytest.raises(binding.CtrError) as ei:
            e.wait(t)
        assert ei.value.code == -1
        t, _ = e.predict_packed_async(lay, *e.pack_candidates(lay, ids, wts))     # and the engine is clean afterwards
        assert np.array_equal(e.wait(t), e.predict(ids, wts))
    finally:
        e.close()


@pytest.mark.parametrize("kind_name,table_dtype", [("wide_deep", 0), ("deepfm", 1), ("lr", 0)])
def test_host_narrowing_of_the_drop_in_path(binding, om, weight_file, kind_name, table_dtype):
    """ctr_config.host_pack: the drop-in request (the reference's int64 id + float weight lists) narrowed on the worker
    thread before it crosses the host link must score what the untouched lists score, bit for bit -- also when a field
    that so far always weighed 1.0 suddenly carries a weight (the layout grows), and when an id does not fit 32 bits
    (the shard travels as it is and fails like it always did)."""
    kind = {"lr": om.LR, "wide_deep": om.WIDE_DEEP, "deepfm": om.DEEPFM}[kind_name]
    w = make_case(kind, hidden=() if kind == om.LR else (256, 128), table_dtype=table_dtype)
    path = weight_file(w)
    plain = binding.Engine(device_ids=(0,), max_candidates=4096, host_pack=1)
    narrow = binding.Engine(device_ids=(0,), max_candidates=4096, host_pack=2)
    plain.load_weights(path)
    narrow.load_weights(path)
    try:
        for n, seed in ((1, 1), (777, 2), (4096, 3)):
            ids, wts = om.criteo_inputs(n, seed=seed, rows_per_field=1 << 10)
            assert np.array_equal(narrow.predict(ids, wts), plain.predict(ids, wts)), n
        ids, wts = om.criteo_inputs(1000, seed=4, rows_per_field=1 << 10)
        wts[17, 30] = 0.25                                            # a sparse field with a weight: the learned layout widens
        wts[500:, 20] = 2.0
        want = plain.predict(ids, wts).copy()
        assert np.array_equal(narrow.predict(ids, wts), want)
        assert np.array_equal(narrow.predict(ids, wts), want)        # and stays right once widened
        ids2, wts2 = om.criteo_inputs(1000, seed=5, rows_per_field=1 << 10)
        assert np.array_equal(narrow.predict(ids2, wts2), plain.predict(ids2, wts2))
        s_n, i_n = narrow.predict_sorted(ids2, wts2)
        s_p, i_p = plain.predict_sorted(ids2, wts2)
        assert np.array_equal(s_n, s_p) and np.array_equal(i_n, i_p)
        ids[3, 25] = (1 << 40) + 5                                    # beyond 32 bits: out of range for the table, must FAIL, not wrap
        for e in (plain, narrow):
            with pytest.raises(binding.CtrError) as ei:
                e.predict(ids, wts)
            assert ei.value.code == -1
        assert np.array_equal(narrow.predict(ids2, wts2), plain.predict(ids2, wts2))
    finally:
        plain.close()
        narrow.close()
