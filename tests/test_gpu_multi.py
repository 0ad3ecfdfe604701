"""Multi-GPU: candidates sharded over the GPUs of one engine by the
reference rule, tables replicated -- and the same with tables row-sharded (row mod G), where the
gather kernel reads peer rows over NVLink.  Scores must equal the 1-GPU scores bit for bit: the
kernels are the same and every candidate's reduction order is fixed."""
import numpy as np
import pytest

from conftest import make_case

pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.fixture(scope="module")
def binding(lib_built):
    from dtfs_b200 import binding
    return binding


@pytest.mark.parametrize("kind_name,n", [("wide_deep", 10000), ("deepfm", 1001), ("dcn_v2", 7), ("lr", 1500)])
def test_candidate_sharding_equals_single_gpu(binding, weight_file, kind_name, n):
    from oracle import model as om
    G = _n_gpus()
    if G < 2:
        pytest.skip("needs >= 2 GPUs")
    kind = {"wide_deep": om.WIDE_DEEP, "deepfm": om.DEEPFM, "dcn_v2": om.DCN_V2, "lr": om.LR}[kind_name]
    hidden = {"wide_deep": (1024, 512, 256), "deepfm": (512, 512, 512), "dcn_v2": (512, 256), "lr": ()}[kind_name]
    w = make_case(kind, hidden=hidden, n_cross=2 if kind == om.DCN_V2 else 0)
    path = weight_file(w)
    ids, wts = om.criteo_inputs(n, seed=4321, rows_per_field=1 << 10)
    with binding.Engine(device_ids=(0,), max_candidates=16384) as e1:
        e1.load_weights(path)
        one = e1.predict(ids, wts).copy()
    for g in sorted({2, G}):
        # min_shard_candidates=1: cut over all g GPUs whatever the size, as the reference does (DC:148)
        with binding.Engine(device_ids=tuple(range(g)), max_candidates=16384, min_shard_candidates=1) as eg:
            eg.load_weights(path)
            many = eg.predict(ids, wts).copy()
        assert np.array_equal(many, one), "%d GPUs: max |d| = %g" % (g, np.abs(many - one).max())
        # engine default: shards of >= 1024 candidates, first GPU rotating -- consecutive small requests
        # land on different GPUs and every one of them must give the same bits
        with binding.Engine(device_ids=tuple(range(g)), max_candidates=16384) as eg:
            eg.load_weights(path)
            for rep in range(g + 1):
                assert np.array_equal(eg.predict(ids, wts), one), "%d GPUs, default sharding, request %d" % (g, rep)
    tol = 1e-5 if kind == om.LR else 2e-3
    assert np.abs(one - om.forward(w, ids, wts, "fp64")).max() <= tol


@pytest.mark.parametrize("hashed", [False, True])
def test_row_sharded_tables_equal_replicated(binding, weight_file, hashed):
    """Rows striped over the GPUs (row mod G), fetched through peer mappings inside the gather
    kernel: the gather is a copy, so results equal the replicated ones bit for bit (config 4)."""
    from oracle import model as om
    G = _n_gpus()
    if G < 2:
        pytest.skip("needs >= 2 GPUs")
    for procedural in (True, False):
        w = make_case(om.DCN_V2, hidden=(512, 256), n_cross=2, id_mode=om.ID_HASHED if hashed else om.ID_DIRECT,
                      procedural=procedural, rows_per_field=1 << 8)
        path = weight_file(w)
        ids, wts = om.criteo_inputs(3000, seed=7, rows_per_field=1 << 8)
        with binding.Engine(device_ids=(0,), max_candidates=4096) as e1:
            e1.load_weights(path)
            one = e1.predict(ids, wts).copy()
        with binding.Engine(device_ids=tuple(range(G)), max_candidates=4096, shard_mode=binding.TABLES_ROW_SHARDED,
                            min_shard_candidates=1) as es:
            info = es.load_weights(path)
            sharded = es.predict(ids, wts).copy()
            rows = np.arange(0, min(w.spec.num_rows, 4096))
            for g in range(G):
                mine = rows[rows % G == g]
                emb, w1 = es.read_rows(mine, device_index=g)
                oe, ow = w.rows(mine)
                assert np.array_equal(emb, oe) and np.array_equal(w1, ow), "shard %d holds the wrong rows" % g
        assert info.table_bytes_per_device * G <= (w.spec.num_rows + G) * 128 + 128 * G
        assert np.array_equal(sharded, one), "row-sharded vs replicated: max |d| = %g" % np.abs(sharded - one).max()
        assert np.abs(one - om.forward(w, ids, wts, "fp64")).max() <= 2e-3


@pytest.mark.parametrize("kind_name,table_dtype,id_mode", [("dcn_v2", 0, 1), ("deepfm", 1, 0), ("wide_deep", 0, 0), ("lr", 0, 0)])
def test_nccl_exchange_equals_replicated_tables(binding, weight_file, kind_name, table_dtype, id_mode):
    """BASELINE config 4 as the north star words it: rows sharded over the GPUs (row mod G), looked up with an NCCL
    all-to-all of row indices and a second one of the rows (ctr_score_devices_exchange).  Every device scores its own
    batch; the scores must be the bits a replicated single-GPU engine gives, and the bits of the peer-load variant."""
    import torch
    from oracle import model as om
    G = _n_gpus()
    if G < 2:
        pytest.skip("needs >= 2 GPUs")
    kind = {"wide_deep": om.WIDE_DEEP, "deepfm": om.DEEPFM, "dcn_v2": om.DCN_V2, "lr": om.LR}[kind_name]
    hidden = {"wide_deep": (256, 128), "deepfm": (256, 128), "dcn_v2": (256, 128), "lr": ()}[kind_name]
    w = make_case(kind, hidden=hidden, n_cross=2 if kind == om.DCN_V2 else 0, table_dtype=table_dtype, id_mode=id_mode)
    path = weight_file(w)
    sizes = [3001 + 517 * g for g in range(G)]                         # a different, ragged batch on every device
    sizes[-1] = 0 if G > 2 else sizes[-1]                              # and one device with nothing to score
    batches = [om.criteo_inputs(max(n, 1), seed=50 + g, rows_per_field=1 << 10) for g, n in enumerate(sizes)]
    with binding.Engine(device_ids=(0,), max_candidates=16384) as e1:
        e1.load_weights(path)
        want = [e1.predict(ids[:n], wts[:n]).copy() if n else np.empty(0, np.float32) for (ids, wts), n in zip(batches, sizes)]
    with binding.Engine(device_ids=tuple(range(G)), max_candidates=16384, shard_mode=binding.TABLES_ROW_SHARDED) as eg:
        eg.load_weights(path)
        bufs = []
        for g, ((ids, wts), n) in enumerate(zip(batches, sizes)):
            with torch.cuda.device(g):
                bufs.append((torch.from_numpy(ids[:max(n, 1)]).cuda(), torch.from_numpy(wts[:max(n, 1)]).cuda(),
                             torch.full((max(n, 1),), -1.0, dtype=torch.float32, device="cuda")))
        for g in range(G):
            torch.cuda.synchronize(g)
        for rep in range(2):                                           # the second call reuses every exchange buffer
            eg.score_devices_exchange([b[0].data_ptr() for b in bufs], [b[1].data_ptr() for b in bufs], sizes, [b[2].data_ptr() for b in bufs])
            for g, n in enumerate(sizes):
                got = bufs[g][2][:n].cpu().numpy()
                assert np.array_equal(got, want[g]), "device %d, call %d: max |d| = %g" % (g, rep, np.abs(got - want[g]).max() if n else 0)
        # the peer-load variant on the same engine: same bits
        for g, n in enumerate(sizes):
            if n:
                o = torch.empty(n, dtype=torch.float32, device="cuda:%d" % g)
                eg.score_device(bufs[g][0].data_ptr(), bufs[g][1].data_ptr(), n, o.data_ptr(), device_index=g)
                assert np.array_equal(o.cpu().numpy(), want[g])
    ids, wts = batches[0]
    tol = 1e-5 if kind == om.LR else 2e-3
    assert np.abs(want[0] - om.forward(w, ids[:sizes[0]], wts[:sizes[0]], "fp64")).max() <= tol


def test_sorted_request_over_several_gpus(binding, weight_file):
    """processARequest including its sort (DCNClient.java:195) when the request is cut over several GPUs: the shards
    merge their scores in order on the first GPU (peer copies) and the sort runs there -- same result as one GPU."""
    from oracle import dcn_client as oc, model as om
    G = _n_gpus()
    if G < 2:
        pytest.skip("needs >= 2 GPUs")
    w = make_case(om.WIDE_DEEP, hidden=(256, 128))
    path = weight_file(w)
    with binding.Engine(device_ids=tuple(range(G)), max_candidates=40000, min_shard_candidates=1) as eg:
        eg.load_weights(path)
        for n in (G, 10000, 70001):                                      # one candidate per GPU; block sort; multi-block sort
            ids, wts = om.criteo_inputs(n, seed=n, rows_per_field=1 << 10)
            base = eg.predict(ids, wts).copy()
            for _ in range(2):
                s, idx = eg.predict_sorted(ids, wts)
                assert np.array_equal(np.sort(idx), np.arange(n))
                assert np.array_equal(base[idx], s)
                assert np.array_equal(s, oc.java_float_sort(base))
