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
