"""CTR_SM_SPLIT: the gather on its own set of SMs and the tower on the rest (green contexts, csrc/sm_split.cuh) is a scheduling
choice only -- the same kernels on the same data, so the scores must be the unsplit engine's bits, through the host path
(sendRequest, DCNClient.java:91-115) and through the device-resident entry point, with one and with two engine streams."""
import os

import numpy as np
import pytest

from conftest import make_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kind_name", ["wide_deep", "deepfm", "dcn_v2"])
def test_sm_split_scores_the_same_bits(lib_built, weight_file, kind_name):
    import torch
    from dtfs_b200 import binding
    from oracle import model as om
    kind = {"wide_deep": om.WIDE_DEEP, "deepfm": om.DEEPFM, "dcn_v2": om.DCN_V2}[kind_name]
    w = make_case(kind, hidden=(256, 128), n_cross=2 if kind == om.DCN_V2 else 0)
    path = weight_file(w)
    n = 70001                                                    # several chunks with CTR_CHUNK_ROWS below, ragged tail
    ids, wts = om.criteo_inputs(n, seed=77, rows_per_field=1 << 10)
    d_ids, d_wts = torch.from_numpy(ids).cuda(), torch.from_numpy(wts).cuda()

    def run(env):
        saved = {k: os.environ.get(k) for k in env}
        os.environ.update(env)
        try:
            e = binding.Engine(device_ids=(0,), max_candidates=32768)
            e.load_weights(path)
            try:
                host = e.predict(ids[:5003], wts[:5003]).copy()
                out = torch.empty(n, dtype=torch.float32, device="cuda")
                e.score_device(d_ids.data_ptr(), d_wts.data_ptr(), n, out.data_ptr())
                torch.cuda.synchronize()
                return host, out.cpu().numpy()
            finally:
                e.close()
        finally:
            for k, v in saved.items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = v

    want_host, want_dev = run({"CTR_CHUNK_ROWS": "16384"})
    assert np.array_equal(want_host, want_dev[:5003])
    for streams in ("1", "2"):
        host, dev = run({"CTR_SM_SPLIT": "48", "CTR_DEVICE_STREAMS": streams, "CTR_CHUNK_ROWS": "16384"})
        assert np.array_equal(host, want_host), "host path, %s stream(s)" % streams
        assert np.array_equal(dev, want_dev), "device path, %s stream(s)" % streams
    assert np.abs(want_host - om.forward(w, ids[:5003], wts[:5003], "fp64")).max() <= 2e-3
