"""ctr_score_device_static -- the entry point for callers that may not allocate.

Some users of this library cannot let a scoring call allocate GPU memory: a serving process
whose HBM is already fully owned by another pool, or a latency-critical request thread for
which a cudaMalloc (it can stall every stream of the device) is unacceptable.  For them the
engine offers a variant that keeps its intermediate activations in a fixed array inside the
library image.  This test calls it through the C ABI, twice in a row with different requests
(the second call finds the first call's leftovers in the shared array), and checks both
results against the ordinary entry point, ctr_score_device.  Both run the same kernels on
the same rows, so the comparison is exact, like the other path-against-path checks in
test_gpu_parity.py.

The file name sorts last on purpose: a run that stops at the first failure has then already
run every other test.
"""
import numpy as np
import pytest

from conftest import make_case

pytestmark = pytest.mark.gpu


def _score(fn, torch, ids, wts):
    d_ids = torch.from_numpy(ids).cuda()
    d_wts = torch.from_numpy(wts).cuda()
    d_out = torch.full((ids.shape[0],), -1.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    fn(d_ids.data_ptr(), d_wts.data_ptr(), ids.shape[0], d_out.data_ptr())
    torch.cuda.synchronize()
    return d_out.cpu().numpy()


@pytest.mark.parametrize("kind_name", ["wide_deep", "deepfm", "dcn_v2"])
def test_static_scratch_matches_regular_entry_point(lib_built, weight_file, kind_name):
    import torch
    from dtfs_b200 import binding
    from oracle import model as om
    kind = {"wide_deep": om.WIDE_DEEP, "deepfm": om.DEEPFM, "dcn_v2": om.DCN_V2}[kind_name]
    hidden = {"wide_deep": (1024, 512, 256), "deepfm": (512, 512, 512), "dcn_v2": (512, 256)}[kind_name]
    w = make_case(kind, hidden=hidden, n_cross=2 if kind == om.DCN_V2 else 0)
    e = binding.Engine(device_ids=(0,), max_candidates=2048)
    e.load_weights(weight_file(w))
    try:
        # two calls in a row, different sizes and inputs; the second sees the first one's leftovers
        for n, seed in ((binding.STATIC_MAX_CANDIDATES, 11), (333, 12)):
            ids, wts = om.criteo_inputs(n, seed=seed, rows_per_field=1 << 10)
            ref = _score(e.score_device, torch, ids, wts)
            got = _score(e.score_device_static, torch, ids, wts)
            assert np.all(np.isfinite(got)) and got.min() >= 0.0 and got.max() <= 1.0
            assert np.array_equal(got, ref), "call with n=%d: max |d| = %g" % (n, np.abs(got - ref).max())
            # and the ordinary entry point agrees with the oracle, so this is not two wrongs
            assert np.abs(ref - om.forward(w, ids, wts, "fp64")).max() <= 2e-3
        # and it really allocates nothing: the device's free memory is the same before and after further calls
        ids, wts = om.criteo_inputs(700, seed=15, rows_per_field=1 << 10)
        d_ids, d_wts = torch.from_numpy(ids).cuda(), torch.from_numpy(wts).cuda()
        d_out = torch.empty(700, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        for _ in range(3):
            e.score_device_static(d_ids.data_ptr(), d_wts.data_ptr(), 700, d_out.data_ptr())
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free0
        # a caller-provided stream works the same
        ids, wts = om.criteo_inputs(500, seed=13, rows_per_field=1 << 10)
        st = torch.cuda.Stream()
        ref = _score(e.score_device, torch, ids, wts)
        with torch.cuda.stream(st):
            got = _score(lambda *a: e.score_device_static(*a, stream=st.cuda_stream), torch, ids, wts)
        assert np.array_equal(got, ref)
        # beyond the fixed scratch: a clear refusal, not a silent overrun
        ids, wts = om.criteo_inputs(binding.STATIC_MAX_CANDIDATES + 1, seed=14, rows_per_field=1 << 10)
        with pytest.raises(binding.CtrError) as ei:
            _score(e.score_device_static, torch, ids, wts)
        assert ei.value.code == -6 and "CTR_STATIC_MAX_CANDIDATES" in str(ei.value)
    finally:
        e.close()


def test_static_scratch_lr_and_bad_ids(lib_built, weight_file):
    import torch
    from dtfs_b200 import binding
    from oracle import model as om
    w = make_case(om.LR, F=43, procedural=False)
    e = binding.Engine(device_ids=(0,), max_candidates=2048)
    e.load_weights(weight_file(w))
    try:
        rng = np.random.default_rng(4321)
        ids = rng.integers(1, 44, (777, 43)).astype(np.int64)
        wts = rng.uniform(0, 1, (777, 43)).astype(np.float32)
        got = _score(e.score_device_static, torch, ids, wts)
        assert np.array_equal(got, _score(e.score_device, torch, ids, wts))
        assert np.abs(got - om.forward(w, ids, wts, "fp64")).max() <= 1e-5
        ids[5, 7] = w.spec.num_rows                      # out of range for a direct-indexed table
        with pytest.raises(binding.CtrError) as ei:
            _score(e.score_device_static, torch, ids, wts)
        assert ei.value.code == -1
        ids[5, 7] = 1                                    # the flag was cleared: the next call is clean
        assert np.all(np.isfinite(_score(e.score_device_static, torch, ids, wts)))
    finally:
        e.close()
