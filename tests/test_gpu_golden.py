"""The committed golden fixtures (tests/golden/*.ctrw + *_io.npz) through the CUDA path.

The fixtures hold weights, inputs and the fp64 master's logits / probabilities as they were when
tests/golden/make_golden.py last ran.  tests/test_oracle.py pins today's oracle to them on the CPU;
this file pins the KERNELS to the same stored numbers, so a change to oracle/model.py cannot move
the oracle and the CUDA path together unnoticed.  (The reference itself holds no golden vectors --
its model runs in an external TF-Serving -- so parity stays "unpinned" in the sense of DESIGN.md
section 2; these are the strongest fixed points available.)"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = {"lr": 1e-5, "wide_deep": 2e-3, "deepfm": 2e-3, "dcn_v2": 2e-3}     # fp32 arithmetic / 16-bit tower (north star)


@pytest.mark.parametrize("name", ["lr", "wide_deep", "deepfm", "dcn_v2"])
def test_golden_fixture_through_the_c_abi(lib_built, name):
    from dtfs_b200 import binding
    io = np.load(os.path.join(GOLD, name + "_io.npz"))
    ids, wts = np.ascontiguousarray(io["ids"]), np.ascontiguousarray(io["wts"])
    with binding.Engine(device_ids=(0,), max_candidates=1024) as e:
        info = e.load_weights(os.path.join(GOLD, name + ".ctrw"))
        assert info.num_fields == ids.shape[1]
        got = e.predict(ids, wts).copy()
        # the other routes into the same kernels give the same bits
        import torch
        d_ids, d_wts = torch.from_numpy(ids).cuda(), torch.from_numpy(wts).cuda()
        d_out = torch.empty(ids.shape[0], dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        e.score_device(d_ids.data_ptr(), d_wts.data_ptr(), ids.shape[0], d_out.data_ptr())
        assert np.array_equal(d_out.cpu().numpy(), got)
        lay = e.packed_layout(0, weighted_fields=range(13))
        ci, cw, rows, pw = e.pack_candidates(lay, ids, wts)
        t, out = e.predict_packed_async(lay, ci, cw, rows, pw)
        assert np.array_equal(e.wait(t), got)
    err = np.abs(got - io["p_fp64"])
    assert err.max() <= TOL[name], "%s: max |dp| = %g against the stored fp64 probabilities" % (name, err.max())
    # the same in logit space, where a saturated probability cannot hide an error
    p = np.clip(got.astype(np.float64), 1e-7, 1 - 1e-7)
    z = np.log(p / (1 - p))
    keep = np.abs(io["logits_fp64"]) < 8
    assert np.abs(z - io["logits_fp64"])[keep].max() <= (1e-4 if name == "lr" else 2.5e-2)
