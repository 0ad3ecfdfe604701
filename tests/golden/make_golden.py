"""Generates tests/golden/*.npz.  Run from the repository root:  python tests/golden/make_golden.py

The reference (neuzxy/Distributed-TF-Serving) holds no golden vectors and cannot be run or
imported here (Java + external TF-Serving), so these fixtures are produced by THIS project's
oracle (oracle/model.py, fp64 master) and pin it against accidental change; the part of each
fixture that IS reference-defined (split bounds, demo inputs, sort order) is restated from
DCNClient.java and checked in tests/test_oracle.py against independent closed forms.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import dcn_client as odc, model as om  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def small_case(kind, hidden, n_cross, seed):
    n_sparse, rpf = 5, 64                     # F = 13 dense + 5 sparse = 18 fields: small files
    F = 13 + n_sparse
    spec = om.ModelSpec(kind, F, 16, om.criteo_num_rows(13, n_sparse, rpf), hidden=hidden, n_cross=n_cross,
                        table_mode=om.TABLE_RAW, seed=seed, scale_e=0.25)
    calib = om.criteo_inputs(128, seed=99, n_sparse=n_sparse, rows_per_field=rpf)
    w = om.make_weights(spec, rng_seed=seed, calib_inputs=calib if kind != om.LR else None)
    ids, wts = om.criteo_inputs(48, seed=4321, n_sparse=n_sparse, rows_per_field=rpf)
    return w, ids, wts


def main():
    # 1. split rule + demo inputs + sort order (reference-defined)
    cases = [(64500, 3), (43000, 2), (1000, 2), (1500, 3), (7, 3), (5, 8), (0, 4), (301, 43)]
    bounds = {"%d_%d" % c: np.array(odc.partition_bounds(*c), np.int64) for c in cases}
    np.savez(os.path.join(OUT, "partition_bounds.npz"), **bounds)
    srt_in = np.array([0.5, -0.0, 0.0, np.nan, -1.5, np.inf, -np.inf, 1e-45, -1e-45, 0.25], np.float32)
    np.savez(os.path.join(OUT, "java_sort.npz"), x=srt_in, sorted=odc.java_float_sort(srt_in), key=odc.java_float_sort_key(srt_in))
    # 2. model math (builder-defined), fp64 master outputs on tiny raw-table models
    for name, kind, hidden, nc in [("lr", om.LR, (), 0), ("wide_deep", om.WIDE_DEEP, (64, 32), 0),
                                   ("deepfm", om.DEEPFM, (64, 64), 0), ("dcn_v2", om.DCN_V2, (64, 32), 2)]:
        w, ids, wts = small_case(kind, hidden, nc, 1234)
        path = os.path.join(OUT, "%s.ctrw" % name)
        om.save_weights(w, path)
        np.savez(os.path.join(OUT, "%s_io.npz" % name), ids=ids, wts=wts,
                 logits_fp64=om.forward_logits(w, ids, wts, "fp64"), p_fp64=om.forward(w, ids, wts, "fp64"))
    # 3. integer pieces
    z = np.array([0, 1, 2, 43, 2 ** 31, 2 ** 63, 2 ** 64 - 1, 0x9E3779B97F4A7C15], np.uint64)
    np.savez(os.path.join(OUT, "mix64.npz"), z=z, h=om.mix64(z), unit=om.proc_unit(1234, 1, z))


if __name__ == "__main__":
    main()
