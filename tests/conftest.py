import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def _has_gpu() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def lib_built():
    """The product library, built in-tree (nvcc cross-compiles without a GPU)."""
    import dtfs_b200
    from dtfs_b200 import build
    return build.build_library()


def make_case(kind, F=39, hidden=(), n_cross=0, rows_per_field=1 << 10, table_dtype=0, id_mode=0,
              procedural=True, n_calib=256, seed=1234, scale_e=0.25, num_rows=None):
    """Oracle weights + a weight file the engine can load.  Returns (oracle Weights, path builder)."""
    from oracle import model as om
    if num_rows is None:
        num_rows = om.criteo_num_rows(13, F - 13, rows_per_field) if F > 13 else F + 1
    if id_mode == om.ID_HASHED:
        num_rows = 1 << int(np.ceil(np.log2(num_rows)))
    spec = om.ModelSpec(kind, F, 16, num_rows, id_mode, table_dtype, tuple(hidden), n_cross,
                        om.TABLE_PROCEDURAL if procedural else om.TABLE_RAW, seed, scale_e, 0.5)
    calib = None
    if kind != om.LR and F > 13:
        calib = om.criteo_inputs(n_calib, seed=99, n_dense=13, n_sparse=F - 13, rows_per_field=rows_per_field)
    return om.make_weights(spec, rng_seed=seed, calib_inputs=calib)


@pytest.fixture
def weight_file(tmp_path):
    from oracle import model as om

    def _save(w):
        p = str(tmp_path / "weights.ctrw")
        om.save_weights(w, p)
        return p
    return _save
