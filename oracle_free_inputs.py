"""Synthetic Criteo-shaped request inputs in numpy for bench.py's host-side legs (the layout of SURVEY.md
section 8(d) config 2: dense field j -> id j+1 with a log1p weight, sparse field s -> id 14 + s*R + h with weight 1.0).
Lives outside oracle/ on purpose: the product's bench legs must not import the checker."""
import numpy as np


def criteo_inputs(n: int, rows_per_field: int, seed: int = 4321, n_dense: int = 13, n_sparse: int = 26):
    rng = np.random.Generator(np.random.PCG64(seed))
    F = n_dense + n_sparse
    ids = np.empty((n, F), np.int64)
    wts = np.ones((n, F), np.float32)
    ids[:, :n_dense] = np.arange(1, n_dense + 1)
    wts[:, :n_dense] = np.log1p(np.exp(rng.standard_normal((n, n_dense)))).astype(np.float32)
    h = rng.integers(0, rows_per_field, (n, n_sparse), dtype=np.int64)
    ids[:, n_dense:] = (n_dense + 1) + np.arange(n_sparse, dtype=np.int64) * rows_per_field + h
    return ids, wts
