"""numpy restatement of the reference client (test infrastructure, see oracle/__init__.py).

Every function cites the lines of
``/root/reference/src/main/java/tensorflow/serving/DCNClient.java`` (``DC``) it follows.
The "server" (TF-Serving + SavedModel, not in the reference repository) is replaced by a
callable ``model(ids[n,F] int64, wts[n,F] float32) -> float32[n]`` from ``oracle.model``.
"""
from __future__ import annotations

from concurrent.futures import ThreadPoolExecutor
from typing import Callable, List, Sequence

import numpy as np

FIELD_NUM = 43          # DC:25
CANDIDATE_NUM = 1500    # DC:29
REQUEST_NUM = 1000      # DC:30
CONCURRENT_NUM = 6      # DC:31
N_HOSTS = 3             # DC:38 (three literal LAN hosts)
POOL_THREADS = 16       # DC:42


def partition_bounds(n: int, part_num: int) -> List[tuple]:
    """DC:46-55.  ``range = N / partNum`` (Java int division); parts 0..P-2 are
    ``[i*range, (i+1)*range)``, the last part is ``[(P-1)*range, N)``."""
    if part_num <= 0:
        raise ZeroDivisionError("partNum must be positive (Java would throw ArithmeticException)")
    rng = n // part_num
    bounds = [(i * rng, (i + 1) * rng) for i in range(part_num - 1)]
    bounds.append(((part_num - 1) * rng, n))
    return bounds


def partition_list(lst: Sequence, part_num: int) -> list:
    """DC:46-55 on a FLAT element list (the reference splits elements, not candidates).
    Returns views (numpy slices), like ``List.subList``."""
    return [lst[a:b] for a, b in partition_bounds(len(lst), part_num)]


def partition_candidates(n_cand: int, part_num: int) -> List[tuple]:
    """The same rule applied to candidates: identical to the
    reference wherever the reference is well-formed (``part_num`` divides ``n_cand``)."""
    return partition_bounds(n_cand, part_num)


def get_feature_ids(candidate_num: int = CANDIDATE_NUM, field_num: int = FIELD_NUM,
                    n_hosts: int = N_HOSTS) -> list:
    """DC:57-65: ids 1..FIELD_NUM repeated per candidate, flat, then element-partitioned."""
    one = np.arange(1, field_num + 1, dtype=np.int64)
    flat = np.tile(one, candidate_num)
    return partition_list(flat, n_hosts)


def get_feature_wts(candidate_num: int = CANDIDATE_NUM, field_num: int = FIELD_NUM,
                    n_hosts: int = N_HOSTS) -> list:
    """DC:67-74: weights 1.0f for every slot, flat, then element-partitioned."""
    flat = np.ones(candidate_num * field_num, dtype=np.float32)
    return partition_list(flat, n_hosts)


def get_tensor_shape(n_elems: int, field_num: int = FIELD_NUM) -> tuple:
    """DC:76-81 with the argument computed at DC:97: ``[size / FIELD_NUM, FIELD_NUM]``
    (Java int division truncates)."""
    return (n_elems // field_num, field_num)


def send_request(model: Callable, feature_ids: np.ndarray, feature_wts: np.ndarray,
                 field_num: int = FIELD_NUM) -> np.ndarray:
    """DC:91-115.  Declares shape ``[len/F, F]`` for both tensors and sends every value.
    TF-Serving rejects a TensorProto whose value count disagrees with its declared shape,
    so an ill-formed shard (``len % F != 0``) raises, as the RPC would."""
    shape = get_tensor_shape(len(feature_ids), field_num)
    if shape[0] * shape[1] != len(feature_ids) or len(feature_wts) != len(feature_ids):
        raise ValueError("declared shape %r disagrees with %d ids / %d weights"
                         % (shape, len(feature_ids), len(feature_wts)))
    ids = np.asarray(feature_ids, dtype=np.int64).reshape(shape)
    wts = np.asarray(feature_wts, dtype=np.float32).reshape(shape)
    out = np.asarray(model(ids, wts), dtype=np.float32).reshape(-1)
    assert out.shape[0] == shape[0]
    return out


def java_float_sort_key(x: np.ndarray) -> np.ndarray:
    """Total order of ``Float.compareTo`` (used by ``Collections.sort`` at DC:195) as an
    int32 key: -0.0 < +0.0, every NaN canonicalised (floatToIntBits) and last."""
    x = np.asarray(x, dtype=np.float32)
    bits = x.view(np.int32).copy()
    bits[np.isnan(x)] = np.int32(0x7FC00000)
    neg = bits < 0
    bits[neg] ^= np.int32(0x7FFFFFFF)
    return bits


def java_float_sort(x: np.ndarray) -> np.ndarray:
    """DC:195 ``Collections.sort(fullCtrList)``: ascending, candidate identity discarded."""
    x = np.asarray(x, dtype=np.float32)
    return x[np.argsort(java_float_sort_key(x), kind="stable")]


def process_a_request(models: Sequence[Callable], feat_ids_list: Sequence[np.ndarray],
                      feat_wts_list: Sequence[np.ndarray], field_num: int = FIELD_NUM,
                      sort: bool = True, pool: ThreadPoolExecutor | None = None) -> np.ndarray:
    """DC:137-203, full-async branch (DC:146-164): one task per host on the shared pool,
    join in index order, concatenate per-shard scores in shard order, then (DC:195) sort.
    ``sort=False`` returns the ordered merge (output index j <-> candidate j)."""
    task_num = len(models)
    own = pool is None
    if own:
        pool = ThreadPoolExecutor(POOL_THREADS)
    try:
        futures = [pool.submit(send_request, models[i], feat_ids_list[i], feat_wts_list[i], field_num)
                   for i in range(task_num)]
        responses = [f.result() for f in futures]          # DC:158-159 join in index order
    finally:
        if own:
            pool.shutdown()
    full = np.concatenate(responses) if responses else np.zeros(0, np.float32)   # DC:161-164
    return java_float_sort(full) if sort else full
