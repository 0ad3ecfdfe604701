"""ctypes loader for oracle/_build/libctr_oracle.so (test infrastructure)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libctr_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "ctr_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _SO


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.oracle_mix64.restype = C.c_uint64
        _lib.oracle_mix64.argtypes = [C.c_uint64]
        for n in ("oracle_encode_int64_packed", "oracle_decode_int64_packed", "oracle_encode_predict_request",
                  "oracle_decode_predict_request", "oracle_encode_predict_response",
                  "oracle_decode_predict_response"):
            getattr(_lib, n).restype = C.c_int64
        _lib.oracle_num_threads.restype = C.c_int32
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def partition_bounds(n: int, p: int):
    b = np.zeros(p, np.int64); e = np.zeros(p, np.int64)
    lib().oracle_partition_bounds(C.c_int64(n), C.c_int32(p), _p(b), _p(e))
    return list(zip(b.tolist(), e.tolist()))


def mix64(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, np.uint64)
    return np.array([lib().oracle_mix64(int(v)) for v in x.ravel()], np.uint64).reshape(x.shape)


def proc_unit(seed: int, tag: int, idx: np.ndarray) -> np.ndarray:
    idx = np.ascontiguousarray(idx, np.uint64)
    out = np.empty(idx.shape, np.float32)
    lib().oracle_proc_unit(C.c_uint64(seed), C.c_uint64(tag), _p(idx), C.c_int64(idx.size), _p(out))
    return out


def fill_table(seed: int, scale_e: float, scale_w1: float, num_rows: int, K: int):
    emb = np.empty((num_rows, K), np.float32); w1 = np.empty(num_rows, np.float32)
    lib().oracle_fill_table(C.c_uint64(seed), C.c_float(scale_e), C.c_float(scale_w1),
                            C.c_int64(num_rows), C.c_int32(K), _p(emb), _p(w1))
    return emb, w1


def gather_fp32(ids, wts, num_rows, hashed, emb, w1, b0, with_fm):
    ids = np.ascontiguousarray(ids, np.int64); wts = np.ascontiguousarray(wts, np.float32)
    n, F = ids.shape; K = emb.shape[1]
    x0 = np.empty((n, F * K), np.float32); wide = np.empty(n, np.float32)
    lib().oracle_gather_fp32(_p(ids), _p(wts), C.c_int64(n), C.c_int32(F), C.c_int32(K), C.c_int64(num_rows),
                             C.c_int32(hashed), _p(emb), _p(w1), C.c_float(b0), C.c_int32(with_fm), _p(x0), _p(wide))
    return x0, wide


def encode_predict_request(ids: np.ndarray, wts: np.ndarray, model="DCN", sig="serving_default") -> bytes:
    ids = np.ascontiguousarray(ids, np.int64); wts = np.ascontiguousarray(wts, np.float32)
    rows, cols = ids.shape; n = rows * cols
    out = np.empty(16 * n + 512, np.uint8); tmp = np.empty(2 * (14 * n + 128) + 64, np.uint8)
    nb = lib().oracle_encode_predict_request(model.encode(), sig.encode(), _p(ids), _p(wts),
                                             C.c_int64(rows), C.c_int64(cols), _p(out), _p(tmp))
    return out[:nb].tobytes()


def decode_predict_request(buf: bytes, cap: int):
    a = np.frombuffer(buf, np.uint8)
    ids = np.empty(cap, np.int64); wts = np.empty(cap, np.float32); shape = np.zeros(2, np.int64)
    r = lib().oracle_decode_predict_request(_p(a), C.c_int64(len(buf)), _p(ids), _p(wts), C.c_int64(cap), _p(shape))
    if r < 0:
        raise ValueError("malformed PredictRequest")
    n = int(shape[0] * shape[1])
    return ids[:n].reshape(tuple(shape)), wts[:n].reshape(tuple(shape))


def encode_predict_response(scores: np.ndarray) -> bytes:
    s = np.ascontiguousarray(scores, np.float32)
    out = np.empty(4 * s.size + 128, np.uint8); tmp = np.empty(4 * s.size + 128, np.uint8)
    nb = lib().oracle_encode_predict_response(_p(s), C.c_int64(s.size), _p(out), _p(tmp))
    return out[:nb].tobytes()


def decode_predict_response(buf: bytes, cap: int) -> np.ndarray:
    a = np.frombuffer(buf, np.uint8); s = np.empty(cap, np.float32)
    n = lib().oracle_decode_predict_response(_p(a), C.c_int64(len(buf)), _p(s), C.c_int64(cap))
    if n < 0:
        raise ValueError("malformed PredictResponse")
    return s[:n]


def java_sort_keys(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, np.float32); k = np.empty(x.shape, np.int32)
    lib().oracle_java_sort_keys(_p(x), C.c_int64(x.size), _p(k))
    return k


def lr_scores(ids, wts, w1, b0):
    ids = np.ascontiguousarray(ids, np.int64); wts = np.ascontiguousarray(wts, np.float32)
    w1 = np.ascontiguousarray(w1, np.float32); out = np.empty(ids.shape[0], np.float32)
    lib().oracle_lr_scores(_p(ids), _p(wts), C.c_int64(ids.shape[0]), C.c_int32(ids.shape[1]), _p(w1),
                           C.c_float(b0), _p(out))
    return out


def num_threads() -> int:
    return int(lib().oracle_num_threads())
