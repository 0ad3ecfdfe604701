"""ctypes binding of include/ctr_engine.h.  No fallback: if libctr_b200.so is missing or a
call fails, this raises."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libctr_b200.so")

CTR_OK = 0
ERR_NAMES = {-1: "CTR_ERR_INVALID", -2: "CTR_ERR_CUDA", -3: "CTR_ERR_IO", -4: "CTR_ERR_STATE",
             -5: "CTR_ERR_TIMEOUT", -6: "CTR_ERR_CAPACITY"}
CTR_MAX_DEVICES, CTR_MAX_HIDDEN = 8, 4
TABLES_REPLICATED, TABLES_ROW_SHARDED = 0, 1
STATIC_MAX_CANDIDATES = 1024

# every symbol include/ctr_engine.h declares (tests check the library exports exactly these)
EXPORTS = ["ctr_engine_create", "ctr_engine_destroy", "ctr_engine_load_weights", "ctr_engine_model_info",
           "ctr_predict_async", "ctr_wait", "ctr_predict", "ctr_alloc_pinned", "ctr_free_pinned",
           "ctr_score_device", "ctr_partition_bounds", "ctr_last_error", "ctr_abi_version",
           "ctr_engine_launch_count", "ctr_engine_set_timing", "ctr_engine_stage_ms", "ctr_debug_read_rows",
           "ctr_debug_gemm", "ctr_debug_gather", "ctr_score_device_static", "ctr_predict_sorted", "ctr_predict_wire",
           "ctr_wire_parse_request", "ctr_wire_build_response", "ctr_predict_shared_async", "ctr_predict_packed_async",
           "ctr_pack_candidates", "ctr_score_devices_exchange", "ctr_engine_h2d_bytes", "ctr_engine_d2h_bytes"]


class CtrConfig(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("n_devices", C.c_int32), ("device_ids", C.c_int32 * CTR_MAX_DEVICES),
                ("max_candidates", C.c_int32), ("workers_per_device", C.c_int32), ("shard_mode", C.c_int32),
                ("min_shard_candidates", C.c_int32), ("host_pack", C.c_int32), ("reserved", C.c_int32 * 6)]


class CtrModelInfo(C.Structure):
    _fields_ = [("model_kind", C.c_int32), ("num_fields", C.c_int32), ("emb_dim", C.c_int32), ("id_mode", C.c_int32),
                ("table_dtype", C.c_int32), ("n_hidden", C.c_int32), ("hidden", C.c_int32 * CTR_MAX_HIDDEN),
                ("n_cross", C.c_int32), ("d_padded", C.c_int32), ("num_rows", C.c_int64),
                ("table_bytes_per_device", C.c_int64), ("chunk_rows", C.c_int32), ("reserved", C.c_int32)]


class CtrPackedLayout(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("n_ctx_fields", C.c_int32), ("weighted_mask", C.c_uint64), ("fixed_id_mask", C.c_uint64)]


CALLBACK = C.CFUNCTYPE(None, C.c_void_p, C.c_int64, C.c_int32)


class CtrError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__("%s (%d): %s" % (ERR_NAMES.get(code, "CTR_ERR"), code, msg))
        self.code = code


_lib: Optional[C.CDLL] = None


def load_library() -> C.CDLL:
    """Loads libctr_b200.so.  Raises if it has not been built: the engine has no other path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("%s is missing: build it with __graft_entry__.build() "
                           "(nvcc, sm_100a). There is no CPU or PyTorch fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    lib.ctr_engine_create.argtypes = [C.POINTER(CtrConfig), C.POINTER(vp)]
    lib.ctr_engine_destroy.argtypes = [vp]
    lib.ctr_engine_load_weights.argtypes = [vp, C.c_char_p]
    lib.ctr_engine_model_info.argtypes = [vp, C.POINTER(CtrModelInfo)]
    lib.ctr_predict_async.argtypes = [vp, vp, vp, i64, vp, vp, vp, C.POINTER(i64)]
    lib.ctr_predict_shared_async.argtypes = [vp, vp, vp, i32, vp, vp, i64, vp, vp, vp, C.POINTER(i64)]
    lib.ctr_predict_packed_async.argtypes = [vp, C.POINTER(CtrPackedLayout), vp, vp, vp, vp, i64, vp, vp, vp, C.POINTER(i64)]
    lib.ctr_pack_candidates.argtypes = [vp, C.POINTER(CtrPackedLayout), vp, vp, i64, vp, vp, vp, vp]
    lib.ctr_score_devices_exchange.argtypes = [vp, vp, vp, vp, vp]
    lib.ctr_wait.argtypes = [vp, i64, i32]
    lib.ctr_predict.argtypes = [vp, vp, vp, i64, vp]
    lib.ctr_alloc_pinned.argtypes = [vp, i64, C.POINTER(vp)]
    lib.ctr_free_pinned.argtypes = [vp, vp]
    lib.ctr_score_device.argtypes = [vp, i32, vp, vp, i64, vp, vp]
    lib.ctr_predict_sorted.argtypes = [vp, vp, vp, i64, vp, vp]
    lib.ctr_predict_wire.argtypes = [vp, vp, i64, vp, i64, C.POINTER(i64)]
    lib.ctr_wire_parse_request.argtypes = [vp, i64, i32, vp, vp, i64, C.POINTER(i64)]
    lib.ctr_wire_build_response.argtypes = [vp, i64, vp, i64, C.POINTER(i64)]
    lib.ctr_score_device_static.argtypes = [vp, i32, vp, vp, i64, vp, vp]
    lib.ctr_partition_bounds.argtypes = [i64, i32, vp, vp]
    lib.ctr_last_error.restype = C.c_char_p
    lib.ctr_engine_launch_count.argtypes = [vp]
    lib.ctr_engine_launch_count.restype = i64
    for fn in (lib.ctr_engine_h2d_bytes, lib.ctr_engine_d2h_bytes):
        fn.argtypes = [vp]
        fn.restype = i64
    lib.ctr_engine_set_timing.argtypes = [vp, i32]
    lib.ctr_engine_stage_ms.argtypes = [vp, i32, i32]
    lib.ctr_engine_stage_ms.restype = C.c_double
    lib.ctr_debug_read_rows.argtypes = [vp, i32, vp, i64, vp, vp]
    lib.ctr_debug_gemm.argtypes = [vp, i32, vp, vp, vp, i32, i32, i32, i32, vp]
    lib.ctr_debug_gather.argtypes = [vp, i32, vp, vp, i64, vp, vp]
    _lib = lib
    return lib


def _check(rc: int) -> None:
    if rc != CTR_OK:
        raise CtrError(rc, (load_library().ctr_last_error() or b"").decode("utf-8", "replace"))


def partition_bounds(n: int, part_num: int):
    """DCNClient.java:46-55 through the C ABI."""
    b = np.zeros(max(part_num, 1), np.int64)
    e = np.zeros(max(part_num, 1), np.int64)
    _check(load_library().ctr_partition_bounds(n, part_num, b.ctypes.data, e.ctypes.data))
    return list(zip(b.tolist(), e.tolist()))


def wire_parse_request(request: bytes, field_num: int):
    """PredictRequest bytes -> (ids [n, F] int64, wts [n, F] float32); host only."""
    lib = load_library()
    n = C.c_int64(0)
    cap = max(len(request), 16)
    ids = np.empty(cap, np.int64)
    wts = np.empty(cap, np.float32)
    _check(lib.ctr_wire_parse_request(request, len(request), field_num, ids.ctypes.data, wts.ctypes.data, cap, C.byref(n)))
    m = n.value * field_num
    return ids[:m].reshape(n.value, field_num).copy(), wts[:m].reshape(n.value, field_num).copy()


def wire_build_response(scores: np.ndarray) -> bytes:
    """scores -> PredictResponse bytes (outputs["prediction_node"], DT_FLOAT [n,1]); host only."""
    lib = load_library()
    s = np.ascontiguousarray(scores, np.float32)
    need = C.c_int64(0)
    cap = 4 * s.size + 128
    buf = C.create_string_buffer(cap)
    _check(lib.ctr_wire_build_response(s.ctypes.data, s.size, buf, cap, C.byref(need)))
    return buf.raw[:need.value]


class PinnedBuffer:
    """numpy view over memory from ctr_alloc_pinned (the zero-copy packer's slab)."""

    def __init__(self, engine: "Engine", shape, dtype):
        self.engine = engine
        dtype = np.dtype(dtype)
        nbytes = int(np.prod(shape)) * dtype.itemsize
        p = C.c_void_p()
        _check(engine.lib.ctr_alloc_pinned(engine.handle, max(nbytes, 16), C.byref(p)))
        self.ptr = p.value
        buf = (C.c_char * max(nbytes, 16)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self):
        if self.ptr is not None and self.engine.handle:
            self.array = None
            _check(self.engine.lib.ctr_free_pinned(self.engine.handle, self.ptr))
        self.ptr = None


class Engine:
    """One engine = the reference's channel list (getChannels, DCNClient.java:118-125): one
    "channel" per GPU."""

    def __init__(self, device_ids=(0,), max_candidates: int = 16384, workers_per_device: int = 0,
                 shard_mode: int = TABLES_REPLICATED, min_shard_candidates: int = 0, host_pack: int = 0):
        self.lib = load_library()
        cfg = CtrConfig()
        cfg.struct_size = C.sizeof(CtrConfig)
        cfg.n_devices = len(device_ids)
        for i, d in enumerate(device_ids):
            cfg.device_ids[i] = d
        cfg.max_candidates = max_candidates
        cfg.workers_per_device = workers_per_device
        cfg.shard_mode = shard_mode
        cfg.min_shard_candidates = min_shard_candidates       # 0: engine default; 1: always cut over all GPUs (DC:148)
        cfg.host_pack = host_pack                             # 0: narrow drop-in shards on the host when the GPU is busy; 1: never; 2: always
        h = C.c_void_p()
        _check(self.lib.ctr_engine_create(C.byref(cfg), C.byref(h)))
        self.handle = h.value
        self.n_devices = len(device_ids)
        self.num_fields = None
        self._keep = {}

    def close(self):
        if self.handle:
            self.lib.ctr_engine_destroy(self.handle)
            self.handle = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def load_weights(self, path: str) -> CtrModelInfo:
        _check(self.lib.ctr_engine_load_weights(self.handle, path.encode()))
        info = CtrModelInfo()
        _check(self.lib.ctr_engine_model_info(self.handle, C.byref(info)))
        self.num_fields = info.num_fields
        self.info = info
        return info

    # -- host request path -------------------------------------------------------------
    def _check_inputs(self, ids, wts):
        ids = np.asarray(ids)
        wts = np.asarray(wts)
        if ids.dtype != np.int64 or wts.dtype != np.float32:
            raise TypeError("ids must be int64 and wts float32 (the reference's DT_INT64 / DT_FLOAT tensors)")
        if not (ids.flags.c_contiguous and wts.flags.c_contiguous):
            raise ValueError("ids / wts must be C-contiguous")
        if ids.size != wts.size or ids.size % self.num_fields:
            raise ValueError("ids / wts need n_cand * %d elements each" % self.num_fields)
        return ids, wts, ids.size // self.num_fields

    def predict_async(self, ids, wts, out: Optional[np.ndarray] = None, callback=None):
        """callback(ticket, status) runs on an engine thread when the request completes (the Java shim
        completes its CompletableFuture there)."""
        ids, wts, n = self._check_inputs(ids, wts)
        if out is None:
            out = np.empty(n, np.float32)
        t = C.c_int64()
        cb = CALLBACK(lambda user, ticket, status: callback(ticket, status)) if callback else None
        _check(self.lib.ctr_predict_async(self.handle, ids.ctypes.data, wts.ctypes.data, n, out.ctypes.data,
                                          C.cast(cb, C.c_void_p) if cb else None, None, C.byref(t)))
        self._keep[t.value] = (ids, wts, cb, out)
        return t.value, out

    def predict_shared_async(self, ctx_ids, ctx_wts, cand_ids, cand_wts, out: Optional[np.ndarray] = None):
        """Shared-context request: the first len(ctx_ids) fields of every candidate given once."""
        ctx_ids = np.ascontiguousarray(ctx_ids, np.int64)
        ctx_wts = np.ascontiguousarray(ctx_wts, np.float32)
        cand_ids = np.asarray(cand_ids)
        cand_wts = np.asarray(cand_wts)
        fk = self.num_fields - ctx_ids.size
        if cand_ids.dtype != np.int64 or cand_wts.dtype != np.float32 or not (cand_ids.flags.c_contiguous and cand_wts.flags.c_contiguous):
            raise TypeError("cand_ids must be C-contiguous int64 and cand_wts C-contiguous float32")
        if fk < 1 or cand_ids.size % fk or cand_ids.size != cand_wts.size or ctx_wts.size != ctx_ids.size:
            raise ValueError("candidate lists need n_cand * %d elements each" % fk)
        n = cand_ids.size // fk
        if out is None:
            out = np.empty(n, np.float32)
        t = C.c_int64()
        _check(self.lib.ctr_predict_shared_async(self.handle, ctx_ids.ctypes.data, ctx_wts.ctypes.data, ctx_ids.size,
                                                 cand_ids.ctypes.data, cand_wts.ctypes.data, n, out.ctypes.data, None, None, C.byref(t)))
        self._keep[t.value] = (ctx_ids, ctx_wts, cand_ids, cand_wts, out)
        return t.value, out

    @staticmethod
    def packed_layout(n_ctx_fields: int = 0, weighted_fields=(), fixed_id_fields=()) -> CtrPackedLayout:
        """Layout of a packed request: which leading fields are shared context, which fields carry weights, which
        per-candidate fields have one id for the whole request (no id column: the id travels once)."""
        lay = CtrPackedLayout()
        lay.struct_size = C.sizeof(CtrPackedLayout)
        lay.n_ctx_fields = n_ctx_fields
        lay.weighted_mask = sum(1 << int(f) for f in weighted_fields)
        lay.fixed_id_mask = sum(1 << int(f) for f in fixed_id_fields if int(f) >= n_ctx_fields)
        return lay

    def packed_shape(self, layout: CtrPackedLayout):
        """(id columns, weight columns, fixed ids) per candidate of a layout for the loaded model."""
        nc = layout.n_ctx_fields
        cand = ((1 << self.num_fields) - 1) & ~((1 << nc) - 1)
        n_fixed = bin(layout.fixed_id_mask & cand).count("1")
        return (self.num_fields - nc - n_fixed, bin(layout.weighted_mask & cand).count("1"), n_fixed)

    def pack_candidates(self, layout: CtrPackedLayout, ids, wts, rows_out=None, wts_out=None):
        """ctr_pack_candidates: reference lists [n, F] -> (ctx_ids, ctx_wts, cand_rows uint32 [n, Fk], cand_wts [n, Wk])."""
        ids, wts, n = self._check_inputs(ids, wts)
        nc = layout.n_ctx_fields
        fk, wk, n_fixed = self.packed_shape(layout)
        rows = rows_out if rows_out is not None else np.empty((n, fk), np.uint32)
        cw = wts_out if wts_out is not None else np.empty((n, wk), np.float32)
        ctx_i = np.zeros(max(nc + n_fixed, 1), np.int64)
        ctx_w = np.zeros(max(nc, 1), np.float32)
        _check(self.lib.ctr_pack_candidates(self.handle, C.byref(layout), ids.ctypes.data, wts.ctypes.data, n, rows.ctypes.data,
                                            cw.ctypes.data if wk else None, ctx_i.ctypes.data, ctx_w.ctypes.data))
        return ctx_i[:nc + n_fixed], ctx_w[:nc], rows, cw

    def predict_packed_async(self, layout: CtrPackedLayout, ctx_ids, ctx_wts, cand_rows, cand_wts, out: Optional[np.ndarray] = None):
        """Packed request: uint32 table rows, weights only for the layout's weighted fields, optional shared context."""
        ctx_ids = np.ascontiguousarray(ctx_ids, np.int64)
        ctx_wts = np.ascontiguousarray(ctx_wts, np.float32)
        cand_rows = np.asarray(cand_rows)
        cand_wts = np.asarray(cand_wts)
        fk, _, n_fixed = self.packed_shape(layout)
        if cand_rows.dtype != np.uint32 or not cand_rows.flags.c_contiguous or cand_rows.size % fk:
            raise TypeError("cand_rows must be C-contiguous uint32 [n, %d]" % fk)
        if cand_wts.size and (cand_wts.dtype != np.float32 or not cand_wts.flags.c_contiguous):
            raise TypeError("cand_wts must be C-contiguous float32")
        if ctx_ids.size != layout.n_ctx_fields + n_fixed or ctx_wts.size != layout.n_ctx_fields:
            raise ValueError("ctx_ids needs n_ctx_fields + fixed-id entries, ctx_wts n_ctx_fields")
        n = cand_rows.size // fk
        if out is None:
            out = np.empty(n, np.float32)
        t = C.c_int64()
        _check(self.lib.ctr_predict_packed_async(self.handle, C.byref(layout), ctx_ids.ctypes.data if ctx_ids.size else None,
                                                 ctx_wts.ctypes.data if ctx_wts.size else None, cand_rows.ctypes.data,
                                                 cand_wts.ctypes.data if cand_wts.size else None, n, out.ctypes.data, None, None, C.byref(t)))
        self._keep[t.value] = (layout, ctx_ids, ctx_wts, cand_rows, cand_wts, out)
        return t.value, out

    def wait(self, ticket: int, timeout_ms: int = -1) -> np.ndarray:
        rc = self.lib.ctr_wait(self.handle, ticket, timeout_ms)
        if rc == -5:
            raise CtrError(rc, "timed out")
        kept = self._keep.pop(ticket, None)
        _check(rc)
        return kept[-1] if kept else None

    def predict(self, ids, wts, out: Optional[np.ndarray] = None) -> np.ndarray:
        t, out = self.predict_async(ids, wts, out)
        return self.wait(t)

    def predict_sorted(self, ids, wts):
        """processARequest including its sort (DCNClient.java:195): (ascending scores, candidate index of each)."""
        ids, wts, n = self._check_inputs(ids, wts)
        out = np.empty(n, np.float32)
        order = np.empty(n, np.int32)
        _check(self.lib.ctr_predict_sorted(self.handle, ids.ctypes.data, wts.ctypes.data, n, out.ctypes.data, order.ctypes.data))
        return out, order

    def predict_wire(self, request: bytes) -> bytes:
        """Serialised PredictRequest -> serialised PredictResponse (the RPC of DCNClient.java:112)."""
        need = C.c_int64(0)
        cap = 64 + len(request)
        buf = C.create_string_buffer(cap)
        rc = self.lib.ctr_predict_wire(self.handle, request, len(request), buf, cap, C.byref(need))
        if rc == -6 and need.value > cap:
            cap = need.value
            buf = C.create_string_buffer(cap)
            rc = self.lib.ctr_predict_wire(self.handle, request, len(request), buf, cap, C.byref(need))
        _check(rc)
        return buf.raw[:need.value]

    def pinned(self, shape, dtype) -> PinnedBuffer:
        return PinnedBuffer(self, shape, dtype)

    # -- device-resident path ----------------------------------------------------------
    def score_device(self, d_ids: int, d_wts: int, n_cand: int, d_scores: int, stream: int = 0, device_index: int = 0):
        """Raw device pointers (ints) and a cudaStream_t handle (0 = engine stream, blocking)."""
        _check(self.lib.ctr_score_device(self.handle, device_index, d_ids, d_wts, n_cand, d_scores, stream or None))

    def score_device_static(self, d_ids: int, d_wts: int, n_cand: int, d_scores: int, stream: int = 0, device_index: int = 0):
        """ctr_score_device_static: same as score_device, but the engine allocates nothing; the
        activations live in the library's fixed scratch (<= STATIC_MAX_CANDIDATES per call)."""
        _check(self.lib.ctr_score_device_static(self.handle, device_index, d_ids, d_wts, n_cand, d_scores, stream or None))

    def score_devices_exchange(self, d_ids, d_wts, n_cand, d_scores):
        """ctr_score_devices_exchange: per-device lists of raw device pointers / counts (row-sharded engine, NCCL all-to-all)."""
        G = self.n_devices
        pi = (C.c_void_p * G)(*d_ids)
        pw = (C.c_void_p * G)(*d_wts)
        po = (C.c_void_p * G)(*d_scores)
        pn = (C.c_int64 * G)(*n_cand)
        _check(self.lib.ctr_score_devices_exchange(self.handle, pi, pw, pn, po))

    def launch_count(self) -> int:
        return int(self.lib.ctr_engine_launch_count(self.handle))

    def link_bytes(self):
        """(host -> device, device -> host) bytes the request path has copied so far."""
        return int(self.lib.ctr_engine_h2d_bytes(self.handle)), int(self.lib.ctr_engine_d2h_bytes(self.handle))

    def set_timing(self, on: bool):
        _check(self.lib.ctr_engine_set_timing(self.handle, 1 if on else 0))

    def stage_ms(self, stage: int, device_index: int = 0) -> float:
        return float(self.lib.ctr_engine_stage_ms(self.handle, device_index, stage))

    def read_rows(self, rows, device_index: int = 0):
        rows = np.ascontiguousarray(rows, np.int64)
        emb = np.empty((rows.size, 16), np.float32)
        w1 = np.empty(rows.size, np.float32)
        _check(self.lib.ctr_debug_read_rows(self.handle, device_index, rows.ctypes.data, rows.size,
                                            emb.ctypes.data, w1.ctypes.data))
        return emb, w1

    def debug_gemm(self, A, W, bias, device_index: int = 0) -> np.ndarray:
        """relu(half(A) @ half(W).T + bias) through the tcgen05 kernel alone; A [M,K], W [N,K]."""
        A = np.ascontiguousarray(A, np.float32); W = np.ascontiguousarray(W, np.float32)
        bias = np.ascontiguousarray(bias, np.float32)
        M, K = A.shape; N = W.shape[0]
        out = np.empty((M, N), np.float32)
        _check(self.lib.ctr_debug_gemm(self.handle, device_index, A.ctypes.data, W.ctypes.data, bias.ctypes.data,
                                       M, N, K, 1, out.ctypes.data))
        return out

    def debug_gather(self, ids, wts, device_index: int = 0):
        ids, wts, n = self._check_inputs(ids, wts)
        d = self.num_fields * 16
        x = np.zeros((n, d), np.float32); wide = np.zeros(n, np.float32)
        _check(self.lib.ctr_debug_gather(self.handle, device_index, ids.ctypes.data, wts.ctypes.data, n,
                                         x.ctypes.data, wide.ctypes.data))
        return x, wide
