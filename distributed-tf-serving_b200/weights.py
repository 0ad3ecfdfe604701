"""Flat weight file of the engine (what ``ctr_engine_load_weights`` reads).

Stands in for the SavedModel TF-Serving loads for the reference client
(model "DCN", DCNClient.java:33-34); the reference repository holds no model, so the format
is this project's own.

Layout (little endian)::

    0    8   magic "CTRW0001"
    8   48   int32 x12: model_kind, num_fields, emb_dim, id_mode, table_dtype, table_mode,
                        n_hidden, hidden[4], n_cross
    56   8   int64  num_rows
    64   8   uint64 table seed           (procedural tables)
    72   8   float32 scale_e, scale_w1   (procedural tables)
    ...      zero padding to 128 bytes
    128      float32 tensors, in order:
             b0[1]
             W1[num_rows], E[num_rows*emb_dim]                   only when table_mode == 0 (raw)
             per hidden layer l: W_l[hidden_l * fan_in_l], b_l[hidden_l]      (row-major [out, in])
             per cross layer  l: Wc_l[d*d], bc_l[d]                           (DCN-v2 only)
             w_out[head_in], b_out[1]                                         (not for LR)

``table_mode == 1`` (procedural) stores no table: the engine generates row r on the device as
``E[r,k] = u(seed, 1, r*emb_dim+k) * scale_e`` and ``W1[r] = u(seed, 2, r) * scale_w1`` with
``u(seed, tag, i) = (mix64(mix64(seed*0x100000001B3 + tag) + i) >> 40) * 2^-23 - 1`` -- this is how
a 7 GB (config 2) or 137 GB (config 4) table exists without a file of that size.
"""
from __future__ import annotations

import struct
from typing import Optional, Sequence

import numpy as np

MAGIC = b"CTRW0001"
HEADER_BYTES = 128
MODEL_LR, MODEL_WIDE_DEEP, MODEL_DEEPFM, MODEL_DCN_V2 = 0, 1, 2, 3
ID_DIRECT, ID_HASHED = 0, 1
TABLE_FP32, TABLE_F16 = 0, 1
TABLE_RAW, TABLE_PROCEDURAL = 0, 1


def write_weight_file(path: str, *, model_kind: int, num_fields: int, num_rows: int, b0: float,
                      emb_dim: int = 16, id_mode: int = ID_DIRECT, table_dtype: int = TABLE_FP32,
                      w1: Optional[np.ndarray] = None, emb: Optional[np.ndarray] = None,
                      seed: int = 0, scale_e: float = 0.0, scale_w1: float = 0.0,
                      mlp: Sequence = (), cross: Sequence = (), w_out: Optional[np.ndarray] = None,
                      b_out: float = 0.0) -> None:
    """``mlp`` / ``cross``: sequences of ``(W[out, in], b[out])`` float32 arrays.  Pass ``w1`` and
    ``emb`` for a raw table, or leave them ``None`` for a procedural one (seed / scales)."""
    hidden = [int(W.shape[0]) for W, _ in mlp]
    if len(hidden) > 4:
        raise ValueError("at most 4 hidden layers")
    raw = w1 is not None
    if raw and (emb is None or emb.shape != (num_rows, emb_dim) or w1.shape != (num_rows,)):
        raise ValueError("raw table needs w1[num_rows] and emb[num_rows, emb_dim]")
    d = num_fields * emb_dim
    fan_in = d
    for W, b in mlp:
        if W.shape != (b.shape[0], fan_in):
            raise ValueError("hidden layer weight has shape %r, expected (%d, %d)" % (W.shape, b.shape[0], fan_in))
        fan_in = W.shape[0]
    for W, b in cross:
        if W.shape != (d, d) or b.shape != (d,):
            raise ValueError("cross layers are [d, d] with bias [d]")
    head_in = 0 if model_kind == MODEL_LR else (hidden[-1] if hidden else 0) + (d if model_kind == MODEL_DCN_V2 else 0)
    if model_kind != MODEL_LR and (w_out is None or w_out.shape != (head_in,)):
        raise ValueError("w_out must have %d entries" % head_in)
    hdr = MAGIC + struct.pack("<12i", model_kind, num_fields, emb_dim, id_mode, table_dtype,
                              TABLE_RAW if raw else TABLE_PROCEDURAL, len(hidden), *(hidden + [0] * (4 - len(hidden))),
                              len(cross))
    hdr += struct.pack("<qQff", num_rows, seed, scale_e, scale_w1)
    hdr += b"\0" * (HEADER_BYTES - len(hdr))
    f32 = lambda a: np.ascontiguousarray(a, dtype="<f4").tobytes()
    with open(path, "wb") as f:
        f.write(hdr)
        f.write(f32(np.float32(b0)))
        if raw:
            f.write(f32(w1))
            f.write(f32(emb))
        for W, b in list(mlp) + list(cross):
            f.write(f32(W))
            f.write(f32(b))
        if model_kind != MODEL_LR:
            f.write(f32(w_out))
            f.write(f32(np.float32(b_out)))


def read_header(path: str) -> dict:
    with open(path, "rb") as f:
        h = f.read(HEADER_BYTES)
    if h[:8] != MAGIC:
        raise ValueError("not a CTRW0001 weight file")
    v = struct.unpack_from("<12i", h, 8)
    num_rows, seed, se, sw = struct.unpack_from("<qQff", h, 56)
    return dict(model_kind=v[0], num_fields=v[1], emb_dim=v[2], id_mode=v[3], table_dtype=v[4], table_mode=v[5],
                hidden=list(v[7:7 + v[6]]), n_cross=v[11], num_rows=num_rows, seed=seed, scale_e=se, scale_w1=sw)
