"""Builder-defined model math for the scorer behind ``sendRequest`` (test infrastructure).

The reference pins only the tensor contract of the served model (``feat_ids`` int64 [B,F],
``feat_wts`` float32 [B,F] -> ``prediction_node`` float32 [B]; DCNClient.java:97-108, :162).
The arithmetic itself lives in an external SavedModel that is not in the reference
repository, so the definitions below are NOT from the reference: they are the ones
SURVEY.md section 8(a-M) fixes for this project.  PARITY UNPINNED (see oracle/__init__.py).

  x[b,f,:] = E[row(ids[b,f])] * wts[b,f]            x0[b] = concat_f x[b,f,:]   (d = F*K)
  lin[b]   = b0 + sum_f W1[row(ids[b,f])] * wts[b,f]
  LR        logit = lin
  WIDE_DEEP logit = lin + MLP(x0)
  DEEPFM    logit = lin + 0.5*sum_k[(sum_f x_fk)^2 - sum_f x_fk^2] + MLP(x0)
  DCN_V2    x_{l+1} = x0 * (W_l x_l + b_l) + x_l ; logit = w . concat(x_L, MLPhidden(x0)) + b
  p = sigmoid(logit)

``row(id) = id`` (direct) or ``mix64(id) & (V-1)`` (hashed, V a power of two).

Three arithmetic modes:
  "fp64"  master: float64 everywhere from the stored fp32 (or f16) parameters
  "fp32"  float32 everywhere
  "f16"   the CUDA path's contract: x0, hidden activations and tower weights rounded to
          IEEE half (RNE), products accumulated wide, bias/ReLU/head/sigmoid in fp32.
          (Half, not bfloat16: same tensor-core rate, 3 more mantissa bits.  With logits of
          std 1.5 a bfloat16 tower measured |dp| up to 6.5e-3 against the fp64 master, above
          the 2e-3 the north star allows a 16-bit path; half stays below it.)
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

import numpy as np

LR, WIDE_DEEP, DEEPFM, DCN_V2 = 0, 1, 2, 3
KIND_NAMES = {LR: "lr", WIDE_DEEP: "wide_deep", DEEPFM: "deepfm", DCN_V2: "dcn_v2"}
ID_DIRECT, ID_HASHED = 0, 1
TABLE_FP32, TABLE_F16 = 0, 1
TABLE_RAW, TABLE_PROCEDURAL = 0, 1

TAG_E, TAG_W1 = 1, 2

MAGIC = b"CTRW0001"
HEADER_BYTES = 128

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


# ----------------------------------------------------------------------------- integer pieces
def mix64(z) -> np.ndarray:
    """splitmix64 finaliser on uint64 (wrapping).  Integer work: must be bit-exact with
    ``ctr_mix64`` in the CUDA path and in oracle/ctr_oracle.c."""
    z = np.asarray(z).astype(np.uint64, copy=True)
    with np.errstate(over="ignore"):
        z += np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def id_to_row(ids: np.ndarray, num_rows: int, id_mode: int) -> np.ndarray:
    ids = np.asarray(ids, dtype=np.int64)
    if id_mode == ID_DIRECT:
        if ids.size and (ids.min() < 0 or ids.max() >= num_rows):
            raise IndexError("id out of range for a direct table of %d rows" % num_rows)
        return ids.astype(np.uint64)
    assert num_rows & (num_rows - 1) == 0, "hashed tables need a power-of-two row count"
    return mix64(ids.view(np.uint64)) & np.uint64(num_rows - 1)


def proc_stream_key(seed: int, tag: int) -> np.uint64:
    with np.errstate(over="ignore"):
        return mix64(np.uint64(seed) * np.uint64(0x100000001B3) + np.uint64(tag))[()]


def proc_unit(seed: int, tag: int, idx) -> np.ndarray:
    """Procedural value in [-1, 1) on a 2^-23 grid, exactly representable in fp32:
    u24 = top 24 bits of mix64(key + idx);  t = u24 * 2^-23 - 1."""
    key = proc_stream_key(seed, tag)
    with np.errstate(over="ignore"):
        u = mix64(np.asarray(idx, dtype=np.uint64) + key)
    u24 = (u >> np.uint64(40)).astype(np.float32)
    return u24 * np.float32(2.0 ** -23) - np.float32(1.0)


def round_f16(x: np.ndarray) -> np.ndarray:
    """float32 -> nearest IEEE half (ties to even) -> float32, saturating at +-65504 instead of overflowing to
    infinity: what the kernels' cvt.rn.satfinite.f16x2.f32 does wherever an activation is stored as half."""
    x = np.clip(np.asarray(x, dtype=np.float32), np.float32(-65504.0), np.float32(65504.0))      # NaN stays NaN
    return x.astype(np.float16).astype(np.float32)


# ----------------------------------------------------------------------------- spec / weights
@dataclass
class ModelSpec:
    kind: int
    num_fields: int
    emb_dim: int = 16
    num_rows: int = 64
    id_mode: int = ID_DIRECT
    table_dtype: int = TABLE_FP32
    hidden: Tuple[int, ...] = ()
    n_cross: int = 0
    table_mode: int = TABLE_RAW
    seed: int = 1234
    scale_e: float = 0.5
    scale_w1: float = 0.5

    @property
    def d(self) -> int:
        return self.num_fields * self.emb_dim

    @property
    def head_in(self) -> int:
        if self.kind == LR:
            return 0
        h = self.hidden[-1] if self.hidden else 0
        return h + (self.d if self.kind == DCN_V2 else 0)


@dataclass
class Weights:
    spec: ModelSpec
    b0: np.float32 = np.float32(0)
    w1: Optional[np.ndarray] = None              # [V] fp32 (raw tables only)
    emb: Optional[np.ndarray] = None             # [V,K] fp32 (raw tables only)
    mlp: List[Tuple[np.ndarray, np.ndarray]] = field(default_factory=list)     # (W[out,in], b[out])
    cross: List[Tuple[np.ndarray, np.ndarray]] = field(default_factory=list)   # (W[d,d], b[d])
    w_out: Optional[np.ndarray] = None           # [head_in]
    b_out: np.float32 = np.float32(0)

    def rows(self, row: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        """(E[row] as float32 [...,K], W1[row] float32 [...]) with the table dtype applied."""
        s = self.spec
        row = np.asarray(row, dtype=np.uint64)
        if s.table_mode == TABLE_RAW:
            e = self.emb[row.astype(np.int64)]
            w = self.w1[row.astype(np.int64)]
        else:
            k = np.arange(s.emb_dim, dtype=np.uint64)
            with np.errstate(over="ignore"):
                eidx = row[..., None] * np.uint64(s.emb_dim) + k
            e = proc_unit(s.seed, TAG_E, eidx) * np.float32(s.scale_e)
            w = proc_unit(s.seed, TAG_W1, row) * np.float32(s.scale_w1)
        e = e.astype(np.float32)
        if s.table_dtype == TABLE_F16:
            e = round_f16(e)
        return e, w.astype(np.float32)


def make_weights(spec: ModelSpec, rng_seed: int = 1234, target_logit_std: float = 1.5,
                 calib_inputs: Optional[Tuple[np.ndarray, np.ndarray]] = None) -> Weights:
    """Random-init parameters.  Tower: Glorot-uniform; table: U(-scale, scale).  When
    ``calib_inputs`` is given the head is rescaled so the logits over that sample have a
    standard deviation near ``target_logit_std`` -- otherwise every probability sits at 0.5
    and an absolute tolerance on p would be vacuous (hard part 7)."""
    rng = np.random.Generator(np.random.PCG64(rng_seed))
    w = Weights(spec)
    w.b0 = np.float32(rng.uniform(-0.1, 0.1))
    if spec.table_mode == TABLE_RAW:
        w.w1 = rng.uniform(-spec.scale_w1, spec.scale_w1, spec.num_rows).astype(np.float32)
        w.emb = rng.uniform(-spec.scale_e, spec.scale_e, (spec.num_rows, spec.emb_dim)).astype(np.float32)
    d = spec.d
    if spec.kind != LR:
        fan_in = d
        for h in spec.hidden:
            lim = np.sqrt(6.0 / (fan_in + h))
            w.mlp.append((rng.uniform(-lim, lim, (h, fan_in)).astype(np.float32),
                          rng.uniform(-0.05, 0.05, h).astype(np.float32)))
            fan_in = h
        for _ in range(spec.n_cross if spec.kind == DCN_V2 else 0):
            lim = np.sqrt(6.0 / (2 * d)) * 0.5
            w.cross.append((rng.uniform(-lim, lim, (d, d)).astype(np.float32),
                            rng.uniform(-0.05, 0.05, d).astype(np.float32)))
        hin = spec.head_in
        lim = np.sqrt(6.0 / (hin + 1))
        w.w_out = rng.uniform(-lim, lim, hin).astype(np.float32)
        w.b_out = np.float32(rng.uniform(-0.1, 0.1))
        if calib_inputs is not None:
            ids, wts = calib_inputs
            w_probe = w
            logit = forward_logits(w_probe, ids, wts, mode="fp64")
            deep = logit - _wide_part(w_probe, ids, wts)
            sd = float(np.std(deep))
            if sd > 0:
                w.w_out = (w.w_out * np.float32(target_logit_std / sd)).astype(np.float32)
    return w


# ----------------------------------------------------------------------------- weight file
def _header(spec: ModelSpec) -> bytes:
    hid = list(spec.hidden) + [0] * (4 - len(spec.hidden))
    h = MAGIC + struct.pack("<12i", spec.kind, spec.num_fields, spec.emb_dim, spec.id_mode,
                            spec.table_dtype, spec.table_mode, len(spec.hidden), *hid, spec.n_cross)
    h += struct.pack("<qQff", spec.num_rows, spec.seed, spec.scale_e, spec.scale_w1)
    assert len(h) <= HEADER_BYTES
    return h + b"\0" * (HEADER_BYTES - len(h))


def save_weights(w: Weights, path: str) -> None:
    """Flat weight file: 128-byte header, then little-endian fp32 tensors in a fixed order
    (b0; [W1, E if raw]; per hidden layer W[out,in], b; per cross layer W[d,d], b; w_out, b_out).
    An independent restatement of the product's format (distributed-tf-serving_b200/weights.py)."""
    s = w.spec
    with open(path, "wb") as f:
        f.write(_header(s))
        f.write(np.float32(w.b0).tobytes())
        if s.table_mode == TABLE_RAW:
            f.write(np.ascontiguousarray(w.w1, np.float32).tobytes())
            f.write(np.ascontiguousarray(w.emb, np.float32).tobytes())
        for W, b in w.mlp:
            f.write(np.ascontiguousarray(W, np.float32).tobytes())
            f.write(np.ascontiguousarray(b, np.float32).tobytes())
        for W, b in w.cross:
            f.write(np.ascontiguousarray(W, np.float32).tobytes())
            f.write(np.ascontiguousarray(b, np.float32).tobytes())
        if s.kind != LR:
            f.write(np.ascontiguousarray(w.w_out, np.float32).tobytes())
            f.write(np.float32(w.b_out).tobytes())


def load_weights(path: str) -> Weights:
    with open(path, "rb") as f:
        raw = f.read()
    assert raw[:8] == MAGIC, "bad magic"
    (kind, F, K, id_mode, tdt, tmode, nh, h0, h1, h2, h3, ncross) = struct.unpack_from("<12i", raw, 8)
    V, seed, se, sw = struct.unpack_from("<qQff", raw, 8 + 48)
    spec = ModelSpec(kind, F, K, V, id_mode, tdt, tuple([h0, h1, h2, h3][:nh]), ncross, tmode, seed, se, sw)
    a = np.frombuffer(raw, dtype="<f4", offset=HEADER_BYTES)
    pos = 0

    def take(n, shape=None):
        nonlocal pos
        out = a[pos:pos + n].copy()
        pos += n
        return out.reshape(shape) if shape else out

    w = Weights(spec)
    w.b0 = np.float32(take(1)[0])
    if tmode == TABLE_RAW:
        w.w1 = take(V)
        w.emb = take(V * K, (V, K))
    d = spec.d
    if kind != LR:
        fan_in = d
        for h in spec.hidden:
            W = take(h * fan_in, (h, fan_in))
            w.mlp.append((W, take(h)))
            fan_in = h
        for _ in range(ncross if kind == DCN_V2 else 0):
            W = take(d * d, (d, d))
            w.cross.append((W, take(d)))
        w.w_out = take(spec.head_in)
        w.b_out = np.float32(take(1)[0])
    assert pos == a.shape[0], "trailing bytes in weight file"
    return w


# ----------------------------------------------------------------------------- forward
def _gather(w: Weights, ids, wts):
    s = w.spec
    ids = np.asarray(ids, dtype=np.int64).reshape(-1, s.num_fields)
    wts = np.asarray(wts, dtype=np.float32).reshape(-1, s.num_fields)
    row = id_to_row(ids, s.num_rows, s.id_mode)
    e, w1 = w.rows(row)                      # [B,F,K] fp32, [B,F] fp32
    return e, w1, wts


def _wide_part(w: Weights, ids, wts) -> np.ndarray:
    """lin (+ FM for DeepFM), float64; zero for DCN-v2 (its logit has no first-order term)."""
    s = w.spec
    e, w1, wt = _gather(w, ids, wts)
    if s.kind == DCN_V2:
        return np.zeros(e.shape[0])
    lin = np.float64(w.b0) + (w1.astype(np.float64) * wt.astype(np.float64)).sum(1)
    if s.kind == DEEPFM:
        x = e.astype(np.float64) * wt.astype(np.float64)[..., None]
        lin = lin + 0.5 * ((x.sum(1) ** 2).sum(1) - (x ** 2).sum((1, 2)))
    return lin


def forward_logits(w: Weights, ids, wts, mode: str = "fp64") -> np.ndarray:
    """Logits [B].  ``mode`` as in the module docstring."""
    s = w.spec
    e, w1, wt = _gather(w, ids, wts)
    B = e.shape[0]
    if mode == "fp64":
        ft = np.float64
    elif mode in ("fp32", "f16"):
        ft = np.float32
    else:
        raise ValueError(mode)

    # wide / first-order / FM terms: fp32 products of fp32 values in every mode but fp64
    x = (e.astype(ft) * wt.astype(ft)[..., None])            # [B,F,K]
    if s.kind == DCN_V2:
        wide = np.zeros(B, ft)
    else:
        wide = ft(w.b0) + (w1.astype(ft) * wt.astype(ft)).sum(1, dtype=ft)
        if s.kind == DEEPFM:
            wide = wide + ft(0.5) * ((x.sum(1, dtype=ft) ** 2).sum(1, dtype=ft) - (x ** 2).sum((1, 2), dtype=ft))
    if s.kind == LR:
        return wide.astype(ft)

    q = round_f16 if mode == "f16" else (lambda a: a)
    acc = np.float64 if mode != "fp32" else np.float32      # "f16": exact products, wide sum

    x0 = q(x.reshape(B, s.d).astype(np.float32)).astype(ft) if mode == "f16" else x.reshape(B, s.d)

    def dense(a, W, b, relu):
        Wq = q(W) if mode == "f16" else W
        y = (a.astype(acc) @ Wq.astype(acc).T).astype(ft) + b.astype(ft)
        return np.maximum(y, 0) if relu else y

    h = x0
    for li, (W, b) in enumerate(w.mlp):
        h = dense(h, W, b, relu=True)
        last = li == len(w.mlp) - 1
        if mode == "f16" and not (last and head_is_fused(s)):
            h = q(h.astype(np.float32)).astype(ft)           # hidden activations stored as f16

    if s.kind == DCN_V2:
        xl = x0
        for ci, (W, b) in enumerate(w.cross):
            y = dense(xl, W, b, relu=False)
            xl = x0 * y + xl
            if mode == "f16" and ci != len(w.cross) - 1:     # the last x_L feeds the head in fp32
                xl = q(xl.astype(np.float32)).astype(ft)
        feat = np.concatenate([xl, h], axis=1)
    else:
        feat = h
    deep = (feat.astype(np.float64) @ w.w_out.astype(np.float64)).astype(ft) if mode != "fp32" \
        else feat @ w.w_out
    return (wide + deep + ft(w.b_out)).astype(ft)


def head_is_fused(spec: ModelSpec) -> bool:
    """Kernel-contract detail for the "f16" mode: the CUDA path applies the output layer to the
    fp32 accumulators of the last hidden layer (and of the last cross layer), so those
    activations are never rounded to half."""
    return bool(spec.hidden)


def sigmoid(z: np.ndarray) -> np.ndarray:
    z = np.asarray(z)
    return (1.0 / (1.0 + np.exp(-z.astype(np.float64)))).astype(np.float32 if z.dtype != np.float64 else np.float64)


def forward(w: Weights, ids, wts, mode: str = "fp64") -> np.ndarray:
    """Probabilities [B]; float32 except in "fp64" mode."""
    return sigmoid(forward_logits(w, ids, wts, mode))


def as_model(w: Weights, mode: str = "fp32"):
    """A ``model(ids, wts) -> float32[n]`` callable for ``oracle.dcn_client.send_request``."""
    return lambda ids, wts: forward(w, ids, wts, mode).astype(np.float32)


# ----------------------------------------------------------------------------- synthetic inputs
def criteo_inputs(n_cand: int, seed: int = 4321, n_dense: int = 13, n_sparse: int = 26,
                  rows_per_field: int = 1 << 22, dist: str = "uniform") -> Tuple[np.ndarray, np.ndarray]:
    """a global id space.  Dense field j -> id j+1 (1..13),
    wt = log1p(lognormal(0,1)); sparse field s -> id = 14 + s*R + h, wt = 1.0, h uniform
    (roofline runs) or Zipf(1.05) over R."""
    rng = np.random.Generator(np.random.PCG64(seed))
    F = n_dense + n_sparse
    ids = np.empty((n_cand, F), np.int64)
    wts = np.ones((n_cand, F), np.float32)
    ids[:, :n_dense] = np.arange(1, n_dense + 1)
    wts[:, :n_dense] = np.log1p(rng.lognormal(0.0, 1.0, (n_cand, n_dense))).astype(np.float32)
    if dist == "uniform":
        h = rng.integers(0, rows_per_field, (n_cand, n_sparse), dtype=np.int64)
    else:
        h = (rng.zipf(1.05, (n_cand, n_sparse)) - 1) % rows_per_field
    ids[:, n_dense:] = (n_dense + 1) + np.arange(n_sparse, dtype=np.int64) * rows_per_field + h
    return ids, wts


def criteo_num_rows(n_dense: int = 13, n_sparse: int = 26, rows_per_field: int = 1 << 22) -> int:
    return n_dense + 1 + n_sparse * rows_per_field
