"""Weight import from a TensorFlow SavedModel / checkpoint.

The reference addresses its model only by name: TF-Serving loads a SavedModel called "DCN" and serves its
``serving_default`` signature with inputs ``feat_ids`` (DT_INT64 [B, F]) and ``feat_wts`` (DT_FLOAT [B, F]) and the
output ``prediction_node`` (DCNClient.java:33-35, :97-108, :162; the signature lives in the SavedModel's
``SignatureDef``, meta_graph.proto:297-311).  The repository holds no such model.  ``tools/export_savedmodel.py``
builds one of the four architectures of SURVEY.md section 8(a-M) in TensorFlow with exactly that signature and
these variable names; this module turns the variables of such a model into the engine's CTRW0001 weight file, so
that the same weights can be served by TF-Serving (off this box) and scored here, and parity can be pinned
against the real thing.

Variable names (one flat name space, float32):

    first_order/w        [V]          W1
    first_order/b0       []           b0
    embedding/table      [V, 16]      E
    mlp/dense_<l>/kernel [in, out]    TF layout; transposed to the engine's [out, in]
    mlp/dense_<l>/bias   [out]
    cross/layer_<l>/kernel [d, d]     applied as x0 * (W x_l + b) + x_l, W stored [out, in] already
    cross/layer_<l>/bias   [d]
    head/kernel          [head_in]    order: x_L (DCN-v2 only), then the last hidden layer
    head/bias            []

Two sources: a directory with a TensorFlow checkpoint (needs tensorflow, absent from this image -- source only), or
the ``variables.npz`` the exporter writes next to the SavedModel (numpy only; this is what the tests use).
"""
from __future__ import annotations

import os
from typing import Dict

import numpy as np

from . import weights as W

KIND_BY_NAME = {"lr": W.MODEL_LR, "wide_deep": W.MODEL_WIDE_DEEP, "deepfm": W.MODEL_DEEPFM, "dcn_v2": W.MODEL_DCN_V2}


def load_variables(path: str) -> Dict[str, np.ndarray]:
    """``path``: a ``variables.npz`` file, a SavedModel directory that holds one, or a TF checkpoint prefix / directory."""
    if os.path.isdir(path) and os.path.exists(os.path.join(path, "variables.npz")):
        path = os.path.join(path, "variables.npz")
    if path.endswith(".npz"):
        with np.load(path) as z:
            return {k: z[k] for k in z.files}
    try:
        import tensorflow as tf  # noqa: PLC0415  (not in this image: this branch is source only)
    except ImportError as ex:
        raise RuntimeError("reading a TensorFlow checkpoint needs tensorflow; pass the exporter's variables.npz instead") from ex
    ckpt = os.path.join(path, "variables", "variables") if os.path.isdir(os.path.join(path, "variables")) else path
    reader = tf.train.load_checkpoint(ckpt)
    out = {}
    for name in reader.get_variable_to_shape_map():
        # object-graph checkpoints name a variable "<attribute path>/.ATTRIBUTES/VARIABLE_VALUE"
        key = name.replace("/.ATTRIBUTES/VARIABLE_VALUE", "")
        out[key] = reader.get_tensor(name)
    return out


def convert(variables: Dict[str, np.ndarray], model: str, out_path: str, id_mode: int = W.ID_DIRECT, table_dtype: int = W.TABLE_FP32) -> dict:
    """Writes ``out_path`` (CTRW0001) from the named variables.  Returns the header fields it wrote."""
    kind = KIND_BY_NAME[model]
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    emb = f32(variables["embedding/table"])
    w1 = f32(variables["first_order/w"]).reshape(-1)
    V, K = emb.shape
    if K != 16 or w1.shape != (V,):
        raise ValueError("embedding/table must be [V, 16] and first_order/w [V]")
    mlp, l = [], 0
    while "mlp/dense_%d/kernel" % l in variables:
        mlp.append((f32(variables["mlp/dense_%d/kernel" % l]).T.copy(), f32(variables["mlp/dense_%d/bias" % l])))
        l += 1
    cross, l = [], 0
    while "cross/layer_%d/kernel" % l in variables:
        cross.append((f32(variables["cross/layer_%d/kernel" % l]), f32(variables["cross/layer_%d/bias" % l])))
        l += 1
    if kind == W.MODEL_LR:
        mlp, cross = [], []
    if kind != W.MODEL_DCN_V2:
        cross = []
    if kind != W.MODEL_LR and not mlp:
        raise ValueError("model %r needs mlp/dense_<l>/kernel variables" % model)
    d = mlp[0][0].shape[1] if mlp else 0
    num_fields = d // K if mlp else int(variables.get("meta/num_fields", 0))
    if not num_fields:
        raise ValueError("an LR model carries no layer to infer the field count from: add meta/num_fields")
    w_out = f32(variables["head/kernel"]).reshape(-1) if kind != W.MODEL_LR else None
    b_out = float(np.asarray(variables.get("head/bias", 0.0)).reshape(-1)[0]) if kind != W.MODEL_LR else 0.0
    W.write_weight_file(out_path, model_kind=kind, num_fields=num_fields, num_rows=V, b0=float(np.asarray(variables["first_order/b0"]).reshape(-1)[0]),
                        id_mode=id_mode, table_dtype=table_dtype, w1=w1, emb=emb, mlp=mlp, cross=cross, w_out=w_out, b_out=b_out)
    return W.read_header(out_path)


def main(argv=None):
    import argparse
    ap = argparse.ArgumentParser(description="TensorFlow SavedModel / checkpoint / variables.npz -> CTRW0001 weight file")
    ap.add_argument("source")
    ap.add_argument("out")
    ap.add_argument("--model", required=True, choices=sorted(KIND_BY_NAME))
    ap.add_argument("--hashed", action="store_true")
    ap.add_argument("--table-dtype", default="fp32", choices=["fp32", "f16"])
    a = ap.parse_args(argv)
    h = convert(load_variables(a.source), a.model, a.out, W.ID_HASHED if a.hashed else W.ID_DIRECT, W.TABLE_F16 if a.table_dtype == "f16" else W.TABLE_FP32)
    print(h)


if __name__ == "__main__":
    main()
