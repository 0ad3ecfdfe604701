#!/usr/bin/env python
"""Exports one of the four model families as a TensorFlow SavedModel that TF-Serving can load for the UNMODIFIED
reference client: model name "DCN", signature "serving_default", inputs "feat_ids"
(DT_INT64 [B, F]) / "feat_wts" (DT_FLOAT [B, F]), output "prediction_node" (DT_FLOAT [B, 1]) -- DCNClient.java:33-35,
:97-108, :162; SignatureDef: meta_graph.proto:297-311.  Next to the SavedModel it writes variables.npz, which
distributed-tf-serving_b200/savedmodel_import.py turns into the engine's weight file: the same weights served by
TF-Serving and scored by the engine is what would pin parity against the real reference path.

SOURCE ONLY on this image: TensorFlow is not installed here (nor is tensorflow_model_server), so this file has not
been run; the importer half is tested against the npz layout.  The math is SURVEY.md section 8(a-M), written with
plain TF ops so that what TF-Serving executes is exactly what oracle/model.py restates.

  python tools/export_savedmodel.py --weights model.ctrw --out /models/DCN/1        (then: tensorflow_model_server
         --port=9999 --model_name=DCN --model_base_path=/models/DCN)
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def variables_from_ctrw(path: str) -> dict:
    """CTRW0001 (raw table) -> the flat variable dictionary of savedmodel_import.py."""
    import struct
    from dtfs_b200 import weights as W
    h = W.read_header(path)
    if h["table_mode"] != W.TABLE_RAW:
        raise ValueError("export needs a raw-table weight file (a procedural table has no tensor to export)")
    F, K, V = h["num_fields"], h["emb_dim"], h["num_rows"]
    data = np.fromfile(path, dtype="<f4", offset=W.HEADER_BYTES)
    pos = 0

    def take(n, shape=None):
        nonlocal pos
        a = data[pos:pos + n]
        pos += n
        return a.reshape(shape) if shape else a
    v = {"first_order/b0": take(1)[0], "first_order/w": take(V), "embedding/table": take(V * K, (V, K)), "meta/num_fields": np.int64(F)}
    fan_in = F * K
    for l, hdim in enumerate(h["hidden"]):
        v["mlp/dense_%d/kernel" % l] = take(hdim * fan_in, (hdim, fan_in)).T.copy()       # TF layout [in, out]
        v["mlp/dense_%d/bias" % l] = take(hdim)
        fan_in = hdim
    d = F * K
    for l in range(h["n_cross"]):
        v["cross/layer_%d/kernel" % l] = take(d * d, (d, d))
        v["cross/layer_%d/bias" % l] = take(d)
    if h["model_kind"] != W.MODEL_LR:
        head_in = h["hidden"][-1] + (d if h["model_kind"] == W.MODEL_DCN_V2 else 0)
        v["head/kernel"] = take(head_in)
        v["head/bias"] = take(1)[0]
    assert pos == data.size
    return v, h


def build_and_save(v: dict, h: dict, out_dir: str, hashed: bool):
    import tensorflow as tf  # noqa: PLC0415

    class CtrModel(tf.Module):
        def __init__(self):
            super().__init__(name="DCN")
            self.vars = {k: tf.Variable(np.asarray(a, np.float32), name=k) for k, a in v.items() if k != "meta/num_fields"}
            self.kind, self.hidden, self.n_cross, self.V = h["model_kind"], h["hidden"], h["n_cross"], h["num_rows"]

        @tf.function(input_signature=[tf.TensorSpec([None, h["num_fields"]], tf.int64, name="feat_ids"),
                                      tf.TensorSpec([None, h["num_fields"]], tf.float32, name="feat_wts")])
        def serve(self, feat_ids, feat_wts):
            rows = feat_ids
            if hashed:                                                # splitmix64 finaliser, as common.cuh::mix64
                z = tf.bitcast(feat_ids, tf.uint64) + tf.constant(0x9E3779B97F4A7C15, tf.uint64)
                z = (z ^ (z >> 30)) * tf.constant(0xBF58476D1CE4E5B9, tf.uint64)
                z = (z ^ (z >> 27)) * tf.constant(0x94D049BB133111EB, tf.uint64)
                rows = tf.cast((z ^ (z >> 31)) & tf.constant(self.V - 1, tf.uint64), tf.int64)
            e = tf.gather(self.vars["embedding/table"], rows) * feat_wts[..., None]           # [B, F, K]
            lin = self.vars["first_order/b0"] + tf.reduce_sum(tf.gather(self.vars["first_order/w"], rows) * feat_wts, axis=1)
            if self.kind == 0:
                return {"prediction_node": tf.sigmoid(lin)[:, None]}
            x0 = tf.reshape(e, [tf.shape(e)[0], -1])
            hcur = x0
            for l in range(len(self.hidden)):
                hcur = tf.nn.relu(tf.matmul(hcur, self.vars["mlp/dense_%d/kernel" % l]) + self.vars["mlp/dense_%d/bias" % l])
            feats, wide = hcur, lin
            if self.kind == 2:                                        # DeepFM second-order term
                s = tf.reduce_sum(e, axis=1)
                wide = lin + 0.5 * tf.reduce_sum(s * s - tf.reduce_sum(e * e, axis=1), axis=1)
            if self.kind == 3:                                        # DCN-v2: no wide term, head over [x_L, h]
                xl = x0
                for l in range(self.n_cross):
                    xl = x0 * (tf.matmul(xl, self.vars["cross/layer_%d/kernel" % l], transpose_b=True) + self.vars["cross/layer_%d/bias" % l]) + xl
                feats, wide = tf.concat([xl, hcur], axis=1), 0.0
            logit = tf.linalg.matvec(feats, self.vars["head/kernel"]) + self.vars["head/bias"] + wide
            return {"prediction_node": tf.sigmoid(logit)[:, None]}

    m = CtrModel()
    tf.saved_model.save(m, out_dir, signatures={"serving_default": m.serve})


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--weights", required=True, help="CTRW0001 file with a raw table")
    ap.add_argument("--out", required=True, help="SavedModel version directory, e.g. /models/DCN/1")
    a = ap.parse_args()
    v, h = variables_from_ctrw(a.weights)
    os.makedirs(a.out, exist_ok=True)
    np.savez(os.path.join(a.out, "variables.npz"), **v)
    build_and_save(v, h, a.out, hashed=h["id_mode"] == 1)
    print("SavedModel written to", a.out)


if __name__ == "__main__":
    main()
