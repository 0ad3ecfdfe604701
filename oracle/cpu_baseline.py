"""CPU restatement of one whole reference request, timed as the CPU baseline (test
infrastructure; see oracle/__init__.py).  "CPU restatement, not TF-Serving": the reference
client and its TF-Serving backend cannot run on this image (no JDK, no tensorflow; BASELINE.md
section 2), so this restates the same steps on the host cores:

  split into P shards (DCNClient.java:46-55, on candidates)  ->  per shard, on a pool thread
  (DC:148-156): protobuf-wire encode of feat_ids / feat_wts (DC:97-108; tensor.proto:48,64)
  -> decode (the server's parse) -> model forward in fp32 (embedding gather in C/OpenMP from a
  table held in host RAM, dense tower with torch CPU matmul = the MKL/oneDNN path a CPU
  TF-Serving would use) -> encode / decode the scores (DC:162) -> concatenate in shard order
  (DC:161-164) -> ascending sort (DC:195).
"""
from __future__ import annotations

import os
import time
from concurrent.futures import ThreadPoolExecutor
from typing import Optional

import numpy as np

from . import clib, dcn_client as odc, model as om


class CpuServer:
    """One "TF-Serving replica" on the host: weights + an in-RAM table."""

    def __init__(self, w: om.Weights, threads: int, table_rows_cap: Optional[int] = None):
        import torch
        self.torch = torch
        self.w = w
        s = w.spec
        self.threads = threads
        self.table_note = "full table in host RAM"
        if s.table_mode == om.TABLE_PROCEDURAL:
            rows = s.num_rows
            self.row_mask = None
            if table_rows_cap is not None and rows > table_rows_cap:
                # not enough host RAM for the real table: fold rows into a smaller one (favours the CPU)
                rows = table_rows_cap
                self.row_mask = rows
                self.table_note = "table folded to %d rows (host RAM)" % rows
            self.emb, self.w1 = clib.fill_table(s.seed, s.scale_e, s.scale_w1, rows, s.emb_dim)
        else:
            self.emb, self.w1 = w.emb, w.w1
            self.row_mask = None
        self.mlp = [(torch.from_numpy(np.ascontiguousarray(W.T)), torch.from_numpy(b)) for W, b in w.mlp]
        self.cross = [(torch.from_numpy(np.ascontiguousarray(W.T)), torch.from_numpy(b)) for W, b in w.cross]
        self.w_out = torch.from_numpy(w.w_out) if w.w_out is not None else None

    def forward(self, ids: np.ndarray, wts: np.ndarray) -> np.ndarray:
        torch = self.torch
        s = self.w.spec
        if self.row_mask is not None:
            ids = ids % self.row_mask
        x0, wide = clib.gather_fp32(ids, wts, self.emb.shape[0], s.id_mode == om.ID_HASHED, self.emb, self.w1,
                                    float(self.w.b0), s.kind == om.DEEPFM)
        if s.kind == om.LR:
            z = torch.from_numpy(wide)
        else:
            x0t = torch.from_numpy(x0)
            h = x0t
            for Wt, b in self.mlp:
                h = torch.relu(torch.addmm(b, h, Wt))
            if s.kind == om.DCN_V2:
                xl = x0t
                for Wt, b in self.cross:
                    xl = x0t * torch.addmm(b, xl, Wt) + xl
                h = torch.cat([xl, h], dim=1)
                z = h @ self.w_out + float(self.w.b_out)
            else:
                z = torch.from_numpy(wide) + h @ self.w_out + float(self.w.b_out)
        return torch.sigmoid(z).numpy().astype(np.float32)

    def serve(self, request_bytes: bytes, cap: int) -> bytes:
        """The server side of DC:112: parse -> run -> serialise."""
        ids, wts = clib.decode_predict_request(request_bytes, cap)
        return clib.encode_predict_response(self.forward(ids, wts))


class CpuClient:
    """The client side: DC:137-203 with P in-process replicas."""

    def __init__(self, w: om.Weights, n_shards: int = 2, with_wire: bool = True, table_rows_cap: Optional[int] = None):
        import torch
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(max(1, self.cores // n_shards))
        self.P = n_shards
        self.with_wire = with_wire
        self.server = CpuServer(w, max(1, self.cores // n_shards), table_rows_cap)   # replicas share read-only weights
        self.pool = ThreadPoolExecutor(odc.POOL_THREADS)
        self.F = w.spec.num_fields

    def _send(self, ids: np.ndarray, wts: np.ndarray) -> np.ndarray:
        if not self.with_wire:
            return self.server.forward(ids, wts)
        req = clib.encode_predict_request(ids, wts)                                # DC:97-109
        resp = self.server.serve(req, ids.size)                                    # DC:112
        return clib.decode_predict_response(resp, ids.shape[0])                    # DC:162

    def process_a_request(self, ids: np.ndarray, wts: np.ndarray, sort: bool = True) -> np.ndarray:
        bounds = odc.partition_candidates(ids.shape[0], self.P)
        futs = [self.pool.submit(self._send, ids[a:b], wts[a:b]) for a, b in bounds if b > a]
        full = np.concatenate([f.result() for f in futs]) if futs else np.zeros(0, np.float32)
        return odc.java_float_sort(full) if sort else full

    def close(self):
        self.pool.shutdown()


def time_requests(client: CpuClient, ids: np.ndarray, wts: np.ndarray, n_requests: int, warmup: int = 1):
    """Seconds per request (mean) over n_requests sequential requests of the given inputs."""
    for _ in range(warmup):
        client.process_a_request(ids, wts)
    t0 = time.perf_counter()
    for _ in range(n_requests):
        client.process_a_request(ids, wts)
    return (time.perf_counter() - t0) / max(n_requests, 1)
