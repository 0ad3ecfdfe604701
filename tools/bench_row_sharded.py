#!/usr/bin/env python
"""BASELINE.json config 4 (shape): DCN-v2 with hashed ids into ONE 2^k-row table that is row-sharded
over the GPUs of the engine (row mod G); the gather kernel fetches remote rows through NVLink peer
mappings.  One engine over G devices, every device scores its own batch concurrently (weak scaling,
like bench.py), device-resident inputs.  Prints candidates/s for replicated vs row-sharded tables.

  python tools/bench_row_sharded.py [--log2-rows 27] [--candidates 320000] [--steps 20]
"""
import argparse
import os
import sys
import tempfile
import threading

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log2-rows", type=int, default=27)
    ap.add_argument("--candidates", type=int, default=320000)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--model", default="dcn_v2")
    ap.add_argument("--skip-replicated", action="store_true", help="only the row-sharded engine (a 2^30-row fp32 table is 137 GB replicated)")
    ap.add_argument("--table-dtype", default="fp32", choices=["fp32", "f16"])
    args = ap.parse_args()
    import torch
    import dtfs_b200  # noqa: F401
    from dtfs_b200 import binding, weights as W
    G = torch.cuda.device_count()
    tmp = tempfile.mkdtemp(prefix="ctr_rs_")
    wpath = os.path.join(tmp, "m.ctrw")
    rpf = ((1 << args.log2_rows) - 14) // bench.N_SPARSE
    num_rows = bench.write_model(wpath, rpf, model=args.model, hashed=True, table_dtype=1 if args.table_dtype == "f16" else 0)
    assert num_rows == 1 << args.log2_rows, num_rows
    n = args.candidates
    out = {}
    modes = ((binding.TABLES_REPLICATED, "replicated"), (binding.TABLES_ROW_SHARDED, "row_sharded"))
    for mode, name in modes[1:] if args.skip_replicated else modes:
        eng = binding.Engine(device_ids=tuple(range(G)), max_candidates=32768, shard_mode=mode)
        info = eng.load_weights(wpath)
        bufs = []
        for g in range(G):
            with torch.cuda.device(g):
                ids, wts = bench.device_inputs(torch, n, rpf, 100 + g)
                bufs.append((ids, wts, torch.empty(n, dtype=torch.float32, device="cuda")))
        def run(g, k):
            ids, wts, o = bufs[g]
            for _ in range(k):
                eng.score_device(ids.data_ptr(), wts.data_ptr(), n, o.data_ptr(), device_index=g)
        for k in (3, args.steps):                       # warm-up, then timed
            ths = [threading.Thread(target=run, args=(g, k)) for g in range(G)]
            for g in range(G):
                torch.cuda.synchronize(g)
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(G)]
            import time
            t0 = time.perf_counter()
            [t.start() for t in ths]
            [t.join() for t in ths]
            for g in range(G):
                torch.cuda.synchronize(g)
            dt = time.perf_counter() - t0
        out[name] = {"candidates_per_s": G * n * args.steps / dt, "table_bytes_per_gpu": int(info.table_bytes_per_device),
                     "scores_head": bufs[0][2][:4].cpu().numpy().tolist()}
        if mode == binding.TABLES_ROW_SHARDED:
            # the same batches through the NCCL all-to-all variant (one host call drives all the devices)
            import time
            args_x = ([b[0].data_ptr() for b in bufs], [b[1].data_ptr() for b in bufs], [n] * G, [b[2].data_ptr() for b in bufs])
            keep = [b[2].clone() for b in bufs]
            for _ in range(2):
                eng.score_devices_exchange(*args_x)
            t0 = time.perf_counter()
            for _ in range(args.steps):
                eng.score_devices_exchange(*args_x)
            dt = time.perf_counter() - t0
            out["row_sharded_nccl_exchange"] = {"candidates_per_s": G * n * args.steps / dt,
                                                "same_bits_as_peer_loads": all(bool(torch.equal(k, b[2])) for k, b in zip(keep, bufs))}
        eng.close()
    out["gpus"] = G
    out["rows"] = num_rows
    if "replicated" in out:
        out["same_scores"] = out["replicated"]["scores_head"] == out["row_sharded"]["scores_head"]
    import json
    print(json.dumps(out))


if __name__ == "__main__":
    main()
