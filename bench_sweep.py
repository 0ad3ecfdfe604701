#!/usr/bin/env python
"""BASELINE.json config 5: ranking request stream sweep -- request size 64 ... 65 536 candidates,
closed loop with 6 client threads (the reference's concurrentNum, DCNClient.java:31), p50 / p99
latency and candidates/s through the host C ABI (pinned buffers, H2D + kernels + D2H inside the
window: DCNClient.java:141 -> :198 without the sort).  Writes a markdown table.

  python bench_sweep.py [--out profiles/r01_sweep.md] [--requests 400] [--threads 6]
"""
from __future__ import annotations

import argparse
import os
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (model writer / input generator of the headline benchmark)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep.md"))
    ap.add_argument("--requests", type=int, default=400, help="timed requests per thread and size")
    ap.add_argument("--threads", type=int, default=6)
    ap.add_argument("--rows-per-field", type=int, default=1 << 22)
    args = ap.parse_args()
    import torch
    import dtfs_b200  # noqa: F401
    from dtfs_b200 import binding

    tmp = tempfile.mkdtemp(prefix="ctr_sweep_")
    wpath = os.path.join(tmp, "wd.ctrw")
    bench.write_model(wpath, args.rows_per_field)
    eng = binding.Engine(device_ids=(0,), max_candidates=65536, workers_per_device=4)
    eng.load_weights(wpath)
    sizes = [64 << i for i in range(11)]
    d_ids, d_wts = bench.device_inputs(torch, 65536 * args.threads, args.rows_per_field, 4321)
    h_ids, h_wts = d_ids.cpu().numpy(), d_wts.cpu().numpy()
    del d_ids, d_wts
    pins = []
    for t in range(args.threads):
        pi, pw, po = eng.pinned((65536, bench.F), np.int64), eng.pinned((65536, bench.F), np.float32), eng.pinned((65536,), np.float32)
        pi.array[:] = h_ids[t * 65536:(t + 1) * 65536]
        pw.array[:] = h_wts[t * 65536:(t + 1) * 65536]
        pins.append((pi, pw, po))
    lines = ["| candidates / request | p50 ms | p99 ms | requests/s | candidates/s |", "|---:|---:|---:|---:|---:|"]
    for C in sizes:
        n_req = max(50, min(args.requests, (1 << 24) // C))
        lat = [[] for _ in range(args.threads)]

        def drive(t):
            pi, pw, po = pins[t]
            for j in range(n_req + 20):
                t0 = time.perf_counter()
                eng.predict(pi.array[:C], pw.array[:C], out=po.array[:C])
                if j >= 20:
                    lat[t].append((time.perf_counter() - t0) * 1e3)
        ths = [threading.Thread(target=drive, args=(t,)) for t in range(args.threads)]
        w0 = time.perf_counter()
        [t.start() for t in ths]
        [t.join() for t in ths]
        wall = time.perf_counter() - w0
        all_lat = np.sort(np.concatenate([np.array(x) for x in lat]))
        rps = args.threads * (n_req + 20) / wall
        lines.append("| %d | %.3f | %.3f | %.0f | %.3g |" % (C, np.percentile(all_lat, 50), np.percentile(all_lat, 99), rps, rps * C))
        print(lines[-1], flush=True)
    eng.close()
    hdr = ("# Request-size sweep (BASELINE config 5), 1 x B200, Wide&Deep Criteo-shaped\n\n"
           "Closed loop, %d client threads x sequential requests through `ctr_predict` from pinned host buffers\n"
           "(reference layout, 468 B/candidate over PCIe); latency window = inputs exist -> ordered scores on host.\n\n" % args.threads)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    open(args.out, "w").write(hdr + "\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
