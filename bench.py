#!/usr/bin/env python
"""bench.py -- candidates scored per second on BASELINE.json config 2:
Wide&Deep, Criteo-shaped (13 dense + 26 sparse fields, dim 16), 10k-candidate requests.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run)
  python bench.py --impl reference ...                     the CPU restatement of the reference request

A step = one pass of the hot path over one batch: R requests of C candidates each
(defaults R=64, C=10 000 -> 640 000 candidates, 300 MB of ids+weights: larger than L2).

  value    device-resident: ids/weights already in HBM, one ctr_score_device call per step
           (the engine coalesces the R requests and walks them in chunks of up to 151 552), CUDA events
  e2e      through the host C ABI (ctr_predict_async / ctr_wait) from pinned host buffers,
           H2D and D2H inside the timed region, R requests in flight per step
  roofline dominant kernel by device time (CUDA events around every kernel of the first chunk
           of each timed step, averaged), against MEASURED_PEAKS.json
  cpu_baseline   the oracle's CPU restatement of the same request on the host cores (N=1 only)

Multi-GPU: candidates are independent, tables replicated, no data-path collective: every rank
scores its own batch (weak scaling); value = all ranks' candidates / max-over-ranks time.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_DENSE, N_SPARSE, EMB = 13, 26, 16
F = N_DENSE + N_SPARSE
HIDDEN = (1024, 512, 256)
# algorithmic figures per candidate
GATHER_BYTES_PER_CAND = F * 12 + N_SPARSE * (EMB * 4 + 4)          # 468 + 1768 = 2236 (gather kernel's own bytes)
FLOP_PER_CAND_LAYER = [2 * F * EMB * HIDDEN[0], 2 * HIDDEN[0] * HIDDEN[1], 2 * HIDDEN[1] * HIDDEN[2] + 2 * HIDDEN[2]]


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=float(d["hbm_gbs"]), tc_burst=float(d["bf16_tflops"]), tc_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), src="measured")
    return dict(hbm=6650.0, tc_burst=1590.0, tc_sustained=1400.0, src="fallback")


# the other model families of BASELINE.json (configs 3 and 4): secondary runs, `--model`
MODELS = {"wide_deep": dict(kind=1, hidden=(1024, 512, 256), n_cross=0),       # SURVEY.md section 8(a-M) shapes
          "deepfm": dict(kind=2, hidden=(512, 512, 512), n_cross=0),
          "dcn_v2": dict(kind=3, hidden=(1024, 512, 256), n_cross=3)}


def model_flop_per_candidate(model: str) -> int:
    m = MODELS[model]
    d, fl, fan_in = F * EMB, 0, F * EMB
    for h in m["hidden"]:
        fl += 2 * fan_in * h
        fan_in = h
    fl += m["n_cross"] * 2 * d * d
    return fl + 2 * (m["hidden"][-1] + (d if m["kind"] == 3 else 0))


def write_model(path: str, rows_per_field: int, seed: int = 1234, model: str = "wide_deep", table_dtype: int = 0, hashed: bool = False):
    """Random-init model of the named architecture; procedural table (no multi-GB file)."""
    from dtfs_b200 import weights as W
    m = MODELS[model]
    rng = np.random.Generator(np.random.PCG64(seed))
    d = F * EMB
    mlp, fan_in = [], d
    for h in m["hidden"]:
        lim = np.sqrt(6.0 / (fan_in + h))
        mlp.append((rng.uniform(-lim, lim, (h, fan_in)).astype(np.float32), rng.uniform(-0.05, 0.05, h).astype(np.float32)))
        fan_in = h
    cross = [(rng.uniform(-0.03, 0.03, (d, d)).astype(np.float32), rng.uniform(-0.05, 0.05, d).astype(np.float32))
             for _ in range(m["n_cross"])]
    w_out = rng.uniform(-1, 1, m["hidden"][-1] + (d if m["kind"] == 3 else 0)).astype(np.float32)
    num_rows = N_DENSE + 1 + N_SPARSE * rows_per_field
    if hashed:
        num_rows = 1 << int(np.ceil(np.log2(num_rows)))
    W.write_weight_file(path, model_kind=m["kind"], num_fields=F, num_rows=num_rows, b0=0.01, table_dtype=table_dtype,
                        id_mode=W.ID_HASHED if hashed else W.ID_DIRECT, seed=seed, scale_e=0.25, scale_w1=0.5,
                        mlp=mlp, cross=cross, w_out=w_out, b_out=-0.02)
    return num_rows


def device_inputs(torch, n: int, rows_per_field: int, seed: int):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    ids = torch.empty((n, F), dtype=torch.int64, device="cuda")
    wts = torch.ones((n, F), dtype=torch.float32, device="cuda")
    ids[:, :N_DENSE] = torch.arange(1, N_DENSE + 1, device="cuda")
    wts[:, :N_DENSE] = torch.log1p(torch.exp(torch.randn((n, N_DENSE), device="cuda", generator=g)))
    h = torch.randint(0, rows_per_field, (n, N_SPARSE), device="cuda", generator=g)      # uniform: roofline run
    ids[:, N_DENSE:] = (N_DENSE + 1) + torch.arange(N_SPARSE, device="cuda") * rows_per_field + h
    return ids, wts


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons during the timed region (NVML, every 5 ms)."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
                 0x80: "hw_power_brake"}
        while not self.stop_flag:
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                r = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.005)

    def result(self):
        self.stop_flag = True
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def cpu_restatement(weight_path: str, rows_per_field: int, n_cand: int, n_requests: int, warmup: int):
    """The oracle's CPU port of one reference request (the only place bench.py executes oracle/)."""
    import psutil
    from oracle import cpu_baseline as cb, model as om
    w = om.load_weights(weight_path)
    need = w.spec.num_rows * 68
    cap = None
    if psutil.virtual_memory().available < need * 1.5 + (8 << 30):
        cap = N_DENSE + 1 + N_SPARSE * (1 << 18)
    client = cb.CpuClient(w, n_shards=2, with_wire=True, table_rows_cap=cap)
    ids, wts = om.criteo_inputs(n_cand, seed=4321, rows_per_field=rows_per_field)
    sec = cb.time_requests(client, ids, wts, n_requests, warmup)
    note = client.server.table_note
    cores = client.cores
    client.close()
    return sec, cores, note


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path (oracle port; the reference itself is Java +
    an external TF-Serving, neither present on this image) with all host threads."""
    if rank != 0:
        return
    os.environ.pop("OMP_NUM_THREADS", None)        # torchrun pins it to 1; the reference arm may use every host core
    tmp = tempfile.mkdtemp(prefix="ctr_bench_")
    path = os.path.join(tmp, "wd.ctrw")
    write_model(path, args.rows_per_field)
    C = args.candidates
    reqs_per_step = max(1, args.ref_requests_per_step)
    t_build = time.perf_counter()
    import psutil
    from oracle import cpu_baseline as cb, model as om
    w = om.load_weights(path)
    cap = None
    if psutil.virtual_memory().available < w.spec.num_rows * 68 * 1.5 + (8 << 30):
        cap = N_DENSE + 1 + N_SPARSE * (1 << 18)
    client = cb.CpuClient(w, n_shards=2, with_wire=True, table_rows_cap=cap)
    ids, wts = om.criteo_inputs(C, seed=4321, rows_per_field=args.rows_per_field)
    t_build = time.perf_counter() - t_build
    for _ in range(args.warmup):
        for _ in range(reqs_per_step):
            client.process_a_request(ids, wts)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        for _ in range(reqs_per_step):
            client.process_a_request(ids, wts)
    dt = time.perf_counter() - t0
    value = args.steps * reqs_per_step * C / dt
    sample = "%d requests of %d candidates per step, P=2 shards, protobuf wire encode/decode + fp32 forward + merge + sort; %s; table build %.1fs untimed" % (
        reqs_per_step, C, client.server.table_note, t_build)
    line = {
        "impl": "reference", "metric": "candidates scored/sec", "value": value, "unit": "candidates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "wide_deep criteo 13 dense + 26 sparse dim16, 10k-candidate requests (BASELINE config 2)",
                   "candidates_per_request": C, "requests_per_step": reqs_per_step, "rows_per_field": args.rows_per_field},
        "cpu_baseline": {"value": value, "unit": "candidates/s", "cores": client.cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "candidates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "CPU restatement of the reference request (oracle port), not TF-Serving: no JDK / tensorflow on this image",
    }
    client.close()
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--candidates", type=int, default=10000)
    ap.add_argument("--requests-per-step", type=int, default=64)
    ap.add_argument("--ref-requests-per-step", type=int, default=2)
    ap.add_argument("--rows-per-field", type=int, default=1 << 22)
    ap.add_argument("--workers", type=int, default=6, help="host worker threads (one stream each) of the engine: the reference's threadPool")
    ap.add_argument("--model", default="wide_deep", choices=sorted(MODELS), help="wide_deep is the headline (config 2); others are secondary runs")
    ap.add_argument("--table-dtype", default="fp32", choices=["fp32", "f16"])
    ap.add_argument("--hashed", action="store_true", help="row = mix64(id) & (V-1) (config 4)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--latency-requests", type=int, default=300)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import dtfs_b200  # noqa: F401
    from dtfs_b200 import binding

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU path")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        os.environ["NCCL_DEBUG"] = "WARN"          # keep stdout to the one JSON line
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    tmp = tempfile.mkdtemp(prefix="ctr_bench_%d_" % rank)
    wpath = os.path.join(tmp, "wd.ctrw")
    num_rows = write_model(wpath, args.rows_per_field, model=args.model, table_dtype=1 if args.table_dtype == "f16" else 0, hashed=args.hashed)
    if args.model != "wide_deep" or args.table_dtype != "fp32" or args.hashed:
        args.no_cpu_baseline = True                      # the CPU baseline is for the headline configuration
    C, R = args.candidates, args.requests_per_step
    n = C * R
    eng = binding.Engine(device_ids=(local_rank,), max_candidates=32768, workers_per_device=args.workers)
    info = eng.load_weights(wpath)
    d_ids, d_wts = device_inputs(torch, n, args.rows_per_field, 4321 + rank)
    d_out = torch.empty(n, dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream()
    step = lambda: eng.score_device(d_ids.data_ptr(), d_wts.data_ptr(), n, d_out.data_ptr(), stream=stream.cuda_stream)

    # ---------------- device-resident throughput (value) + per-kernel timing (roofline) ----------------
    for _ in range(args.warmup):
        step()
    barrier()
    eng.set_timing(True)
    clocks = ClockSampler(local_rank)
    clocks.start()
    launches0 = eng.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for _ in range(args.steps):
        step()
    ev1.record(stream)
    barrier()
    t_dev = max_over_ranks(ev0.elapsed_time(ev1) * 1e-3)
    launches = eng.launch_count() - launches0
    clk = clocks.result()
    value = world * n * args.steps / t_dev
    stage_ms = {"gather": eng.stage_ms(0), "tower": eng.stage_ms(1) + max(eng.stage_ms(2), 0.0), "chunk": eng.stage_ms(-1)}   # DCN-v2: cross chain + MLP chain
    eng.set_timing(False)
    chunk_rows = min(n, int(info.chunk_rows))          # candidates per kernel launch (first chunk is a full one)
    assert np.isfinite(d_out[:1000].cpu().numpy()).all()

    peaks = load_peaks()
    try:    # DRAM bytes per candidate measured by ncu (--set full) for each kernel, committed under profiles/
        traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except Exception:
        traffic = {}
    per_cand = lambda k: (traffic.get(k) or {}).get("dram_bytes_per_candidate")
    rl = []
    if stage_ms["gather"] > 0:
        gbytes = F * 12 + N_SPARSE * ((EMB * 2 if args.table_dtype == "f16" else EMB * 4) + 4)
        a = gbytes * chunk_rows / (stage_ms["gather"] * 1e-3) / 1e9
        rl.append({"kernel": "gather_pool_kernel", "bound": "hbm", "achieved": a, "peak": peaks["hbm"], "unit": "GB/s",
                   "frac": a / peaks["hbm"], "traffic": ((per_cand("gather_pool_kernel") or 0) * chunk_rows or None) if (args.table_dtype == "fp32" and args.model == "wide_deep") else None, "ms": stage_ms["gather"],
                   "algorithmic_bytes_per_candidate": gbytes})
    if stage_ms["tower"] > 0:
        flop = model_flop_per_candidate(args.model)
        a = flop * chunk_rows / (stage_ms["tower"] * 1e-3) / 1e12
        rl.append({"kernel": "tower_chain_kernel (624->1024->512->256->1 in one launch)", "bound": "tensor", "achieved": a,
                   "peak": peaks["tc_sustained"], "unit": "TFLOP/s", "frac": a / peaks["tc_sustained"],
                   "traffic": (per_cand("tower_chain_kernel") or 0) * chunk_rows or None, "ms": stage_ms["tower"], "algorithmic_flop_per_candidate": flop,
                   "frac_of_burst_peak": a / peaks["tc_burst"]})
    dominant = max(rl, key=lambda r: r["ms"]) if rl else None
    if dominant is not None:
        dominant = dict(dominant)
        dominant["peak_source"] = "%s (MEASURED_PEAKS.json %s)" % ("hbm_gbs" if dominant["bound"] == "hbm" else "bf16_tflops_sustained", peaks["src"])
        dominant["launch_rows"] = chunk_rows

    # ---------------- end to end through the host C ABI (e2e) ----------------
    e2e = None
    lat = None
    if not args.no_e2e:
        h_ids_t, h_wts_t = d_ids.cpu().numpy(), d_wts.cpu().numpy()
        pin_i, pin_w, pin_o = eng.pinned((n, F), np.int64), eng.pinned((n, F), np.float32), eng.pinned((n,), np.float32)
        pin_i.array[:] = h_ids_t
        pin_w.array[:] = h_wts_t

        def e2e_step():
            tickets = []
            for r in range(R):
                t, _ = eng.predict_async(pin_i.array[r * C:(r + 1) * C], pin_w.array[r * C:(r + 1) * C], out=pin_o.array[r * C:(r + 1) * C])
                tickets.append(t)
            for t in tickets:
                eng.wait(t)
        for _ in range(args.warmup):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        torch.cuda.synchronize()
        t_e2e = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": world * n * args.steps / t_e2e, "unit": "candidates/s", "h2d_bytes_per_step": n * F * 12,
               "d2h_bytes_per_step": n * 4, "ms_per_step": t_e2e / args.steps * 1e3,
               "how": "R ctr_predict_async + ctr_wait per step from pinned host buffers, %d worker streams" % args.workers}
        assert np.array_equal(pin_o.array[:C], d_out[:C].cpu().numpy()), "host path and device path disagree"
        # request latency (reference window DC:141 -> DC:198 without the sort): one 10k request at a time
        ts = []
        for _ in range(args.latency_requests):
            t0 = time.perf_counter()
            eng.predict(pin_i.array[:C], pin_w.array[:C], out=pin_o.array[:C])
            ts.append((time.perf_counter() - t0) * 1e3)
        ts = np.sort(np.array(ts[10:]))
        lat = {"p50_ms": float(np.percentile(ts, 50)), "p99_ms": float(np.percentile(ts, 99)), "n": int(ts.size),
               "what": "one %d-candidate request through ctr_predict (H2D + kernels + D2H), sequential" % C}
        # the same window including the reference's sort (DC:195), done on the device
        ts2 = []
        for _ in range(max(20, args.latency_requests // 4)):
            t0 = time.perf_counter()
            eng.predict_sorted(pin_i.array[:C], pin_w.array[:C])
            ts2.append((time.perf_counter() - t0) * 1e3)
        ts2 = np.sort(np.array(ts2[5:]))
        lat["with_sort_p50_ms"] = float(np.percentile(ts2, 50))
        lat["with_sort_p99_ms"] = float(np.percentile(ts2, 99))
        # additive API: 26 user/context fields given once per request, 13 per-candidate fields shipped
        n_ctx = 26
        fk = F - n_ctx
        pin_ci, pin_cw = eng.pinned((n, fk), np.int64), eng.pinned((n, fk), np.float32)
        pin_ci.array[:] = h_ids_t[:, n_ctx:]
        pin_cw.array[:] = h_wts_t[:, n_ctx:]
        ctx_i = [np.ascontiguousarray(h_ids_t[r * C, :n_ctx]) for r in range(R)]
        ctx_w = [np.ascontiguousarray(h_wts_t[r * C, :n_ctx]) for r in range(R)]

        def shared_step():
            tickets = []
            for r in range(R):
                t, _ = eng.predict_shared_async(ctx_i[r], ctx_w[r], pin_ci.array[r * C:(r + 1) * C], pin_cw.array[r * C:(r + 1) * C],
                                                out=pin_o.array[r * C:(r + 1) * C])
                tickets.append(t)
            for t in tickets:
                eng.wait(t)
        for _ in range(args.warmup):
            shared_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            shared_step()
        torch.cuda.synchronize()
        t_sh = max_over_ranks(time.perf_counter() - t0)
        e2e["shared_context"] = {"value": world * n * args.steps / t_sh, "unit": "candidates/s", "h2d_bytes_per_step": n * fk * 12 + R * n_ctx * 12,
                                 "d2h_bytes_per_step": n * 4, "what": "ctr_predict_shared_async: 26 of 39 fields (13 dense + 13 sparse) given once per request, "
                                 "13 per-candidate fields shipped; additive API beside the drop-in one"}
        for b in (pin_i, pin_w, pin_o, pin_ci, pin_cw):
            b.free()

    # ---------------- CPU baseline (rank 0, N=1 only) ----------------
    cpu = None
    if world == 1 and rank == 0 and not args.no_cpu_baseline:
        try:
            sec, cores, note = cpu_restatement(wpath, args.rows_per_field, C, n_requests=4, warmup=1)
            cpu = {"value": C / sec, "unit": "candidates/s", "cores": cores, "kind": "port",
                   "sample": "4 sequential requests of %d candidates, P=2 shards, wire encode/decode + fp32 forward + merge + sort; %s" % (C, note)}
        except Exception as ex:  # noqa: BLE001
            cpu = {"value": None, "unit": "candidates/s", "cores": os.cpu_count(), "kind": "port", "sample": "failed: %r" % (ex,)}

    eng.close()
    if rank == 0:
        line = {
            "metric": "candidates scored/sec", "value": value, "unit": "candidates/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_dev / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": {"workload": "wide_deep criteo 13 dense + 26 sparse dim16, 10k-candidate requests (BASELINE config 2)",
                       "model": args.model, "tower": "624->" + "->".join(map(str, MODELS[args.model]["hidden"])) + "->1" + (" + %d cross layers 624x624" % MODELS[args.model]["n_cross"] if MODELS[args.model]["n_cross"] else ""),
                       "id_mode": "hashed" if args.hashed else "direct", "candidates_per_request": C, "requests_per_step": R,
                       "candidates_per_step_per_gpu": n, "table_rows": int(num_rows), "table_dtype": args.table_dtype,
                       "table_bytes_per_gpu": int(info.table_bytes_per_device), "id_distribution": "uniform",
                       "l2": "inputs larger than L2 (%.0f MB ids+weights per step; %.1f GB table, random rows)" % (n * F * 12 / 1e6, info.table_bytes_per_device / 1e9),
                       "parallelism": "candidate-sharded x%d, tables replicated, no collective" % world},
            "roofline": dominant, "rooflines": rl, "stage_ms_first_chunk": stage_ms,
            "cpu_baseline": cpu, "e2e": e2e, "latency": lat, "gpu_launches": int(launches), "clocks": clk,
        }
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
