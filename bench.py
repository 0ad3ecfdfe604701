#!/usr/bin/env python
"""bench.py -- candidates scored per second on BASELINE.json config 2:
Wide&Deep, Criteo-shaped (13 dense + 26 sparse fields, dim 16), 10k-candidate requests.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run)
  python bench.py --impl reference ...                     the CPU restatement of the reference request

A step = one pass of the hot path over one batch: R requests of C candidates each
(defaults R=64, C=10 000 -> 640 000 candidates, 300 MB of ids+weights: larger than L2).

  value    device-resident: ids/weights already in HBM, one ctr_score_device call per step
           (the engine coalesces the R requests and walks them in chunks of up to 303 104), CUDA events
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
import subprocess
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


def tower_name(model: str) -> str:
    m = MODELS[model]
    return "624->" + "->".join(map(str, m["hidden"])) + "->1" + (" + %d cross layers 624x624" % m["n_cross"] if m["n_cross"] else "")


def num_table_rows(rows_per_field: int, hashed: bool) -> int:
    n = N_DENSE + 1 + N_SPARSE * rows_per_field
    return 1 << int(np.ceil(np.log2(n))) if hashed else n


WORKLOAD = "wide_deep criteo 13 dense + 26 sparse dim16, 10k-candidate requests (BASELINE config 2)"


def workload_config(args, world: int) -> dict:
    """The `config` block: identical keys and values in both arms (ours / --impl reference)."""
    rows = num_table_rows(args.rows_per_field, args.hashed)
    row_bytes = 64 if args.table_dtype == "f16" else 128
    C, R = args.candidates, args.requests_per_step
    return {"workload": WORKLOAD, "model": args.model, "tower": tower_name(args.model),
            "id_mode": "hashed" if args.hashed else "direct", "candidates_per_request": C, "requests_per_step": R,
            "candidates_per_step_per_gpu": C * R, "table_rows": int(rows), "table_dtype": args.table_dtype,
            "table_bytes_per_gpu": int(rows * row_bytes), "id_distribution": "uniform",
            "l2": "inputs larger than L2 (%.0f MB ids+weights per step; %.1f GB table, random rows)" % (C * R * F * 12 / 1e6, rows * row_bytes / 1e9),
            "parallelism": "candidate-sharded x%d, tables replicated, no collective" % world}


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
    num_rows = num_table_rows(rows_per_field, hashed)
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


def cpu_baseline_subprocess(args):
    """The CPU baseline of the repo arm = the reference arm on a bounded sample, in its own process (so that this
    process never loads anything under oracle/)."""
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "12", "--warmup", "1",
           "--ref-requests-per-step", "20", "--candidates", str(args.candidates), "--rows-per-field", str(args.rows_per_field),
           "--requests-per-step", str(args.requests_per_step)]
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "OMP_NUM_THREADS"):
        env.pop(k, None)
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
        for ln in reversed(out.stdout.splitlines()):
            if ln.startswith("{"):
                return json.loads(ln)["cpu_baseline"]
        return {"value": None, "unit": "candidates/s", "cores": os.cpu_count(), "kind": "port", "sample": "failed: " + out.stderr[-300:]}
    except Exception as ex:  # noqa: BLE001
        return {"value": None, "unit": "candidates/s", "cores": os.cpu_count(), "kind": "port", "sample": "failed: %r" % (ex,)}


def random_row_rate(device_index: int):
    """Random 128-byte rows per second this GPU's HBM delivers to a kernel that does nothing else
    (tools/probes/random_rows_probe.cu, run live; 36.2 G rows/s on this pool's B200s for rows of 32 to 128 bytes alike, with
    16 to 64 warps per SM and 2 to 16 loads in flight per lane: profiles/r02_random_rows_probe.txt).  The gather kernel's time
    follows the rows it looks up, not the bytes: this is the roofline that bounds it."""
    exe = os.path.join(ROOT, "tools", "probes", "random_rows_probe")
    out = {"peak_rows_per_s": 36.2e9, "source": "profiles/r02_random_rows_probe.txt (committed measurement: the probe did not run here)"}
    if not os.path.exists(exe):
        return out
    try:
        env = dict(os.environ)
        visible = env.get("CUDA_VISIBLE_DEVICES")
        env["CUDA_VISIBLE_DEVICES"] = visible.split(",")[device_index] if visible else str(device_index)
        txt = subprocess.run([exe, "8", "quick"], capture_output=True, text=True, timeout=120, env=env).stdout
        import re
        m = re.search(r"([0-9.]+) G rows/s", txt)
        if m:
            out = {"peak_rows_per_s": float(m.group(1)) * 1e9,
                   "source": "tools/probes/random_rows_probe.cu run on this GPU after the timed region: " + " ".join(txt.split())}
    except Exception as ex:  # noqa: BLE001
        out["probe_error"] = repr(ex)
    return out


def closed_loop_latency(wpath: str, args, workers: int):
    """p50 / p99 of a 10 000-candidate request under the reference's load shape (DCNClient.java:30-31, :213-224):
    6 closed-loop client threads, 10 000 timed requests after 1 000 warm-up, C++ over the C ABI."""
    from dtfs_b200 import build
    exe = build.build_stream_sweep()
    C = args.candidates
    cmd = [exe, "--weights", wpath, "--lib", build.LIB, "--threads", "6", "--requests", str(args.closed_loop_requests), "--warmup", "1000",
           "--min", str(C), "--max", str(C), "--rows-per-field", str(args.rows_per_field), "--workers", str(workers)]
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        for ln in out.stdout.splitlines():
            cells = [c.strip() for c in ln.strip().strip("|").split("|")]
            if len(cells) >= 9 and cells[0] == str(C):
                return {"p50_ms": float(cells[1]), "p99_ms": float(cells[2]), "requests_per_s": float(cells[3]), "clients": 6,
                        "timed_requests": args.closed_loop_requests, "warmup_requests": 1000,
                        "open_loop_70pct": {"requests_per_s": float(cells[6]), "p50_ms": float(cells[7]), "p99_ms": float(cells[8])},
                        "what": "%d-candidate requests, 6 closed-loop client threads (DCNClient.java:213-224), ctr_predict_async + ctr_wait from pinned "
                                "buffers; open loop: Poisson arrivals at 70%% of the closed-loop rate, latency from the scheduled arrival" % C}
        return {"error": (out.stderr or out.stdout)[-300:]}
    except Exception as ex:  # noqa: BLE001
        return {"error": repr(ex)}


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
    # a bounded sample of the step: the step is 64 requests, the CPU scores ~40 requests/s
    reqs_per_step = max(1, min(args.ref_requests_per_step, 3000 // max(args.steps, 1)))
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
    sample = "%d of the step's %d requests of %d candidates per step (%d requests timed), P=2 shards, protobuf wire encode/decode + fp32 forward + merge + sort; %s; table build %.1fs untimed" % (
        reqs_per_step, args.requests_per_step, C, reqs_per_step * args.steps, client.server.table_note, t_build)
    line = {
        "impl": "reference", "metric": "candidates scored/sec", "value": value, "unit": "candidates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, max(world, 1)),
        "cpu_baseline": {"value": value, "unit": "candidates/s", "cores": client.cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "candidates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "CPU restatement of the reference request (oracle port), not TF-Serving: no JDK / tensorflow on this image",
    }
    client.close()
    print(json.dumps(line))


def sharded_request(n_dev: int, tmp: str):
    """ONE engine over all N GPUs (what the Java client would hold): processARequest's fan-out (DCNClient.java:146-164)
    inside the engine.  Reports request latency and whether the scores are the bits a 1-GPU engine gives."""
    from dtfs_b200 import binding
    from oracle_free_inputs import criteo_inputs
    rpf = 1 << 20
    out = {"n_devices": n_dev, "rows_per_field": rpf,
           "what": "rank 0 opens ONE engine over all %d GPUs, min_shard_candidates = 1 (always cut N ways, DC:148); sequential requests "
                   "through ctr_predict from pinned buffers; bit_identical_to_1gpu compares with a single-device engine on the same inputs" % n_dev}

    def lat(eng, ids, wts, n_req):
        n = ids.shape[0]
        pi, pw, po = eng.pinned(ids.shape, np.int64), eng.pinned(wts.shape, np.float32), eng.pinned((n,), np.float32)
        pi.array[:] = ids
        pw.array[:] = wts
        ts = []
        for _ in range(n_req + 20):
            t0 = time.perf_counter()
            eng.predict(pi.array, pw.array, out=po.array)
            ts.append((time.perf_counter() - t0) * 1e3)
        ts = np.sort(np.array(ts[20:]))
        res = po.array.copy()
        for b in (pi, pw, po):
            b.free()
        return res, {"p50_ms": float(np.percentile(ts, 50)), "p99_ms": float(np.percentile(ts, 99)), "n": int(ts.size),
                     "candidates_per_s": float(n / (np.percentile(ts, 50) * 1e-3))}

    cases = [("config3_deepfm_100k", "deepfm", 100000, binding.TABLES_REPLICATED, 200),
             ("wide_deep_10k", "wide_deep", 10000, binding.TABLES_REPLICATED, 500),
             ("wide_deep_10k_row_sharded", "wide_deep", 10000, binding.TABLES_ROW_SHARDED, 500)]
    devs = tuple(range(n_dev))
    for name, model, C, mode, n_req in cases:
        try:
            wpath = os.path.join(tmp, "sh_%s.ctrw" % model)
            if not os.path.exists(wpath):
                write_model(wpath, rpf, model=model)
            ids, wts = criteo_inputs(C, rpf, seed=99)
            cap = C // n_dev + n_dev + 1
            with binding.Engine(device_ids=devs, max_candidates=cap, workers_per_device=2, shard_mode=mode, min_shard_candidates=1) as eng:
                eng.load_weights(wpath)
                got, l = lat(eng, ids, wts, n_req)
            with binding.Engine(device_ids=(0,), max_candidates=C, workers_per_device=2) as one:
                one.load_weights(wpath)
                ref, l1 = lat(one, ids, wts, max(50, n_req // 4))
            l["bit_identical_to_1gpu"] = bool(np.array_equal(got, ref))
            l["one_gpu_p50_ms"] = l1["p50_ms"]
            l["tables"] = "row-sharded (row mod N), remote rows over NVLink peer loads" if mode == binding.TABLES_ROW_SHARDED else "replicated"
            out[name] = l
        except Exception as ex:  # noqa: BLE001
            out[name] = {"error": repr(ex)}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--candidates", type=int, default=10000)
    ap.add_argument("--requests-per-step", type=int, default=64)
    ap.add_argument("--ref-requests-per-step", type=int, default=20, help="reference arm: requests timed per step (a bounded sample of the step)")
    ap.add_argument("--rows-per-field", type=int, default=1 << 22)
    ap.add_argument("--workers", type=int, default=0, help="host worker threads (one stream each) of the engine, the reference's threadPool; "
                    "0 = the host cores left per rank, at most 12 (the reference's pool has 16, DC:42)")
    ap.add_argument("--model", default="wide_deep", choices=sorted(MODELS), help="wide_deep is the headline (config 2); others are secondary runs")
    ap.add_argument("--table-dtype", default="fp32", choices=["fp32", "f16"])
    ap.add_argument("--hashed", action="store_true", help="row = mix64(id) & (V-1) (config 4)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--latency-requests", type=int, default=300)
    ap.add_argument("--closed-loop-requests", type=int, default=10000, help="timed requests of the 6-client closed-loop latency leg (0 = skip)")
    ap.add_argument("--no-sharded-request", action="store_true", help="N > 1: skip rank 0's one-engine-over-all-GPUs leg")
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
    # stdout carries exactly one JSON line: everything else that would land there (NCCL_DEBUG=INFO lines, library
    # chatter) is sent to stderr, where it stays visible to whoever runs this
    json_out = os.fdopen(os.dup(1), "w")
    sys.stdout.flush()
    os.dup2(2, 1)
    if args.workers <= 0:
        cores = os.cpu_count() or 8
        # the reference's pool has 16 threads (DC:42); one per host core that is left, at most 12, at least 2
        args.workers = max(2, min(12, cores // world - (4 if world == 1 else 1)))
    dist = None
    if world > 1:
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
        rl.append({"kernel": "gather_tile_kernel", "bound": "hbm", "achieved": a, "peak": peaks["hbm"], "unit": "GB/s",
                   "frac": a / peaks["hbm"], "traffic": ((per_cand("gather_tile_kernel") or 0) * chunk_rows or None) if (args.table_dtype == "fp32" and args.model == "wide_deep") else None, "ms": stage_ms["gather"],
                   "algorithmic_bytes_per_candidate": gbytes})
    if stage_ms["tower"] > 0:
        flop = model_flop_per_candidate(args.model)
        a = flop * chunk_rows / (stage_ms["tower"] * 1e-3) / 1e12
        capped = "sw_power_cap" in (clk.get("reasons") or [])
        pk = peaks["tc_sustained"] if capped else peaks["tc_burst"]
        rl.append({"kernel": "tower_chain_kernel (%s in one launch)" % tower_name(args.model), "bound": "tensor", "achieved": a,
                   "peak": pk, "unit": "TFLOP/s", "frac": a / pk,
                   "peak_is": "bf16_tflops_sustained (the clock record of this run shows sw_power_cap)" if capped else "bf16_tflops (burst: no power cap in this run's clock record)",
                   "traffic": ((per_cand("tower_chain_kernel") or 0) * chunk_rows or None) if args.model == "wide_deep" else None, "ms": stage_ms["tower"],
                   "algorithmic_flop_per_candidate": flop,
                   "frac_of_burst_peak": a / peaks["tc_burst"], "frac_of_sustained_peak": a / peaks["tc_sustained"],
                   "note": "one engine stream: kernels run back to back, the event-to-event time is the kernel's own"})
    dominant = max(rl, key=lambda r: r["ms"]) if rl else None
    if dominant is not None:
        dominant = dict(dominant)
        dominant["peak_source"] = "%s (MEASURED_PEAKS.json %s)" % ("hbm_gbs" if dominant["bound"] == "hbm" else dominant.get("peak_is", "bf16_tflops"), peaks["src"])
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
        lb0 = eng.link_bytes()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        torch.cuda.synchronize()
        t_e2e = max_over_ranks(time.perf_counter() - t0)
        lb1 = eng.link_bytes()
        h2d_step, d2h_step = (lb1[0] - lb0[0]) // args.steps, (lb1[1] - lb0[1]) // args.steps      # what the engine really copied
        e2e = {"value": world * n * args.steps / t_e2e, "unit": "candidates/s", "h2d_bytes_per_step": int(h2d_step),
               "d2h_bytes_per_step": int(d2h_step), "ms_per_step": t_e2e / args.steps * 1e3,
               "caller_bytes_per_step": n * F * 12, "h2d_gbs_per_gpu": h2d_step * args.steps / t_e2e / 1e9,
               "how": "R ctr_predict_async + ctr_wait per step from pinned host buffers (the reference's layout: int64 id + float weight per "
                      "field slot, 468 B per candidate); %d worker streams per GPU; with 5 or more shards in flight the workers narrow their "
                      "shard on the host before it crosses the link (ctr_config.host_pack = 0)" % args.workers}
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
        # additive API: the narrow marshaller -- uint32 table rows, weights only for the 13 dense fields (208 B per candidate),
        # alone and combined with the shared context (13 per-candidate sparse fields: 52 B per candidate).  The caller
        # keeps its features in this layout: the packing below runs once, outside the timed region.
        def packed_leg(n_ctx_p):
            lay = eng.packed_layout(n_ctx_p, weighted_fields=range(N_DENSE), fixed_id_fields=range(N_DENSE))
            fkp, wk, n_fixed = eng.packed_shape(lay)
            # one pinned slab per request: its rows, then its weights (the engine ships such a request with one copy)
            per_req = C * (fkp + wk)
            pin_slab = eng.pinned((R * per_req,), np.uint32)
            reqs = []
            for r in range(R):
                sl = slice(r * C, (r + 1) * C)
                ids_r, wts_r = h_ids_t[sl].copy(), h_wts_t[sl].copy()
                ids_r[:, :n_ctx_p] = ids_r[0, :n_ctx_p]               # one user / context per request, as a real request has
                wts_r[:, :n_ctx_p] = wts_r[0, :n_ctx_p]
                base = pin_slab.array[r * per_req:(r + 1) * per_req]
                rows_v = base[:C * fkp].reshape(C, fkp)
                wts_v = base[C * fkp:].view(np.float32).reshape(C, wk) if wk else np.empty((C, 0), np.float32)
                ci, cw, _, _ = eng.pack_candidates(lay, ids_r, wts_r, rows_out=rows_v, wts_out=wts_v if wk else None)
                reqs.append((ci.copy(), cw.copy(), rows_v, wts_v))

            def pstep():
                tickets = []
                for r in range(R):
                    ci, cw, rows_v, wts_v = reqs[r]
                    t, _ = eng.predict_packed_async(lay, ci, cw, rows_v, wts_v, out=pin_o.array[r * C:(r + 1) * C])
                    tickets.append(t)
                for t in tickets:
                    eng.wait(t)
            for _ in range(args.warmup):
                pstep()
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                pstep()
            torch.cuda.synchronize()
            t_p = max_over_ranks(time.perf_counter() - t0)
            same = bool(np.array_equal(pin_o.array[:C], d_out[:C].cpu().numpy())) if n_ctx_p == 0 else None
            pin_slab.free()
            return {"value": world * n * args.steps / t_p, "unit": "candidates/s", "h2d_bytes_per_step": n * (fkp * 4 + wk * 4) + R * (n_ctx_p * 12 + n_fixed * 8),
                    "d2h_bytes_per_step": n * 4, "bytes_per_candidate": fkp * 4 + wk * 4, "bit_identical_to_drop_in": same}
        e2e["packed"] = packed_leg(0)
        e2e["packed"]["what"] = ("ctr_predict_packed_async: uint32 table rows for the 26 sparse fields + values of the 13 dense fields, whose ids are "
                                 "the same for every candidate and travel once (156 instead of 468 B per candidate); additive API, features kept in "
                                 "this layout by the caller")
        e2e["packed_shared_context"] = packed_leg(26)
        e2e["packed_shared_context"]["what"] = "the same with 26 of 39 fields given once per request: 13 uint32 rows per candidate"
        for b in (pin_i, pin_w, pin_o, pin_ci, pin_cw):
            b.free()
        if world == 1 and args.closed_loop_requests > 0 and args.model == "wide_deep":
            lat["closed_loop_6_clients"] = closed_loop_latency(wpath, args, args.workers)

    # ---------------- CPU baseline (rank 0, N=1 only) ----------------
    cpu = None
    eng.close()
    del d_ids, d_wts, d_out
    torch.cuda.empty_cache()
    if rank == 0 and rl and rl[0]["kernel"].startswith("gather"):
        rr = random_row_rate(local_rank)
        g = rl[0]
        rows = N_SPARSE * chunk_rows / (g["ms"] * 1e-3)
        g["random_rows"] = {"achieved_rows_per_s": rows, "peak_rows_per_s": rr["peak_rows_per_s"], "frac": rows / rr["peak_rows_per_s"],
                            "what": "table rows fetched from HBM per second (%d per candidate: the sparse fields; the dense fields' rows are cache-resident) "
                                    "against the rate of a kernel that only fetches random rows; beside them this kernel streams %d B of lists and "
                                    "%d B of X per candidate" % (N_SPARSE, F * 12, 2 * int(info.d_padded)),
                            "peak_source": rr["source"]}
        if "probe_error" in rr:
            g["random_rows"]["probe_error"] = rr["probe_error"]
    if world == 1 and rank == 0 and not args.no_cpu_baseline:
        cpu = cpu_baseline_subprocess(args)

    # ---------------- N > 1: one engine over all the GPUs (rank 0; the other ranks wait on the store, GPUs idle) ----------------
    sharded = None
    if dist is not None and not args.no_sharded_request:
        store = dist.distributed_c10d._get_default_store()
        barrier()
        if rank == 0:
            try:
                sharded = sharded_request(world, tmp)
            except Exception as ex:  # noqa: BLE001
                sharded = {"error": repr(ex)}
            store.set("ctr_sharded_done", "1")
        else:
            store.wait(["ctr_sharded_done"])
    if rank == 0:
        line = {
            "metric": "candidates scored/sec", "value": value, "unit": "candidates/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_dev / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": workload_config(args, world),
            "roofline": dominant, "rooflines": rl, "stage_ms_first_chunk": stage_ms,
            "cpu_baseline": cpu, "e2e": e2e, "latency": lat, "gpu_launches": int(launches), "clocks": clk,
            "host": {"cores": os.cpu_count(), "worker_streams_per_gpu": args.workers},
        }
        if sharded is not None:
            line["sharded_request"] = sharded
        assert int(info.table_bytes_per_device) == line["config"]["table_bytes_per_gpu"]
        json_out.write(json.dumps(line) + "\n")
        json_out.flush()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
