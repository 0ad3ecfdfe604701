# One 8-GPU session: topology, multi-GPU parity tests, torchrun bench at N=8,
# request stream sweep at G = 1, 2, 4, 8.   gpurun --gpus 8 -- 'bash tools/gpurun_n8_session.sh'
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
lscpu | grep -i "numa\|socket\|^CPU(s)" | head -8
for d in /sys/bus/pci/devices/*; do if [ -e $d/local_cpulist ] && grep -q 0x10de $d/vendor 2>/dev/null; then echo "$(basename $d) $(cat $d/local_cpulist) numa $(cat $d/numa_node)"; fi; done | head -12
timeout 600 python -m pytest tests/test_gpu_multi.py -q -m gpu -x --timeout 300 2>&1 | tail -4
N=8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/r01_e_bench_n$N.json 2> gpurun_out/bench_n$N.err; tail -c 1200 gpurun_out/r01_e_bench_n$N.json; tail -3 gpurun_out/bench_n$N.err
python - <<'PY'
import json
for f in ("gpurun_out/r01_e_bench_n8.json",):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, "value %.0fM e2e %.0fM shared %.0fM" % (d["value"] / 1e6, d["e2e"]["value"] / 1e6, d["e2e"]["shared_context"]["value"] / 1e6), d.get("latency"))
    except Exception as ex:
        print(f, "unreadable", ex)
PY
bash tools/gpurun_sweep.sh "1 2 4 8"
# config 3: DeepFM, 100 000 candidates per request cut over the GPUs of one engine (through the C ABI, pinned host buffers)
python - <<'PY'
import dtfs_b200, bench
bench.write_model("/tmp/deepfm_sweep.ctrw", 1 << 22, model="deepfm")
PY
for g in 1 8; do echo "== config 3 (DeepFM, 100k candidates/request) on $g GPU(s)"; timeout 600 ./distributed-tf-serving_b200/stream_sweep_host --lib distributed-tf-serving_b200/libctr_b200.so --weights /tmp/deepfm_sweep.ctrw --devices $g --min 100000 --max 100000 --requests 3000 --out gpurun_out/r01_config3_g$g.md; done
# config 4: DCN-v2, ids hashed into ONE 2^30-row fp32 table (137 GB), rows striped over the 8 GPUs, remote rows over NVLink
echo "== config 4, 2^27 rows (replicated vs row-sharded)"; timeout 900 python tools/bench_row_sharded.py --log2-rows 27 2>&1 | tail -1
echo "== config 4, 2^30 rows (row-sharded only)"; timeout 900 python tools/bench_row_sharded.py --log2-rows 30 --skip-replicated 2>&1 | tail -1
# host link check: pinned host -> device copies on all GPUs at once (what bounds the end-to-end number at N=8)
python - <<'PY'
import threading, time, torch
G = torch.cuda.device_count()
n = 1 << 28
host = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(G)]
dev = [torch.empty(n, dtype=torch.uint8, device="cuda:%d" % g) for g in range(G)]
def run(g, k):
    with torch.cuda.device(g):
        for _ in range(k):
            dev[g].copy_(host[g], non_blocking=True)
        torch.cuda.synchronize(g)
for gs in ([0], list(range(G))):
    for k in (2, 10):
        ths = [threading.Thread(target=run, args=(g, k)) for g in gs]
        t0 = time.perf_counter(); [t.start() for t in ths]; [t.join() for t in ths]; dt = time.perf_counter() - t0
    print("H2D from pinned memory, %d GPU(s) at once: %.1f GB/s aggregate" % (len(gs), len(gs) * 10 * n / dt / 1e9))
PY
