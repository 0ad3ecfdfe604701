# end-to-end drop-in rate against worker threads and how a worker waits for its stream (CTR_WAIT=sync|yield)
set -x
nproc
for W in 12 16 20; do for M in sync yield; do
  CTR_WAIT=$M python bench.py --workers $W --steps 60 --warmup 3 --no-cpu-baseline --closed-loop-requests 2000 \
     > gpurun_out/r02_w${W}_${M}.json 2> gpurun_out/r02_w${W}_${M}.err
  python - gpurun_out/r02_w${W}_${M}.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
e = d["e2e"]
print(sys.argv[1], "value %.1f M" % (d["value"] / 1e6), "e2e %.1f M" % (e["value"] / 1e6),
      {k: (round(v["value"] / 1e6, 1) if isinstance(v, dict) and "value" in v else v) for k, v in e.items() if k in ("shared_context", "packed", "packed_shared_context")},
      "lat", {k: v for k, v in d.get("latency", {}).items() if k != "closed_loop_6_clients"}, "closed", {k: v for k, v in d.get("latency", {}).get("closed_loop_6_clients", {}).items() if "p" in k or "rps" in k})
PY
done; done
CTR_WAIT=yield python bench.py --workers 28 --steps 60 --warmup 3 --no-cpu-baseline --closed-loop-requests 0 > gpurun_out/r02_w28_yield.json 2> gpurun_out/r02_w28_yield.err
tail -c 1500 gpurun_out/r02_w28_yield.json
