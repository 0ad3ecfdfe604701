# the other models / table formats with the round's final code: one bench line each (device-resident + e2e legs, no CPU baseline)
for a in "deepfm --model deepfm" "dcn_v2 --model dcn_v2" "f16 --table-dtype f16" "hashed --hashed"; do
  set -- $a; tag=$1; shift
  python bench.py "$@" --steps 100 --warmup 3 --no-cpu-baseline --closed-loop-requests 0 > gpurun_out/r02_bench_$tag.json 2> gpurun_out/r02_bench_$tag.err
  python - gpurun_out/r02_bench_$tag.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("RESULT", sys.argv[1], "value %.1f M" % (d["value"] / 1e6), "e2e %.1f M" % (d["e2e"]["value"] / 1e6), [(r["kernel"][:18], round(r["frac"], 3), (r.get("random_rows") or {}).get("frac")) for r in d["rooflines"]], d["clocks"]["sm_mhz"])
PY
done
