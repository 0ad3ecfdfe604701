# the driver's scaling run at N ranks with the default flags: reference arm first, then ours; N from $1
set -x
N=${1:-2}
nproc; nvidia-smi -L | wc -l
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --impl reference --gpus $N --steps 200 --warmup 3 > gpurun_out/r02_scale_ref_n$N.json 2> gpurun_out/r02_scale_ref_n$N.err
cut -c1-300 gpurun_out/r02_scale_ref_n$N.json
(time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus $N --steps 200 --warmup 3) > gpurun_out/r02_scale_n$N.json 2> gpurun_out/r02_scale_n$N.err
tail -4 gpurun_out/r02_scale_n$N.err
python - gpurun_out/r02_scale_n$N.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
e = d["e2e"]
print("value %.1f M" % (d["value"] / 1e6), "e2e %.1f M" % (e["value"] / 1e6), "h2d GB/s per gpu", e.get("h2d_gbs_per_gpu"),
      {k: round(v["value"] / 1e6, 1) for k, v in e.items() if isinstance(v, dict) and "value" in v}, d["host"], d["clocks"])
print("sharded_request", json.dumps(d.get("sharded_request"))[:900])
PY
