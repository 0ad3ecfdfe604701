# round 2 profiling session (1 GPU): launch list of the bench command, full captures of the kernels, secondary bench lines
set -x
mkdir -p gpurun_out
P="python bench.py --no-cpu-baseline --steps 2 --warmup 3 --requests-per-step 32 --closed-loop-requests 0 --latency-requests 40"
$P > gpurun_out/r02_prof_plain.json 2> gpurun_out/r02_prof_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches.csv $P > /dev/null 2> gpurun_out/ncu_l.err
Q="python bench.py --no-cpu-baseline --no-e2e --steps 2 --warmup 3 --requests-per-step 32"
CTR_DEVICE_STREAMS=1 ncu --set full --clock-control none --import-source on -k regex:"gather_tile|tower_chain" -s 6 -c 2 -f -o gpurun_out/r02_wd $Q > /dev/null 2> gpurun_out/ncu_f1.err
CTR_DEVICE_STREAMS=1 ncu --set full --clock-control none --import-source on -k regex:"gather_tile|tower_chain" -s 6 -c 3 -f -o gpurun_out/r02_dcn $Q --model dcn_v2 > /dev/null 2> gpurun_out/ncu_f2.err
# the request path: compact gather + block sort (one 10k request with sort, one packed request)
cat > /tmp/req.py <<'PY'
import sys, os, tempfile, numpy as np
sys.path.insert(0, os.getcwd())
import bench, dtfs_b200
from dtfs_b200 import binding
from oracle_free_inputs import criteo_inputs
tmp = tempfile.mkdtemp(); w = os.path.join(tmp, "m.ctrw"); bench.write_model(w, 1 << 22)
e = binding.Engine(device_ids=(0,), max_candidates=16384); e.load_weights(w)
ids, wts = criteo_inputs(10000, 1 << 22)
for _ in range(3):
    e.predict_sorted(ids, wts)
    lay = e.packed_layout(0, weighted_fields=range(13)); t, _ = e.predict_packed_async(lay, *e.pack_candidates(lay, ids, wts)); e.wait(t)
e.close()
PY
ncu --set full --clock-control none --import-source on -k regex:"sort_block|gather_tile" -s 2 -c 2 -f -o gpurun_out/r02_req python /tmp/req.py > /dev/null 2> gpurun_out/ncu_f3.err
tail -2 gpurun_out/ncu_f1.err gpurun_out/ncu_f2.err gpurun_out/ncu_f3.err
B="python bench.py --no-e2e --no-cpu-baseline --steps 100 --warmup 5"
for m in deepfm dcn_v2; do $B --model $m > gpurun_out/r02_bench_$m.json 2>/dev/null; done
$B --table-dtype f16 > gpurun_out/r02_bench_f16.json 2>/dev/null
$B --hashed > gpurun_out/r02_bench_hashed.json 2>/dev/null
ls -la gpurun_out | head -40
