set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu --timeout 600 2>&1 | tail -4
cat > /tmp/small.py <<'PY'
import sys, os, tempfile, numpy as np
sys.path.insert(0, os.getcwd())
import bench, dtfs_b200
from dtfs_b200 import binding
from oracle_free_inputs import criteo_inputs
tmp = tempfile.mkdtemp()
for model in ("wide_deep", "deepfm", "dcn_v2"):
    w = os.path.join(tmp, model + ".ctrw"); bench.write_model(w, 1 << 12, model=model)
    e = binding.Engine(device_ids=(0,), max_candidates=4096); e.load_weights(w)
    ids, wts = criteo_inputs(1003, 1 << 12)
    a = e.predict(ids, wts).copy()
    s, idx = e.predict_sorted(ids, wts)
    lay = e.packed_layout(13, weighted_fields=range(13)); ids[:, :13] = ids[0, :13]; wts[:, :13] = wts[0, :13]
    b = e.predict(ids, wts).copy()
    t, _ = e.predict_packed_async(lay, *e.pack_candidates(lay, ids, wts)); c = e.wait(t)
    assert np.array_equal(b, c) and np.array_equal(a[idx], s)
    e.close()
print("small ok")
PY
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python /tmp/small.py > gpurun_out/r02_sanitizer_memcheck.txt 2>&1; echo "memcheck rc=$?" >> gpurun_out/r02_sanitizer_memcheck.txt
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python /tmp/small.py > gpurun_out/r02_sanitizer_racecheck.txt 2>&1; echo "racecheck rc=$?" >> gpurun_out/r02_sanitizer_racecheck.txt
tail -4 gpurun_out/r02_sanitizer_memcheck.txt gpurun_out/r02_sanitizer_racecheck.txt
python bench.py --workers 12 --no-cpu-baseline --steps 100 --closed-loop-requests 0 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); e=d['e2e']
print('workers 12: e2e %.1fM shared %.1fM packed %.1fM packed+shared %.1fM'%(e['value']/1e6,e['shared_context']['value']/1e6,e['packed']['value']/1e6,e['packed_shared_context']['value']/1e6))"
