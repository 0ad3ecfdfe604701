timeout 1500 python -m pytest tests -q -m gpu -x --timeout 600 2>&1 | tail -3
python - <<'PY'
import dtfs_b200, bench
from dtfs_b200 import build
build.build_stream_sweep()
bench.write_model("/tmp/wd_sweep.ctrw", 1 << 22)
PY
for sp in 0 30 100; do echo "== CTR_SPIN_US=$sp"; CTR_SPIN_US=$sp timeout 600 ./distributed-tf-serving_b200/stream_sweep_host --lib distributed-tf-serving_b200/libctr_b200.so --weights /tmp/wd_sweep.ctrw --devices 1 --max 16384 --requests 6000; done
for sp in 0 30; do echo "== bench latency CTR_SPIN_US=$sp"; CTR_SPIN_US=$sp python bench.py --steps 40 --warmup 5 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.1fM e2e %.1fM shared %.1fM'%(d['value']/1e6,d['e2e']['value']/1e6,d['e2e']['shared_context']['value']/1e6), d['latency'])
"; done
