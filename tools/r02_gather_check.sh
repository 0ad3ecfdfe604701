# gather / parity tests, then the device-resident bench (stage times of the first chunk) on fp32 and f16 rows
set -x
python -m pytest tests/test_gpu_parity.py tests/test_gpu_packed.py tests/test_gpu_golden.py tests/test_gpu_full_size.py -q -m gpu 2>&1 | tail -3
for T in fp32 f16; do
  python bench.py --steps 40 --warmup 3 --no-cpu-baseline --no-e2e --closed-loop-requests 0 --table-dtype $T 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('RESULT $T', d['stage_ms_first_chunk'], round(d['value']/1e6,1), d['clocks']['sm_mhz'])"
done
