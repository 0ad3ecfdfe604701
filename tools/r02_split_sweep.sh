# gather and tower on disjoint SMs (CTR_SM_SPLIT = SMs of the gather's partition): device-resident rate against the split, the chunk size
# and the number of engine streams
set -x
run() {
  env "$@" python bench.py --steps 40 --warmup 3 --no-cpu-baseline --no-e2e --closed-loop-requests 0 2>gpurun_out/_split.err | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('RESULT $*', d['stage_ms_first_chunk'], round(d['value']/1e6,1), d['clocks']['sm_mhz'])
except Exception as ex:
    print('RESULT $* FAILED', ex)"
  tail -2 gpurun_out/_split.err | cut -c1-300
}
run CTR_DEVICE_STREAMS=1
run CTR_DEVICE_STREAMS=2 CTR_CHUNK_ROWS=75776
run CTR_SM_SPLIT=32 CTR_DEVICE_STREAMS=2 CTR_CHUNK_ROWS=75776
run CTR_SM_SPLIT=32 CTR_DEVICE_STREAMS=2 CTR_CHUNK_ROWS=151552
run CTR_SM_SPLIT=24 CTR_DEVICE_STREAMS=2 CTR_CHUNK_ROWS=75776
run CTR_SM_SPLIT=40 CTR_DEVICE_STREAMS=2 CTR_CHUNK_ROWS=75776
run CTR_SM_SPLIT=48 CTR_DEVICE_STREAMS=2 CTR_CHUNK_ROWS=75776
run CTR_SM_SPLIT=32 CTR_DEVICE_STREAMS=3 CTR_CHUNK_ROWS=75776
run CTR_SM_SPLIT=32 CTR_DEVICE_STREAMS=1
