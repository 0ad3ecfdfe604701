set -x
B="python bench.py --no-e2e --no-cpu-baseline"
for s in 1 2 3 4; do echo "== streams $s"; CTR_DEVICE_STREAMS=$s $B --steps 50 --warmup 5 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.1fM ms/step %.3f'%(d['value']/1e6,d['ms_per_step']), {k:round(v*1000,1) for k,v in d['stage_ms_first_chunk'].items()}, d['clocks'])
"; done
for c in 9472 37888; do echo "== chunk $c streams 2"; CTR_CHUNK_ROWS=$c $B --steps 50 --warmup 5 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.1fM ms/step %.3f'%(d['value']/1e6,d['ms_per_step']), {k:round(v*1000,1) for k,v in d['stage_ms_first_chunk'].items()})
"; done
P="python bench.py --no-e2e --no-cpu-baseline --steps 3 --warmup 3 --requests-per-step 8"
CTR_DEVICE_STREAMS=1 $P > gpurun_out/plain.log 2>&1 && \
CTR_DEVICE_STREAMS=1 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"gather_pool|gemm_f16|head_kernel" -c 200 --csv --log-file gpurun_out/r01_launches.csv $P > gpurun_out/ncu_l.log 2>&1
CTR_DEVICE_STREAMS=1 ncu --set full --clock-control none --import-source on -k regex:"gather_pool|gemm_f16" -s 40 -c 4 -o gpurun_out/r01_prof $P > gpurun_out/ncu_f.log 2>&1
tail -3 gpurun_out/ncu_l.log gpurun_out/ncu_f.log
ls -la gpurun_out
