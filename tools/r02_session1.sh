# round 2, GPU session 1: correctness of the new gather / shared scratch / setmaxnreg tower, then the overlap sweep
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv,noheader; nproc
timeout 900 python -m pytest tests -q -m gpu -x --timeout 600 2>&1 | tail -15
python __graft_entry__.py smoke 2>&1 | tail -3
B="python bench.py --no-e2e --no-cpu-baseline --steps 40 --warmup 5"
show() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.1fM ms/step %.3f'%(d['value']/1e6,d['ms_per_step']), {k:round(v*1000,1) for k,v in d['stage_ms_first_chunk'].items()}, d['clocks']['reasons'], d['clocks']['sm_mhz'])
    elif 'rror' in l or 'Trace' in l: print(l.rstrip())
"; }
for s in 1 2 3; do echo "== streams $s (default chunk)"; CTR_DEVICE_STREAMS=$s $B 2>&1 | show; done
for c in 37888 75776 303104; do echo "== chunk $c streams 2"; CTR_CHUNK_ROWS=$c $B 2>&1 | show; done
echo "== streams 2, no L2 persistence"; CTR_NO_L2_PERSIST=1 $B 2>&1 | show
echo "== streams 1, chunk 303104 (round-1 configuration)"; CTR_DEVICE_STREAMS=1 CTR_CHUNK_ROWS=303104 $B 2>&1 | show
echo "== f16 table, streams 2"; $B --table-dtype f16 2>&1 | show
echo "== f16 table, streams 1"; CTR_DEVICE_STREAMS=1 $B --table-dtype f16 2>&1 | show
echo "== deepfm"; $B --model deepfm 2>&1 | show
echo "== dcn_v2"; $B --model dcn_v2 2>&1 | show
echo "== full default bench"; python bench.py 2>&1 | tail -2 > gpurun_out/r02_a_bench.json; cat gpurun_out/r02_a_bench.json | cut -c1-1500
