set -x
B="python bench.py --no-e2e --no-cpu-baseline --steps 40 --warmup 5"
show() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.1fM ms/step %.3f'%(d['value']/1e6,d['ms_per_step']), {k:round(v*1000,1) for k,v in d['stage_ms_first_chunk'].items()}, d['clocks']['reasons'], d['clocks']['sm_mhz'])
    elif 'rror' in l or 'Trace' in l: print(l.rstrip())
"; }
echo "== streams 1 chunk 303104, max carveout"; CTR_DEVICE_STREAMS=1 CTR_CHUNK_ROWS=303104 $B 2>&1 | show
echo "== streams 1 chunk 303104, default carveout for the gather"; CTR_GATHER_DEFAULT_CARVEOUT=1 CTR_DEVICE_STREAMS=1 CTR_CHUNK_ROWS=303104 $B 2>&1 | show
echo "== streams 2, default carveout for the gather"; CTR_GATHER_DEFAULT_CARVEOUT=1 $B 2>&1 | show
echo "== streams 2, max carveout"; $B 2>&1 | show
echo "== streams 1 chunk 303104, gather 5 CTAs"; CTR_GATHER_CTAS=5 CTR_GATHER_DEFAULT_CARVEOUT=1 CTR_DEVICE_STREAMS=1 CTR_CHUNK_ROWS=303104 $B 2>&1 | show
