set -x
B="python bench.py --no-e2e --no-cpu-baseline --steps 40 --warmup 5"
show() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.1fM ms/step %.3f'%(d['value']/1e6,d['ms_per_step']), {k:round(v*1000,1) for k,v in d['stage_ms_first_chunk'].items()}, d['clocks']['reasons'], d['clocks']['sm_mhz'])
    elif 'rror' in l or 'Trace' in l: print(l.rstrip())
"; }
echo "== small table (L2-resident rows), streams 1"; CTR_DEVICE_STREAMS=1 $B --rows-per-field 4096 2>/dev/null | show
echo "== small table, streams 2"; $B --rows-per-field 4096 2>/dev/null | show
echo "== small table, streams 2, co-resident"; CTR_GATHER_MAX_CARVEOUT=1 $B --rows-per-field 4096 2>/dev/null | show
echo "== full table, streams 2, co-resident"; CTR_GATHER_MAX_CARVEOUT=1 $B 2>/dev/null | show
echo "== full table, streams 2"; $B 2>/dev/null | show
echo "== full table, streams 2, gather 1 CTA/SM"; CTR_GATHER_CTAS=1 $B 2>/dev/null | show
echo "== full table, streams 2, gather 2 CTA/SM"; CTR_GATHER_CTAS=2 $B 2>/dev/null | show
