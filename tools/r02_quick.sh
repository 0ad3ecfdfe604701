set -x
B="python bench.py --no-e2e --no-cpu-baseline --steps 300 --warmup 5"
show() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.1fM ms/step %.3f'%(d['value']/1e6,d['ms_per_step']), {k:round(v*1000,1) for k,v in d['stage_ms_first_chunk'].items()}, d['clocks']['reasons'], d['clocks']['sm_mhz'])
    elif 'rror' in l or 'Trace' in l: print(l.rstrip())
"; }
for rep in 1 2; do
echo "== default"; $B 2>/dev/null | show
echo "== 1 tile in flight"; CTR_TOWER_TILES=1 $B 2>/dev/null | show
echo "== 1 tile in flight + L2 persist"; CTR_TOWER_TILES=1 CTR_L2_PERSIST=1 $B 2>/dev/null | show
echo "== 1 tile, streams 1"; CTR_TOWER_TILES=1 CTR_DEVICE_STREAMS=1 $B 2>/dev/null | show
done
