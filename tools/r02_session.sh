# round 2 GPU session: tests, then device-resident bench at 1 and 2 engine streams (fp32 and f16 rows)
set -x
mkdir -p gpurun_out
if [ "$1" != "notest" ]; then timeout 1200 python -m pytest tests -q -m gpu --timeout 600 2>&1 | tail -6; fi
B="python bench.py --no-e2e --no-cpu-baseline --steps 40 --warmup 5"
show() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.1fM ms/step %.3f'%(d['value']/1e6,d['ms_per_step']), {k:round(v*1000,1) for k,v in d['stage_ms_first_chunk'].items()}, d['clocks']['reasons'], d['clocks']['sm_mhz'])
    elif 'rror' in l or 'Trace' in l: print(l.rstrip())
"; }
echo "== streams 2"; $B 2>&1 | show
echo "== streams 1"; CTR_DEVICE_STREAMS=1 $B 2>&1 | show
echo "== streams 1 chunk 303104"; CTR_DEVICE_STREAMS=1 CTR_CHUNK_ROWS=303104 $B 2>&1 | show
echo "== streams 2 f16"; $B --table-dtype f16 2>&1 | show
echo "== streams 1 f16"; CTR_DEVICE_STREAMS=1 $B --table-dtype f16 2>&1 | show
echo "== deepfm"; $B --model deepfm 2>&1 | show
echo "== dcn_v2"; $B --model dcn_v2 2>&1 | show
