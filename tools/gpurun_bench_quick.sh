# quick device-resident bench at 1 and 2 engine streams (+ gpu tests when "test" is passed)
if [ "$1" = "test" ]; then timeout 600 python -m pytest tests -q -m gpu -x --timeout 300 2>&1 | tail -4; fi
B="python bench.py --no-e2e --no-cpu-baseline --steps 50 --warmup 5"
for s in 1 2; do echo "== streams $s"; CTR_DEVICE_STREAMS=$s $B 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.1fM ms/step %.3f'%(d['value']/1e6,d['ms_per_step']), {k:round(v*1000,1) for k,v in d['stage_ms_first_chunk'].items()}, d['clocks']['reasons'])
    elif 'rror' in l: print(l.rstrip())
"; done
