# secondary runs: the other model families and the f16 table (device-resident + e2e, no CPU baseline)
for a in "--model deepfm" "--model dcn_v2 --hashed" "--model wide_deep --table-dtype f16"; do
  echo "== $a"; python bench.py $a --steps 50 --warmup 5 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(json.dumps({'value':d['value'],'ms_per_step':d['ms_per_step'],'e2e':d['e2e']['value'],'stage_ms':d['stage_ms_first_chunk'],'rooflines':[(r['kernel'][:24],round(r['achieved'],1),r['unit'],round(r['frac'],3)) for r in d['rooflines']],'config':d['config']['tower'],'clocks':d['clocks']['reasons']}))
    elif 'rror' in l: print(l.rstrip())
"; done
