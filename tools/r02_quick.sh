set -x
show() { python -c "
import sys,json
d=json.loads(sys.stdin.read()); e=d['e2e']; l=d['latency']
print('value %.1fM e2e %.1fM shared %.1fM packed %.1fM (same bits %s) packed+shared %.1fM | seq p50 %.3f p99 %.3f'%(d['value']/1e6,e['value']/1e6,e['shared_context']['value']/1e6,e['packed']['value']/1e6,e['packed']['bit_identical_to_drop_in'],e['packed_shared_context']['value']/1e6,l['p50_ms'],l['p99_ms']))"; }
B="python bench.py --no-cpu-baseline --steps 100 --closed-loop-requests 0"
timeout 900 python -m pytest tests -q -m gpu --timeout 600 2>&1 | tail -2
echo "== compact layouts on the block-tile kernel"; $B 2>/dev/null | show
echo "== compact layouts on the warp-tile kernel"; CTR_COMPACT_WARP_TILE=1 $B 2>/dev/null | show
