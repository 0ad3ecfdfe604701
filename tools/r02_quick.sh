set -x
show() { python -c "
import sys,json
d=json.loads(sys.stdin.read()); e=d['e2e']; l=d['latency']
print('value %.1fM e2e %.1fM shared %.1fM packed %.1fM packed+shared %.1fM | seq p50 %.3f p99 %.3f | closed loop p50 %.3f p99 %.3f rps %.0f'%(d['value']/1e6,e['value']/1e6,e['shared_context']['value']/1e6,e['packed']['value']/1e6,e['packed_shared_context']['value']/1e6,l['p50_ms'],l['p99_ms'],l['closed_loop_6_clients']['p50_ms'],l['closed_loop_6_clients']['p99_ms'],l['closed_loop_6_clients']['requests_per_s']))"; }
timeout 600 python -m pytest tests/test_gpu_packed.py -q -m gpu 2>&1 | tail -2
B="python bench.py --no-cpu-baseline --steps 100"
echo "== host_pack off"; CTR_HOST_PACK=1 $B 2>/dev/null | show
echo "== host_pack auto (default)"; $B 2>/dev/null | show
echo "== host_pack always"; CTR_HOST_PACK=2 $B 2>/dev/null | show
echo "== host_pack auto, 12 workers"; $B --workers 12 2>/dev/null | show
