set -x
for CFG in "1 0" "2 75776" "2 37888"; do
  set -- $CFG
  CTR_DEVICE_STREAMS=$1 CTR_CHUNK_ROWS=$2 timeout 600 python tools/bench_row_sharded.py --log2-rows 27 --candidates 160000 --steps 10 --skip-replicated 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('STREAMS=$1 CHUNK=$2 gpus', d['gpus'], 'row_sharded %.1f M' % (d['row_sharded']['candidates_per_s']/1e6), 'exchange %.1f M' % (d['row_sharded_nccl_exchange']['candidates_per_s']/1e6))"
done
