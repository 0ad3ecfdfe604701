# BASELINE config 5 through the C++ driver: bash tools/gpurun_sweep.sh "1 2 4 8"   (engine over G GPUs, in-process sharding)
GS=${1:-1}
python - <<'PY'
import dtfs_b200, bench
from dtfs_b200 import build
build.build_stream_sweep()
bench.write_model("/tmp/wd_sweep.ctrw", 1 << 22)
PY
for g in $GS; do
  echo "== $g GPU(s)"
  timeout 900 ./distributed-tf-serving_b200/stream_sweep_host --lib distributed-tf-serving_b200/libctr_b200.so --weights /tmp/wd_sweep.ctrw --devices $g --workers ${WORKERS:-12} --out gpurun_out/${TAG:-r02}_stream_sweep_g$g.md
done
