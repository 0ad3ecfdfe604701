# the C++ mirror of DCNClient.main over the C ABI: reference defaults (1500 candidates x 43 fields, 6 threads x 1000 requests)
python - <<'PY'
import numpy as np, sys, os
sys.path.insert(0, ".")
import dtfs_b200
from dtfs_b200 import weights as W, build
build.build_host_demo()
rng = np.random.default_rng(0)
W.write_weight_file("/tmp/lr43.ctrw", model_kind=0, num_fields=43, num_rows=44, b0=0.1,
                    w1=rng.uniform(-.5, .5, 44).astype(np.float32), emb=rng.uniform(-.5, .5, (44, 16)).astype(np.float32))
d = 43 * 16
mlp, fan = [], d
for h in (1024, 512, 256):
    mlp.append((rng.uniform(-.05, .05, (h, fan)).astype(np.float32), np.zeros(h, np.float32))); fan = h
W.write_weight_file("/tmp/wd43.ctrw", model_kind=1, num_fields=43, num_rows=44, b0=0.1,
                    w1=rng.uniform(-.5, .5, 44).astype(np.float32), emb=rng.uniform(-.5, .5, (44, 16)).astype(np.float32),
                    mlp=mlp, w_out=rng.uniform(-1, 1, 256).astype(np.float32))
PY
L=distributed-tf-serving_b200/libctr_b200.so
echo "== LR (BASELINE config 1 family), reference defaults"; ./distributed-tf-serving_b200/dcn_client_host --lib $L --weights /tmp/lr43.ctrw
echo "== Wide&Deep F=43, reference defaults"; ./distributed-tf-serving_b200/dcn_client_host --lib $L --weights /tmp/wd43.ctrw
echo "== config 1: 1000 candidates"; ./distributed-tf-serving_b200/dcn_client_host --lib $L --weights /tmp/lr43.ctrw --candidates 1000
