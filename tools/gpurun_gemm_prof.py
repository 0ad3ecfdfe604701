import os, sys, tempfile
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dtfs_b200
from dtfs_b200 import binding, weights as W
os.environ["CTR_GEMM_PROF"] = "1"
tmp = tempfile.mkdtemp()
path = os.path.join(tmp, "lr.ctrw")
W.write_weight_file(path, model_kind=0, num_fields=39, num_rows=64, b0=0.0, w1=np.zeros(64, np.float32), emb=np.zeros((64, 16), np.float32))
e = binding.Engine(device_ids=(0,), max_candidates=1024)
e.load_weights(path)
rng = np.random.default_rng(0)
for (M, N, K) in [(18944, 1024, 640), (18944, 512, 1024), (18944, 256, 512), (18944, 640, 640), (18944, 768, 640)]:
    A = rng.uniform(-1, 1, (M, K)).astype(np.float32); Wt = rng.uniform(-.1, .1, (N, K)).astype(np.float32)
    e.debug_gemm(A, Wt, np.zeros(N, np.float32))
e.close()
