"""CPU tests of the drop-in boundary: the C-ABI library builds, loads, exports every symbol
include/ctr_engine.h declares, refuses to run without a GPU (no fallback), and the host
mirror of the reference client's interface behaves like the reference."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ctr_engine.h")


@pytest.fixture(scope="module")
def binding(lib_built):
    from dtfs_b200 import binding
    return binding


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ctr_[a-z0-9_]+)\s*\(", src)) - {"ctr_callback"})


def test_library_exports_every_declared_symbol(binding):
    lib = binding.load_library()
    declared = _declared_symbols()
    assert len(declared) >= 17
    for sym in declared:
        assert hasattr(lib, sym), "libctr_b200.so does not export %s" % sym
    assert sorted(binding.EXPORTS) == declared, "binding.EXPORTS is out of date with the header"
    assert lib.ctr_abi_version() == 1
    out = subprocess.check_output(["nm", "-D", "--defined-only", binding.LIB_PATH], text=True)
    exported = set(re.findall(r" T (ctr_[a-z0-9_]+)", out))
    assert exported == set(declared), "exported but undeclared: %r" % (exported - set(declared),)


def test_header_is_plain_c_and_struct_layout_matches(binding, tmp_path):
    """No CUDA / C++ / torch types in the ABI: the header compiles as C99, and the ctypes
    mirrors of its structs have the same size."""
    src = tmp_path / "t.c"
    src.write_text('#include "ctr_engine.h"\n#include <stdio.h>\nint main(void){printf("%zu %zu %d\\n", sizeof(ctr_config),'
                   ' sizeof(ctr_model_info), CTR_STATIC_MAX_CANDIDATES);return 0;}\n')
    exe = tmp_path / "t"
    cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
    subprocess.check_call([cc, "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.dirname(HEADER), str(src), "-o", str(exe)])
    a, b, c = map(int, subprocess.check_output([str(exe)], text=True).split())
    assert a == ctypes.sizeof(binding.CtrConfig) and b == ctypes.sizeof(binding.CtrModelInfo)
    assert c == binding.STATIC_MAX_CANDIDATES
    code = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    assert not re.search(r"cudaStream_t|CUstream|torch|at::|#include\s*<cuda", code), "no CUDA / torch types in signatures"


def test_no_cpu_fallback(binding):
    import torch
    if torch.cuda.is_available():
        pytest.skip("this check is for a box without a GPU")
    with pytest.raises(binding.CtrError) as ei:
        binding.Engine(device_ids=(0,))
    assert ei.value.code == -2 and "no CPU fallback" in str(ei.value)


def test_missing_library_is_loud(binding, monkeypatch):
    monkeypatch.setattr(binding, "_lib", None)
    monkeypatch.setattr(binding, "LIB_PATH", "/nonexistent/libctr_b200.so")
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        binding.load_library()


def test_bad_arguments(binding):
    lib = binding.load_library()
    assert lib.ctr_engine_create(None, None) == -1 and b"null" in lib.ctr_last_error()
    cfg = binding.CtrConfig()
    cfg.struct_size = 4                                    # ABI mismatch is caught before any CUDA call
    h = ctypes.c_void_p()
    assert lib.ctr_engine_create(ctypes.byref(cfg), ctypes.byref(h)) == -1 and b"size mismatch" in lib.ctr_last_error()
    assert lib.ctr_partition_bounds(10, 0, None, None) == -1
    assert lib.ctr_engine_destroy(None) == 0
    assert lib.ctr_wait(None, 1, 0) == -1


@given(st.integers(0, 10 ** 7), st.integers(1, 64))
@settings(max_examples=200, deadline=None)
def test_partition_bounds_abi_equals_oracle(binding, n, p):
    from oracle import dcn_client as odc
    assert binding.partition_bounds(n, p) == odc.partition_bounds(n, p)


def test_dcn_client_mirror_matches_reference_restatement(binding):
    """The host mirror (product) against the oracle's restatement of DCNClient.java."""
    from dtfs_b200 import dcn_client as dc
    from oracle import dcn_client as odc
    assert (dc.FIELD_NUM, dc.candidateNum, dc.requestNum, dc.concurrentNum) == (43, 1500, 1000, 6)
    for a, b in zip(dc.getFeatureIds(3), odc.get_feature_ids()):
        assert np.array_equal(a, b)
    for a, b in zip(dc.getFeatureWts(3), odc.get_feature_wts()):
        assert np.array_equal(a, b)
    lst = np.arange(1001)
    for p in (1, 2, 3, 7, 1001, 2000):
        got, ref = dc.partitionList(lst, p), odc.partition_list(lst, p)
        assert len(got) == len(ref) and all(np.array_equal(x, y) for x, y in zip(got, ref))

    class FakeEngine:                                        # stands in for a GPU: scores = first weight
        num_fields = 43

        def predict(self, ids, wts):
            return wts.reshape(-1, 43)[:, 0].copy()
    chans = [dc.Channel(FakeEngine()) for _ in range(3)]
    rng = np.random.default_rng(1)
    wts = rng.uniform(size=(1500, 43)).astype(np.float32)
    ids = np.ones((1500, 43), np.int64)
    b = binding.partition_bounds(1500, 3)
    merged = dc.processARequest(chans, [ids[s:e] for s, e in b], [wts[s:e] for s, e in b], sort=False)
    assert np.array_equal(merged, wts[:, 0]), "ordered merge: output j is candidate j (DC:161-164)"
    srt = dc.processARequest(chans, [ids[s:e] for s, e in b], [wts[s:e] for s, e in b], sort=True)
    assert np.array_equal(srt, odc.java_float_sort(wts[:, 0]))
    with pytest.raises(ValueError):                          # DC:97: value count vs declared shape
        dc.sendRequest(chans[0], np.ones(44, np.int64), np.ones(44, np.float32))


def test_host_demo_builds(lib_built):
    from dtfs_b200 import build
    exe = build.build_host_demo()
    assert exe and os.path.exists(exe)
    r = subprocess.run([exe, "--help"], capture_output=True, text=True)
    assert r.returncode == 0 and "driver threads" in r.stdout
    exe = build.build_stream_sweep()
    assert exe and os.path.exists(exe)
    r = subprocess.run([exe, "--help"], capture_output=True, text=True)
    assert r.returncode == 0 and "open loop" in r.stdout


# ---------------------------------------------------------------- N > 1, one process per GPU (gloo on CPU)
def _dist_worker(rank, world, port, q):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    import dtfs_b200  # noqa: F401
    from dtfs_b200 import dcn_client as dc
    from oracle import model as om
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    spec = om.ModelSpec(om.DEEPFM, 39, 16, om.criteo_num_rows(rows_per_field=64), hidden=(32,), scale_e=0.25)
    w = om.make_weights(spec)                                # same seed on every rank = replicated tables
    score = om.as_model(w, "fp32")                           # CPU stand-in for engine.predict
    for n in (1001, 2, 0):
        ids, wts = (om.criteo_inputs(n, rows_per_field=64) if rank == 0 else (None, None))
        out = dc.processARequestDistributed(score, ids, wts, 39)
        tag = lambda i, w_: ((i[:, 13] * 31 + i[:, 20]) % 1000003).astype(np.float32)   # row-local and exact
        out_tag = dc.processARequestDistributed(tag, ids, wts, 39)
        if rank == 0:
            ref = score(ids, wts) if n else np.zeros(0, np.float32)
            ok = out.shape == (n,) and np.allclose(out, ref, rtol=0, atol=1e-6)     # BLAS blocking differs per shard size
            ok = ok and np.array_equal(out_tag, tag(ids, wts) if n else np.zeros(0, np.float32))   # ordered merge, exact
            q.put((n, bool(ok)))
        else:
            assert out is None and out_tag is None
        a, b = dc.shard_for_rank(n, world, rank)
        assert (b - a) == (n // world if rank < world - 1 else n - (world - 1) * (n // world))
    dist.destroy_process_group()


def test_one_process_per_gpu_split_and_ordered_merge():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_dist_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    [p.join(180) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    results = [q.get(timeout=5) for _ in range(3)]
    assert sorted(r[0] for r in results) == [0, 2, 1001] and all(r[1] for r in results)


# ---------------------------------------------------------------- wire-compat front door (host half)
def test_wire_front_door_matches_oracle_codec(binding):
    """The product's hand-written PredictRequest parser / PredictResponse writer against the oracle's
    independent codec (both restate tensor.proto:48,64 and predict.proto:12-40)."""
    from oracle import clib, model as om
    ids, wts = om.criteo_inputs(37, n_sparse=30, rows_per_field=1 << 20)        # F = 43, the reference's FIELD_NUM
    ids[3, 5] = -7                                                             # 10-byte varint
    req = clib.encode_predict_request(ids, wts)                                # int64_val / float_val, as DCNClient.java:101,107
    i2, w2 = binding.wire_parse_request(req, 43)
    assert np.array_equal(i2, ids) and np.array_equal(w2, wts)
    scores = np.linspace(0, 1, 37, dtype=np.float32)
    resp = binding.wire_build_response(scores)
    assert np.array_equal(clib.decode_predict_response(resp, 37), scores)
    assert b"prediction_node" in resp and b"DCN" in resp
    # tensor_content variant (raw little-endian bytes), built by hand from the field numbers
    def varint(v):
        out = b""
        while v >= 0x80:
            out += bytes([v & 0x7F | 0x80]); v >>= 7
        return out + bytes([v])
    def ld(field, payload):
        return varint(field << 3 | 2) + varint(len(payload)) + payload
    shape = b"".join(ld(2, varint(1 << 3) + varint(d)) for d in (37, 43))
    t_ids = varint(1 << 3) + varint(9) + ld(2, shape) + ld(4, ids.astype("<i8").tobytes())
    t_wts = varint(1 << 3) + varint(1) + ld(2, shape) + ld(4, wts.astype("<f4").tobytes())
    req2 = ld(2, ld(1, b"feat_wts") + ld(2, t_wts)) + ld(2, ld(1, b"feat_ids") + ld(2, t_ids)) + ld(1, ld(1, b"DCN"))
    i3, w3 = binding.wire_parse_request(req2, 43)
    assert np.array_equal(i3, ids) and np.array_equal(w3, wts)
    # INVALID_ARGUMENT cases: wrong FIELD_NUM, value count vs shape, missing input, truncated message
    for bad, f in ((req, 39), (req[:len(req) - 9], 43), (ld(2, ld(1, b"feat_ids") + ld(2, t_ids)), 43), (req2[:50], 43)):
        with pytest.raises(binding.CtrError) as ei:
            binding.wire_parse_request(bad, f)
        assert ei.value.code == -1
    short = varint(1 << 3) + varint(9) + ld(2, shape) + ld(10, b"".join(varint(1) for _ in range(10)))   # 10 values, shape 37x43
    with pytest.raises(binding.CtrError):
        binding.wire_parse_request(ld(2, ld(1, b"feat_ids") + ld(2, short)) + ld(2, ld(1, b"feat_wts") + ld(2, t_wts)), 43)


# ---------------------------------------------------------------- f4: SavedModel variables -> weight file
@pytest.mark.parametrize("name", ["lr", "wide_deep", "deepfm", "dcn_v2"])
def test_savedmodel_variables_round_trip(tmp_path, name):
    """tools/export_savedmodel.py's variable dictionary (TF layout: kernels [in, out]) converted back by
    savedmodel_import.py reproduces the golden weight file byte for byte (but for the
    procedural-table seed / scales in the header, which a raw table does not use).  (The TensorFlow half of the exporter is
    source only on this image; this pins the naming and layout contract between the two halves.)"""
    import importlib.util
    import dtfs_b200  # noqa: F401
    from dtfs_b200 import savedmodel_import as imp
    spec = importlib.util.spec_from_file_location("export_savedmodel", os.path.join(ROOT, "tools", "export_savedmodel.py"))
    exp = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(exp)
    src = os.path.join(ROOT, "tests", "golden", name + ".ctrw")
    v, h = exp.variables_from_ctrw(src)
    npz = str(tmp_path / "variables.npz")
    np.savez(npz, **v)
    out = str(tmp_path / "back.ctrw")
    imp.convert(imp.load_variables(npz), name, out)
    a, b = bytearray(open(out, "rb").read()), bytearray(open(src, "rb").read())
    a[64:80] = b[64:80] = bytes(16)        # table seed / scales: only meaningful for procedural tables
    assert a == b
