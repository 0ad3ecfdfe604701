"""Builds libctr_b200.so (CUDA kernels + C ABI + host runtime) for sm_100a, in-tree."""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libctr_b200.so")
HOST_DEMO = os.path.join(HERE, "dcn_client_host")
STREAM_SWEEP = os.path.join(HERE, "stream_sweep_host")
SOURCES = ["engine.cu", "engine_base.cuh", "weight_file.cuh", "engine_state.cuh", "device_setup.cuh", "device_path.cuh", "workers.cuh",
           "sm_split.cuh", "debug_hooks.cuh", "static_scratch.cuh", "exchange.cuh", "exchange_host.cuh", "common.cuh", "gather.cuh", "tower.cuh", "sort.cuh", "wire.hpp",
           os.path.join("..", "..", "include", "ctr_engine.h")]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC,-Wall,-pthread", "-shared", "-ldl"]


def _nvcc() -> str:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: the engine has no prebuilt or CPU path")
    return nvcc


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build_library(force: bool = False, verbose: bool = False) -> str:
    deps = [os.path.join(CSRC, s) for s in SOURCES]
    if force or _stale(LIB, deps):
        cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB, os.path.join(CSRC, "engine.cu")]
        subprocess.check_call(cmd, cwd=CSRC)
    return LIB


def build_host_demo(force: bool = False) -> str:
    """C++ host mirror of DCNClient.main (6 closed-loop driver threads over the C ABI)."""
    src = os.path.join(CSRC, "dcn_client_host.cpp")
    if not os.path.exists(src):
        return ""
    if force or _stale(HOST_DEMO, [src, os.path.join(CSRC, "..", "..", "include", "ctr_engine.h")]):
        cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
        subprocess.check_call([cxx, "-O2", "-std=c++17", "-pthread", "-o", HOST_DEMO, src, "-ldl"], cwd=CSRC)
    return HOST_DEMO


def build_stream_sweep(force: bool = False) -> str:
    """C++ driver of BASELINE config 5 (request stream sweep: closed loop + open-loop Poisson arrivals)."""
    src = os.path.join(CSRC, "stream_sweep_host.cpp")
    if force or _stale(STREAM_SWEEP, [src, os.path.join(CSRC, "..", "..", "include", "ctr_engine.h")]):
        cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
        subprocess.check_call([cxx, "-O2", "-std=c++17", "-pthread", "-Wall", "-o", STREAM_SWEEP, src, "-ldl"], cwd=CSRC)
    return STREAM_SWEEP


ROW_PROBE = os.path.join(HERE, "..", "tools", "probes", "random_rows_probe")


def build_row_probe(force: bool = False) -> str:
    """tools/probes/random_rows_probe.cu: the random-row rate of HBM that bench.py quotes beside the gather kernel's."""
    src = ROW_PROBE + ".cu"
    if not os.path.exists(src):
        return ""
    if force or _stale(ROW_PROBE, [src]):
        subprocess.check_call([_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-o", ROW_PROBE, src])
    return ROW_PROBE


def build_all(force: bool = False) -> None:
    build_library(force)
    build_host_demo(force)
    build_stream_sweep(force)
    build_row_probe(force)
