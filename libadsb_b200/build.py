"""Build libadsb_b200/libadsbcu.so (CUDA kernels + C ABI) in-tree with nvcc for sm_100a."""
from __future__ import annotations

import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libadsbcu.so")
SOURCES = ["adsbcu_front.cu", "adsbcu_msgs.cu", "adsbcu_finish.cu", "adsbcu_kernels.cu", "adsbcu_pathb.cu", "adsbcu_uat.cu", "adsbcu_modeac.cu", "adsbcu_fields.cu", "adsbcu_api.cpp", "adsbcu_partition.cpp", "adsbcu_shard.cpp", "adsbcu_wire.cpp", "adsbcu_tables.cpp"]
HOST_SRC = [os.path.join("..", "host", "adsbcu_replay.cpp"), os.path.join("..", "host", "adsbcu_handlers.hpp")]
HEADERS = ["adsbcu_internal.h", "adsbcu_device.cuh", os.path.join("..", "..", "include", "adsbcu.h")]


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS + HOST_SRC) or not os.path.exists(REPLAY)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
           "-Xcompiler", "-fPIC", "-Xcompiler", "-pthread", "-shared", "-o", LIB] + SOURCES
    cmd[1:1] = os.environ.get("ADSBCU_CFLAGS", "").split()
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.run(cmd, cwd=CSRC, check=True)
    build_replay()
    return LIB


REPLAY = os.path.join(HERE, "host", "adsbcu_replay")


def build_replay() -> str:
    """The C++ handler shim + file replay CLI (host/), linked against libadsbcu.so."""
    subprocess.run(["g++", "-O2", "-std=c++20", "-Wall", "-o", REPLAY, os.path.join(HERE, "host", "adsbcu_replay.cpp"),
                    "-L" + HERE, "-ladsbcu", "-Wl,-rpath,$ORIGIN/.."], check=True)
    return REPLAY


if __name__ == "__main__":
    print(build(force=True, verbose=True))
