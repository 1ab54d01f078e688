"""TEST INFRASTRUCTURE ONLY.  Compiles libadsb_b200/csrc/* with g++ against the SIMT emulator
(tests/emu/include/cuda_runtime.h shadows the CUDA runtime header) into tests/emu/_build/libadsbcu_emu.so.
Only tests/test_emu_*.py load that library; the product binding (libadsb_b200/adsbcu.py) never does."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "libadsb_b200", "csrc")
OUT = os.path.join(HERE, "_build")
LIB = os.path.join(OUT, "libadsbcu_emu.so")


def sources():
    sys.path.insert(0, ROOT)
    from libadsb_b200 import build as b
    return [os.path.join(CSRC, s) for s in b.SOURCES]


def build(force: bool = False) -> str:
    os.makedirs(OUT, exist_ok=True)
    srcs = sources() + [os.path.join(HERE, "emu_runtime.cpp")]
    deps = srcs + [os.path.join(HERE, "include", "cuda_runtime.h")] + \
        [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))] + [os.path.join(ROOT, "include", "adsbcu.h")]
    if not force and os.path.exists(LIB) and all(os.path.getmtime(d) <= os.path.getmtime(LIB) for d in deps):
        return LIB
    objs = []
    procs = []
    for s in srcs:
        o = os.path.join(OUT, os.path.basename(s) + ".o")
        objs.append(o)
        cmd = ["g++", "-x", "c++", "-std=c++17", "-O1", "-g", "-fPIC", "-ffp-contract=off", "-pthread", "-w",
               "-I" + os.path.join(HERE, "include"), "-I" + CSRC, "-c", s, "-o", o]
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    bad = False
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode:
            bad = True
            sys.stderr.write(f"--- {s}\n{out}\n")
    if bad:
        raise RuntimeError("emulator build failed")
    subprocess.run(["g++", "-shared", "-pthread", "-o", LIB] + objs, check=True)
    return LIB


if __name__ == "__main__":
    print(build(force=True))
