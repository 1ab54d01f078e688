"""TEST INFRASTRUCTURE ONLY: point the ctypes binding of THIS process at the emulator build of the kernel sources.
Imported by tests/emu/run_case.py (a subprocess of tests/test_emu_*.py), never by product code."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def load_emulated():
    sys.path.insert(0, HERE)
    import build_emu
    lib = build_emu.build()
    from libadsb_b200 import adsbcu
    adsbcu._lib = None
    adsbcu.LIB_PATH = lib
    adsbcu.load()
    return adsbcu
