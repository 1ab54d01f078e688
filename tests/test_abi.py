"""CPU suite: the C-ABI library builds, loads, exports every symbol include/adsbcu.h declares,
and refuses to run without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    src = open(os.path.join(ROOT, "include", "adsbcu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(adsbcu_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(adsbcu_mod):
    L = adsbcu_mod.load()
    names = _declared_functions()
    assert len(names) >= 11
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/adsbcu.h but not exported"
    assert sorted(adsbcu_mod.EXPORTS) == names
    assert L.adsbcu_abi_version() == 2


def test_struct_layouts(adsbcu_mod):
    assert adsbcu_mod.MSG_DTYPE.itemsize == 56
    assert C.sizeof(adsbcu_mod.Config) == 24
    assert C.sizeof(adsbcu_mod.Stats) == 72


def test_argument_errors_without_touching_the_gpu(adsbcu_mod):
    L = adsbcu_mod.load()
    assert L.adsbcu_create(None, None) == -1
    bad = adsbcu_mod.Config(C.sizeof(adsbcu_mod.Config), 0, 7, 1, 0, 0)
    h = C.c_void_p()
    assert L.adsbcu_create(C.byref(bad), C.byref(h)) == -1
    assert b"nfix_crc" in L.adsbcu_last_error(None)
    assert L.adsbcu_1090_feed(None, None, 0, None, 0, None) == -1
    assert L.adsbcu_get_stats(None, None) == -1


def test_no_cpu_fallback(adsbcu_mod):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(adsbcu_mod.AdsbcuError, match="no usable CUDA device"):
        adsbcu_mod.Demod1090()


def test_product_does_not_import_oracle():
    """Only tests/, smoke() and bench.py's CPU-baseline legs may touch oracle/."""
    for d, _, files in os.walk(os.path.join(ROOT, "libadsb_b200")):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".cuh", ".h", ".hpp")):
                txt = open(os.path.join(d, f)).read()
                assert "oracle1090" not in txt and "liboracle" not in txt, os.path.join(d, f)
