"""CPU: the CUDA kernel SOURCES, compiled by g++ against the SIMT emulator of tests/emu (test infrastructure, never the
product), run a selection of the -m gpu parity tests in a subprocess with ADSBCU_TEST_EMU=1.  What this proves on a box
without a GPU: the algorithm in the .cu files (front / message / order / resolve / level kernels, Mode A/C incl. the
static-scratch entry point) gives the reference's message set.  What it cannot show (memory ordering, real concurrency,
speed) is what the -m gpu run on the B200 is for."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_selection(args, extra_env=None, timeout=900):
    env = dict(os.environ, ADSBCU_TEST_EMU="1", **(extra_env or {}))
    p = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-p", "no:cacheprovider", *args], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=timeout)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-2000:]
    return p.stdout


def test_emulated_kernels_match_reference_goldens():
    out = run_selection(["tests/test_gpu_parity.py", "-k", "golden"])
    assert " passed" in out


def test_emulated_resolve_kernel_small_batches_and_sequential_replay():
    """The ordered pass in its three modes: default batches, 40-record batches (many batch boundaries, skip state carried
    across them), and the one-thread replay."""
    for env in ({"ADSBCU_RESOLVE_BATCH": "40"}, {"ADSBCU_RESOLVE_SEQ": "1"}):
        out = run_selection(["tests/test_gpu_parity.py", "-k", "golden and c3_dense"], env)
        assert " passed" in out


def test_emulated_filter_expiry_and_static_scratch():
    out = run_selection(["tests/test_gpu_parity.py", "tests/test_zz_no_alloc.py", "-k", "expiry or static_scratch"])
    assert " passed" in out


def test_emulated_es_payload_decode_matches_reference_fixture():
    """fields_es_kernel's source against what the reference binary's decodeModesMessage produced (tests/golden/fields_es.npz)."""
    out = run_selection(["tests/test_fields_es.py", "-k", "gpu_es_equals_reference_fixture"])
    assert " passed" in out


def test_emulated_commb_payload_decode_matches_reference_fixture():
    """fields_commb_kernel's source against what the reference binary's decodeCommB produced (tests/golden/fields_mb.npz)."""
    out = run_selection(["tests/test_fields_mb.py", "-k", "gpu_commb_equals_reference_fixture"])
    assert " passed" in out
