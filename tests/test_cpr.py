"""SURVEY 8f rank 2, Path A side: the CPR decoders of dump1090/cpr.c (:162-369).
Golden vectors: the tables of the reference's own dump1090/cprtests.c (tests/make_golden_cpr.py), with the exact
doubles the reference's cpr.c returns.  Results are compared as bit patterns."""
import os

import numpy as np
import pytest

from oracle import oraclecpr as P

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "cpr_vectors.npz"))


def same_bits(a, b):
    return a.tobytes() == b.tobytes()


def first_diff(a, b, rec):
    k = next(i for i in range(a.size) if a[i].tobytes() != b[i].tobytes())
    return rec[k], a[k], b[k]


# ---- CPU -------------------------------------------------------------------------------------------------------

def test_oracle_on_reference_test_vectors():
    """cprtests.c: result codes equal, positions within the 1e-6 degrees that test allows ..."""
    got = P.decode(G["vectors"])
    want = G["expected"]
    assert G["vectors"].size == 80
    assert np.array_equal(got["result"], want["result"])
    assert np.all(np.abs(got["lat"] - want["lat"]) <= 1e-6) and np.all(np.abs(got["lon"] - want["lon"]) <= 1e-6)
    # ... and the very doubles the reference's cpr.c produced for them
    assert same_bits(got, G["reference"]), first_diff(got, G["reference"], G["vectors"])


def test_oracle_on_reference_random_outputs():
    got = P.decode(G["random_in"])
    assert same_bits(got, G["random_reference"]), first_diff(got, G["random_reference"], G["random_in"])
    assert (got["result"] == 0).sum() > got.size // 2


@pytest.mark.skipif(not os.path.exists(P.REF_PATH), reason="oracle/_ref/libcpr_ref.so not built (needs /root/reference)")
def test_oracle_against_compiled_reference():
    rec = P.random_records(30000, 77)
    got, ref = P.decode(rec), P.decode_ref(rec)
    assert same_bits(got, ref), first_diff(got, ref, rec)
    for mode in range(4):
        assert ((rec["mode"] == mode) & (got["result"] == 0)).sum() > 1000


@pytest.mark.skipif(not os.path.exists(os.path.join(P.HERE, "_ref", "cprtests_ref")), reason="reference test program not built")
def test_reference_passes_its_own_cprtests():
    import subprocess
    r = subprocess.run([os.path.join(P.HERE, "_ref", "cprtests_ref")], capture_output=True, text=True)
    assert r.returncode == 0 and "FAIL" not in r.stdout + r.stderr


def test_nl_zone_edges():
    assert P.nl(0.0) == 59 and P.nl(10.47047129) == 59 and P.nl(10.47047130) == 58 and P.nl(-10.47047130) == 58
    assert P.nl(86.9999) == 2 and P.nl(87.0) == 1 and P.nl(-90.0) == 1


# ---- GPU -------------------------------------------------------------------------------------------------------

@pytest.mark.gpu
def test_gpu_cpr_reference_vectors():
    from libadsb_b200 import adsbcu
    got = adsbcu.Demod1090().cpr_decode_a(G["vectors"])
    assert same_bits(got, G["reference"]), first_diff(got, G["reference"], G["vectors"])
    got = adsbcu.Demod1090().cpr_decode_a(G["random_in"])
    assert same_bits(got, G["random_reference"]), first_diff(got, G["random_reference"], G["random_in"])


@pytest.mark.gpu
def test_gpu_cpr_random_matches_oracle():
    from libadsb_b200 import adsbcu
    d = adsbcu.Demod1090()
    for seed, n in ((1, 400000), (2, 1), (3, 257)):
        rec = P.random_records(n, seed) if n > 8 else P.random_records(64, seed)[:n]
        got, want = d.cpr_decode_a(rec), P.decode(rec)
        assert same_bits(got, want), first_diff(got, want, rec)


@pytest.mark.gpu
def test_gpu_cpr_dense_grid_airborne():
    """Every 37th/41st raw value against each other: zone edges, poles and the date line all occur."""
    from libadsb_b200 import adsbcu
    v = np.arange(0, 131072, 37 * 41)
    a, b, c, e = np.meshgrid(v, v[::3], v, v[::3], indexing="ij")
    rec = np.zeros(a.size * 2, dtype=P.IN_DTYPE)
    for f, x in (("even_lat", a), ("even_lon", b), ("odd_lat", c), ("odd_lon", e)):
        rec[f] = np.tile(x.ravel(), 2)
    rec["fflag"][a.size:] = 1
    got, want = adsbcu.Demod1090().cpr_decode_a(rec), P.decode(rec)
    assert same_bits(got, want), first_diff(got, want, rec)


@pytest.mark.gpu
def test_gpu_cpr_empty_and_bad_mode():
    from libadsb_b200 import adsbcu
    d = adsbcu.Demod1090()
    assert d.cpr_decode_a(np.zeros(0, dtype=P.IN_DTYPE)).size == 0
    rec = P.random_records(16, 4)
    rec["mode"][5] = 9
    with pytest.raises(Exception):
        d.cpr_decode_a(rec)
