"""CPU suite for Path C (978 MHz UAT).  PARITY UNPINNED: the dump978 sources are not in /root/reference, so
the oracle restatement (oracle/oracle978.c) is checked through what the domain offers: DO-282B frame geometry,
Reed-Solomon encode -> corrupt -> decode round trips with exact correction counts against an INDEPENDENT
encoder (libadsb_b200/synth978.py), the phase table formula of UAT978.cpp:76-100, and
modulate -> demodulate round trips."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def O():
    from oracle import oracle978 as O
    O.lib()
    return O


@pytest.fixture(scope="module")
def S():
    from libadsb_b200 import synth978 as S
    return S


def test_phase_table(O):
    t = O.Oracle978().phase_table()
    for i, q in ((0, 0), (255, 255), (128, 128), (127, 128), (200, 30), (13, 250)):   # UAT978.cpp:86-96
        ang = np.arctan2(q - 127.5, i - 127.5) + np.pi
        want = min(max(np.round(32768 * ang / np.pi), 0), 65535)
        assert t[i | (q << 8)] == int(want)
    assert t.min() >= 0 and len(np.unique(t)) > 20000


@pytest.mark.parametrize("nroots,k,t", [(12, 18, 6), (14, 34, 7), (20, 72, 10)])
def test_rs_round_trip(O, S, nroots, k, t):
    rng = np.random.default_rng(nroots)
    for _ in range(40):
        data = rng.integers(0, 256, size=k, dtype=np.uint8).tobytes()
        par = S.rs_parity(data, nroots)
        assert par == O.rs_parity(data, nroots)                 # independent encoder == libfec-style encoder
        block = data + par
        for e in range(t + 1):                                  # up to t symbol errors: corrected, exact count
            n, fixed = O.rs_decode(S.corrupt(block, e, rng), nroots)
            assert (n, fixed) == (e, block)
        bad = S.corrupt(block, t + 1 + int(rng.integers(0, 3)), rng)
        n, fixed = O.rs_decode(bad, nroots)                     # beyond t: reported, or miscorrected to another code word
        assert n == -1 and fixed == bad or n >= 0 and O.rs_decode(fixed, nroots)[0] == 0


def test_demod_round_trip(O, S):
    iq, sent = S.config5(seconds=3.0, seed=5, sigma=1.0)
    frames = O.Oracle978().handle_data(iq)
    got = {bytes(f["data"][:f["len"]]): f for f in frames}
    hit = [s for s in sent if s[2] in got]
    assert len(hit) >= 0.97 * len(sent)
    for t0, kind, payload, nerr in hit:
        f = got[payload]
        assert chr(f["updown"]) == kind and f["len"] == len(payload)
        assert abs(int(f["offset"]) - t0) <= 2 and f["rs_errors"] >= nerr      # channel noise may add symbol errors
    assert any(chr(f["updown"]) == "+" for f in frames) and {int(f["len"]) for f in frames} == {18, 34, 432}
    assert np.all(np.diff(frames["offset"].astype(np.int64)) > 0)


def test_call_segmentation_invariance(O, S):
    """HandleData-sized feeds (262144 bytes, RTLSDR.hpp:55) and ragged feeds decode the same frames."""
    iq, _ = S.config5(seconds=1.5, seed=8)
    whole = O.Oracle978().handle_data(iq)
    o = O.Oracle978()
    parts = [o.handle_data(iq[k:k + 262144]) for k in range(0, iq.size, 262144)]
    parts = np.concatenate(parts)
    assert len(whole) > 250 and whole.tobytes() == parts.tobytes()
