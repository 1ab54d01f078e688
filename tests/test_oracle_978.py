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
    got = {}
    for f in frames:             # first sighting: a frame lying in the stale stretch is reported again on later calls
        got.setdefault(bytes(f["data"][:f["len"]]), f)
    hit = [s for s in sent if s[2] in got]
    # the reference's staging loses the frames that reach into the stretch its short memmove leaves stale (UAT978.cpp:57,
    # about 4500 of every 56560 samples): roughly one frame in twelve
    assert 0.85 * len(sent) <= len(hit) <= 0.97 * len(sent)
    for t0, kind, payload, nerr in hit:
        f = got[payload]
        assert chr(f["updown"]) == kind and f["len"] == len(payload)
        assert abs(int(f["offset"]) - t0) <= 2 and f["rs_errors"] >= nerr      # channel noise may add symbol errors
    assert any(chr(f["updown"]) == "+" for f in frames) and {int(f["len"]) for f in frames} == {18, 34, 432}
    assert np.all(np.diff(frames["offset"].astype(np.int64)) > 0)


def test_call_segmentation_matters_as_in_the_reference(O, S):
    """HandleData presents what is left of a feed at once (UAT978.cpp:49-52), and what its short memmove leaves stale
    depends on that length: one call with the whole capture and 262144-byte feeds (RTLSDR.hpp:55) see different streams.
    (tests below pin both flows to the compiled reference.)"""
    iq, _ = S.config5(seconds=1.5, seed=8)
    whole = O.Oracle978().handle_data(iq)
    o = O.Oracle978()
    parts = np.concatenate([o.handle_data(iq[k:k + 262144]) for k in range(0, iq.size, 262144)])
    assert len(whole) > 200 and len(parts) > 200
    common = set(map(int, whole["offset"])) & set(map(int, parts["offset"]))
    assert len(common) > 0.8 * min(len(whole), len(parts))


# ---- stages pinned to the reference's own UAT978.cpp (oracle/_ref/libref978.so, compiled where it lies) -----------
import os  # noqa: E402

GOLD978 = os.path.join(os.path.dirname(__file__), "golden", "uat_ref.npz")


def feeds_of(iq, sizes):
    k, i = 0, 0
    while k < iq.size:
        step = sizes[i % len(sizes)]
        yield iq[k:k + step]
        k += step
        i += 1


def test_phase_table_equals_reference_fixture(O):
    """C1: every entry of the table the compiled InitATan2Table produced (UAT978.cpp:76-100)."""
    with np.load(GOLD978) as z:
        assert np.array_equal(O.Oracle978().phase_table(), z["phase_table"])


def test_staging_equals_reference_fixture(O, S):
    """C2: the process_buffer(len, offset) -> consumed sequence HandleData made in the reference (UAT978.cpp:43-60),
    and the frames found on the way, for a capture fed in RTLSDR.hpp's 262144-byte buffers."""
    with np.load(GOLD978) as z:
        calls, offs, rs, ln = z["calls"], z["frame_offsets"], z["frame_rs"], z["frame_len"]
        iq = S.config5(seconds=float(z["seconds"]), seed=int(z["seed"]))[0]
        buf = int(z["buf"])
    o = O.Oracle978()
    frames = np.concatenate([o.handle_data(part) for part in feeds_of(iq, [buf])])
    assert np.array_equal(o.calls(), calls)
    assert np.array_equal(frames["offset"], offs) and np.array_equal(frames["rs_errors"], rs) and np.array_equal(frames["len"], ln)
    assert len(frames) > 80


@pytest.mark.parametrize("sizes", [[262144], [65536 * 2], [100000, 7, 262144, 4096, 30], [2, 8190, 600000]])
def test_staging_vs_reference_live(O, S, sizes):
    """The compiled reference handler with the oracle's process_buffer plugged in == the oracle's own HandleData
    restatement, call by call, for regular and ragged feeds; the table it built equals the oracle's."""
    from oracle import ref978 as R
    if not R.have_ref():
        pytest.skip("reference UAT978.cpp not built here")
    iq = S.config5(seconds=0.4, seed=sum(sizes) % 97)[0]
    inner = O.Oracle978()
    r = R.Ref978(lambda phi, offset: inner.process_buffer(phi, offset)[0])
    assert np.array_equal(r.phase_table(), inner.phase_table())
    o = O.Oracle978()
    for part in feeds_of(iq, sizes):
        r.handle_data(part)
        o.handle_data(part)
        assert r.state() == o.state()
    assert np.array_equal(np.array(r.calls, np.int64).reshape(-1, 3), o.calls())
    assert len(r.calls) >= 10
