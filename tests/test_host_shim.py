"""The C++ handler shim (libadsb_b200/host/): HandleData(span) -> C ABI -> sink, and the wire formats
of the decoded set (AVR raw line net_io.c:529-545, Beast binary net_io.c:440-487)."""
import os
import subprocess

import numpy as np
import pytest


def _replay():
    from libadsb_b200 import build
    build.build()
    return build.REPLAY


def beast_reference(recs) -> bytes:
    """writeBeastMessage (net_io.c:440-487) restated on the reference-generated golden records."""
    out = bytearray()

    def put(b):
        out.append(b)
        if b == 0x1A:
            out.append(b)
    for r in recs:
        n = int(r["msgbits"]) // 8
        out.append(0x1A)
        out.append(ord("2") if n == 7 else ord("3"))
        ts = int(r["timestamp"])
        for s in range(40, -1, -8):
            put((ts >> s) & 0xFF)
        sl = float(r["signal_level"])
        sig = int(np.floor(np.sqrt(sl) * 255 + 0.5))
        if sl > 0 and sig < 1:
            sig = 1
        put(min(sig, 255))
        for b in r["verbatim"][:n]:
            put(int(b))
    return bytes(out)


def test_beast_restatement_equals_reference_bytes(golden):
    """tests/golden/beast_ref.npz: the bytes the reference's own dump1090 sent on its --net-verbatim Beast port for the golden
    capture (tests/make_golden_beast.py).  The restatement above, used by the GPU test below, must reproduce them."""
    with np.load(os.path.join(os.path.dirname(__file__), "golden", "beast_ref.npz")) as z:
        for nfix in (0, 2):
            assert beast_reference(golden["c3_dense"][f"ref_fix{nfix}_df1"]) == z[f"fix{nfix}"].tobytes()


def test_abi_encoders_equal_reference_bytes(golden):
    """adsbcu_1090_encode_beast / _avr (host code behind the C ABI, no GPU needed) on the reference-generated records ==
    the reference's Beast bytes / its raw-port lines."""
    from libadsb_b200 import adsbcu
    with np.load(os.path.join(os.path.dirname(__file__), "golden", "beast_ref.npz")) as z:
        for nfix in (0, 2):
            ref = golden["c3_dense"][f"ref_fix{nfix}_df1"]
            msgs = np.zeros(len(ref), adsbcu.MSG_DTYPE)
            for f in adsbcu.MSG_DTYPE.names:
                msgs[f] = ref[f]
            assert adsbcu.encode_beast(msgs, verbatim=True) == z[f"fix{nfix}"].tobytes()
            want = "".join("@%012X%s;\n" % (int(r["timestamp"]), bytes(r["msg"][:int(r["msgbits"]) // 8]).hex().upper()) for r in ref)
            assert adsbcu.encode_avr(msgs).decode() == want
            cooked = adsbcu.encode_beast(msgs, verbatim=False)
            keep = ref[ref["correctedbits"] < 2]
            assert cooked.count(b"\x1a") >= len(keep) and len(cooked) <= len(z[f"fix{nfix}"])


def test_replay_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    p = subprocess.run([_replay(), "--path", "a", os.devnull], capture_output=True, text=True)
    assert p.returncode == 3
    assert "no CPU fallback" in p.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("nfix,flag", [(0, []), (2, ["--fix-2bit"])])
def test_replay_avr_and_beast_match_reference(golden, tmp_path, nfix, flag):
    g = golden["c3_dense"]
    cap = tmp_path / "cap.uc8"
    g["iq"].tofile(cap)
    beast = tmp_path / "out.beast"
    p = subprocess.run([_replay(), "--path", "a", *flag, "--beast", str(beast), str(cap)], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    ref = g[f"ref_fix{nfix}_df1"]
    want = ["@%012X%s;" % (int(r["timestamp"]), bytes(r["msg"][:int(r["msgbits"]) // 8]).hex()) for r in ref]
    assert p.stdout.split() == want
    with np.load(os.path.join(os.path.dirname(__file__), "golden", "beast_ref.npz")) as z:
        assert beast.read_bytes() == z[f"fix{nfix}"].tobytes()      # the bytes the reference itself sent on its Beast port


@pytest.mark.gpu
def test_replay_path_b_matches_reference_fixture(tmp_path):
    """b_mixed.npz was produced by the reference's own ADSB1090.cpp fed 262144-byte buffers (make_golden_b.py)."""
    with np.load(os.path.join(os.path.dirname(__file__), "golden", "b_mixed.npz")) as z:
        iq, msgs = z["iq"], z["oracle_msgs"]
    cap = tmp_path / "b.bin"
    iq.tofile(cap)
    p = subprocess.run([_replay(), "--path", "b", str(cap)], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    want = ["*%s;" % bytes(r["msg"][:int(r["msgbits"]) // 8]).hex() for r in msgs]
    assert p.stdout.split() == want


@pytest.mark.gpu
def test_replay_path_b_aircraft_callbacks_match_reference(tmp_path):
    """The reference's IListener::OnChanged callbacks for the same capture (b_mixed.npz, `ref_callbacks`, recorded from
    the reference's own ADSB1090Handler + TrafficManager): field decode and CPR run on the device
    (adsbcu_1090b_decode_fields / adsbcu_1090b_cpr_decode), the aircraft table on the host."""
    with np.load(os.path.join(os.path.dirname(__file__), "golden", "b_mixed.npz")) as z:
        iq, calls = z["iq"], z["ref_callbacks"]
    cap = tmp_path / "b.bin"
    iq.tofile(cap)
    p = subprocess.run([_replay(), "--path", "b", "--aircraft", str(cap)], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    want = ["%06X,%d,%d,%d,%d,%d,%s" % (int(c["addr"]), int(c["altitude"]), int(c["speed"]), int(c["track"]), int(c["lat1e7"]),
                                        int(c["lon1e7"]), c["callsign"].decode("ascii")) for c in calls]
    got = p.stdout.splitlines()
    assert len(got) == len(want)
    assert got == want
    assert sum(1 for c in calls if c["lat1e7"] != 0) > 10 and sum(1 for c in calls if c["speed"] != 0) > 10


@pytest.mark.gpu
def test_replay_decode_lines_equal_the_oracles_fields(golden, tmp_path, oracle_mod):
    """`adsbcu_replay --path a --fix-2bit --decode`: the handler's decoded sink (three batched decode calls per buffer) prints what
    decodeModesMessage fills in; the lines must be those the oracle's field / ES / Comm-B restatements (each pinned to the
    reference binary) give for the reference's own messages of the capture."""
    g = golden["c3_dense"]
    cap = tmp_path / "cap.uc8"
    g["iq"].tofile(cap)
    p = subprocess.run([_replay(), "--path", "a", "--fix-2bit", "--decode", str(cap)], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    got = [ln for ln in p.stdout.splitlines() if ln.startswith("# ")]
    ref = g["ref_fix2_df1"]
    f = oracle_mod.decode_fields(ref)
    e = oracle_mod.decode_es(ref["msg"], ref["correctedbits"])
    b = oracle_mod.decode_commb(ref["msg"], ref["correctedbits"])
    assert len(got) == len(ref) > 200

    def text(s):
        return bytes(s).split(b"\0")[0].decode("latin-1")

    for k in range(len(ref)):
        want = ("# DF%u airground=%u alt=%d/%u squawk=%04x/%u es=%u metype=%u.%u addr=%07X addrtype=%u valid=%08X callsign=%s "
                "cpr=%u,%u,%u alt_baro=%d alt_geom=%d gs=%.3f heading=%.3f commb=%u mbvalid=%05X mbcallsign=%s" % (
                    ref["msgtype"][k], e["airground"][k] if e["es"][k] else f["airground"][k], f["altitude_baro"][k], f["altitude_baro_valid"][k],
                    f["squawk"][k], f["squawk_valid"][k], e["es"][k], e["metype"][k], e["mesub"][k], e["addr"][k], e["addrtype"][k], e["valid"][k],
                    text(e["callsign"][k]), e["cpr_odd"][k], e["cpr_lat"][k], e["cpr_lon"][k], e["altitude_baro"][k], e["altitude_geom"][k],
                    float(e["gs_selected"][k]), float(e["heading"][k]), b["commb_format"][k], b["valid"][k], text(b["callsign"][k])))
        assert got[k] == want, (k, got[k], want)
    assert sum(" es=1 " in ln for ln in got) > 100 and sum(" commb=2 " in ln for ln in got) > 3
