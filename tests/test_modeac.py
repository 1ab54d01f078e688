"""Mode A/C (dump1090 --modeac, demod_2400.c:539-726)."""
import os

import numpy as np
import pytest

BLOCK = 131072      # MODES_MAG_BUF_SAMPLES (dump1090.h:92-94)
GOLD = os.path.join(os.path.dirname(__file__), "golden", "ac_mixed.npz")


def oracle_with_modeac(O, nfix, iq):
    o = O.Oracle1090(nfix, 1)
    O.lib().oracle1090_set_mode_ac(o.h, 1)
    return o.feed(iq, cap=1 << 18)


def test_oracle_modeac_matches_reference_fixture(oracle_mod):
    with np.load(GOLD) as z:
        iq, ref = z["iq"], z["ref_fix1_modeac"]
    got = oracle_with_modeac(oracle_mod, 1, iq)
    assert (ref["msgtype"] == 32).sum() > 50
    assert oracle_mod.same_messages(ref, got) == []


def test_oracle_modeac_vs_reference_binary(oracle_mod):
    if not oracle_mod.have_ref():
        pytest.skip("reference binary not built here")
    from libadsb_b200 import synth
    for seed, sigma in ((5, 2.0), (6, 5.0)):
        iq = synth.config_ac(seconds=0.25, seed=seed, sigma=sigma)[0]
        ref = oracle_mod.run_ref(iq, 0, 1, extra=("--modeac",))
        assert oracle_mod.same_messages(ref, oracle_with_modeac(oracle_mod, 0, iq)) == []


def dump1090_loop(d, iq, block=BLOCK):
    """dump1090.c:895-905: per mag_buf demodulate2400, then demodulate2400AC on the same buffer."""
    out, n = [], iq.size // 2
    for s in range(0, n, block):
        e = min(n, s + block)
        out.append(d.feed(iq[2 * s:2 * e]))
        out.append(d.modeac_block(iq[2 * s:2 * e], iq[2 * max(0, s - 326):2 * s], 5 * s))
    return np.concatenate(out)


@pytest.mark.gpu
def test_gpu_modeac_matches_reference_fixture(adsbcu_mod, oracle_mod):
    with np.load(GOLD) as z:
        iq, ref = z["iq"], z["ref_fix1_modeac"]
    got = dump1090_loop(adsbcu_mod.Demod1090(1, 1), iq)
    assert oracle_mod.same_messages(ref, got) == []


@pytest.mark.gpu
@pytest.mark.parametrize("seed,sigma,block", [(11, 2.0, BLOCK), (12, 6.0, BLOCK), (13, 2.0, 40000)])
def test_gpu_modeac_vs_oracle(adsbcu_mod, oracle_mod, seed, sigma, block):
    from libadsb_b200 import synth
    iq = synth.config_ac(seconds=0.4, seed=seed, sigma=sigma)[0]
    o = oracle_mod.Oracle1090(0, 1, block)
    oracle_mod.lib().oracle1090_set_mode_ac(o.h, 1)
    want = o.feed(iq, cap=1 << 18)
    got = dump1090_loop(adsbcu_mod.Demod1090(0, 1, block_samples=block), iq, block)
    assert (want["msgtype"] == 32).sum() > 50
    assert oracle_mod.same_messages(want, got) == []


@pytest.mark.gpu
def test_gpu_modeac_edges(adsbcu_mod):
    d = adsbcu_mod.Demod1090()
    assert len(d.modeac_block(np.zeros(0, np.uint8), None, 0)) == 0
    assert len(d.modeac_block(np.full(2, 127, np.uint8), None, 0)) == 0
    assert len(d.modeac_block(np.full(2000, 127, np.uint8), np.full(10, 127, np.uint8), 0)) == 0
    with pytest.raises(adsbcu_mod.AdsbcuError):
        d.modeac_block(np.zeros(100, np.uint8), np.zeros(2 * 327, np.uint8), 0)


@pytest.mark.gpu
def test_replay_cli_modeac_matches_reference(tmp_path):
    """host shim + CLI: `adsbcu_replay --path a --fix --modeac --block 262144` == `dump1090 --ifile X --fix --modeac`."""
    import subprocess
    from libadsb_b200 import build
    build.build()
    with np.load(GOLD) as z:
        iq, ref = z["iq"], z["ref_fix1_modeac"]
    cap = tmp_path / "cap.uc8"
    iq.tofile(cap)
    p = subprocess.run([build.REPLAY, "--path", "a", "--fix", "--modeac", "--block", "262144", str(cap)], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    want = ["@%012X%s;" % (int(r["timestamp"]), bytes(r["msg"][:int(r["msgbits"]) // 8]).hex()) for r in ref]
    assert p.stdout.split() == want
