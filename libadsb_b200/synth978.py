"""Synthetic 978 MHz UAT captures (uint8 I/Q at 2.083334 MS/s, RTL-SDR "UC8") for tests and bench.

The reference vendors no UAT capture (tests/test_1090.cpp:127-146 reads them from an environment
directory), so inputs are produced here from fixed seeds.

Signal model (RTCA DO-282B): continuous-phase FSK, 1.041667 Mbit/s, modulation index 0.6 (a one
advances the phase by +0.6 pi per bit, a zero by -0.6 pi), 2 samples per bit.  A frame is a 36-bit
sync word (ADS-B 0xEACDDA4E2, ground uplink 0x153225B1D) followed by
  * basic ADS-B: 18 payload bytes + 12 RS parity bytes          RS(30,18)
  * long ADS-B:  34 payload bytes + 14 RS parity bytes          RS(48,34)
  * uplink:      6 blocks of 72 payload + 20 parity bytes, byte-interleaved (byte i of block b is
                 transmitted as byte 6 i + b)                                  6 x RS(92,72)
all over GF(256) with field polynomial 0x187 and generator roots alpha^120 ...  The encoder below is
an independent implementation (polynomial division), not shared with the oracle or the CUDA decoder.
"""
from __future__ import annotations

import numpy as np

FS_978 = 2_083_334
ADSB_SYNC = 0xEACDDA4E2
UPLINK_SYNC = 0x153225B1D

_EXP = np.zeros(512, dtype=np.int64)
_LOG = np.zeros(256, dtype=np.int64)
_x = 1
for _i in range(255):
    _EXP[_i] = _x
    _LOG[_x] = _i
    _x <<= 1
    if _x & 0x100:
        _x ^= 0x187
_EXP[255:510] = _EXP[0:255]


def _gmul(a: int, b: int) -> int:
    return 0 if a == 0 or b == 0 else int(_EXP[_LOG[a] + _LOG[b]])


def _genpoly(nroots: int, fcr: int = 120):
    g = [1]
    for r in range(fcr, fcr + nroots):
        a = int(_EXP[r % 255])
        ng = [0] * (len(g) + 1)
        for i, c in enumerate(g):          # g(x) * (x + a), coefficients highest degree first
            ng[i] ^= c
            ng[i + 1] ^= _gmul(c, a)
        g = ng
    return g


_GEN = {n: _genpoly(n) for n in (12, 14, 20)}


def rs_parity(data: bytes, nroots: int) -> bytes:
    """Remainder of data(x) * x^nroots modulo the generator polynomial (systematic RS parity)."""
    g = _GEN[nroots]
    rem = [0] * nroots
    for d in data:
        fb = d ^ rem[0]
        rem = rem[1:] + [0]
        if fb:
            for j in range(nroots):
                rem[j] ^= _gmul(fb, g[j + 1])
    return bytes(rem)


def adsb_frame(payload: bytes) -> bytes:
    """18-byte (basic, payload type 0) or 34-byte (long, payload type 1..) payload -> transmitted frame bytes."""
    assert len(payload) in (18, 34)
    return bytes(payload) + rs_parity(payload, 12 if len(payload) == 18 else 14)


def uplink_frame(payload: bytes) -> bytes:
    assert len(payload) == 432
    blocks = [bytes(payload[72 * b:72 * b + 72]) for b in range(6)]
    blocks = [blk + rs_parity(blk, 20) for blk in blocks]
    out = bytearray(552)
    for b in range(6):
        for i in range(92):
            out[6 * i + b] = blocks[b][i]
    return bytes(out)


def frame_bits(sync: int, frame: bytes) -> np.ndarray:
    s = np.array([(sync >> (35 - i)) & 1 for i in range(36)], dtype=np.int8)
    return np.concatenate([s, np.unpackbits(np.frombuffer(frame, dtype=np.uint8)).astype(np.int8)])


def corrupt(frame: bytes, nerr: int, rng: np.random.Generator, lo: int = 0, hi: int | None = None) -> bytes:
    """Replace `nerr` distinct bytes in [lo, hi) by different values."""
    b = bytearray(frame)
    hi = len(b) if hi is None else hi
    for k in rng.choice(np.arange(lo, hi), size=nerr, replace=False):
        b[int(k)] ^= int(rng.integers(1, 256))
    return bytes(b)


def synth_capture(n_samples: int, bit_arrays, starts, amps, sigma: float, rng: np.random.Generator, dc: float = 127.5) -> np.ndarray:
    """Render CPFSK bursts: bit_arrays[k] starts at (fractional) sample starts[k] with amplitude amps[k] LSB."""
    sig = np.zeros(n_samples, dtype=np.complex64)
    for bits, t0, a in zip(bit_arrays, starts, amps):
        nb = len(bits)
        first = int(np.ceil(t0))
        last = min(n_samples, int(np.floor(t0 + 2 * nb)))
        if last <= first:
            continue
        tau = (np.arange(first, last) - t0) / 2.0                     # time in bit periods since the burst began
        k = np.minimum(tau.astype(np.int64), nb - 1)
        nrz = 2.0 * bits.astype(np.float64) - 1.0
        cum = np.concatenate([[0.0], np.cumsum(nrz)])
        phase = 0.6 * np.pi * (cum[k] + (tau - k) * nrz[k]) + rng.uniform(0, 2 * np.pi)
        sig[first:last] += (a * np.exp(1j * phase)).astype(np.complex64)
    out = np.empty(2 * n_samples, dtype=np.uint8)
    chunk = 1 << 22
    for s in range(0, n_samples, chunk):
        e = min(n_samples, s + chunk)
        nz = rng.standard_normal((e - s, 2), dtype=np.float32) * np.float32(sigma)
        nz[:, 0] += sig[s:e].real + np.float32(dc)
        nz[:, 1] += sig[s:e].imag + np.float32(dc)
        np.rint(nz, out=nz)
        np.clip(nz, 0, 255, out=nz)
        out[2 * s:2 * e] = nz.astype(np.uint8).reshape(-1)
    return out


def config5(seconds: float = 60.0, seed: int = 0x09780005, adsb_per_s: float = 200.0, uplink_per_s: float = 1.0,
            sigma: float = 2.0, max_err_frac: float = 1.0):
    """200 ADS-B frames/s (80 % long, 20 % basic) + 1 ground uplink frame/s, each with
    0..t corrupted bytes (t = 6 / 7 / 10 per block), amplitudes 12..60 LSB over sigma = 2 noise.
    Returns (iq, sent) where sent is a list of (start_sample, kind, payload, n_corrupted)."""
    rng = np.random.default_rng(seed)
    n = int(round(seconds * FS_978))
    m = int(round(seconds * (adsb_per_s + uplink_per_s)))
    slot = n / max(m, 1)
    bit_arrays, starts, amps, sent = [], [], [], []
    up_every = max(1, int(round((adsb_per_s + uplink_per_s) / max(uplink_per_s, 1e-9))))
    for k in range(m):
        if uplink_per_s > 0 and k % up_every == up_every // 2 and slot > 9100:
            payload = rng.integers(0, 256, size=432, dtype=np.uint8).tobytes()
            frame = bytearray(uplink_frame(payload))
            nerr = 0
            for b in range(6):                                        # up to 10 byte errors in every block
                e = int(rng.integers(0, int(10 * max_err_frac) + 1))
                for i in rng.choice(92, size=e, replace=False):
                    frame[6 * int(i) + b] ^= int(rng.integers(1, 256))
                nerr += e
            bits = frame_bits(UPLINK_SYNC, bytes(frame))
            kind = "+"
        else:
            long_ = rng.random() < 0.8
            payload = bytearray(rng.integers(0, 256, size=34 if long_ else 18, dtype=np.uint8).tobytes())
            payload[0] = (payload[0] & 7) | ((int(rng.integers(1, 11)) if long_ else 0) << 3)
            payload = bytes(payload)
            nerr = int(rng.integers(0, int((7 if long_ else 6) * max_err_frac) + 1))
            bits = frame_bits(ADSB_SYNC, corrupt(adsb_frame(payload), nerr, rng))
            kind = "-"
        room = slot - 2 * len(bits) - 8
        t0 = k * slot + (rng.random() * room if room > 0 else 0.0)
        bit_arrays.append(bits); starts.append(t0); amps.append(rng.uniform(12.0, 60.0))
        sent.append((t0, kind, payload, nerr))
    iq = synth_capture(n, bit_arrays, np.array(starts), np.array(amps), sigma, rng)
    return iq, sent
