"""Synthetic uint8 I/Q capture generator (RTL-SDR "UC8" format) for the 1090 MHz path.

The reference vendors no raw captures (its test inputs are downloaded at CMake
configure time, CMakeLists.txt:141; tests/test_1090.cpp:127-146), so every
input used by tests/ and bench.py is produced here from a fixed seed
(SURVEY.md section 8d, configs 2 and 3).

Signal model: Mode S pulse-position modulation sampled at 2.4 MS/s.  A frame is
an 8 us preamble (0.5 us pulses at 0, 1.0, 3.5, 4.5 us) followed by 56 or 112
data bits of 1 us each (bit 1 = pulse in the first half, bit 0 = pulse in the
second half).  Each sample integrates the pulse envelope over its 416.7 ns
window; the carrier has a random phase; additive white Gaussian noise with
standard deviation `sigma` LSB is added to I and Q around the RTL-SDR DC point
127.4; the result is rounded and clipped to [0, 255].
"""
from __future__ import annotations

import numpy as np

FS_1090 = 2_400_000
SAMPLES_PER_US = 2.4
_POLY = 0xFFF409


def _crc_table():
    t = np.zeros(256, dtype=np.uint32)
    for i in range(256):
        c = i << 16
        for _ in range(8):
            c = ((c << 1) ^ _POLY) if (c & 0x800000) else (c << 1)
        t[i] = c & 0xFFFFFF
    return t


_CRC_TABLE = _crc_table()


def modes_crc(data: bytes) -> int:
    """CRC-24 remainder of `data` (generator 0xFFF409), MSB first."""
    rem = 0
    for b in data:
        rem = ((rem << 8) & 0xFFFFFF) ^ int(_CRC_TABLE[b ^ ((rem >> 16) & 0xFF)])
    return rem


def make_frame(df: int, icao: int, rng: np.random.Generator, iid: int = 0) -> bytes:
    """Build one valid Mode S frame (bytes) for the given downlink format.

    DF11/17/18 carry the address in bytes 1..3 and parity (xor IID for DF11)
    in the last three bytes; the Address/Parity formats DF0/4/5/16/20/21
    overlay the address on the parity field.
    """
    n = 14 if (df & 16) else 7
    body = bytearray(rng.integers(0, 256, size=n - 3, dtype=np.uint8).tobytes())
    body[0] = ((df & 31) << 3) | (body[0] & 7)
    if df in (11, 17, 18):
        body[1] = (icao >> 16) & 0xFF
        body[2] = (icao >> 8) & 0xFF
        body[3] = icao & 0xFF
        par = modes_crc(bytes(body))
        if df == 11:
            par ^= iid & 0x7F
    else:
        par = modes_crc(bytes(body)) ^ (icao & 0xFFFFFF)
    return bytes(body) + bytes([(par >> 16) & 0xFF, (par >> 8) & 0xFF, par & 0xFF])


def flip_bits(frame: bytes, bits) -> bytes:
    b = bytearray(frame)
    for k in bits:
        b[k >> 3] ^= 0x80 >> (k & 7)
    return bytes(b)


def _pulse_starts_us(frame: bytes) -> np.ndarray:
    bits = np.unpackbits(np.frombuffer(frame, dtype=np.uint8))
    data = 8.0 + np.arange(bits.size) + np.where(bits == 1, 0.0, 0.5)
    return np.concatenate([np.array([0.0, 1.0, 3.5, 4.5]), data])


AC_BITS = {"F1": 0, "C1": 1, "A1": 2, "C2": 3, "A2": 4, "C4": 5, "A4": 6, "B1": 8, "D1": 9, "B2": 10, "D2": 11, "B4": 12,
           "D4": 13, "F2": 14, "SPI": 17}   # pulse slots of a Mode A/C reply, 1.45 us apart (demod_2400.c:556-577)


def modeac_pulses_us(code: int) -> np.ndarray:
    """Pulse start times (us) of the Mode A/C reply whose 13 data bits are `code` in the reference's layout
    00 A4 A2 A1  00 B4 B2 B1  SPI C4 C2 C1  00 D4 D2 D1 (demod_2400.c:690-705)."""
    m = {"C1": 0x0010, "A1": 0x1000, "C2": 0x0020, "A2": 0x2000, "C4": 0x0040, "A4": 0x4000, "B1": 0x0100, "D1": 0x0001,
         "B2": 0x0200, "D2": 0x0002, "B4": 0x0400, "D4": 0x0004, "SPI": 0x0080}
    slots = [0, 14] + [AC_BITS[k] for k, v in m.items() if code & v]
    return 1.45 * np.array(sorted(slots), dtype=np.float64)


def synth_capture(n_samples: int,
                  frames,
                  starts,
                  amps,
                  carrier_phase,
                  sigma: float,
                  rng: np.random.Generator,
                  dc: float = 127.4,
                  samples_per_us: float = SAMPLES_PER_US) -> np.ndarray:
    """Render `frames` (list of bytes; an int item is a Mode A/C reply code, 0.45 us pulses) into an
    interleaved uint8 I/Q array.

    starts: message start times in (fractional) samples; amps: peak amplitude
    in LSB; carrier_phase: radians.  Returns uint8[2*n_samples].
    """
    sig = np.zeros(n_samples + 8, dtype=np.complex64)
    if len(frames):
        idx_all, val_all = [], []
        for f, t0, a, th in zip(frames, starts, amps, carrier_phase):
            is_ac = isinstance(f, (int, np.integer))
            ps = t0 + (modeac_pulses_us(int(f)) if is_ac else _pulse_starts_us(f)) * samples_per_us   # pulse start, samples
            pe = ps + (0.45 if is_ac else 0.5) * samples_per_us
            base = np.floor(ps).astype(np.int64)
            c = a * np.exp(1j * th)
            for k in range(3):                                       # a 1.2-sample pulse spans <= 3 samples
                lo = np.maximum(ps, base + k)
                hi = np.minimum(pe, base + k + 1)
                ov = np.clip(hi - lo, 0.0, None)
                idx_all.append(base + k)
                val_all.append(ov * c)
        idx = np.concatenate(idx_all)
        val = np.concatenate(val_all)
        ok = (idx >= 0) & (idx < n_samples)
        idx, val = idx[ok], val[ok]
        sig[:n_samples] += (np.bincount(idx, weights=val.real, minlength=n_samples)[:n_samples]
                            + 1j * np.bincount(idx, weights=val.imag, minlength=n_samples)[:n_samples]
                            ).astype(np.complex64)
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


def config2(seconds: float = 60.0, seed: int = 0x10900002, msgs_per_s: float = 1000.0,
            sigma: float = 2.0, pool: int = 512, snr_db=(6.0, 30.0)):
    """non-overlapping DF17 (70 %) / DF11 IID0 (30 %)
    frames at 6..30 dB SNR over sigma=2 LSB noise.  Returns (iq uint8, frames, starts)."""
    rng = np.random.default_rng(seed)
    n = int(round(seconds * FS_1090))
    m = int(round(seconds * msgs_per_s))
    icaos = rng.integers(1, 1 << 24, size=pool)
    slot = n / max(m, 1)
    frames, starts = [], []
    for k in range(m):
        df = 17 if rng.random() < 0.7 else 11
        frames.append(make_frame(df, int(icaos[rng.integers(0, pool)]), rng))
        # one frame per slot, uniformly placed so that frames never overlap
        room = slot - 300.0
        starts.append(k * slot + (rng.random() * room if room > 0 else 0.0))
    starts = np.array(starts)
    snr = rng.uniform(snr_db[0], snr_db[1], size=m)
    amps = np.minimum(sigma * np.sqrt(2.0) * 10.0 ** (snr / 20.0), 90.0)
    ph = rng.uniform(0, 2 * np.pi, size=m)
    iq = synth_capture(n, frames, starts, amps, ph, sigma, rng)
    return iq, frames, starts


def config3(seconds: float = 1.0, seed: int = 0x10900003, msgs_per_s: float = 2000.0,
            sigma: float = 2.0, pool: int = 2000):
    """Poisson arrivals (overlaps allowed), mixed DFs,
    10 % frames with one flipped bit and 5 % with two."""
    rng = np.random.default_rng(seed)
    n = int(round(seconds * FS_1090))
    m = int(rng.poisson(seconds * msgs_per_s))
    icaos = rng.integers(1, 1 << 24, size=pool)
    dfs = rng.choice([17, 11, 4, 5, 20, 21, 18, 0, 16], size=m,
                     p=[0.40, 0.15, 0.10, 0.08, 0.08, 0.07, 0.06, 0.03, 0.03])
    frames = []
    for k in range(m):
        f = make_frame(int(dfs[k]), int(icaos[rng.integers(0, pool)]), rng,
                       iid=int(rng.integers(0, 16)) if rng.random() < 0.2 else 0)
        u = rng.random()
        nb = len(f) * 8
        if u < 0.10:
            f = flip_bits(f, [int(rng.integers(0, nb))])
        elif u < 0.15:
            f = flip_bits(f, [int(x) for x in rng.choice(nb, size=2, replace=False)])
        frames.append(f)
    starts = np.sort(rng.uniform(0, max(n - 400, 1), size=m))
    snr = rng.uniform(6.0, 30.0, size=m)
    amps = np.minimum(sigma * np.sqrt(2.0) * 10.0 ** (snr / 20.0), 90.0)
    ph = rng.uniform(0, 2 * np.pi, size=m)
    iq = synth_capture(n, frames, starts, amps, ph, sigma, rng)
    return iq, frames, starts


FS_1090B = 2_000_000


def config_b(seconds: float = 1.0, seed: int = 0x1090B002, msgs_per_s: float = 1000.0, sigma: float = 1.5,
             pool: int = 300, flip_frac: float = 0.15, ap_frac: float = 0.25):
    """Capture for the 2.0 MS/s path libadsb compiles (ADSB1090.cpp): sample-aligned and
    half-sample-offset DF17/DF11 frames (the latter exercise the phase-correction retry), a
    share of single-bit errors (FixSingleBitErrors) and Address/Parity replies (BruteForceAp)
    from aircraft of the same pool.  DC point 127 as the handler assumes (ADSB1090.cpp:165-166)."""
    rng = np.random.default_rng(seed)
    n = int(round(seconds * FS_1090B))
    m = int(round(seconds * msgs_per_s))
    icaos = rng.integers(1, 1 << 24, size=pool)
    slot = n / max(m, 1)
    frames, starts = [], []
    for k in range(m):
        u = rng.random()
        if u < ap_frac:
            df = int(rng.choice([0, 4, 5, 16, 20, 21]))
        else:
            df = 17 if rng.random() < 0.7 else 11
        f = make_frame(df, int(icaos[rng.integers(0, pool)]), rng)
        if rng.random() < flip_frac:
            f = flip_bits(f, [int(rng.integers(0, len(f) * 8))])
        frames.append(f)
        room = slot - 260.0
        t0 = k * slot + (rng.random() * room if room > 0 else 0.0)
        frac = rng.choice([0.0, 0.0, 0.5, rng.random()])           # aligned, half-sample, arbitrary
        starts.append(np.floor(t0) + frac)
    starts = np.array(starts)
    amps = rng.uniform(25.0, 100.0, size=m)
    ph = rng.uniform(0, 2 * np.pi, size=m)
    iq = synth_capture(n, frames, starts, amps, ph, sigma, rng, dc=127.0, samples_per_us=2.0)
    return iq, frames, starts


def config_ac(seconds: float = 1.0, seed: int = 0x1090AC01, msgs_per_s: float = 1500.0, ac_per_s: float = 3000.0,
              sigma: float = 2.0, pool: int = 300):
    """Mode S traffic (as config 2) plus Mode A/C replies (random 12-bit codes, some with SPI), for --modeac
    (demod_2400.c:539-726).  Returns (iq, frames, starts); a Mode A/C item of `frames` is its int code."""
    rng = np.random.default_rng(seed)
    n = int(round(seconds * FS_1090))
    m, k = int(round(seconds * msgs_per_s)), int(round(seconds * ac_per_s))
    icaos = rng.integers(1, 1 << 24, size=pool)
    frames = [make_frame(17 if rng.random() < 0.7 else 11, int(icaos[rng.integers(0, pool)]), rng) for _ in range(m)]
    for _ in range(k):
        code = int(rng.integers(0, 1 << 12))
        code = ((code & 0x007) << 0) | ((code & 0x038) << 1) | ((code & 0x1C0) << 2) | ((code & 0xE00) << 3)   # D, C, B, A octal digits
        if rng.random() < 0.1:
            code |= 0x0080
        frames.append(code)
    order = rng.permutation(len(frames))
    frames = [frames[i] for i in order]
    starts = np.sort(rng.uniform(400.0, n - 400.0, size=len(frames)))
    snr = rng.uniform(10.0, 30.0, size=len(frames))
    amps = sigma * np.sqrt(2.0) * 10.0 ** (snr / 20.0)
    phases = rng.uniform(0, 2 * np.pi, size=len(frames))
    iq = synth_capture(n, frames, starts, amps, phases, sigma, rng)
    return iq, frames, starts


def to_sc16(iq_uc8: np.ndarray, q11: bool = False) -> np.ndarray:
    """The same capture as interleaved int16 I/Q (dump1090 --iformat SC16 / SC16Q11): 8-bit samples around the
    127 DC point scaled to the 16-bit (or 12-bit Q11) range, with a little sub-LSB dither so that the converter's
    float rounding is exercised on more than 256 distinct values per axis."""
    rng = np.random.default_rng(int(iq_uc8[:64].sum()) + int(q11))
    x = (iq_uc8.astype(np.int32) - 127) * (16 if q11 else 256) + rng.integers(-7, 8, size=iq_uc8.size)
    lim = 2047 if q11 else 32767
    return np.clip(x, -lim - 1, lim).astype("<i2")
