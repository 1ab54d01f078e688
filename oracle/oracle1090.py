"""TEST INFRASTRUCTURE ONLY: ctypes binding of oracle/liboracle1090.so (the CPU
restatement) and a runner for oracle/_ref/dump1090_ref (the reference's own code).
Imported only by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle1090.so")
REF_BIN = os.path.join(HERE, "_ref", "dump1090_ref")

MSG_DTYPE = np.dtype([
    ("timestamp", "<u8"), ("msg", "u1", 14), ("verbatim", "u1", 14), ("msgbits", "u1"),
    ("correctedbits", "u1"), ("msgtype", "u1"), ("score", "u1"), ("crc", "<u4"), ("addr", "<u4"),
    ("signal_level", "<f8")])
assert MSG_DTYPE.itemsize == 56

_lib = None


def build():
    subprocess.run(["make", "-s", "-C", HERE, "all"], check=True)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.oracle1090_create.restype = C.c_void_p
        L.oracle1090_create.argtypes = [C.c_int, C.c_int, C.c_uint]
        L.oracle1090_destroy.argtypes = [C.c_void_p]
        L.oracle1090_set_mode_ac.argtypes = [C.c_void_p, C.c_int]
        L.oracle1090_set_input_format.argtypes = [C.c_void_p, C.c_int]
        L.oracle1090_feed.restype = C.c_size_t
        L.oracle1090_feed.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
        L.oracle1090_mag_table.argtypes = [C.c_void_p]
        L.oracle1090_crc.restype = C.c_uint32
        L.oracle1090_crc.argtypes = [C.c_void_p, C.c_int]
        L.oracle1090_score.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle1090_syndrome_table.restype = C.c_size_t
        L.oracle1090_syndrome_table.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
        L.oracle1090_filter_add.argtypes = [C.c_void_p, C.c_uint32]
        L.oracle1090_filter_test.argtypes = [C.c_void_p, C.c_uint32]
        L.oracle1090_filter_expire.argtypes = [C.c_void_p]
        L.oracle1090_preamble.argtypes = [C.c_void_p]
        L.oracle1090_slice.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        _lib = L
    return _lib


class Oracle1090:
    def __init__(self, nfix_crc=0, fix_df=1, block_samples=0):
        self.h = lib().oracle1090_create(nfix_crc, fix_df, block_samples)
        if not self.h:
            raise MemoryError("oracle1090_create failed")

    def close(self):
        if self.h:
            lib().oracle1090_destroy(self.h)
            self.h = None

    __del__ = close

    def set_input_format(self, fmt: int):
        """0 UC8, 1 SC16, 2 SC16Q11; feed() then takes int16 arrays for 1 and 2."""
        lib().oracle1090_set_input_format(self.h, fmt)
        self._fmt = fmt

    def feed(self, iq: np.ndarray, cap: int | None = None) -> np.ndarray:
        iq = np.ascontiguousarray(iq, dtype="<i2" if getattr(self, "_fmt", 0) else np.uint8)
        n = iq.size // 2
        cap = cap if cap is not None else max(1024, n // 100)
        out = np.zeros(cap, dtype=MSG_DTYPE)
        got = lib().oracle1090_feed(self.h, iq.ctypes.data, n, out.ctypes.data, cap)
        if got > cap:
            raise OverflowError(f"oracle produced {got} messages, cap {cap}")
        return out[:got]

    def score(self, msg: bytes) -> int:
        b = (C.c_uint8 * 14)(*bytes(msg).ljust(14, b"\0"))
        return lib().oracle1090_score(self.h, b)

    def syndrome_table(self, bits: int) -> np.ndarray:
        out = np.zeros((8192, 4), dtype=np.uint32)
        n = lib().oracle1090_syndrome_table(self.h, bits, out.ctypes.data, 8192)
        return out[:n]

    def filter_add(self, a): lib().oracle1090_filter_add(self.h, a)
    def filter_test(self, a): return bool(lib().oracle1090_filter_test(self.h, a))
    def filter_expire(self): lib().oracle1090_filter_expire(self.h)


def mag_table() -> np.ndarray:
    t = np.zeros(65536, dtype=np.uint16)
    lib().oracle1090_mag_table(t.ctypes.data)
    return t


def crc(msg: bytes, bits: int) -> int:
    b = (C.c_uint8 * 14)(*bytes(msg).ljust(14, b"\0"))
    return lib().oracle1090_crc(b, bits)


def have_ref() -> bool:
    return os.path.exists(REF_BIN) and os.access(REF_BIN, os.X_OK)


FIELDS_A_DTYPE = np.dtype([("AC", "<u2"), ("ID", "<u2")] + [(k, "u1") for k in (
    "CA", "CC", "CF", "DR", "FS", "KE", "ND", "RI", "SL", "UM", "VS", "airground", "alert_valid", "alert", "spi_valid", "spi",
    "altitude_baro_valid", "altitude_baro_unit", "squawk_valid", "es_pending")] + [("squawk", "<u2"), ("pad", "<u2"), ("altitude_baro", "<i4")])
assert FIELDS_A_DTYPE.itemsize == 32


def decode_fields(msgs: np.ndarray) -> np.ndarray:
    """oracle1090_decode_fields on every record of a MSG_DTYPE array."""
    L = lib()
    L.oracle1090_decode_fields.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    out = np.zeros(len(msgs), FIELDS_A_DTYPE)
    for k in range(len(msgs)):
        m = np.ascontiguousarray(msgs["msg"][k])
        L.oracle1090_decode_fields(m.ctypes.data, int(msgs["msgtype"][k]), out[k:k + 1].ctypes.data)
    return out


FIELDS_ES_DTYPE = np.dtype([("addr", "<u4"), ("valid", "<u4")] + [(k, "u1") for k in (
    "es", "metype", "mesub", "addrtype", "source", "reliable", "airground", "CA_CF", "cpr_type", "cpr_odd", "heading_type", "nav_heading_type",
    "alert", "spi", "emergency", "category", "altitude_baro_unit", "altitude_geom_unit", "nav_altitude_source", "nav_modes")] +
    [("callsign", "S8")] + [(k, "<i4") for k in ("altitude_baro", "altitude_geom", "geom_delta", "baro_rate", "geom_rate")] +
    [("cpr_lat", "<u4"), ("cpr_lon", "<u4"), ("ias", "<u2"), ("tas", "<u2"), ("squawk", "<u2"), ("ops_flags", "<u2")] +
    [(k, "<f4") for k in ("heading", "gs_v0", "gs_v2", "gs_selected", "nav_heading", "nav_qnh")] +
    [("nav_fms_altitude", "<i4"), ("nav_mcp_altitude", "<i4")] + [(k, "u1") for k in (
        "nic_a", "nic_b", "nic_c", "nic_baro", "nac_p", "nac_v", "sil", "sil_type", "gva", "sda",
        "ops_version", "ops_tah", "ops_hrd", "ops_cc_tc", "ops_cc_lw", "ops_cc_antenna_offset")])
assert FIELDS_ES_DTYPE.itemsize == 120


def decode_es(frames: np.ndarray, corrected=None) -> np.ndarray:
    """oracle1090_decode_es on n corrected 14-byte frames (uint8 [n, 14]); corrected: bits repaired per frame."""
    frames = np.ascontiguousarray(frames, np.uint8).reshape(-1, 14)
    L = lib()
    L.oracle1090_decode_es_batch.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
    L.oracle1090_decode_es_batch.restype = None
    out = np.zeros(len(frames), FIELDS_ES_DTYPE)
    cor = None if corrected is None else np.ascontiguousarray(corrected, np.uint8)
    L.oracle1090_decode_es_batch(frames.ctypes.data, len(frames), None if cor is None else cor.ctypes.data, out.ctypes.data)
    return out


FIELDS_MB_DTYPE = np.dtype([("valid", "<u4")] + [(k, "u1") for k in (
    "commb_format", "heading_type", "nav_modes", "nav_altitude_source", "mrar_source", "turbulence")] + [("reserved", "u1", 2), ("callsign", "S8")] +
    [(k, "<f4") for k in ("roll", "heading", "track_rate", "gs", "mach", "nav_qnh", "wind_speed", "wind_dir", "temperature", "pressure", "humidity")] +
    [(k, "<i4") for k in ("baro_rate", "geom_rate", "nav_mcp_altitude", "nav_fms_altitude")] + [("ias", "<u2"), ("tas", "<u2")])
assert FIELDS_MB_DTYPE.itemsize == 84


def decode_commb(frames: np.ndarray, corrected=None) -> np.ndarray:
    """oracle1090_decode_commb on n corrected 14-byte frames (uint8 [n, 14])."""
    frames = np.ascontiguousarray(frames, np.uint8).reshape(-1, 14)
    L = lib()
    L.oracle1090_decode_commb_batch.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
    L.oracle1090_decode_commb_batch.restype = None
    out = np.zeros(len(frames), FIELDS_MB_DTYPE)
    cor = None if corrected is None else np.ascontiguousarray(corrected, np.uint8)
    L.oracle1090_decode_commb_batch(frames.ctypes.data, len(frames), None if cor is None else cor.ctypes.data, out.ctypes.data)
    return out


def run_ref_decode(frames: np.ndarray, commb=False, header=False):
    """The reference's own decodeModesMessage on 14-byte frames as received (third tap of ref_harness.c): the ES
    record of each frame (commb=True: the Comm-B record; header=True: the surveillance-field record) and
    decodeModesMessage's return code."""
    frames = np.ascontiguousarray(frames, np.uint8).reshape(-1, 14)
    rec = np.dtype([("es", FIELDS_ES_DTYPE), ("mb", FIELDS_MB_DTYPE), ("hdr", FIELDS_A_DTYPE), ("rc", "<i4")])
    tmpdir = "/dev/shm" if os.path.isdir("/dev/shm") else None
    with tempfile.TemporaryDirectory(dir=tmpdir) as d:
        fin, fout = os.path.join(d, "frames.bin"), os.path.join(d, "es.bin")
        frames.tofile(fin)
        subprocess.run([REF_BIN], env=dict(os.environ, REF1090_DECODE_IN=fin, REF1090_DECODE_OUT=fout), check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        r = np.fromfile(fout, dtype=rec)
    assert len(r) == len(frames)
    return r["hdr" if header else "mb" if commb else "es"].copy(), r["rc"].copy()


def run_ref_fields(iq: np.ndarray, nfix_crc=0, fix_df=1):
    """The reference's accepted messages AND the surveillance fields its decodeModesMessage filled in for them."""
    flags = {0: [], 1: ["--fix"], 2: ["--fix-2bit"]}[nfix_crc] + ([] if fix_df else ["--no-fix-df"])
    tmpdir = "/dev/shm" if os.path.isdir("/dev/shm") else None
    with tempfile.TemporaryDirectory(dir=tmpdir) as d:
        cap, rec, fld = os.path.join(d, "cap.uc8"), os.path.join(d, "out.rec"), os.path.join(d, "out.fields")
        np.ascontiguousarray(iq).tofile(cap)
        env = dict(os.environ, REF1090_OUT=rec, REF1090_FIELDS=fld)
        subprocess.run([REF_BIN, "--ifile", cap, "--quiet", *flags], env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=False)
        return np.fromfile(rec, dtype=MSG_DTYPE), np.fromfile(fld, dtype=FIELDS_A_DTYPE)


def run_ref(iq: np.ndarray, nfix_crc=0, fix_df=1, extra=()) -> np.ndarray:
    """Run the reference's own dump1090 (--ifile) on `iq`; returns its accepted messages."""
    flags = {0: [], 1: ["--fix"], 2: ["--fix-2bit"]}[nfix_crc] + ([] if fix_df else ["--no-fix-df"])
    tmpdir = "/dev/shm" if os.path.isdir("/dev/shm") else None
    with tempfile.TemporaryDirectory(dir=tmpdir) as d:
        cap, rec = os.path.join(d, "cap.uc8"), os.path.join(d, "out.rec")
        np.ascontiguousarray(iq).tofile(cap)     # uint8 (UC8) or int16 (with --iformat SC16 / SC16Q11 in `extra`)
        env = dict(os.environ, REF1090_OUT=rec)
        subprocess.run([REF_BIN, "--ifile", cap, "--quiet", *flags, *extra], env=env,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=False)
        if not os.path.exists(rec):
            return np.zeros(0, dtype=MSG_DTYPE)
        return np.fromfile(rec, dtype=MSG_DTYPE)


def same_messages(a: np.ndarray, b: np.ndarray, check_signal=True) -> list[str]:
    """Compare two message arrays on the parity tuple; returns a list of differences."""
    diffs = []
    if len(a) != len(b):
        diffs.append(f"count {len(a)} != {len(b)}")
    for k in range(min(len(a), len(b))):
        x, y = a[k], b[k]
        nb = int(x["msgbits"]) // 8
        for f in ("timestamp", "msgbits", "correctedbits", "msgtype", "score", "crc", "addr"):
            if x[f] != y[f]:
                diffs.append(f"#{k} {f}: {x[f]} != {y[f]}")
        if bytes(x["msg"][:nb]) != bytes(y["msg"][:nb]):
            diffs.append(f"#{k} msg differs")
        if bytes(x["verbatim"][:nb]) != bytes(y["verbatim"][:nb]):
            diffs.append(f"#{k} verbatim differs")
        if check_signal and x["signal_level"] != y["signal_level"]:
            diffs.append(f"#{k} signal_level {x['signal_level']} != {y['signal_level']}")
        if len(diffs) > 20:
            break
    return diffs
