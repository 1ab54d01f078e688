"""TEST INFRASTRUCTURE ONLY: ctypes bindings of oracle/liboracle1090b.so (CPU restatement of
libadsb's own 2.0 MS/s handler, ADSB1090.cpp) and of oracle/_ref/libref1090b.so (the reference's
ADSB1090.cpp itself, compiled in place).  Imported only by tests/, smoke() and bench.py's CPU legs."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle1090b.so")
REF_PATH = os.path.join(HERE, "_ref", "libref1090b.so")

MSG_DTYPE = np.dtype([("pos", "<u4"), ("crc", "<u4"), ("addr", "<u4"), ("msg", "u1", 14), ("msgbits", "u1"),
                      ("msgtype", "u1"), ("errorbit", "i1"), ("corrected", "u1"), ("pad", "<u2")])
assert MSG_DTYPE.itemsize == 32
CB_DTYPE = np.dtype([("addr", "<u4"), ("altitude", "<i4"), ("speed", "<u4"), ("track", "<u4"), ("lat1e7", "<i4"),
                     ("lon1e7", "<i4"), ("callsign", "S8")])
assert CB_DTYPE.itemsize == 32
STAT_NAMES = ["valid_preamble", "demodulated", "goodcrc", "badcrc", "fixed", "single_bit_fix", "out_of_phase"]


def have_ref() -> bool:
    return os.path.exists(REF_PATH)


def _bind(path, prefix):
    L = C.CDLL(path)
    f = lambda n: getattr(L, prefix + n)
    f("create").restype = C.c_void_p
    f("destroy").argtypes = [C.c_void_p]
    f("handle_data").restype = C.c_size_t
    f("handle_data").argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
    f("stats").argtypes = [C.c_void_p, C.c_void_p]
    f("lut").argtypes = [C.c_void_p, C.c_void_p]
    f("checksum").restype = C.c_uint32
    f("checksum").argtypes = [C.c_void_p, C.c_int]
    f("fix_single").argtypes = [C.c_void_p, C.c_int]
    f("decode").argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    f("fields").argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    f("cpr").argtypes = [C.c_int] * 5 + [C.c_void_p]
    f("cache").restype = C.c_size_t
    f("cache").argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    return L

# == adsbcu_fields_b (include/adsbcu.h)
FIELDS_DTYPE = np.dtype([("ca", "<i4"), ("metype", "<i4"), ("mesub", "<i4"), ("heading", "<i4"), ("fflag", "<i4"),
                         ("raw_latitude", "<i4"), ("raw_longitude", "<i4"), ("flight", "S8"), ("ew_dir", "<i4"),
                         ("ew_velocity", "<i4"), ("ns_dir", "<i4"), ("ns_velocity", "<i4"), ("velocity", "<i4"), ("fs", "<i4"),
                         ("dr", "<i4"), ("um", "<i4"), ("identity", "<i4"), ("altitude", "<i4"), ("unit", "<i4")])
assert FIELDS_DTYPE.itemsize == 80


class _Base:
    prefix = ""
    rec_dtype = MSG_DTYPE

    def __init__(self, lib):
        self.L = lib
        self.h = getattr(lib, self.prefix + "create")()

    def _f(self, n):
        return getattr(self.L, self.prefix + n)

    def handle_data(self, iq: np.ndarray, cap: int = 1 << 16) -> np.ndarray:
        iq = np.ascontiguousarray(iq, dtype=np.uint8)
        out = np.zeros(cap, dtype=self.rec_dtype)
        n = self._f("handle_data")(self.h, iq.ctypes.data, iq.size, out.ctypes.data, cap)
        assert n <= cap
        return out[:n]

    def stats(self) -> dict:
        v = np.zeros(7, dtype=np.uint64)
        self._f("stats")(self.h, v.ctypes.data)
        return dict(zip(STAT_NAMES, (int(x) for x in v)))

    def lut(self) -> np.ndarray:
        v = np.zeros(129 * 129, dtype=np.uint16)
        self._f("lut")(self.h, v.ctypes.data)
        return v

    def checksum(self, msg: bytes, bits: int) -> int:
        b = np.frombuffer(bytes(msg).ljust(14, b"\0"), dtype=np.uint8).copy()
        return int(self._f("checksum")(b.ctypes.data, bits))

    def fix_single(self, msg: bytes, bits: int):
        b = np.frombuffer(bytes(msg).ljust(14, b"\0"), dtype=np.uint8).copy()
        r = self._f("fix_single")(b.ctypes.data, bits)
        return int(r), bytes(b)

    def decode(self, msg: bytes):
        b = np.frombuffer(bytes(msg).ljust(14, b"\0"), dtype=np.uint8).copy()
        o = np.zeros(6, dtype=np.int32)
        self._f("decode")(self.h, b.ctypes.data, o.ctypes.data)
        return dict(msg=bytes(b), msgbits=int(o[0]), msgtype=int(o[1]), crcok=int(o[2]), errorbit=int(o[3]),
                    crc=int(o[4]) & 0xFFFFFF, addr=int(o[5]) & 0xFFFFFF)

    def fields(self, msg: bytes) -> np.ndarray:
        """Every Message field DecodeModesMessage fills beyond the CRC stage, as one FIELDS_DTYPE record."""
        b = np.frombuffer(bytes(msg).ljust(14, b"\0"), dtype=np.uint8).copy()
        o = np.zeros(1, dtype=FIELDS_DTYPE)
        self._f("fields")(self.h, b.ctypes.data, o.ctypes.data)
        return o[0]

    def cpr(self, even_lat, even_lon, odd_lat, odd_lon, even_is_newer):
        """DecodeCpr on one even/odd pair -> (updated, lat1E7, lon1E7)."""
        o = np.zeros(2, dtype=np.int32)
        ok = self._f("cpr")(int(even_lat), int(even_lon), int(odd_lat), int(odd_lon), int(bool(even_is_newer)), o.ctypes.data)
        return int(ok), int(o[0]), int(o[1])

    def fields_bulk(self, msgs14: np.ndarray, msgtype: np.ndarray) -> np.ndarray:
        """Fields of frames that are already past the CRC stage (msgtype = the DF as sliced)."""
        msgs14 = np.ascontiguousarray(msgs14, dtype=np.uint8).reshape(-1, 14)
        msgtype = np.ascontiguousarray(msgtype, dtype=np.uint8)
        out = np.zeros(msgs14.shape[0], dtype=FIELDS_DTYPE)
        f = self.L.oracle1090b_fields_bulk
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        f(msgs14.ctypes.data, msgtype.ctypes.data, msgs14.shape[0], out.ctypes.data)
        return out

    def cpr_bulk(self, pairs5: np.ndarray) -> np.ndarray:
        pairs5 = np.ascontiguousarray(pairs5, dtype=np.int32).reshape(-1, 5)
        out = np.zeros((pairs5.shape[0], 3), dtype=np.int32)
        f = self.L.oracle1090b_cpr_bulk
        f.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        f(pairs5.ctypes.data, pairs5.shape[0], out.ctypes.data)
        return out

    def cache(self) -> set:
        v = np.zeros(1 << 16, dtype=np.uint32)
        n = self._f("cache")(self.h, v.ctypes.data, v.size)
        return set(int(x) for x in v[:n])


_lib = _ref = None


class Oracle1090B(_Base):
    prefix = "oracle1090b_"

    def __init__(self):
        global _lib
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                subprocess.run(["make", "-s", "-C", HERE, "liboracle1090b.so"], check=True)
            _lib = _bind(LIB_PATH, self.prefix)
        super().__init__(_lib)


class Ref1090B(_Base):
    """The reference's own ADSB1090Handler (records are IListener::OnChanged callbacks)."""
    prefix = "ref1090b_"
    rec_dtype = CB_DTYPE

    def __init__(self):
        global _ref
        if _ref is None:
            _ref = _bind(REF_PATH, self.prefix)
        super().__init__(_ref)
