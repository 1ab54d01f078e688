"""TEST INFRASTRUCTURE ONLY: ctypes binding of oracle/liboracle978.so (CPU restatement of the UAT path;
PARITY UNPINNED, see the header of oracle978.c).  Imported only by tests/, smoke() and bench.py's CPU legs."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle978.so")
FRAME_DTYPE = np.dtype([("offset", "<u8"), ("rs_errors", "<i2"), ("len", "<u2"), ("updown", "u1"), ("pad", "u1", 3),
                        ("data", "u1", 432)])
assert FRAME_DTYPE.itemsize == 448
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            subprocess.run(["make", "-s", "-C", HERE, "liboracle978.so"], check=True)
        L = C.CDLL(LIB_PATH)
        L.oracle978_create.restype = C.c_void_p
        L.oracle978_destroy.argtypes = [C.c_void_p]
        L.oracle978_phase_table.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle978_handle_data.restype = C.c_size_t
        L.oracle978_handle_data.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
        L.oracle978_process_buffer.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_uint64, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.oracle978_calls.restype = C.c_size_t
        L.oracle978_calls.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.oracle978_state.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle978_rs_encode.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.oracle978_rs_decode.argtypes = [C.c_int, C.c_void_p]
        L.oracle978_correct_adsb.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        L.oracle978_correct_uplink.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_int)]
        _lib = L
    return _lib


class Oracle978:
    def __init__(self):
        self.h = lib().oracle978_create()

    def close(self):
        if self.h:
            lib().oracle978_destroy(self.h)
            self.h = None

    __del__ = close

    def phase_table(self) -> np.ndarray:
        t = np.zeros(65536, dtype=np.uint16)
        lib().oracle978_phase_table(self.h, t.ctypes.data)
        return t

    def handle_data(self, iq: np.ndarray, cap: int = 1 << 14) -> np.ndarray:
        iq = np.ascontiguousarray(iq, dtype=np.uint8)
        out = np.zeros(cap, dtype=FRAME_DTYPE)
        n = lib().oracle978_handle_data(self.h, iq.ctypes.data, iq.size, out.ctypes.data, cap)
        assert n <= cap
        return out[:n]

    def calls(self) -> np.ndarray:
        """(len, offset, returned) of every process_buffer() call HandleData made so far."""
        buf = np.zeros((4096, 3), np.int64)
        n = lib().oracle978_calls(self.h, buf.ctypes.data, 4096)
        assert n <= 4096
        return buf[:n]

    def state(self):
        s = np.zeros(2, np.uint64)
        lib().oracle978_state(self.h, s.ctypes.data)
        return int(s[0]), int(s[1])

    def process_buffer(self, phi: np.ndarray, offset: int = 0, cap: int = 1 << 14):
        phi = np.ascontiguousarray(phi, dtype=np.uint16)
        out = np.zeros(cap, dtype=FRAME_DTYPE)
        n = C.c_size_t(0)
        done = lib().oracle978_process_buffer(self.h, phi.ctypes.data, phi.size, offset, out.ctypes.data, cap, C.byref(n))
        return done, out[:n.value]


KIND = {12: 0, 14: 1, 20: 2}


def rs_parity(data: bytes, nroots: int) -> bytes:
    d = np.frombuffer(bytes(data), dtype=np.uint8).copy()
    p = np.zeros(nroots, dtype=np.uint8)
    lib().oracle978_rs_encode(KIND[nroots], d.ctypes.data, p.ctypes.data)
    return p.tobytes()


def rs_decode(block: bytes, nroots: int):
    b = np.frombuffer(bytes(block), dtype=np.uint8).copy()
    n = lib().oracle978_rs_decode(KIND[nroots], b.ctypes.data)
    return int(n), b.tobytes()
