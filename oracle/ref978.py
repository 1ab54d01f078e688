"""TEST INFRASTRUCTURE ONLY: ctypes binding of oracle/_ref/libref978.so -- the reference's own UAT978.cpp (phase table
and HandleData staging loop) compiled where it lies, with dump978's process_buffer() supplied as a callback."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libref978.so")
PROCESS_FN = C.CFUNCTYPE(C.c_int, C.POINTER(C.c_uint16), C.c_int, C.c_uint64)
_lib = None


def have_ref() -> bool:
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.ref978_create.restype = C.c_void_p
        L.ref978_destroy.argtypes = [C.c_void_p]
        L.ref978_phase_table.argtypes = [C.c_void_p, C.c_void_p]
        L.ref978_handle_data.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.ref978_state.argtypes = [C.c_void_p, C.c_void_p]
        L.ref978_set_process_buffer.argtypes = [PROCESS_FN]
        _lib = L
    return _lib


class Ref978:
    """UAT978Handler of the reference; process(phi: np.ndarray[uint16], offset) -> samples consumed."""

    def __init__(self, process):
        self.calls = []                     # (len, offset, returned)

        def cb(phi, n, offset):
            done = int(process(np.ctypeslib.as_array(phi, shape=(n,)), int(offset)))
            self.calls.append((int(n), int(offset), done))
            return done
        self._cb = PROCESS_FN(cb)
        lib().ref978_set_process_buffer(self._cb)
        self.h = lib().ref978_create()

    def close(self):
        if self.h:
            lib().ref978_destroy(self.h)
            self.h = None

    def phase_table(self) -> np.ndarray:
        t = np.zeros(65536, np.uint16)
        lib().ref978_phase_table(self.h, t.ctypes.data)
        return t

    def handle_data(self, iq: np.ndarray):
        iq = np.ascontiguousarray(iq, np.uint8)
        lib().ref978_set_process_buffer(self._cb)
        lib().ref978_handle_data(self.h, iq.ctypes.data, iq.size)

    def state(self):
        s = np.zeros(2, np.uint64)
        lib().ref978_state(self.h, s.ctypes.data)
        return int(s[0]), int(s[1])
