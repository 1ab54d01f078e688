"""TEST INFRASTRUCTURE ONLY.  ctypes front-ends of oracle/liboraclecpr.so (this repo's restatement of
dump1090/cpr.c) and of oracle/_ref/libcpr_ref.so (the reference's cpr.c compiled where it lies)."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboraclecpr.so")
REF_PATH = os.path.join(HERE, "_ref", "libcpr_ref.so")

# == adsbcu_cpr_in / adsbcu_cpr_out (include/adsbcu.h)
IN_DTYPE = np.dtype([("reflat", "<f8"), ("reflon", "<f8"), ("even_lat", "<i4"), ("even_lon", "<i4"), ("odd_lat", "<i4"),
                     ("odd_lon", "<i4"), ("fflag", "<i4"), ("mode", "<i4")])
OUT_DTYPE = np.dtype([("lat", "<f8"), ("lon", "<f8"), ("result", "<i4"), ("reserved", "<i4")])
assert IN_DTYPE.itemsize == 40 and OUT_DTYPE.itemsize == 24
AIRBORNE, SURFACE, RELATIVE_AIRBORNE, RELATIVE_SURFACE = 0, 1, 2, 3

_lib = _ref = None


def decode(rec: np.ndarray) -> np.ndarray:
    """The restatement, one OUT_DTYPE record per IN_DTYPE record."""
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB_PATH)
        _lib.oraclecpr_decode.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
    rec = np.ascontiguousarray(rec, dtype=IN_DTYPE)
    out = np.zeros(rec.size, dtype=OUT_DTYPE)
    _lib.oraclecpr_decode(rec.ctypes.data, rec.size, out.ctypes.data)
    return out


def decode_ref(rec: np.ndarray) -> np.ndarray:
    """The reference's own decodeCPRairborne / decodeCPRsurface / decodeCPRrelative, record by record."""
    global _ref
    if _ref is None:
        _ref = C.CDLL(REF_PATH)
        pd = C.POINTER(C.c_double)
        _ref.decodeCPRairborne.argtypes = [C.c_int] * 5 + [pd, pd]
        _ref.decodeCPRsurface.argtypes = [C.c_double, C.c_double] + [C.c_int] * 5 + [pd, pd]
        _ref.decodeCPRrelative.argtypes = [C.c_double, C.c_double] + [C.c_int] * 4 + [pd, pd]
    out = np.zeros(rec.size, dtype=OUT_DTYPE)
    lat, lon = C.c_double(), C.c_double()
    for k, r in enumerate(rec):
        lat.value = lon.value = 0.0
        a = [int(r[f]) for f in ("even_lat", "even_lon", "odd_lat", "odd_lon", "fflag")]
        mode = int(r["mode"])
        if mode == AIRBORNE:
            res = _ref.decodeCPRairborne(*a, C.byref(lat), C.byref(lon))
        elif mode == SURFACE:
            res = _ref.decodeCPRsurface(float(r["reflat"]), float(r["reflon"]), *a, C.byref(lat), C.byref(lon))
        else:
            res = _ref.decodeCPRrelative(float(r["reflat"]), float(r["reflon"]), a[0], a[1], a[4], int(mode == RELATIVE_SURFACE),
                                         C.byref(lat), C.byref(lon))
        out[k] = (lat.value, lon.value, res, 0)
    return out


def random_records(n: int, seed: int) -> np.ndarray:
    """A mix of all four modes: uniformly random raw values (mostly rejected) and pairs encoded from a true position
    near the reference point, which decode."""
    rng = np.random.default_rng(seed)
    rec = np.zeros(n, dtype=IN_DTYPE)
    rec["mode"] = rng.integers(0, 4, n)
    rec["fflag"] = rng.integers(0, 2, n)
    for f in ("even_lat", "even_lon", "odd_lat", "odd_lon"):
        rec[f] = rng.integers(0, 131072, n)
    rec["reflat"] = rng.uniform(-90, 90, n)
    rec["reflon"] = rng.uniform(-180, 180, n)
    k = (2 * n) // 3                       # encoded from a position
    lat = rng.uniform(-89.9, 89.9, k)
    lon = rng.uniform(-180, 180, k)
    surface = (rec["mode"][:k] == SURFACE) | (rec["mode"][:k] == RELATIVE_SURFACE)
    span = np.where(surface, 90.0, 360.0)
    for odd, (fl, fo) in enumerate((("even_lat", "even_lon"), ("odd_lat", "odd_lon"))):
        # relative modes carry their single frame in even_*; encode it with the frame's own parity
        par = np.where(rec["mode"][:k] >= 2, rec["fflag"][:k], odd) if odd == 0 else np.full(k, 1)
        dlat = span / (60 - par)
        yz = np.floor(131072 * (np.mod(lat, dlat) / dlat) + 0.5)
        rlat = dlat * (yz / 131072 + np.floor(lat / dlat))
        nlv = np.array([nl(v) for v in rlat])
        dlon = span / np.maximum(nlv - par, 1)
        xz = np.floor(131072 * (np.mod(lon, dlon) / dlon) + 0.5)
        rec[fl][:k] = yz.astype(np.int64) & 0x1FFFF
        rec[fo][:k] = xz.astype(np.int64) & 0x1FFFF
    # the reference point: close to the true position (inside half a cell most of the time), sometimes far away
    near = rng.random(k) < 0.8
    rec["reflat"][:k] = np.where(near, np.clip(lat + rng.uniform(-0.6, 0.6, k), -90, 90), rec["reflat"][:k])
    rec["reflon"][:k] = np.where(near, lon + rng.uniform(-0.5, 0.5, k), rec["reflon"][:k])
    # edge values
    rec["even_lat"][k::7] = 0
    rec["odd_lat"][k::7] = 0
    rec["reflat"][k::5] = rng.choice([-90.0, -45.0, 0.0, 45.0, 90.0], rec["reflat"][k::5].size)
    rec["reflon"][k::11] = rng.choice([-180.0, 180.0, 0.0], rec["reflon"][k::11].size)
    return rec


def nl(lat: float) -> int:
    global _lib
    if _lib is None:
        decode(np.zeros(0, dtype=IN_DTYPE))
    _lib.oraclecpr_nl.argtypes = [C.c_double]
    return _lib.oraclecpr_nl(float(lat))
