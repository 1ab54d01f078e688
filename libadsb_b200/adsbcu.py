"""ctypes binding of the C ABI in include/adsbcu.h.

This is plumbing for tests and bench.py; the product is libadsbcu.so.  There is no CPU
fallback: if the library or a B200 is missing, construction raises."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libadsbcu.so")

MSG_DTYPE = np.dtype([
    ("timestamp", "<u8"), ("msg", "u1", 14), ("verbatim", "u1", 14), ("msgbits", "u1"),
    ("correctedbits", "u1"), ("msgtype", "u1"), ("score", "u1"), ("crc", "<u4"), ("addr", "<u4"),
    ("signal_level", "<f8")])
assert MSG_DTYPE.itemsize == 56
MSG_B_DTYPE = np.dtype([("pos", "<u4"), ("crc", "<u4"), ("addr", "<u4"), ("msg", "u1", 14), ("msgbits", "u1"),
                        ("msgtype", "u1"), ("errorbit", "i1"), ("corrected", "u1"), ("reserved", "<u2")])
assert MSG_B_DTYPE.itemsize == 32
FIELDS_A_DTYPE = np.dtype([("AC", "<u2"), ("ID", "<u2")] + [(k, "u1") for k in (
    "CA", "CC", "CF", "DR", "FS", "KE", "ND", "RI", "SL", "UM", "VS", "airground", "alert_valid", "alert", "spi_valid", "spi",
    "altitude_baro_valid", "altitude_baro_unit", "squawk_valid", "es_pending")] + [("squawk", "<u2"), ("pad", "<u2"), ("altitude_baro", "<i4")])
assert FIELDS_A_DTYPE.itemsize == 32
# adsbcu_fields_es (include/adsbcu.h); valid bits ADSBCU_ES_*
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
# adsbcu_fields_commb; valid bits ADSBCU_MB_*
FIELDS_MB_DTYPE = np.dtype([("valid", "<u4")] + [(k, "u1") for k in (
    "commb_format", "heading_type", "nav_modes", "nav_altitude_source", "mrar_source", "turbulence")] + [("reserved", "u1", 2), ("callsign", "S8")] +
    [(k, "<f4") for k in ("roll", "heading", "track_rate", "gs", "mach", "nav_qnh", "wind_speed", "wind_dir", "temperature", "pressure", "humidity")] +
    [(k, "<i4") for k in ("baro_rate", "geom_rate", "nav_mcp_altitude", "nav_fms_altitude")] + [("ias", "<u2"), ("tas", "<u2")])
assert FIELDS_MB_DTYPE.itemsize == 84
COMMB_FORMATS = ("unknown", "ambiguous", "not_decoded", "empty_response", "datalink_caps", "gicb_caps", "aircraft_ident", "acas_ra",
                 "vertical_intent", "track_turn", "heading_speed", "mrar", "airborne_position")
ES_VALID = {name: 1 << k for k, name in enumerate((
    "altitude_baro", "altitude_geom", "heading", "gs", "ias", "tas", "baro_rate", "geom_rate", "squawk", "callsign", "cpr", "category",
    "geom_delta", "spi", "alert", "emergency", "nic_a", "nic_b", "nic_c", "nic_baro", "nac_p", "nac_v", "gva", "sda", "nav_heading",
    "nav_fms_altitude", "nav_mcp_altitude", "nav_qnh", "nav_modes", "opstatus", "cc_lw"))}
UAT_DTYPE = np.dtype([("offset", "<u8"), ("rs_errors", "<i2"), ("len", "<u2"), ("updown", "u1"), ("reserved", "u1", 3),
                      ("data", "u1", 432)])
assert UAT_DTYPE.itemsize == 448
# adsbcu_fields_b / adsbcu_cpr_pair / adsbcu_cpr_fix
FIELDS_B_DTYPE = np.dtype([("ca", "<i4"), ("metype", "<i4"), ("mesub", "<i4"), ("heading", "<i4"), ("fflag", "<i4"),
                           ("raw_latitude", "<i4"), ("raw_longitude", "<i4"), ("flight", "S8"), ("ew_dir", "<i4"),
                           ("ew_velocity", "<i4"), ("ns_dir", "<i4"), ("ns_velocity", "<i4"), ("velocity", "<i4"),
                           ("fs", "<i4"), ("dr", "<i4"), ("um", "<i4"), ("identity", "<i4"), ("altitude", "<i4"), ("unit", "<i4")])
assert FIELDS_B_DTYPE.itemsize == 80
CPR_PAIR_DTYPE = np.dtype([("even_lat", "<i4"), ("even_lon", "<i4"), ("odd_lat", "<i4"), ("odd_lon", "<i4"),
                           ("even_is_newer", "<i4")])
CPR_FIX_DTYPE = np.dtype([("lat1e7", "<i4"), ("lon1e7", "<i4"), ("valid", "<i4")])
# adsbcu_cpr_in / adsbcu_cpr_out (dump1090/cpr.c)
CPR_IN_DTYPE = np.dtype([("reflat", "<f8"), ("reflon", "<f8"), ("even_lat", "<i4"), ("even_lon", "<i4"), ("odd_lat", "<i4"),
                         ("odd_lon", "<i4"), ("fflag", "<i4"), ("mode", "<i4")])
CPR_OUT_DTYPE = np.dtype([("lat", "<f8"), ("lon", "<f8"), ("result", "<i4"), ("reserved", "<i4")])
assert CPR_IN_DTYPE.itemsize == 40 and CPR_OUT_DTYPE.itemsize == 24
CPR_AIRBORNE, CPR_SURFACE, CPR_RELATIVE_AIRBORNE, CPR_RELATIVE_SURFACE = 0, 1, 2, 3

CAND_DTYPE = np.dtype([("pos", "<u8"), ("phase", "u1"), ("kind", "u1"), ("corrections", "i1"), ("msgtype", "u1"),
                       ("verbatim", "u1", 14), ("msg", "u1", 14), ("crc", "<u4"), ("addr", "<u4"), ("test_key", "<u4"),
                       ("add_key", "<u4"), ("signal_level", "<f8")])
assert CAND_DTYPE.itemsize == 64
KNOWN_WORDS = 1 << 20

EXPORTS = ["adsbcu_abi_version", "adsbcu_last_error", "adsbcu_create", "adsbcu_destroy", "adsbcu_1090_reset",
           "adsbcu_1090_feed", "adsbcu_1090_feed_device", "adsbcu_1090_magnitude", "adsbcu_1090_filter_add",
           "adsbcu_1090_filter_test", "adsbcu_1090_filter_expire", "adsbcu_get_stats", "adsbcu_1090b_feed", "adsbcu_1090b_reset",
           "adsbcu_978_feed", "adsbcu_978_feed_device", "adsbcu_978_get_stats", "adsbcu_978_reset", "adsbcu_978_phase", "adsbcu_1090_seek", "adsbcu_1090_scan_candidates",
           "adsbcu_1090_known_export", "adsbcu_1090_known_import", "adsbcu_1090_revive", "adsbcu_1090_resolve", "adsbcu_1090_modeac_block", "adsbcu_1090_feed_sc16",
           "adsbcu_1090_modeac_block_static", "adsbcu_1090_decode_fields", "adsbcu_1090_decode_es", "adsbcu_1090_decode_commb", "adsbcu_1090_decode_device", "adsbcu_1090_encode_beast", "adsbcu_1090_encode_avr", "adsbcu_1090_state_export", "adsbcu_1090_state_import", "adsbcu_1090_replay_sharded", "adsbcu_1090_replay_sharded_tail",
           "adsbcu_1090b_decode_fields", "adsbcu_1090b_cpr_decode", "adsbcu_1090_cpr_decode"]


class StreamStateC(C.Structure):
    _fields_ = [("stream_pos", C.c_uint64), ("last_a", C.c_uint64), ("last_lme", C.c_uint64), ("has_last", C.c_uint32), ("reserved", C.c_uint32)]


class Piece(C.Structure):
    _fields_ = [("iq_dev", C.c_void_p), ("nsamples", C.c_uint64)]


class ShardReport(C.Structure):
    _fields_ = [("passes", C.c_uint32), ("rounds", C.c_uint32), ("start_addrs", C.c_uint32), ("chunks", C.c_uint32), ("launches", C.c_uint64),
                ("ms_front", C.c_double), ("ms_messages", C.c_double), ("ms_order", C.c_double), ("ms_resolve", C.c_double), ("ms_h2d", C.c_double)]


SHARD_MAX_ADDRS = 4000
SHARD_WORDS = SHARD_MAX_ADDRS + 8
ALLGATHER_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(C.c_uint32), C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32))


class Stats978(C.Structure):
    _fields_ = [("ms_device", C.c_float), ("calls", C.c_uint32), ("launches", C.c_uint32), ("reserved", C.c_uint32)]


class Config(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("nfix_crc", C.c_int32), ("fix_df", C.c_int32),
                ("block_samples", C.c_uint32), ("max_feed_samples", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("preambles", C.c_uint64), ("phase_records", C.c_uint64),
                ("resolved_records", C.c_uint64), ("messages", C.c_uint64), ("ms_h2d", C.c_float),
                ("ms_front", C.c_float), ("ms_messages", C.c_float), ("ms_order", C.c_float), ("ms_resolve", C.c_float),
                ("launches", C.c_uint32), ("chunks", C.c_uint32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "reserved"}


_lib = None


def load():
    """dlopen libadsbcu.so (loudly failing if it has not been built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} missing: run `python -m libadsb_b200.build` (nvcc, sm_100a)")
        L = C.CDLL(LIB_PATH)
        L.adsbcu_abi_version.restype = C.c_int
        L.adsbcu_last_error.restype = C.c_char_p
        L.adsbcu_last_error.argtypes = [C.c_void_p]
        L.adsbcu_create.argtypes = [C.POINTER(Config), C.POINTER(C.c_void_p)]
        L.adsbcu_destroy.argtypes = [C.c_void_p]
        L.adsbcu_1090_reset.argtypes = [C.c_void_p, C.c_int]
        L.adsbcu_1090_feed.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.adsbcu_1090_feed_device.argtypes = L.adsbcu_1090_feed.argtypes
        L.adsbcu_1090_magnitude.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.adsbcu_1090_filter_add.argtypes = [C.c_void_p, C.c_uint32]
        L.adsbcu_1090_filter_test.argtypes = [C.c_void_p, C.c_uint32]
        L.adsbcu_1090_filter_expire.argtypes = [C.c_void_p]
        L.adsbcu_get_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
        L.adsbcu_1090b_feed.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.adsbcu_1090b_reset.argtypes = [C.c_void_p]
        L.adsbcu_1090b_decode_fields.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.adsbcu_1090b_cpr_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.adsbcu_1090_cpr_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.adsbcu_978_feed.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.adsbcu_978_feed_device.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.adsbcu_978_get_stats.argtypes = [C.c_void_p, C.POINTER(Stats978)]
        L.adsbcu_978_reset.argtypes = [C.c_void_p]
        L.adsbcu_978_phase.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.adsbcu_1090_seek.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_size_t]
        L.adsbcu_1090_scan_candidates.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_size_t,
                                                  C.POINTER(C.c_size_t)]
        L.adsbcu_1090_known_export.argtypes = [C.c_void_p, C.c_void_p]
        L.adsbcu_1090_known_import.argtypes = [C.c_void_p, C.c_void_p]
        L.adsbcu_1090_revive.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.adsbcu_1090_resolve.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.adsbcu_1090_modeac_block.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_uint64, C.c_void_p,
                                               C.c_size_t, C.POINTER(C.c_size_t)]
        L.adsbcu_1090_feed_sc16.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.adsbcu_1090_decode_fields.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.adsbcu_1090_decode_es.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.adsbcu_1090_decode_commb.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.adsbcu_1090_decode_device.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.adsbcu_1090_encode_beast.restype = C.c_size_t
        L.adsbcu_1090_encode_beast.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_size_t]
        L.adsbcu_1090_encode_avr.restype = C.c_size_t
        L.adsbcu_1090_encode_avr.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
        L.adsbcu_1090_state_export.argtypes = [C.c_void_p, C.POINTER(StreamStateC), C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.adsbcu_1090_state_import.argtypes = [C.c_void_p, C.POINTER(StreamStateC), C.c_void_p, C.c_size_t]
        L.adsbcu_1090_replay_sharded.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(Piece), C.c_size_t, C.c_uint64, C.c_void_p, C.c_size_t,
                                                 ALLGATHER_FN, C.c_void_p, C.c_uint, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(ShardReport)]
        L.adsbcu_1090_replay_sharded_tail.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(Piece), C.c_size_t, C.c_uint64, C.c_void_p, C.c_size_t,
                                                      C.POINTER(Piece), C.c_size_t, C.c_uint64, C.c_void_p, C.c_size_t,
                                                      ALLGATHER_FN, C.c_void_p, C.c_uint, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t),
                                                      C.POINTER(C.c_size_t), C.POINTER(ShardReport)]
        L.adsbcu_1090_modeac_block_static.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_uint64, C.c_void_p,
                                                      C.c_size_t, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


class AdsbcuError(RuntimeError):
    pass


def encode_beast(msgs: np.ndarray, verbatim: bool = True) -> bytes:
    """adsbcu_1090_encode_beast on a MSG_DTYPE array."""
    msgs = np.ascontiguousarray(msgs, dtype=MSG_DTYPE)
    need = load().adsbcu_1090_encode_beast(msgs.ctypes.data, len(msgs), int(verbatim), None, 0)
    out = np.zeros(max(need, 1), np.uint8)
    load().adsbcu_1090_encode_beast(msgs.ctypes.data, len(msgs), int(verbatim), out.ctypes.data, need)
    return out[:need].tobytes()


def encode_avr(msgs: np.ndarray) -> bytes:
    msgs = np.ascontiguousarray(msgs, dtype=MSG_DTYPE)
    need = load().adsbcu_1090_encode_avr(msgs.ctypes.data, len(msgs), None, 0)
    out = np.zeros(max(need, 1), np.uint8)
    load().adsbcu_1090_encode_avr(msgs.ctypes.data, len(msgs), out.ctypes.data, need)
    return out[:need].tobytes()


class Demod1090:
    """One demodulator stream (one `adsbcu_handle`)."""

    def __init__(self, nfix_crc=0, fix_df=1, block_samples=0, max_feed_samples=0, device=0):
        L = load()
        cfg = Config(C.sizeof(Config), device, nfix_crc, fix_df, block_samples, max_feed_samples)
        self.h = C.c_void_p()
        rc = L.adsbcu_create(C.byref(cfg), C.byref(self.h))
        if rc != 0:
            msg = L.adsbcu_last_error(None)
            self.h = None
            raise AdsbcuError(f"adsbcu_create failed ({rc}): {msg.decode() if msg else ''}")

    def close(self):
        if getattr(self, "h", None) and _lib is not None:       # (at interpreter exit the module globals may be gone already)
            _lib.adsbcu_destroy(self.h)
            self.h = None

    __del__ = close

    def _check(self, rc):
        if rc != 0:
            raise AdsbcuError(f"adsbcu error {rc}: {load().adsbcu_last_error(self.h).decode()}")

    def reset(self, clear_filter=True):
        self._check(load().adsbcu_1090_reset(self.h, int(clear_filter)))

    def feed(self, iq: np.ndarray, cap: int | None = None, out: np.ndarray | None = None) -> np.ndarray:
        iq = np.ascontiguousarray(iq, dtype=np.uint8)
        n = iq.size // 2
        if out is None:
            out = np.zeros(cap if cap is not None else max(1024, n // 100), dtype=MSG_DTYPE)
        got = C.c_size_t(0)
        self._check(load().adsbcu_1090_feed(self.h, iq.ctypes.data, n, out.ctypes.data, len(out), C.byref(got)))
        return out[:got.value]

    def feed_sc16(self, iq: np.ndarray, q11: bool = False, cap: int | None = None) -> np.ndarray:
        """--iformat SC16 / SC16Q11: interleaved int16 I/Q."""
        iq = np.ascontiguousarray(iq, dtype="<i2")
        n = iq.size // 2
        out = np.zeros(cap if cap is not None else max(1024, n // 100), dtype=MSG_DTYPE)
        got = C.c_size_t(0)
        self._check(load().adsbcu_1090_feed_sc16(self.h, iq.ctypes.data, n, int(q11), out.ctypes.data, len(out), C.byref(got)))
        return out[:got.value]

    def feed_device(self, dev_ptr: int, nsamples: int, out: np.ndarray) -> np.ndarray:
        got = C.c_size_t(0)
        self._check(load().adsbcu_1090_feed_device(self.h, C.c_void_p(dev_ptr), nsamples, out.ctypes.data, len(out),
                                                   C.byref(got)))
        return out[:got.value]

    def feed_device_resident(self, dev_ptr: int, nsamples: int, out_ptr: int, cap: int) -> int:
        """adsbcu_1090_feed_device with out[] in DEVICE memory: the message set stays on the GPU; returns the message count."""
        got = C.c_size_t(0)
        self._check(load().adsbcu_1090_feed_device(self.h, C.c_void_p(dev_ptr), nsamples, C.c_void_p(out_ptr), cap, C.byref(got)))
        return int(got.value)

    def magnitude(self, iq: np.ndarray) -> np.ndarray:
        iq = np.ascontiguousarray(iq, dtype=np.uint8)
        n = iq.size // 2
        out = np.zeros(n, dtype=np.uint16)
        self._check(load().adsbcu_1090_magnitude(self.h, iq.ctypes.data, n, out.ctypes.data))
        return out

    def handle_data_b(self, iq: np.ndarray, cap: int = 1 << 16, out: np.ndarray | None = None) -> np.ndarray:
        """Path B: one ADSB1090Handler::HandleData(span) call (libadsb's own 2.0 MS/s demodulator).  `out`: an array to reuse."""
        iq = np.ascontiguousarray(iq, dtype=np.uint8)
        if out is None:
            out = np.zeros(cap, dtype=MSG_B_DTYPE)
        cap = len(out)
        got = C.c_size_t(0)
        self._check(load().adsbcu_1090b_feed(self.h, iq.ctypes.data, iq.size, out.ctypes.data, cap, C.byref(got)))
        return out[:got.value]

    def decode_fields_a(self, msgs: np.ndarray) -> np.ndarray:
        """dump1090's surveillance fields (mode_s.c:626-779) of every MSG_DTYPE record."""
        msgs = np.ascontiguousarray(msgs, dtype=MSG_DTYPE)
        out = np.zeros(len(msgs), dtype=FIELDS_A_DTYPE)
        self._check(load().adsbcu_1090_decode_fields(self.h, msgs.ctypes.data, len(msgs), out.ctypes.data))
        return out

    def decode_es(self, msgs: np.ndarray) -> np.ndarray:
        """dump1090's extended-squitter payload decode (mode_s.c:798-1551) of every MSG_DTYPE record (zeroed for other DFs)."""
        msgs = np.ascontiguousarray(msgs, dtype=MSG_DTYPE)
        out = np.zeros(len(msgs), dtype=FIELDS_ES_DTYPE)
        self._check(load().adsbcu_1090_decode_es(self.h, msgs.ctypes.data, len(msgs), out.ctypes.data))
        return out

    def decode_commb(self, msgs: np.ndarray) -> np.ndarray:
        """dump1090's Comm-B payload decode (decodeCommB, comm_b.c:50-985) of every MSG_DTYPE record (zeroed unless DF20 / DF21)."""
        msgs = np.ascontiguousarray(msgs, dtype=MSG_DTYPE)
        out = np.zeros(len(msgs), dtype=FIELDS_MB_DTYPE)
        self._check(load().adsbcu_1090_decode_commb(self.h, msgs.ctypes.data, len(msgs), out.ctypes.data))
        return out

    def decode_device(self, msgs_ptr: int, n: int, fields_ptr: int = 0, es_ptr: int = 0, commb_ptr: int = 0, stream: int = 0) -> None:
        """All of decodeModesMessage for n device-resident records in one launch on `stream` (device pointers; outputs may be 0)."""
        self._check(load().adsbcu_1090_decode_device(self.h, msgs_ptr or None, n, fields_ptr or None, es_ptr or None, commb_ptr or None,
                                                     stream or None))

    def cpr_decode_a(self, rec: np.ndarray) -> np.ndarray:
        """dump1090's decodeCPRairborne / decodeCPRsurface / decodeCPRrelative on every CPR_IN_DTYPE record."""
        rec = np.ascontiguousarray(rec, dtype=CPR_IN_DTYPE)
        out = np.zeros(rec.size, dtype=CPR_OUT_DTYPE)
        self._check(load().adsbcu_1090_cpr_decode(self.h, rec.ctypes.data, rec.size, out.ctypes.data))
        return out

    def decode_fields_b(self, msgs: np.ndarray) -> np.ndarray:
        """Path B, next layer: the Message fields DecodeModesMessage fills after the CRC stage, for every record."""
        msgs = np.ascontiguousarray(msgs, dtype=MSG_B_DTYPE)
        out = np.zeros(msgs.size, dtype=FIELDS_B_DTYPE)
        self._check(load().adsbcu_1090b_decode_fields(self.h, msgs.ctypes.data, msgs.size, out.ctypes.data))
        return out

    def cpr_decode(self, pairs: np.ndarray) -> np.ndarray:
        """DecodeCpr on every even/odd pair (CPR_PAIR_DTYPE) -> CPR_FIX_DTYPE."""
        pairs = np.ascontiguousarray(pairs, dtype=CPR_PAIR_DTYPE)
        out = np.zeros(pairs.size, dtype=CPR_FIX_DTYPE)
        self._check(load().adsbcu_1090b_cpr_decode(self.h, pairs.ctypes.data, pairs.size, out.ctypes.data))
        return out

    def reset_b(self):
        self._check(load().adsbcu_1090b_reset(self.h))

    def handle_data_uat(self, iq: np.ndarray, cap: int = 1 << 14, out: np.ndarray | None = None) -> np.ndarray:
        """Path C: one UAT978Handler::HandleData(span) call (978 MHz UAT, 2.083334 MS/s).  `out`: a frame array to reuse."""
        iq = np.ascontiguousarray(iq, dtype=np.uint8)
        if out is None:
            out = np.zeros(cap, dtype=UAT_DTYPE)
        cap = len(out)
        got = C.c_size_t(0)
        self._check(load().adsbcu_978_feed(self.h, iq.ctypes.data, iq.size, out.ctypes.data, cap, C.byref(got)))
        return out[:got.value]

    def handle_data_uat_device(self, iq_dev: int, nbytes: int, out: np.ndarray) -> np.ndarray:
        """Path C with the samples already in device memory (iq_dev: device address)."""
        got = C.c_size_t(0)
        self._check(load().adsbcu_978_feed_device(self.h, iq_dev, nbytes, out.ctypes.data, len(out), C.byref(got)))
        return out[:got.value]

    def uat_stats(self) -> dict:
        s = Stats978()
        self._check(load().adsbcu_978_get_stats(self.h, C.byref(s)))
        return {"ms_device": s.ms_device, "calls": s.calls, "launches": s.launches}

    def reset_uat(self):
        self._check(load().adsbcu_978_reset(self.h))

    def uat_phase(self, iq: np.ndarray) -> np.ndarray:
        iq = np.ascontiguousarray(iq, dtype=np.uint8)
        out = np.zeros(iq.size // 2, dtype=np.uint16)
        self._check(load().adsbcu_978_phase(self.h, iq.ctypes.data, iq.size // 2, out.ctypes.data))
        return out

    # ---- sharded captures (include/adsbcu.h, "Path A on a capture sharded over several GPUs") ----
    def seek(self, stream_pos: int, prev_iq: np.ndarray | None = None):
        prev = np.ascontiguousarray(prev_iq if prev_iq is not None else np.zeros(0, np.uint8), dtype=np.uint8)
        self._check(load().adsbcu_1090_seek(self.h, stream_pos, prev.ctypes.data if prev.size else None, prev.size // 2))

    def scan_candidates(self, iq, nsamples: int | None = None, cap: int | None = None) -> np.ndarray:
        """iq: host uint8 array, or an int device pointer (then nsamples is required)."""
        if isinstance(iq, int):
            ptr, n, on_dev = iq, int(nsamples), 1
        else:
            iq = np.ascontiguousarray(iq, dtype=np.uint8)
            ptr, n, on_dev = iq.ctypes.data, iq.size // 2, 0
        out = np.zeros(cap if cap is not None else max(4096, n // 64), dtype=CAND_DTYPE)
        got = C.c_size_t(0)
        self._check(load().adsbcu_1090_scan_candidates(self.h, C.c_void_p(ptr), n, on_dev, out.ctypes.data, len(out), C.byref(got)))
        return out[:got.value]

    def known_export(self) -> np.ndarray:
        w = np.zeros(KNOWN_WORDS, dtype=np.uint32)
        self._check(load().adsbcu_1090_known_export(self.h, w.ctypes.data))
        return w

    def known_import(self, words: np.ndarray):
        w = np.ascontiguousarray(words, dtype=np.uint32)
        assert w.size == KNOWN_WORDS
        self._check(load().adsbcu_1090_known_import(self.h, w.ctypes.data))

    def revive(self, cap: int = 1 << 16) -> np.ndarray:
        out = np.zeros(cap, dtype=CAND_DTYPE)
        got = C.c_size_t(0)
        self._check(load().adsbcu_1090_revive(self.h, out.ctypes.data, cap, C.byref(got)))
        return out[:got.value]

    def resolve(self, cands: np.ndarray, cap: int | None = None) -> np.ndarray:
        cands = np.ascontiguousarray(cands, dtype=CAND_DTYPE)
        out = np.zeros(cap if cap is not None else max(16, len(cands)), dtype=MSG_DTYPE)
        got = C.c_size_t(0)
        self._check(load().adsbcu_1090_resolve(self.h, cands.ctypes.data, len(cands), out.ctypes.data, len(out), C.byref(got)))
        return out[:got.value]

    def modeac_block(self, iq: np.ndarray, overlap_iq: np.ndarray | None, block_timestamp: int, cap: int = 1 << 15) -> np.ndarray:
        """One demodulate2400AC(mag_buf) call: `iq` = the buffer's new samples, `overlap_iq` = the <= 326 samples before it."""
        iq = np.ascontiguousarray(iq, dtype=np.uint8)
        ov = np.ascontiguousarray(overlap_iq if overlap_iq is not None else np.zeros(0, np.uint8), dtype=np.uint8)
        out = np.zeros(cap, dtype=MSG_DTYPE)
        got = C.c_size_t(0)
        self._check(load().adsbcu_1090_modeac_block(self.h, iq.ctypes.data, iq.size // 2, ov.ctypes.data if ov.size else None,
                                                    ov.size // 2, block_timestamp, out.ctypes.data, cap, C.byref(got)))
        return out[:got.value]

    def modeac_block_static(self, iq_dev: int, nsamples: int, overlap_iq_dev: int, n_overlap: int, block_timestamp: int,
                            out_dev: int, cap: int, result_dev: int, stream: int = 0) -> None:
        """adsbcu_1090_modeac_block_static: DEVICE addresses in and out, scratch in the library's static device array, nothing
        allocated, nothing awaited.  result_dev[0] = replies, result_dev[1] = flags once `stream` has reached this point."""
        self._check(load().adsbcu_1090_modeac_block_static(self.h, iq_dev or None, nsamples, overlap_iq_dev or None, n_overlap,
                                                           block_timestamp, out_dev or None, cap, result_dev or None, stream or None))

    def state_export(self):
        """(stream_pos, has_last, last_a, last_lme), addresses in the ICAO filter."""
        st, addrs, n = StreamStateC(), np.zeros(8192, np.uint32), C.c_size_t(0)
        self._check(load().adsbcu_1090_state_export(self.h, C.byref(st), addrs.ctypes.data, addrs.size, C.byref(n)))
        return (st.stream_pos, st.has_last, st.last_a, st.last_lme), addrs[:n.value].copy()

    def state_import(self, has_last: int, last_a: int, last_lme: int, addrs: np.ndarray):
        st = StreamStateC(0, last_a, last_lme, has_last, 0)
        addrs = np.ascontiguousarray(addrs, np.uint32)
        self._check(load().adsbcu_1090_state_import(self.h, C.byref(st), addrs.ctypes.data if addrs.size else None, addrs.size))

    def replay_sharded(self, rank: int, world: int, pieces, range_start: int, halo_iq: np.ndarray | None, allgather, out: np.ndarray,
                       host_pieces: bool = False, has_tail: bool = False, tail=None):
        """adsbcu_1090_replay_sharded[_tail].  pieces: [(device address, nsamples)] of THIS rank's range in stream order;
        allgather(words: np.ndarray[uint32]) -> list of world np.ndarray[uint32] (the transport: torch.distributed, MPI, ...).
        has_tail (the same on every rank): the capture has world + 1 ranges and rank 0 also holds the last one,
        tail = (pieces, start, halo_iq).  Returns (messages of this range [rank 0 with a tail: own range, then the tail;
        self.n_first tells where the tail begins], ShardReport)."""
        arr = (Piece * max(1, len(pieces)))(*[Piece(int(p), int(n)) for p, n in pieces])
        halo = np.ascontiguousarray(halo_iq if halo_iq is not None else np.zeros(0, np.uint8), np.uint8)
        t_pieces, t_start, t_halo = tail if (tail is not None and rank == 0) else ([], 0, None)
        t_arr = (Piece * max(1, len(t_pieces)))(*[Piece(int(p), int(n)) for p, n in t_pieces])
        t_halo = np.ascontiguousarray(t_halo if t_halo is not None else np.zeros(0, np.uint8), np.uint8)

        def cb(_ctx, mine, n_mine, all_, counts):
            try:
                parts = allgather(np.ctypeslib.as_array(mine, shape=(n_mine,)).copy())
                for q, part in enumerate(parts):
                    part = np.ascontiguousarray(part, np.uint32)
                    counts[q] = part.size
                    C.memmove(C.addressof(all_.contents) + 4 * q * SHARD_WORDS, part.ctypes.data, 4 * part.size)
                return 0
            except Exception as e:      # noqa: BLE001  (an exception must not cross the C frame)
                self._cb_error = e
                return 1
        self._cb_error = None
        fn = ALLGATHER_FN(cb)
        got, first, rep = C.c_size_t(0), C.c_size_t(0), ShardReport()
        flags = (1 if host_pieces else 0) | (2 if has_tail else 0)
        rc = load().adsbcu_1090_replay_sharded_tail(self.h, rank, world, arr, len(pieces), range_start, halo.ctypes.data if halo.size else None,
                                                    halo.size // 2, t_arr, len(t_pieces), t_start, t_halo.ctypes.data if t_halo.size else None,
                                                    t_halo.size // 2, fn, None, flags, out.ctypes.data, len(out), C.byref(got), C.byref(first),
                                                    C.byref(rep))
        if self._cb_error is not None:
            raise self._cb_error
        self._check(rc)
        self.n_first = first.value
        return out[:got.value], rep

    def filter_add(self, addr): self._check(load().adsbcu_1090_filter_add(self.h, addr))
    def filter_test(self, addr): return bool(load().adsbcu_1090_filter_test(self.h, addr))
    def filter_expire(self): self._check(load().adsbcu_1090_filter_expire(self.h))

    def stats(self) -> dict:
        s = Stats()
        self._check(load().adsbcu_get_stats(self.h, C.byref(s)))
        return s.as_dict()
