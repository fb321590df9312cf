"""ctypes binding of include/smsgate_b200.h.

There is no CPU implementation behind this module: if the CUDA library is
missing or no B200-class GPU is visible, loading / creating a context raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# SMSGATE_B200_LIB: another build of the same library (kernel experiments, smsgate_b200.build.build_variant)
LIB_PATH = os.environ.get("SMSGATE_B200_LIB") or os.path.join(_HERE, "libsmsgate_b200.so")

ABI_VERSION = 2

# flags / status / entry flags — mirror the header
F_WORKER_SKIP, F_PARSER_OTP, F_MODIFIED, F_BODY_DATE = 0x01, 0x02, 0x04, 0x08
(ST_UPSERT, ST_PARSED_NO_MERCHANT, ST_SKIPPED_FILTER, ST_CACHE_MISS, ST_UNMATCHED, ST_BROKEN,
 ST_FUTURE_DATE, ST_FAILED_EXC, ST_ESCAPED) = range(9)
STATUS_NAMES = ("UPSERT", "PARSED_NO_MERCHANT", "SKIPPED_FILTER", "CACHE_MISS", "UNMATCHED", "BROKEN",
                "FUTURE_DATE", "FAILED_EXC", "ESCAPED")
# what actually reaches sms.failed.  FUTURE_DATE and ESCAPED do not: the reference's handler raises there
# (worker.py:177 / :174, observed by running it: tests/golden/worker_flow.json)
DLQ_STATUSES = (ST_CACHE_MISS, ST_UNMATCHED, ST_FAILED_EXC)
HANDLER_RAISES = (ST_FUTURE_DATE, ST_ESCAPED)
E_IS_DICT, E_HAS_TXN, E_SCHEMA_OK, E_CARD_STR, E_MERCHANT = 0x1, 0x2, 0x4, 0x8, 0x10
E_DATE_STRICT, E_DATE_HOST, E_DATE_NOBODY, E_DATE_TZ, E_DATE_KEY, E_DATE_PHONE = 0x20, 0x40, 0x80, 0x100, 0x200, 0x400
DEC_OK, DEC_SYNTAX, DEC_OVERFLOW = 0, 1, 2
LEG_SKIPPED_OTP, LEG_DEBIT, LEG_CREDIT, LEG_FAILED = 0, 1, 2, 3
LEG_F_BUDGET, LEG_F_INEXACT, LEG_GROUPS = 1, 2, 10
LEG_MODE_NO_STRIP, LEG_MODE_NO_OTP, LEG_MODE_NO_DEBIT, LEG_MODE_NO_CREDIT = 1, 2, 4, 8
DEDUP_STATIC_MAX = 1 << 20
MAX_WORLD = 64
RECORD_BYTES = 32
P2P_RECORD_BYTES = 48
E_LIMIT = -5
E_XML, E_UNSUPPORTED = -6, -7
XML_HAS_BODY, XML_HAS_ADDRESS, XML_DATE_BAD, XML_DATE_BIG = 1, 2, 4, 8

# (column name, numpy dtype string, elements per message)
COLUMNS = (
    ("msg_id", "u1", 20), ("cache_key", "u1", 32), ("flags", "u1", 1), ("body_date", "i4", 1),
    ("fixed_len", "i4", 1), ("fixed_bytes", "u1", 0), ("slot", "i4", 1), ("status", "u1", 1),
    ("amount_coef", "i8", 1), ("balance_coef", "i8", 1), ("amount_exp", "i4", 1), ("balance_exp", "i4", 1),
    ("dec_flags", "u1", 1), ("date", "i8", 1), ("card", "u1", 16), ("card_len", "u1", 1),
    ("survivor", "u1", 1), ("first_seen", "u1", 1), ("last_idx", "i8", 1),
    ("amount_cents", "i8", 1), ("balance_cents", "i8", 1),
)


# struct smsgate_record as a numpy structured dtype (64 bytes, no padding)
RECORD_FIELDS = [("msg_id", "u1", (20,)), ("status", "u1"), ("flags", "u1"), ("dec_flags", "u1"), ("card_len", "u1"),
                 ("slot", "<i4"), ("amount_exp", "<i4"), ("balance_exp", "<i4"), ("body_date", "<i4"),
                 ("amount_coef", "<i8"), ("balance_coef", "<i8"), ("date", "<i8")]
RECORD_BYTES_OUT = 64


class Outputs(C.Structure):
    _fields_ = [(name, C.c_void_p) for name, _, _ in COLUMNS]


class LegacyOutputs(C.Structure):
    """struct smsgate_legacy_outputs"""
    _fields_ = [(n, C.c_void_p) for n in ("status", "lflags", "strip", "caps", "amount", "balance", "body_hash")]


class EmitInput(C.Structure):
    """struct smsgate_emit_input"""
    _fields_ = [(n, C.c_void_p) for n in (
        "records", "msg_id", "status", "flags", "dec_flags", "slot", "amount_coef", "balance_coef", "amount_exp",
        "balance_exp", "date", "card", "card_len", "body_offsets", "body_bytes", "fixed_len", "fixed_bytes",
        "sender_offsets", "sender_bytes")] + [("device_id", C.c_char_p)] + [
        (n, C.c_void_p) for n in ("text_offsets", "text_bytes", "text_kind")]


class P2P(C.Structure):
    """struct smsgate_p2p"""
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32),
                ("inbox", C.c_void_p * 64), ("in_count", C.c_void_p * 64),
                ("gfl", C.c_void_p * 64), ("capacity", C.c_int64),
                ("stage", C.c_void_p * 64), ("collide", C.c_void_p * 64), ("bitmap_words", C.c_int64),
                ("flags", C.c_void_p * 64), ("ovr", C.c_void_p * 64), ("ovr_words", C.c_int64), ("gfl_slots", C.c_int64)]


class LibraryMissing(RuntimeError):
    pass


_lib = None


def lib():
    """The loaded shared library.  Raises LibraryMissing when it has not been
    built (`python -m smsgate_b200.build`): there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise LibraryMissing(f"{LIB_PATH} is not built; run `python -m smsgate_b200.build` "
                             "(the bulk-replay path has no CPU implementation)")
    L = C.CDLL(LIB_PATH)
    vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int
    sigs = {
        "smsgate_abi_version": (i32, []),
        "smsgate_create": (i32, [i32, vp, C.POINTER(vp)]),
        "smsgate_destroy": (None, [vp]),
        "smsgate_last_error": (C.c_char_p, [vp]),
        "smsgate_reserve": (i32, [vp, i64, i64]),
        "smsgate_launch_count": (i64, [vp]),
        "smsgate_cache_load": (i32, [vp, i64, vp, vp, vp, vp, vp]),
        "smsgate_cache_size": (i64, [vp]),
        "smsgate_cache_share": (i32, [vp, vp]),
        "smsgate_set_stream": (i32, [vp, vp]),
        "smsgate_set_scan_ctas": (i32, [vp, i32]),
        "smsgate_set_host_pipeline": (i32, [vp, i32, i64, i32]),
        "smsgate_set_phone_dates": (i32, [vp, vp]),
        "smsgate_dcache_open": (i32, [C.c_char_p, C.c_double, C.POINTER(vp), C.c_char_p, i32]),
        "smsgate_dcache_sizes": (i32, [vp, vp, vp, vp, vp, vp]),
        "smsgate_dcache_fill": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
        "smsgate_dcache_close": (None, [vp]),
        "smsgate_strict_date_entries": (i32, [vp, i64, vp, vp, vp]),
        "smsgate_emit_json": (i32, [C.POINTER(EmitInput), i32, i64, vp, vp, vp, i64, vp, vp]),
        "smsgate_decimal_batch": (i32, [vp, i64, vp, vp, vp, vp, vp]),
        "smsgate_strict_date_batch": (i32, [vp, i64, vp, vp, vp]),
        "smsgate_dt_str_batch": (i32, [vp, i64, vp, vp, vp]),
        "smsgate_xml_scan": (i32, [vp, vp, i64, vp, vp, vp, vp]),
        "smsgate_xml_emit": (i32, [vp, vp, vp, vp, vp, vp, vp]),
        "smsgate_xml_release": (i32, [vp]),
        "smsgate_xml_attrs_sizes": (i32, [vp, vp, vp, vp]),
        "smsgate_xml_attrs_emit": (i32, [vp, vp, vp, vp, vp, vp]),
        "smsgate_legacy_parse": (i32, [vp, i64, vp, vp, C.POINTER(LegacyOutputs)]),
        "smsgate_legacy_parse_mode": (i32, [vp, i64, vp, vp, C.c_uint32, C.POINTER(LegacyOutputs)]),
        "smsgate_scan_keys": (i32, [vp, vp, vp, i64, C.POINTER(Outputs)]),
        "smsgate_join_route": (i32, [vp, i64, i64, C.POINTER(Outputs)]),
        "smsgate_dedup": (i32, [vp, i64, i64, C.POINTER(Outputs)]),
        "smsgate_join_route_dedup": (i32, [vp, i64, i64, i64, C.POINTER(Outputs)]),
        "smsgate_device_keys": (i32, [vp, vp, vp, i64, vp]),
        "smsgate_cross_source": (i32, [vp, i64, vp, i64, vp, vp]),
        "smsgate_replay_device": (i32, [vp, vp, vp, i64, i64, i64, C.POINTER(Outputs)]),
        "smsgate_replay_host": (i32, [vp, vp, vp, i64, i64, i64, C.POINTER(Outputs)]),
        "smsgate_replay_host_records": (i32, [vp, vp, vp, i64, i64, i64, vp, vp]),
        "smsgate_replay_host_records_begin": (i32, [vp, vp, vp, i64, i64, i64, vp]),
        "smsgate_replay_host_records_finish": (i32, [vp, vp]),
        "smsgate_replay_host_records8": (i32, [vp, vp, vp, i64, i64, i64, vp, vp]),
        "smsgate_replay_host_records8_finish": (i32, [vp, vp]),
        "smsgate_fetch_columns": (i32, [vp, i64, C.POINTER(Outputs)]),
        "smsgate_replay_host_timeline": (i32, [vp, C.POINTER(C.c_float), i32]),
        "smsgate_replay_host_begin": (i32, [vp, vp, vp, i64, i64, i64, C.POINTER(Outputs)]),
        "smsgate_device_view": (i32, [vp, i64, C.POINTER(Outputs)]),
        "smsgate_replay_host_finish": (i32, [vp, C.POINTER(Outputs)]),
        "smsgate_dedup_export": (i32, [vp, i64, i64, i32, C.POINTER(Outputs), vp, vp, vp]),
        "smsgate_reserve_merge": (i32, [vp, i64]),
        "smsgate_dedup_merge": (i32, [vp, i64, vp, vp]),
        "smsgate_dedup_apply": (i32, [vp, i64, i64, i64, vp, vp, C.POINTER(Outputs)]),
        "smsgate_p2p_reserve": (i32, [vp, i64]),
        "smsgate_p2p_barrier": (i32, [vp, C.POINTER(P2P), i32]),
        "smsgate_p2p_fuse_filter": (i32, [vp, i64]),
        "smsgate_p2p_exchange": (i32, [vp, i64, i64, i32, i32, C.POINTER(Outputs), C.POINTER(P2P)]),
        "smsgate_p2p_reserve_filter": (i32, [vp, i64]),
        "smsgate_p2p_filter_push": (i32, [vp, i64, C.POINTER(Outputs), C.POINTER(P2P)]),
        "smsgate_p2p_filter_reduce": (i32, [vp, C.POINTER(P2P)]),
        "smsgate_p2p_scatter": (i32, [vp, i64, i64, i32, C.POINTER(Outputs), C.POINTER(P2P)]),
        "smsgate_p2p_merge": (i32, [vp, i32, C.POINTER(P2P)]),
        "smsgate_p2p_resolve": (i32, [vp, i64, i64, i32, C.POINTER(Outputs), C.POINTER(P2P), vp]),
        "smsgate_dedup_static": (i32, [i32, vp, i64, i64, vp, vp, vp, vp, vp, C.c_char_p, i32]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(L, name)          # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    if L.smsgate_abi_version() != ABI_VERSION:
        raise LibraryMissing("libsmsgate_b200.so was built from a different header; rebuild it")
    L._exported = tuple(sigs)
    _lib = L
    return L


class SmsgateError(RuntimeError):
    def __init__(self, code: int, text: str):
        super().__init__(f"smsgate_b200 error {code}: {text}")
        self.code = code


def check(ctx, rc: int) -> None:
    if rc != 0:
        msg = lib().smsgate_last_error(ctx)
        raise SmsgateError(rc, msg.decode("utf-8", "replace") if msg else "")
