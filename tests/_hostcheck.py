"""TEST ONLY: ctypes wrapper over tests/hostcheck/libhostcheck.so, the CPU build
of the kernels' __host__ __device__ cores (see hostcheck.cpp).  It exists so the
device logic can be checked in a container without a GPU; the product never
loads it."""
import ctypes as C

import numpy as np

from smsgate_b200 import build
from smsgate_b200.batch import ragged_from_strings

_lib = None
SLACK = 8        # readable bytes past the last byte of a byte buffer: what include/smsgate_b200.h asks callers for


def lib():
    global _lib
    if _lib is None:
        import os
        # tests/test_host_cpp_sanitized.py points this at a build with AddressSanitizer + UBSan
        _lib = C.CDLL(os.environ.get("SMSGATE_HOSTCHECK_LIB") or build.build_hostcheck())
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def scan(bodies):
    data, offsets = ragged_from_strings(bodies)
    n = len(bodies)
    cols = dict(msg_id=np.zeros((n, 20), np.uint8), cache_key=np.zeros((n, 32), np.uint8),
                flags=np.zeros(n, np.uint8), body_date=np.zeros(n, np.int32), fixed_len=np.zeros(n, np.int32),
                fixed_bytes=np.zeros(max(len(data), 1), np.uint8))
    data = np.concatenate([data, np.zeros(SLACK, np.uint8)])
    rc = lib().hc_scan_keys(_p(data), _p(offsets), C.c_int64(n), _p(cols["msg_id"]), _p(cols["cache_key"]),
                            _p(cols["flags"]), _p(cols["body_date"]), _p(cols["fixed_len"]), _p(cols["fixed_bytes"]))
    assert rc > -200000, f"sg_hash_pass disagrees with the plain digests (code {rc})"
    assert rc == 0, f"first pass + full pass disagree with the full pass at body {-100 - rc}: {bodies[-100 - rc]!r}"
    cols["offsets"] = offsets
    return cols


def md5(bodies):
    data, offsets = ragged_from_strings(bodies)
    n = len(bodies)
    out = np.zeros((n, 16), np.uint8)
    data = np.concatenate([data, np.zeros(SLACK, np.uint8)])
    lib().hc_md5(_p(data), _p(offsets), C.c_int64(n), _p(out))
    return out


def decimal(strings):
    data, offsets = ragged_from_strings(strings, np.int32)
    n = len(strings)
    coef, exp, flags = np.zeros(n, np.int64), np.zeros(n, np.int32), np.zeros(n, np.uint8)
    data = np.concatenate([data, np.zeros(SLACK, np.uint8)])
    lib().hc_decimal(C.c_int64(n), _p(offsets), _p(data), _p(coef), _p(exp), _p(flags))
    return coef, exp, flags


def strict_date(strings):
    data, offsets = ragged_from_strings(strings, np.int32)
    n = len(strings)
    out = np.zeros(n, np.int64)
    data = np.concatenate([data, np.zeros(SLACK, np.uint8)])
    lib().hc_strict_date(C.c_int64(n), _p(offsets), _p(data), _p(out))
    return out


def dt_str(strings):
    data, offsets = ragged_from_strings(strings, np.int32)
    out = np.zeros(len(strings), np.int64)
    data = np.concatenate([data, np.zeros(SLACK, np.uint8)])
    lib().hc_dt_str(C.c_int64(len(strings)), _p(offsets), _p(data), _p(out))
    return out


def join_route(scan_cols, table, now_packed, phone_dt=None):
    n = len(scan_cols["flags"])
    cols = dict(slot=np.zeros(n, np.int32), status=np.zeros(n, np.uint8), amount_coef=np.zeros(n, np.int64),
                balance_coef=np.zeros(n, np.int64), amount_exp=np.zeros(n, np.int32),
                balance_exp=np.zeros(n, np.int32), dec_flags=np.zeros(n, np.uint8), date=np.zeros(n, np.int64),
                card=np.zeros((n, 16), np.uint8), card_len=np.zeros(n, np.uint8),
                amount_cents=np.zeros(n, np.int64), balance_cents=np.zeros(n, np.int64))
    keys = np.ascontiguousarray(table.keys)
    vb = np.concatenate([table.val_bytes, np.zeros(SLACK, np.uint8)])
    lib().hc_join_route(C.c_int64(n), _p(scan_cols["cache_key"]), _p(scan_cols["flags"]), _p(scan_cols["body_date"]),
                        C.c_int64(now_packed), C.c_int64(table.m), _p(keys), _p(table.val_offsets), _p(vb),
                        _p(table.eflags), _p(table.date_packed), _p(cols["slot"]), _p(cols["status"]),
                        _p(cols["amount_coef"]), _p(cols["balance_coef"]), _p(cols["amount_exp"]),
                        _p(cols["balance_exp"]), _p(cols["dec_flags"]), _p(cols["date"]), _p(cols["card"]),
                        _p(cols["card_len"]), _p(cols["amount_cents"]), _p(cols["balance_cents"]),
                        _p(np.ascontiguousarray(phone_dt, dtype=np.int64)) if phone_dt is not None else None)
    return cols


def legacy(bodies, mode=0):
    data, offsets = ragged_from_strings(bodies)
    n = len(bodies)
    cols = dict(status=np.zeros(n, np.uint8), lflags=np.zeros(n, np.uint8), strip=np.zeros((n, 2), np.int32),
                caps=np.full((n, 20), -1, np.int32), amount=np.zeros(n, np.float64), balance=np.zeros(n, np.float64),
                body_hash=np.zeros((n, 32), np.uint8))
    data = np.concatenate([data, np.zeros(SLACK, np.uint8)])
    lib().hc_legacy(C.c_int64(n), _p(offsets), _p(data), _p(cols["status"]), _p(cols["lflags"]), _p(cols["strip"]),
                    _p(cols["caps"]), _p(cols["amount"]), _p(cols["balance"]), _p(cols["body_hash"]), C.c_uint32(mode))
    return cols


def xml_extract(data: bytes):
    """(code, bodies, senders, date_ms, mflags) from the CPU build of the XML tokenizer logic."""
    x = np.frombuffer(data + b"\0" * 16, dtype=np.uint8)
    nx = len(data)
    cap = nx + 16
    n = C.c_int64(0)
    bb, sb = np.zeros(cap, np.uint8), np.zeros(cap, np.uint8)
    bo, so, dm = np.zeros(cap // 8 + 2, np.int64), np.zeros(cap // 8 + 2, np.int64), np.zeros(cap // 8 + 2, np.int64)
    mf = np.zeros(cap // 8 + 2, np.uint8)
    rc = lib().hc_xml_extract(_p(x), C.c_int64(nx), C.byref(n), _p(bb), _p(bo), _p(sb), _p(so), _p(dm), _p(mf))
    k = n.value
    bodies = [bb[bo[i]:bo[i + 1]].tobytes().decode() for i in range(k)]
    senders = [sb[so[i]:so[i + 1]].tobytes().decode() for i in range(k)]
    return rc, bodies, senders, dm[:k].tolist(), mf[:k].tolist()


class HostTokenizerEngine:
    """TEST stand-in for ReplayEngine where only the XML host logic is under test: `xml_extract` runs the
    tokenizer compiled for the CPU (same xml_core.cuh), `msg_ids` is hashlib."""

    def __init__(self):
        self.host_fallbacks = 0

    def xml_extract(self, data):
        import xml.etree.ElementTree as ET

        import torch

        from smsgate_b200.batch import ragged_from_strings
        from smsgate_b200.replay import UnsupportedXML
        code, bodies, senders, dates, mflags = xml_extract(bytes(data))
        if code == 2:
            raise UnsupportedXML("outside the tokenizer's dialect")
        if code == 1:
            raise ET.ParseError("not well-formed")
        bb, bo = ragged_from_strings(bodies)
        sb, so = ragged_from_strings(senders)

        def t(a, dtype):
            a = np.array(a, dtype=dtype)
            return torch.from_numpy(a if a.size else np.zeros(1, dtype))
        return {"n": len(bodies), "body_offsets": t(bo, np.int64), "sender_offsets": t(so, np.int64),
                "body_bytes": t(bb, np.uint8), "sender_bytes": t(sb, np.uint8), "mflags": t(mflags, np.uint8),
                "date_ms": t([d if -2 ** 63 <= d < 2 ** 63 else 0 for d in dates], np.int64)}

    def msg_ids(self, batch):
        import hashlib
        return np.array([np.frombuffer(hashlib.sha1(batch.body(i).encode("utf-8")).digest(), np.uint8)
                         for i in range(batch.n)]).reshape(batch.n, 20)
