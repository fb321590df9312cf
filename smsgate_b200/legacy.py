"""Drop-in for the reference's legacy offline replay (read_xml.py, process_cached.py).

`parse_transaction_message` / `parse_credit_message` / `process_sms_from_cache` keep
the reference's names, arguments and return shapes (process_cached.py:150-268);
the per-message work — strip, OTP filter, both regular expressions, the float
conversion and the SHA-256 of the stripped body — runs on the GPU
(smsgate_legacy_parse), whole cache at a time.  DiskCache reads and writes stay
on the host: that is storage, not arithmetic.
"""
from __future__ import annotations

import ctypes as C
from typing import Any, Dict, List, Optional, Sequence

import numpy as np

from . import _lib as L
from .batch import ragged_from_strings

DEBIT_GROUPS = dict(merchant=1, city=2, address=3, date=4, time=5, card=6, amount=7, currency=8, balance=9)
CREDIT_GROUPS = dict(type=1, date=2, time=3, card=4, amount=5, currency=6, balance=7)


class LegacyResult:
    """Columns of smsgate_legacy_parse for a batch of bodies + dict materialisation."""

    def __init__(self, bodies: Sequence[str], data: np.ndarray, offsets: np.ndarray, cols: dict):
        self.bodies, self.data, self.offsets, self.cols = bodies, data, offsets, cols

    def _span(self, i: int, lo: int, hi: int) -> str:
        base = int(self.offsets[i]) + int(self.cols["strip"][i, 0])
        return self.data[base + lo:base + hi].tobytes().decode("utf-8")

    def stripped(self, i: int) -> str:
        a = int(self.offsets[i])
        return self.data[a + int(self.cols["strip"][i, 0]):a + int(self.cols["strip"][i, 1])].tobytes().decode("utf-8")

    def group(self, i: int, g: int) -> Optional[str]:
        lo, hi = int(self.cols["caps"][i, 2 * g]), int(self.cols["caps"][i, 2 * g + 1])
        return None if lo < 0 else self._span(i, lo, hi)

    def body_hash_hex(self, i: int) -> str:
        return self.cols["body_hash"][i].tobytes().hex()

    def parsed(self, i: int) -> Optional[Dict[str, Any]]:
        """The dict parse_transaction_message / parse_credit_message would return."""
        st = int(self.cols["status"][i])
        if int(self.cols["lflags"][i]) & L.LEG_F_BUDGET:
            raise RuntimeError(f"message {i}: the matcher ran out of its step budget")
        if st == L.LEG_DEBIT:
            g = DEBIT_GROUPS
            if int(self.cols["lflags"][i]) & L.LEG_F_INEXACT:
                raise OverflowError(f"message {i}: amount or balance has more than 15 significant digits")
            return {"direction": "debit", "merchant": self.group(i, g["merchant"]),
                    "city": (self.group(i, g["city"]) or "ONLINE").upper(),
                    "address": self.group(i, g["address"]) or "N/A", "date": self.group(i, g["date"]),
                    "time": self.group(i, g["time"]), "card": "***" + self.group(i, g["card"]),
                    "amount": float(self.cols["amount"][i]), "currency": self.group(i, g["currency"]),
                    "balance": float(self.cols["balance"][i])}
        if st == L.LEG_CREDIT:
            g = CREDIT_GROUPS
            if int(self.cols["lflags"][i]) & L.LEG_F_INEXACT:
                raise OverflowError(f"message {i}: amount or balance has more than 15 significant digits")
            return {"direction": "credit", "type": self.group(i, g["type"]).upper(), "date": self.group(i, g["date"]),
                    "time": self.group(i, g["time"]), "card": "***" + self.group(i, g["card"]),
                    "amount": float(self.cols["amount"][i]), "currency": self.group(i, g["currency"]),
                    "balance": float(self.cols["balance"][i])}
        return None


def parse_bodies(engine, bodies: Sequence[Optional[str]], mode: int = 0) -> LegacyResult:
    """Run the legacy per-message path on the GPU for a list of bodies (None counts as "").  mode 0 is what
    process_sms_from_cache does per entry; the LEG_MODE_* bits give the single-message functions."""
    texts = [b or "" for b in bodies]
    data, offsets = ragged_from_strings(texts)
    n = len(texts)
    cols = dict(status=np.zeros(n, np.uint8), lflags=np.zeros(n, np.uint8), strip=np.zeros((n, 2), np.int32),
                caps=np.full((n, 2 * L.LEG_GROUPS), -1, np.int32), amount=np.zeros(n, np.float64),
                balance=np.zeros(n, np.float64), body_hash=np.zeros((n, 32), np.uint8))
    out = L.LegacyOutputs(**{k: v.ctypes.data for k, v in cols.items()})
    L.check(engine._ctx, L.lib().smsgate_legacy_parse_mode(engine._ctx, n, offsets.ctypes.data_as(C.c_void_p),
                                                           data.ctypes.data_as(C.c_void_p) if len(data) else None,
                                                           int(mode), C.byref(out)))
    return LegacyResult(texts, data, offsets, cols)


def parse_one(engine, message: str, credit: bool) -> Optional[Dict[str, Any]]:
    """parse_transaction_message / parse_credit_message (process_cached.py:150-204) for one message: ONE
    pattern applied to the message as given - no strip(), no OTP filter, no falling through to the other
    pattern.  An amount the float conversion cannot carry exactly raises OverflowError (representation
    limit of the kernel: beyond 2^64 / 22 fraction digits)."""
    if not message:
        return None
    mode = L.LEG_MODE_NO_STRIP | L.LEG_MODE_NO_OTP | (L.LEG_MODE_NO_DEBIT if credit else L.LEG_MODE_NO_CREDIT)
    res = parse_bodies(engine, [message], mode)
    want = L.LEG_CREDIT if credit else L.LEG_DEBIT
    return res.parsed(0) if int(res.cols["status"][0]) == want else None


def parse_transaction_message(engine, message: str) -> Optional[Dict[str, Any]]:
    """process_cached.parse_transaction_message (process_cached.py:150-177) for one message."""
    return parse_one(engine, message, credit=False)


def parse_credit_message(engine, message: str) -> Optional[Dict[str, Any]]:
    """process_cached.parse_credit_message (process_cached.py:180-204) for one message."""
    return parse_one(engine, message, credit=True)


def process_sms_from_cache(engine, source_dir: str, purchase_dest_dir: str, credit_dest_dir: str) -> Dict[str, int]:
    """process_cached.process_sms_from_cache (process_cached.py:210-268): same three DiskCache directories,
    same keys, values and status write-back; returns the four counters the reference only logs."""
    from diskcache import Cache
    stats = {"processed_debit": 0, "processed_credit": 0, "failed": 0, "skipped": 0}
    with Cache(source_dir) as source, Cache(purchase_dest_dir) as purchase, Cache(credit_dest_dir) as credit:
        keys, entries = [], []
        for key in list(source):                                          # :227 rowid order
            data = source.get(key)
            if data.get("status") == "processed":                         # :229-231
                stats["skipped"] += 1
                continue
            keys.append(key)
            entries.append(data)
        res = parse_bodies(engine, [e.get("body") for e in entries])
        for i, (key, data) in enumerate(zip(keys, entries)):
            st = int(res.cols["status"][i])
            if st == L.LEG_SKIPPED_OTP:                                   # :236-240
                data["status"] = "skipped_otp"
                stats["skipped"] += 1
            elif st in (L.LEG_DEBIT, L.LEG_CREDIT):                       # :243-257
                parsed = res.parsed(i)
                parsed["msg_id"] = key
                parsed["original_body"] = res.stripped(i)
                (purchase if st == L.LEG_DEBIT else credit).set(res.body_hash_hex(i), parsed)
                stats["processed_debit" if st == L.LEG_DEBIT else "processed_credit"] += 1
                data["status"] = "processed"
            else:                                                         # :259-261
                data["status"] = "failed_to_parse"
                stats["failed"] += 1
            source.set(key, data)                                         # :263
    return stats


# ---------------------------------------------------------------------------
# save_to_pocketbase.py:34-78 — PocketBase record builders over the legacy caches, in bulk
# ---------------------------------------------------------------------------
def dt_str_batch(engine, pairs: Sequence[tuple]) -> list:
    """`_dt_str(date, time_)` (save_to_pocketbase.py:34-43) for a list of (date, time_) pairs: the six
    strptime formats are evaluated on the GPU (smsgate_dt_str_batch), the text is laid out here the way
    `strftime("%Y-%m-%d %H:%M:%S")` does on glibc (the year is not zero-padded)."""
    import ctypes as C

    import numpy as np

    from . import _lib as L
    from .batch import ragged_from_strings
    from .gemini_cache import unpack_datetime
    strings = [f"{d} {t}" for d, t in pairs]
    if not strings:
        return []
    data, offsets = ragged_from_strings(strings, np.int32, errors="surrogatepass")
    packed = np.zeros(len(strings), dtype=np.int64)
    data = np.concatenate([data, np.zeros(16, np.uint8)])
    L.check(engine._ctx, L.lib().smsgate_dt_str_batch(engine._ctx, len(strings), offsets.ctypes.data_as(C.c_void_p),
                                                      data.ctypes.data_as(C.c_void_p),
                                                      packed.ctypes.data_as(C.c_void_p)))
    out = []
    for p in packed:
        if p == 0:
            out.append(None)
        else:
            d = unpack_datetime(int(p))
            out.append(f"{d.year}-{d.month:02d}-{d.day:02d} {d.hour:02d}:{d.minute:02d}:00")
    return out


def build_sms_data(engine, records: Sequence[Dict[str, Any]]) -> list:
    """`_build_sms_data` (save_to_pocketbase.py:46-62) for every record of the purchase cache: the payload
    for the `sms_data` collection, or None where the date does not parse."""
    dts = dt_str_batch(engine, [(r.get("date", ""), r.get("time", "")) for r in records])
    out = []
    for r, pb_dt in zip(records, dts):
        out.append(None if not pb_dt else {
            "merchant": r.get("merchant"), "city": r.get("city"), "address": r.get("address"), "datetime": pb_dt,
            "card": r.get("card"), "amount": str(r.get("amount", 0.0)), "currency": r.get("currency"),
            "balance": str(r.get("balance", 0.0)), "msg_id": r.get("msg_id"), "original_body": r.get("original_body")})
    return out


def build_transactions(engine, records: Sequence[Dict[str, Any]]) -> list:
    """`_build_transactions` (save_to_pocketbase.py:65-78) for every record of the credit cache."""
    dts = dt_str_batch(engine, [(r.get("date", ""), r.get("time", "")) for r in records])
    out = []
    for r, pb_dt in zip(records, dts):
        out.append(None if not pb_dt else {
            "transaction_id": r.get("msg_id"), "transaction_type": r.get("type", r.get("direction")),
            "amount": r.get("amount"), "currency": r.get("currency"), "balance_after": r.get("balance"),
            "timestamp": pb_dt, "status": "parsed"})
    return out
