"""Loader for the Gemini response cache: turns the reference's
`diskcache.Cache(".gemini_cache")` (libs/gemini_parser.py:33) — a mapping
hex(SHA-256(fixed_body)) -> response dict (schema :46-61) — into the flat
arrays `smsgate_cache_load` builds the device hash table from.

What is decided here, once per cache entry at load time, is only what does
not depend on the message: which schema facts hold (the keys the reference
indexes, the types pydantic would accept, libs/llm_core.py:9-19) and, for
date strings outside the strict `dd.mm.yy HH:MM` form the GPU parses, what
the reference's own `strptime` -> `dateutil` chain makes of them
(libs/gemini_parser.py:106-119).  Amounts, balances, cards and strict dates
stay strings and are evaluated per message on the GPU.
"""
from __future__ import annotations

import datetime as _dt
from dataclasses import dataclass, field
from typing import Any, Mapping, Optional

import numpy as np

from . import _lib as L

_TXN_VALUES = ("debit", "credit", "otp", "unknown")
_NINE = ("txn_type", "date", "amount", "currency", "card", "merchant", "city", "address", "balance")
_HEX = set("0123456789abcdef")


def pack_datetime(d: _dt.datetime) -> int:
    return ((d.year << 46) | (d.month << 42) | (d.day << 37) | (d.hour << 32) | (d.minute << 26)
            | (d.second << 20) | d.microsecond)


def unpack_datetime(p: int) -> _dt.datetime:
    p = int(p)
    return _dt.datetime(p >> 46, (p >> 42) & 15, (p >> 37) & 31, (p >> 32) & 31, (p >> 26) & 63,
                        (p >> 20) & 63, p & 0xFFFFF)


_dt_adapter = None


def _host_date(value: Any):
    """The reference's chain for a cached `date` the GPU does not parse:
    strptime, else dateutil, else (date left as it was) what pydantic makes of
    the raw value.  Returns (kind, datetime | None)."""
    global _dt_adapter
    try:
        return "dt", _dt.datetime.strptime(value, "%d.%m.%y %H:%M")
    except Exception:
        pass
    try:
        from dateutil.parser import parse
        return "dt", parse(value)
    except Exception as e:
        if "String does not contain a date" in str(e):
            return "rawdate", None          # falls back to int(raw.date): raises for XML-sourced messages
    try:
        if _dt_adapter is None:
            from pydantic import TypeAdapter
            _dt_adapter = TypeAdapter(_dt.datetime)
        return "nobody", _dt_adapter.validate_python(value)
    except Exception:
        return "bad", None


def _str_or_none_ok(v: Any) -> bool:
    return v is None or type(v) is str or (isinstance(v, str))


@dataclass
class CacheTable:
    """Flat form of the response cache + the dicts for materialising ParsedSMS."""
    keys: np.ndarray                 # uint8 [m, 32]
    val_offsets: np.ndarray          # int32 [4m+1]
    val_bytes: np.ndarray            # uint8
    eflags: np.ndarray               # uint16 [m]
    date_packed: np.ndarray          # int64 [m]
    values: list                     # the cached objects, entry order
    host_dates: dict = field(default_factory=dict)   # entry -> datetime (non-strict dates)
    skipped: int = 0                 # keys that are not 64 lowercase hex chars (can never be probed)

    @property
    def m(self) -> int:
        return len(self.values)


def build_table(mapping: Mapping[str, Any], strict_date_batch) -> CacheTable:
    """`strict_date_batch(list[str]) -> int64 array` is the GPU call
    (ReplayEngine.strict_date_batch); entries it answers -1 for are resolved by
    `_host_date`."""
    keys, values = [], []
    skipped = 0
    for k, v in mapping.items() if hasattr(mapping, "items") else ((k, mapping[k]) for k in mapping):
        if not (isinstance(k, str) and len(k) == 64 and set(k) <= _HEX):
            skipped += 1
            continue
        keys.append(k)
        values.append(v)
    m = len(values)
    eflags = np.zeros(m, dtype=np.uint16)
    date_packed = np.zeros(m, dtype=np.int64)
    parts: list = []
    date_strs: list = []
    date_idx: list = []
    host_dates: dict = {}
    for e, v in enumerate(values):
        f = 0
        amount = balance = date_s = card = ""
        if isinstance(v, dict):
            f |= L.E_IS_DICT
            if "txn_type" in v:
                f |= L.E_HAS_TXN
            ok = all(k in v for k in _NINE)
            if ok:
                t = v["txn_type"]
                ok = (isinstance(t, str) and (t in _TXN_VALUES)) and all(
                    _str_or_none_ok(v[k]) for k in ("currency", "merchant", "city", "address"))
            if ok:
                f |= L.E_SCHEMA_OK
            if "amount" in v:
                amount = str(v["amount"])
            if "balance" in v:
                balance = str(v["balance"])
            c = v.get("card")
            if isinstance(c, str):
                f |= L.E_CARD_STR
                card = c
            mer = v.get("merchant")
            if isinstance(mer, str) and mer:
                f |= L.E_MERCHANT
            if "date" in v:
                f |= L.E_DATE_KEY
                d = v["date"]
                if type(d) is str and d.isascii() and len(d) <= 32:
                    date_s = d
                    date_strs.append(d)
                    date_idx.append(e)
                else:
                    kind, obj = _host_date(d)
                    f |= _apply_host_date(kind, obj, e, date_packed, host_dates)
        eflags[e] = f
        parts.append(amount.encode("utf-8"))
        parts.append(balance.encode("utf-8"))
        parts.append(date_s.encode("utf-8"))
        parts.append(card.encode("utf-8"))
    if date_strs:
        packed = strict_date_batch(date_strs)
        for j, e in enumerate(date_idx):
            if packed[j] > 0:
                eflags[e] |= L.E_DATE_STRICT
                date_packed[e] = packed[j]
            else:
                kind, obj = _host_date(values[e]["date"])
                eflags[e] |= _apply_host_date(kind, obj, e, date_packed, host_dates)
    val_offsets = np.zeros(4 * m + 1, dtype=np.int64)
    if parts:
        np.cumsum(np.fromiter((len(p) for p in parts), dtype=np.int64, count=len(parts)), out=val_offsets[1:])
    if val_offsets[-1] >= 2 ** 31:
        raise ValueError("cache value strings exceed 2 GiB")
    key_arr = (np.frombuffer(bytes.fromhex("".join(keys)), dtype=np.uint8).reshape(m, 32).copy()
               if m else np.zeros((0, 32), dtype=np.uint8))
    t = CacheTable(keys=key_arr, val_offsets=val_offsets.astype(np.int32),
                   val_bytes=np.frombuffer(b"".join(parts), dtype=np.uint8).copy(), eflags=eflags,
                   date_packed=date_packed, values=values, host_dates=host_dates, skipped=skipped)
    t.text = _text_arrays(values)
    return t


def _text_arrays(values) -> tuple:
    """The five host-side strings of every entry (txn_type, currency, merchant, city, address) in the layout
    smsgate_dcache_fill produces: offsets[5m+1], bytes, kind[5m] (0 missing, 1 None, 2 str, 3 other) - what the
    bulk payload emitter (smsgate_emit_json) reads."""
    m = len(values)
    kind = np.zeros(5 * m, np.uint8)
    parts = []
    for e, v in enumerate(values):
        for j, f in enumerate(("txn_type", "currency", "merchant", "city", "address")):
            x = v.get(f, _MISSING) if isinstance(v, dict) else _MISSING
            if x is _MISSING:
                parts.append(b"")
            elif x is None:
                kind[5 * e + j] = 1
                parts.append(b"")
            elif type(x) is str:
                kind[5 * e + j] = 2
                parts.append(x.encode("utf-8", "surrogatepass"))
            else:
                kind[5 * e + j] = 3
                parts.append(b"")
    off = np.zeros(5 * m + 1, np.int64)
    if parts:
        np.cumsum(np.fromiter((len(p) for p in parts), dtype=np.int64, count=len(parts)), out=off[1:])
    return off, np.frombuffer(b"".join(parts), dtype=np.uint8).copy(), kind


_MISSING = object()


def _apply_host_date(kind: str, obj, e: int, date_packed: np.ndarray, host_dates: dict) -> int:
    if kind == "dt" or kind == "nobody":
        date_packed[e] = pack_datetime(obj)
        host_dates[e] = obj
        f = L.E_DATE_HOST if kind == "dt" else L.E_DATE_NOBODY
        if obj.tzinfo is not None:
            f |= L.E_DATE_TZ
        return f
    if kind == "rawdate":
        return L.E_DATE_PHONE
    return 0


def open_diskcache(path: str) -> Mapping[str, Any]:
    """The on-disk cache as the reference opens it (diskcache 5.6.3, pickled dicts)."""
    import diskcache
    return diskcache.Cache(path)


# diskcache 5.6.3 storage modes (core.py: MODE_RAW .. MODE_PICKLE)
_MODE_RAW, _MODE_BINARY, _MODE_TEXT, _MODE_PICKLE = 1, 2, 3, 4


def read_diskcache_sqlite(path: str, now: Optional[float] = None) -> dict:
    """the whole `.gemini_cache` in ONE query instead of two queries + one unpickle per
    probe (libs/gemini_parser.py:208-209 through diskcache.Cache.__contains__ / __getitem__).

    Reads diskcache's SQLite file directly — table `Cache(rowid, key, raw, store_time, expire_time,
    access_time, access_count, tag, size, mode, filename, value)` — and returns `{key: value}` in rowid
    order (diskcache's iteration order) for exactly the entries `key in cache` would answer True for:
    string keys (stored as TEXT with raw = 1) whose expire_time is NULL or in the future.  Values are
    decoded by mode the way `Disk.fetch` does: inline pickles, raw values, and pickles / blobs that
    diskcache moved to files because they exceed 32 KiB.  The file is opened read-only."""
    import os
    import pickle
    import sqlite3
    import time
    db = os.path.join(path, "cache.db")
    if not os.path.exists(db):
        raise FileNotFoundError(db)
    now = time.time() if now is None else now
    con = sqlite3.connect(f"file:{db}?mode=ro", uri=True, timeout=60)
    try:
        rows = con.execute(
            "SELECT key, mode, filename, value FROM Cache "
            "WHERE raw = 1 AND typeof(key) = 'text' AND (expire_time IS NULL OR expire_time > ?) ORDER BY rowid",
            (now,)).fetchall()
    finally:
        con.close()
    out = {}
    loads = pickle.loads
    for key, mode, filename, value in rows:
        if mode == _MODE_PICKLE and filename is None:
            out[key] = loads(value)
        elif mode == _MODE_RAW:
            out[key] = bytes(value) if isinstance(value, (bytes, memoryview)) else value
        elif mode == _MODE_PICKLE:
            with open(os.path.join(path, filename), "rb") as f:
                out[key] = pickle.load(f)
        elif mode == _MODE_BINARY:
            with open(os.path.join(path, filename), "rb") as f:
                out[key] = f.read()
        elif mode == _MODE_TEXT:
            with open(os.path.join(path, filename), "r", encoding="UTF-8") as f:
                out[key] = f.read()
    return out


# --------------------------------------------------------------------------- native loader
class LazyValues:
    """`CacheTable.values` of a natively loaded cache: the response dict of entry e, unpickled from
    `cache.db` by rowid only when somebody materialises a ParsedSMS from it (libs/gemini_parser.py:249-266
    needs currency / merchant / city / address, which never travel to the GPU)."""

    def __init__(self, path: str, rowids: np.ndarray):
        self.path = path
        self.rowids = rowids
        self._con = None
        self._memo: dict = {}

    def __len__(self) -> int:
        return len(self.rowids)

    def _fetch(self, e: int):
        import os
        import pickle
        import sqlite3
        if self._con is None:
            self._con = sqlite3.connect(f"file:{os.path.join(self.path, 'cache.db')}?mode=ro", uri=True, timeout=60)
        mode, filename, value = self._con.execute("SELECT mode, filename, value FROM Cache WHERE rowid = ?",
                                                  (int(self.rowids[e]),)).fetchone()
        if mode == _MODE_PICKLE and filename is None:
            return pickle.loads(value)
        if mode == _MODE_RAW:
            return bytes(value) if isinstance(value, (bytes, memoryview)) else value
        with open(os.path.join(self.path, filename), "rb" if mode != _MODE_TEXT else "r",
                  **({} if mode != _MODE_TEXT else {"encoding": "UTF-8"})) as f:
            return pickle.load(f) if mode == _MODE_PICKLE else f.read()

    def __getitem__(self, e: int):
        e = int(e)
        if e not in self._memo:
            if len(self._memo) > 65536:
                self._memo.clear()
            self._memo[e] = self._fetch(e)
        return self._memo[e]


def read_diskcache_native(path: str, engine, now: Optional[float] = None) -> CacheTable:
    """The whole `.gemini_cache` as a CacheTable WITHOUT a Python object per entry: the C++ reader
    (smsgate_dcache_*, csrc/cache_reader.h) walks diskcache's SQLite file and decodes the pickled response
    dicts into the flat arrays; the date strings are classified on the GPU (smsgate_strict_date_entries).
    Only what the decoder declines - values kept in side files, float amounts, nested containers, dates the
    strict parser does not take - goes through the per-entry Python code of `build_table`, and is counted in
    `table.host_decoded`."""
    import ctypes as C
    import os
    import time
    lib = L.lib()
    if not os.path.exists(os.path.join(path, "cache.db")):
        raise FileNotFoundError(os.path.join(path, "cache.db"))
    h = C.c_void_p()
    err = C.create_string_buffer(512)
    rc = lib.smsgate_dcache_open(os.fsencode(path), float(time.time() if now is None else now), C.byref(h), err, len(err))
    if rc != 0:
        raise L.SmsgateError(rc, err.value.decode("utf-8", "replace"))
    try:
        m, vb, tb, skipped, nhost = (C.c_int64(0) for _ in range(5))
        lib.smsgate_dcache_sizes(h, C.byref(m), C.byref(vb), C.byref(tb), C.byref(skipped), C.byref(nhost))
        m = m.value
        keys = np.zeros((m, 32), np.uint8)
        val_offsets = np.zeros(4 * m + 1, np.int32)
        val_bytes = np.zeros(vb.value + 16, np.uint8)
        eflags = np.zeros(m, np.uint16)
        rowids = np.zeros(m, np.int64)
        needs_host = np.zeros(m, np.uint8)
        text_offsets = np.zeros(5 * m + 1, np.int64)
        text_bytes = np.zeros(tb.value + 16, np.uint8)
        text_kind = np.zeros(5 * m, np.uint8)
        p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
        L.check(None, lib.smsgate_dcache_fill(h, p(keys), p(val_offsets), p(val_bytes), p(eflags), p(rowids), p(needs_host),
                                              p(text_offsets), p(text_bytes), p(text_kind)))
    finally:
        lib.smsgate_dcache_close(h)
    values = LazyValues(path, rowids)
    date_packed = np.zeros(m, np.int64)
    host_dates: dict = {}
    # entries the decoder left to Python entirely: the old per-entry code, spliced into the flat arrays
    whole = np.nonzero(needs_host == 1)[0]
    if len(whole):
        sub = build_table({f"{k:064x}": values[int(e)] for k, e in enumerate(whole)}, engine.strict_date_batch)
        val_offsets, val_bytes = _splice_values(val_offsets, val_bytes[:vb.value], whole, sub)
        val_bytes = np.concatenate([val_bytes, np.zeros(16, np.uint8)])
        for k, e in enumerate(whole):
            eflags[e] = sub.eflags[k]
            date_packed[e] = sub.date_packed[k]
            if k in sub.host_dates:
                host_dates[int(e)] = sub.host_dates[k]
    # dates: the strict form on the GPU for every entry at once ...
    if m:
        packed = np.zeros(m, np.int64)
        L.check(engine._ctx, lib.smsgate_strict_date_entries(engine._ctx, m, p(val_offsets), p(val_bytes), p(packed)))
        native = needs_host != 1
        strict = native & (packed > 0)
        eflags[strict] |= L.E_DATE_STRICT
        date_packed[strict] = packed[strict]
        # ... and the reference's strptime -> dateutil -> pydantic chain for the rest that carry a date key
        rest = np.nonzero(native & ~strict & ((eflags & L.E_DATE_KEY) != 0))[0]
        for e in rest:
            kind, obj = _host_date(values[int(e)]["date"])
            eflags[e] |= _apply_host_date(kind, obj, int(e), date_packed, host_dates)
    else:
        rest = whole
    t = CacheTable(keys=keys, val_offsets=val_offsets, val_bytes=val_bytes[:int(val_offsets[-1])].copy(), eflags=eflags,
                   date_packed=date_packed, values=values, host_dates=host_dates, skipped=int(skipped.value))
    t.host_decoded = int(len(whole))
    t.host_dates_resolved = int(len(rest)) if m else 0
    t.text = (text_offsets, text_bytes[:tb.value], text_kind)
    return t


def _splice_values(val_offsets: np.ndarray, val_bytes: np.ndarray, entries: np.ndarray, sub: CacheTable):
    """Give the entries `entries` (whose four strings are empty in the native arrays) the strings of the
    table `sub` built for them: one vectorised rebuild of offsets + bytes."""
    m = (len(val_offsets) - 1) // 4
    lens = np.diff(val_offsets.astype(np.int64))
    sub_lens = np.diff(sub.val_offsets.astype(np.int64))
    idx = (entries[:, None] * 4 + np.arange(4)[None, :]).ravel()
    lens[idx] = sub_lens
    new_off = np.zeros(4 * m + 1, np.int64)
    np.cumsum(lens, out=new_off[1:])
    if new_off[-1] >= 2 ** 31:
        raise ValueError("cache value strings exceed 2 GiB")
    out = np.zeros(int(new_off[-1]), np.uint8)
    # native strings: copy run by run between the spliced entries (entries are few)
    src = 0
    prev = 0
    for k, e in enumerate(entries):
        a, b = int(val_offsets[4 * prev]), int(val_offsets[4 * e])
        out[int(new_off[4 * prev]):int(new_off[4 * prev]) + (b - a)] = val_bytes[a:b]
        sa, sb = int(sub.val_offsets[4 * k]), int(sub.val_offsets[4 * k + 4])
        out[int(new_off[4 * e]):int(new_off[4 * e]) + (sb - sa)] = sub.val_bytes[sa:sb]
        prev = int(e) + 1
    a, b = int(val_offsets[4 * prev]) if prev <= m else 0, int(val_offsets[-1])
    if prev <= m and b > a:
        out[int(new_off[4 * prev]):int(new_off[4 * prev]) + (b - a)] = val_bytes[a:b]
    del src
    return new_off.astype(np.int32), out
