"""Bulk replay engine: the Python face of libsmsgate_b200.so.

One `ReplayEngine` = one GPU = one C-ABI context.  It does what the reference
does to every message of a backup between `_iter_sms` and the PocketBase /
Postgres upsert (services/xml_watcher/watcher.py:35-54,
services/parser_worker/worker.py:112-189, libs/gemini_parser.py:193-271,
services/pb_writer/writer.py:65-84) for a whole batch at once, and hands the
results back either as flat columns or, lazily, as the reference's own
`ParsedSMS` objects.

torch is used for device memory and streams only.
"""
from __future__ import annotations

import ctypes as C
import datetime as _dt
from decimal import Decimal
from typing import Any, Mapping, Optional, Sequence

import numpy as np

from . import _lib as L
from .batch import Batch, ragged_from_strings
from .gemini_cache import CacheTable, build_table, pack_datetime, unpack_datetime
from .models import BrokenMessage, ParsedSMS


class UnsupportedXML(RuntimeError):
    """The backup uses an XML construct outside the dialect the GPU tokenizer covers
    (DOCTYPE, CDATA, a non-UTF-8 encoding, '<' inside a comment)."""


class CacheMiss(RuntimeError):
    """The body is not in the Gemini response cache; the reference would call
    Gemini here (libs/gemini_parser.py:210-222), which is out of scope offline."""


def _np_ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class _RecordColumns:
    """The `cols` mapping of a ReplayResult that came back as packed records
    (smsgate_replay_host_records / _records8).  Three kinds of column:
      - fields of the 64-byte record: strided views of the record array;
      - the dedup verdicts: decoded on first use from the verdict array - one int64 per message
        (-1, or last_idx << 2 | first_seen << 1 | survivor) or one byte per message (0, or
        0x80 | first_seen << 1 | survivor; `last_idx` then stays on the device);
      - what the record leaves out (card, cache_key, fixed body, cents): still on the device until the
        engine's next replay, fetched on first use (smsgate_fetch_columns)."""

    _VERDICT = ("survivor", "first_seen", "last_idx")
    _ON_DEVICE = ("card", "cache_key", "fixed_len", "fixed_bytes", "amount_cents", "balance_cents")

    def __init__(self, records: np.ndarray, verdict: np.ndarray, engine=None, generation: int = 0, total: int = 0):
        self.records = records
        self.verdict = verdict
        self._engine = engine
        self._generation = generation
        self._total = total
        self._decoded: dict = {}
        self._fetched: dict = {}

    @property
    def _compact(self) -> bool:
        return self.verdict.dtype == np.uint8

    def keys(self):
        return tuple(self.records.dtype.names) + self._VERDICT + self._ON_DEVICE

    def __iter__(self):
        return iter(self.keys())

    def __contains__(self, name) -> bool:
        return name in self.keys()

    def _fetch(self, names) -> None:
        """Bring columns that were left on the device to the host (once)."""
        names = [c for c in names if c not in self._fetched]
        if not names:
            return
        eng = self._engine
        if eng is None or eng._generation != self._generation:
            raise RuntimeError(f"columns {names} of this result were left on the device and a later replay has "
                               "replaced them; read them (or call ReplayResult.detach()) before replaying again, "
                               "or use ReplayEngine.replay(columns=...)")
        if len(self.records) == 0:
            self._fetched.update(eng._host_columns(0, 0, tuple(names), pinned=False))
        else:
            self._fetched.update(eng._fetch_columns(len(self.records), self._total, names))

    def __getitem__(self, name):
        if name in self.records.dtype.names:
            return self.records[name]
        if name in self._ON_DEVICE:
            self._fetch(self._ON_DEVICE)                    # all of them in one call: they are wanted together
            return self._fetched[name]
        if name not in self._VERDICT:
            raise KeyError(name)
        if name == "last_idx" and self._compact:
            self._fetch(["last_idx"])
            return self._fetched["last_idx"]
        if name not in self._decoded:
            v = self.verdict
            if self._compact:
                self._decoded["survivor"] = (v & 1).astype(np.uint8)
                self._decoded["first_seen"] = ((v >> 1) & 1).astype(np.uint8)
            else:
                up = v >= 0
                self._decoded["survivor"] = (up & ((v & 1) != 0)).astype(np.uint8)
                self._decoded["first_seen"] = (up & ((v & 2) != 0)).astype(np.uint8)
                self._decoded["last_idx"] = np.where(up, v >> 2, -1).astype(np.int64)
        return self._decoded[name]

    @property
    def nbytes(self) -> int:
        """Bytes that crossed PCIe for this result (records + verdicts)."""
        return int(self.records.nbytes + self.verdict.nbytes)

    def detach(self) -> None:
        """Own copies of everything: the records and verdicts alias the engine's pinned staging buffer and
        the remaining columns live on the device, both only until the engine's next replay."""
        self._fetch(self._ON_DEVICE + (("last_idx",) if self._compact else ()))
        self.records = self.records.copy()
        self.verdict = self.verdict.copy()
        self._decoded = {}


class ReplayResult:
    """Columns of one replayed batch (numpy, host) + lazy ParsedSMS views."""

    def __init__(self, batch: Batch, cols: dict, table: Optional[CacheTable], base_index: int):
        self.batch = batch
        self.cols = cols
        self.table = table
        self.base_index = base_index

    def __getattr__(self, name):
        cols = self.__dict__.get("cols", {})
        if name in cols:
            return cols[name]
        raise AttributeError(name)

    @property
    def n(self) -> int:
        return self.batch.n

    def msg_id_hex(self, i: int) -> str:
        return self.cols["msg_id"][i].tobytes().hex()

    def cache_key_hex(self, i: int) -> str:
        return self.cols["cache_key"][i].tobytes().hex()

    def fixed_body(self, i: int) -> str:
        """fixed_body of message i (libs/gemini_parser.py:202-203) = ParsedSMS.raw_body."""
        a = int(self.batch.body_offsets[i])
        if self.cols["flags"][i] & L.F_MODIFIED:
            return self.cols["fixed_bytes"][a:a + int(self.cols["fixed_len"][i])].tobytes().decode("utf-8")
        return self.batch.body(i)

    def decimal(self, which: str, i: int) -> Decimal:
        fl = int(self.cols["dec_flags"][i])
        err = (fl & 3) if which == "amount" else ((fl >> 2) & 3)
        sign = (fl >> 4) & 1 if which == "amount" else (fl >> 5) & 1
        if err == L.DEC_SYNTAX:
            raise ValueError(f"{which} of message {i} is not a number")
        if err == L.DEC_OVERFLOW:
            raise OverflowError(f"{which} of message {i} has more than 18 significant digits")
        coef = int(self.cols[which + "_coef"][i])
        return Decimal((sign, tuple(int(ch) for ch in str(coef)), int(self.cols[which + "_exp"][i])))

    def datetime(self, i: int) -> Optional[_dt.datetime]:
        p = int(self.cols["date"][i])
        if p == 0:
            return None
        e = int(self.cols["slot"][i])
        if self.table is not None and e in self.table.host_dates and not (self.cols["flags"][i] & L.F_BODY_DATE):
            return self.table.host_dates[e]        # keeps tzinfo / exact object of non-strict dates
        return unpack_datetime(p)

    def parsed(self, i: int, sender: Optional[str] = None, device_id: Optional[str] = "xml_backup",
               msg_id: Optional[str] = None) -> Optional[ParsedSMS]:
        """What `parse_sms_llm` returns (or raises) for message i."""
        st = int(self.cols["status"][i])
        if st == L.ST_CACHE_MISS:
            raise CacheMiss(self.cache_key_hex(i))
        if st == L.ST_BROKEN:
            raise BrokenMessage("no card number in the message")
        if st in (L.ST_SKIPPED_FILTER, L.ST_UNMATCHED):
            return None
        e = int(self.cols["slot"][i])
        v = self.table.values[e]
        if st == L.ST_FAILED_EXC and not isinstance(v, dict):
            raise TypeError("the cached response is not a mapping")       # resp_data['txn_type'] on a non-dict (:239)
        if st == L.ST_FAILED_EXC and "txn_type" not in v:
            raise KeyError("txn_type")
        if sender is None:
            sender = self.batch.senders[i] if self.batch.senders is not None else ""
        address = v.get("address")
        return ParsedSMS(
            msg_id=msg_id if msg_id is not None else self.msg_id_hex(i), device_id=device_id, sender=sender,
            date=self.datetime(i), raw_body=self.fixed_body(i), txn_type=v["txn_type"],
            amount=self.decimal("amount", i), currency=v.get("currency"),
            card=self.cols["card"][i, :int(self.cols["card_len"][i])].tobytes().decode("utf-8"),
            merchant=v.get("merchant"), city=v.get("city"), address="" if address == "null" else address,
            balance=self.decimal("balance", i), parser_version="llm-0.2.0")

    def detach(self) -> "ReplayResult":
        """Make the result independent of the engine (see _RecordColumns.detach); a no-op for `replay`
        without pinned buffers."""
        if hasattr(self.cols, "detach"):
            self.cols.detach()
        return self

    def status_counts(self) -> dict:
        c = np.bincount(self.cols["status"], minlength=9)
        return {L.STATUS_NAMES[k]: int(c[k]) for k in range(9)}

    def dlq_indices(self) -> np.ndarray:
        """Messages the reference would publish to sms.failed (libs/nats_utils.py:28)."""
        return np.nonzero(np.isin(self.cols["status"], L.DLQ_STATUSES))[0]

    def survivor_indices(self) -> np.ndarray:
        return np.nonzero(self.cols["survivor"])[0]

    def upserts(self) -> dict:
        """msg_id hex -> ParsedSMS of the record that remains after the replay."""
        return {self.msg_id_hex(int(i)): self.parsed(int(i)) for i in self.survivor_indices()}

    def emit_json(self, indices: np.ndarray, mode: int = 0, device_id: Optional[str] = "xml_backup"):
        """Bytes for the messages `indices`, assembled natively from the columns and the cache texts
        (smsgate_emit_json): mode 0 = sms.parsed payloads (`ParsedSMS.model_dump_json()`), mode 1 = PocketBase
        records as compact JSON.  Returns (offsets int64[k+1], bytes uint8, fallback uint8[k]); a row flagged in
        `fallback` has an empty span and is rendered by the caller through pydantic."""
        t = self.table
        idx = np.ascontiguousarray(indices, dtype=np.int64)
        text = getattr(t, "text", None)
        if text is None:
            raise RuntimeError("the loaded cache carries no text arrays (load it with ReplayEngine.load_cache)")
        cols = self.cols
        inp = L.EmitInput()
        keep = []

        def ptr(a):
            a = np.ascontiguousarray(a)
            keep.append(a)
            return a.ctypes.data
        if hasattr(cols, "records"):
            inp.records = ptr(cols.records)
        else:
            for name in ("msg_id", "status", "flags", "dec_flags", "slot", "amount_coef", "balance_coef", "amount_exp",
                         "balance_exp", "date"):
                setattr(inp, name, ptr(cols[name]))
        inp.card, inp.card_len = ptr(cols["card"]), ptr(cols["card_len"])
        inp.body_offsets = ptr(np.asarray(self.batch.body_offsets, dtype=np.int64))
        inp.body_bytes = ptr(self.batch.body_bytes) if self.batch.body_bytes.size else None
        inp.fixed_len, inp.fixed_bytes = ptr(cols["fixed_len"]), (ptr(cols["fixed_bytes"]) if cols["fixed_bytes"].size else None)
        if getattr(self, "_sender_ragged", None) is None:
            self._sender_ragged = ragged_from_strings(self.batch.senders if self.batch.senders is not None else [""] * self.n)
        sb, so = self._sender_ragged
        inp.sender_offsets, inp.sender_bytes = ptr(so), (ptr(sb) if sb.size else None)
        inp.device_id = device_id.encode("utf-8") if device_id is not None else None
        inp.text_offsets, inp.text_bytes, inp.text_kind = ptr(text[0]), (ptr(text[1]) if text[1].size else None), ptr(text[2])
        k = len(idx)
        offsets = np.zeros(k + 1, np.int64)
        fallback = np.zeros(max(k, 1), np.uint8)
        needed = C.c_int64(0)
        lib = L.lib()
        rc = lib.smsgate_emit_json(C.byref(inp), mode, k, idx.ctypes.data, offsets.ctypes.data, None, 0, None, C.byref(needed))
        if rc:
            raise L.SmsgateError(rc, "emit_json: bad input")
        out = np.zeros(max(needed.value, 1), np.uint8)
        rc = lib.smsgate_emit_json(C.byref(inp), mode, k, idx.ctypes.data, offsets.ctypes.data, out.ctypes.data, out.size,
                                   fallback.ctypes.data, C.byref(needed))
        if rc:
            raise L.SmsgateError(rc, "emit_json failed")
        return offsets, out[:needed.value], fallback[:k]

    def parsed_payloads(self) -> list:
        """The `sms.parsed` payloads of the surviving records, in replay order: what `publish_parsed_sms`
        puts on the wire, `sms.model_dump_json().encode('utf-8')` (libs/nats_utils.py:127, worker.py:183).
        Assembled as bytes by smsgate_emit_json; only rows it declines go through pydantic."""
        idx = self.survivor_indices()
        if getattr(self.table, "text", None) is None:
            return [self.parsed(int(i)).model_dump_json().encode("utf-8") for i in idx]
        off, data, fb = self.emit_json(idx, 0)
        raw = data.tobytes()
        return [raw[off[k]:off[k + 1]] if not fb[k] else self.parsed(int(i)).model_dump_json().encode("utf-8")
                for k, i in enumerate(idx)]

    def pocketbase_records_json(self) -> list:
        """The same records as `pocketbase_records`, as compact JSON bytes assembled natively
        (smsgate_emit_json mode 1): what a bulk import into PocketBase is fed."""
        idx = self.survivor_indices()
        if getattr(self.table, "text", None) is None:
            import json
            return [json.dumps(r, ensure_ascii=False, separators=(",", ":")).encode("utf-8") for r in self.pocketbase_records()]
        off, data, fb = self.emit_json(idx, 1)
        raw = data.tobytes()
        import json
        rec = None
        out = []
        for k in range(len(idx)):
            if fb[k]:
                rec = rec if rec is not None else self.pocketbase_records()
                out.append(json.dumps(rec[k], ensure_ascii=False, separators=(",", ":")).encode("utf-8"))
            else:
                out.append(raw[off[k]:off[k + 1]])
        return out

    def pocketbase_records(self) -> list:
        """The record dict `upsert_parsed_sms` sends to PocketBase for every surviving record
        (libs/pocketbase.py:296-309), in replay order."""
        out = []
        for i in self.survivor_indices():
            p = self.parsed(int(i))
            out.append({
                "msg_id": p.msg_id, "original_body": p.raw_body, "sender": p.sender,
                "datetime": p.date.isoformat(), "card": p.card, "amount": str(p.amount), "currency": p.currency,
                "balance": str(p.balance) if p.balance is not None else None, "merchant": p.merchant,
                "address": p.address, "city": p.city, "txn_type": p.txn_type})
        return out


class ReplayEngine:
    def __init__(self, device: int = 0, use_torch_stream: bool = True):
        self._lib = L.lib()
        self._ctx = C.c_void_p()
        self.device = device
        stream = None
        if use_torch_stream:
            import torch
            if not torch.cuda.is_available():
                raise RuntimeError("smsgate_b200 needs a CUDA device (B200, sm_100a); there is no CPU path")
            torch.cuda.set_device(device)
            stream = C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
        rc = self._lib.smsgate_create(device, stream, C.byref(self._ctx))
        if rc != 0:
            raise L.SmsgateError(rc, (self._lib.smsgate_last_error(None) or b"").decode())
        self.table: Optional[CacheTable] = None
        self._pinned: dict = {}
        self._outputs_cache: dict = {}
        self._generation = 0             # bumped by every host-buffer replay (stale-result detection)
        self.host_fallbacks = 0          # XML files handed to ElementTree because of their dialect (xml_ingest)

    # -- lifecycle ---------------------------------------------------------
    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx.value:
            self._lib.smsgate_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def reserve(self, max_msgs: int, max_bytes: int) -> None:
        L.check(self._ctx, self._lib.smsgate_reserve(self._ctx, int(max_msgs), int(max_bytes)))

    @property
    def launch_count(self) -> int:
        return int(self._lib.smsgate_launch_count(self._ctx))

    # -- cache -------------------------------------------------------------
    def load_cache(self, cache: "Mapping[str, Any] | CacheTable") -> CacheTable:
        if isinstance(cache, (str, bytes)) or hasattr(cache, "__fspath__"):
            # a `.gemini_cache` directory: diskcache's SQLite file and its pickles read natively, no Python
            # object per entry (gemini_cache.read_diskcache_native; the pure-Python reader stays as
            # gemini_cache.read_diskcache_sqlite for comparison)
            import os
            from .gemini_cache import read_diskcache_native
            cache = read_diskcache_native(os.fspath(cache), self)
        table = cache if isinstance(cache, CacheTable) else build_table(cache, self.strict_date_batch)
        keys = np.ascontiguousarray(table.keys)
        L.check(self._ctx, self._lib.smsgate_cache_load(
            self._ctx, table.m, _np_ptr(keys), _np_ptr(table.val_offsets), _np_ptr(table.val_bytes),
            _np_ptr(table.eflags), _np_ptr(table.date_packed)))
        self.table = table
        return table

    # -- batch helpers -----------------------------------------------------
    def decimal_batch(self, strings: Sequence[str]):
        """parse_ambiguous_decimal over many strings -> (coef int64, exp int32, flags uint8)."""
        data, offsets = ragged_from_strings(strings, np.int32)
        n = len(strings)
        coef, exp, flags = np.zeros(n, np.int64), np.zeros(n, np.int32), np.zeros(n, np.uint8)
        L.check(self._ctx, self._lib.smsgate_decimal_batch(self._ctx, n, _np_ptr(offsets), _np_ptr(data),
                                                           _np_ptr(coef), _np_ptr(exp), _np_ptr(flags)))
        return coef, exp, flags

    def strict_date_batch(self, strings: Sequence[str]) -> np.ndarray:
        data, offsets = ragged_from_strings(strings, np.int32)
        n = len(strings)
        out = np.zeros(n, np.int64)
        L.check(self._ctx, self._lib.smsgate_strict_date_batch(self._ctx, n, _np_ptr(offsets), _np_ptr(data),
                                                               _np_ptr(out)))
        return out

    # -- host-buffer replay (the end-to-end call) ----------------------------
    def _host_columns(self, n: int, total: int, want: Sequence[str], pinned: bool) -> dict:
        cols = {}
        for name, dt, width in L.COLUMNS:
            if name not in want:
                continue
            shape = (total,) if name == "fixed_bytes" else ((n, width) if width > 1 else (n,))
            if pinned:
                import torch
                tdt = {"u1": torch.uint8, "i4": torch.int32, "i8": torch.int64}[dt]
                # one grow-only page-locked buffer per column: valid until the next pinned replay
                need = int(np.prod(shape))
                t = self._pinned.get(name)
                if t is None or t.numel() < need:
                    t = torch.empty(max(need, 1), dtype=tdt, pin_memory=True)
                    self._pinned[name] = t
                cols[name] = t.numpy()[:need].reshape(shape)
            else:
                cols[name] = np.zeros(shape, dtype=np.dtype(dt))
        return cols

    ALL_COLUMNS = tuple(name for name, _, _ in L.COLUMNS)

    def replay_records(self, batch: Batch, now: _dt.datetime, base_index: int = 0, pinned: bool = True,
                       exchange=None, compact_verdict: bool = True) -> ReplayResult:
        """`replay` with the results returned as one 64-byte record per message plus one verdict integer
        (smsgate_replay_host_records): one device->host transfer per chunk instead of one per column, which
        is what the end-to-end time depends on.  The ReplayResult reads the same either way: the columns the
        record leaves out (card, cache_key, fixed_len, fixed_bytes, *_cents) stay on the device and are
        fetched on first use.  With `pinned` the arrays alias this engine's page-locked staging buffer: the
        result is valid until the next replay on this engine; `result.detach()` makes it independent."""
        n = batch.n
        total = int(batch.body_offsets[-1]) if n else 0
        self.reserve(max(n, 1), max(total, 1))
        dtype = np.dtype(L.RECORD_FIELDS)
        assert dtype.itemsize == L.RECORD_BYTES_OUT
        self._generation += 1
        if pinned:
            # ONE grow-only page-locked staging buffer per engine: the returned arrays alias it and stay
            # valid until the next replay_records(pinned=True) on this engine (ReplayResult.detach() copies)
            import torch
            t = self._pinned.get("records")
            if t is None or t[1].numel() < max(n, 1):
                cap = max(n, 1) if t is None else max(n, 2 * t[1].numel())
                t = (torch.empty(cap * L.RECORD_BYTES_OUT, dtype=torch.uint8, pin_memory=True),
                     torch.empty(cap, dtype=torch.int64, pin_memory=True))
                self._pinned["records"] = t
            rec_buf, verdict_buf = t[0].numpy().view(dtype), t[1].numpy()
        else:
            rec_buf, verdict_buf = np.zeros(max(n, 1), dtype=dtype), np.zeros(max(n, 1), dtype=np.int64)
        if compact_verdict:
            # one byte of verdict per message (smsgate_replay_host_records8); last_idx stays on the device
            verdict_buf = verdict_buf.view(np.uint8)
        rec, verdict = rec_buf[:n], verdict_buf[:n]
        body = np.ascontiguousarray(batch.body_bytes)
        offs = np.ascontiguousarray(batch.body_offsets, dtype=np.int64)
        rp, vp = rec_buf.ctypes.data, verdict_buf.ctypes.data      # never NULL, also for an empty shard
        call = self._lib.smsgate_replay_host_records8 if compact_verdict else self._lib.smsgate_replay_host_records
        finish = self._lib.smsgate_replay_host_records8_finish if compact_verdict else self._lib.smsgate_replay_host_records_finish
        if exchange is None:
            L.check(self._ctx, call(self._ctx, _np_ptr(body), _np_ptr(offs), n, pack_datetime(now), int(base_index), rp, vp))
        else:
            L.check(self._ctx, self._lib.smsgate_replay_host_records_begin(self._ctx, _np_ptr(body), _np_ptr(offs), n,
                                                                           pack_datetime(now), int(base_index), rp))
            view = L.Outputs()
            L.check(self._ctx, self._lib.smsgate_device_view(self._ctx, n, C.byref(view)))
            exchange.run_outputs(view, n, int(base_index))
            L.check(self._ctx, finish(self._ctx, vp))
        return ReplayResult(batch, _RecordColumns(rec, verdict, self, self._generation, total), self.table, base_index)

    def _fetch_columns(self, n: int, total: int, names: Sequence[str]) -> dict:
        """Columns of the last host-buffer batch that are still on the device (smsgate_fetch_columns)."""
        cols = self._host_columns(n, total, tuple(names), pinned=False)
        out = L.Outputs(**{k: (v.ctypes.data if v.size else None) for k, v in cols.items()})
        L.check(self._ctx, self._lib.smsgate_fetch_columns(self._ctx, n, C.byref(out)))
        return cols

    def set_stream(self, stream) -> None:
        """Queue this engine's work on another torch stream from now on (smsgate_set_stream)."""
        L.check(self._ctx, self._lib.smsgate_set_stream(self._ctx, C.c_void_p(stream.cuda_stream)))

    def set_phone_dates(self, d_phone_dt) -> None:
        """Per-message phone timestamps (int64 device tensor, packed datetimes, 0 = none) for the next
        join_route calls; None resets (smsgate_set_phone_dates)."""
        self._phone_keepalive = d_phone_dt
        L.check(self._ctx, self._lib.smsgate_set_phone_dates(self._ctx, d_phone_dt.data_ptr() if d_phone_dt is not None else None))

    def set_host_pipeline(self, divisor: int = 3, min_chunk: int = 131072, offsets_first: bool = True) -> None:
        """How the host-buffer calls cut a batch into chunks (smsgate_set_host_pipeline)."""
        L.check(self._ctx, self._lib.smsgate_set_host_pipeline(self._ctx, int(divisor), int(min_chunk), int(offsets_first)))

    def set_scan_ctas(self, ctas: int) -> None:
        L.check(self._ctx, self._lib.smsgate_set_scan_ctas(self._ctx, int(ctas)))

    def share_cache_of(self, other: "ReplayEngine") -> None:
        """Probe `other`'s loaded response cache instead of loading a second copy (smsgate_cache_share)."""
        L.check(self._ctx, self._lib.smsgate_cache_share(self._ctx, other._ctx))
        self.table = other.table

    def replay(self, batch: Batch, now: _dt.datetime, base_index: int = 0,
               columns: Optional[Sequence[str]] = None, pinned: bool = False, exchange=None) -> ReplayResult:
        """Replay a host-resident batch end to end (host->device copy, three
        kernel stages, device->host copy of the requested columns).  With
        `exchange` (a dist.GlobalDedup) the batch is this rank's shard and the
        dedup verdicts are made global before they are copied out."""
        n = batch.n
        total = int(batch.body_offsets[-1]) if n else 0
        self.reserve(max(n, 1), max(total, 1))
        self._generation += 1
        want = tuple(columns) if columns is not None else self.ALL_COLUMNS
        cols = self._host_columns(n, total, want, pinned)
        out = L.Outputs(**{k: (v.ctypes.data if v.size else None) for k, v in cols.items()})
        body = np.ascontiguousarray(batch.body_bytes)
        offs = np.ascontiguousarray(batch.body_offsets, dtype=np.int64)
        if exchange is None:
            L.check(self._ctx, self._lib.smsgate_replay_host(self._ctx, _np_ptr(body), _np_ptr(offs), n,
                                                             pack_datetime(now), int(base_index), C.byref(out)))
        else:
            L.check(self._ctx, self._lib.smsgate_replay_host_begin(self._ctx, _np_ptr(body), _np_ptr(offs), n,
                                                                   pack_datetime(now), int(base_index), C.byref(out)))
            view = L.Outputs()
            L.check(self._ctx, self._lib.smsgate_device_view(self._ctx, n, C.byref(view)))
            exchange.run_outputs(view, n, int(base_index))
            L.check(self._ctx, self._lib.smsgate_replay_host_finish(self._ctx, C.byref(out)))
        return ReplayResult(batch, cols, self.table, base_index)

    def xml_extract(self, data, attrs: bool = False) -> dict:
        """Tokenize an XML backup (bytes / uint8 array) on the GPU.  Returns device tensors
        body_bytes, body_offsets, sender_bytes, sender_offsets, date_ms, mflags (+ n); with `attrs`
        also every attribute of every <sms>: attr_offsets[n+1], name_offsets / name_bytes,
        value_offsets / value_bytes (smsgate_xml_attrs_sizes / _emit).
        Raises xml.etree.ElementTree.ParseError where ElementTree would, UnsupportedXML for
        constructs outside the dialect."""
        import torch
        import xml.etree.ElementTree as ET
        buf = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
        n, nb, ns, pos = C.c_int64(0), C.c_int64(0), C.c_int64(0), C.c_int64(-1)
        rc = self._lib.smsgate_xml_scan(self._ctx, _np_ptr(buf) if buf.size else None, int(buf.size), C.byref(n),
                                        C.byref(nb), C.byref(ns), C.byref(pos))
        if rc == L.E_XML:
            raise ET.ParseError(f"not well-formed near byte {pos.value}")
        if rc == L.E_UNSUPPORTED:
            raise UnsupportedXML((self._lib.smsgate_last_error(self._ctx) or b"").decode())
        L.check(self._ctx, rc)
        dev = torch.device("cuda", self.device)
        out = dict(n=n.value,
                   body_bytes=torch.empty(nb.value + 64, dtype=torch.uint8, device=dev),
                   body_offsets=torch.empty(n.value + 1, dtype=torch.int64, device=dev),
                   sender_bytes=torch.empty(ns.value + 64, dtype=torch.uint8, device=dev),
                   sender_offsets=torch.empty(n.value + 1, dtype=torch.int64, device=dev),
                   date_ms=torch.empty(max(n.value, 1), dtype=torch.int64, device=dev),
                   mflags=torch.empty(max(n.value, 1), dtype=torch.uint8, device=dev))
        if attrs:
            na, nnb, nvb = C.c_int64(0), C.c_int64(0), C.c_int64(0)
            L.check(self._ctx, self._lib.smsgate_xml_attrs_sizes(self._ctx, C.byref(na), C.byref(nnb), C.byref(nvb)))
            out.update(attr_offsets=torch.empty(n.value + 1, dtype=torch.int64, device=dev),
                       name_offsets=torch.empty(na.value + 1, dtype=torch.int64, device=dev),
                       name_bytes=torch.empty(nnb.value + 64, dtype=torch.uint8, device=dev),
                       value_offsets=torch.empty(na.value + 1, dtype=torch.int64, device=dev),
                       value_bytes=torch.empty(nvb.value + 64, dtype=torch.uint8, device=dev))
            L.check(self._ctx, self._lib.smsgate_xml_attrs_emit(
                self._ctx, out["attr_offsets"].data_ptr(), out["name_offsets"].data_ptr(), out["name_bytes"].data_ptr(),
                out["value_offsets"].data_ptr(), out["value_bytes"].data_ptr()))
        L.check(self._ctx, self._lib.smsgate_xml_emit(self._ctx, out["body_bytes"].data_ptr(), out["body_offsets"].data_ptr(),
                                                      out["sender_bytes"].data_ptr(), out["sender_offsets"].data_ptr(),
                                                      out["date_ms"].data_ptr(), out["mflags"].data_ptr()))
        return out

    def xml_attribs(self, data) -> list:
        """`[elem.attrib for elem in root.findall("sms")]` (read_xml.py:32-34) with the file tokenized and
        every value decoded on the GPU: one dict per top-level <sms>, attributes in document order."""
        d = self.xml_extract(data, attrs=True)
        ao = d["attr_offsets"].cpu().numpy()
        no, vo = d["name_offsets"].cpu().numpy(), d["value_offsets"].cpu().numpy()
        nb = d["name_bytes"].cpu().numpy().tobytes()
        vb = d["value_bytes"].cpu().numpy().tobytes()
        out = []
        for i in range(d["n"]):
            out.append({nb[no[a]:no[a + 1]].decode("utf-8"): vb[vo[a]:vo[a + 1]].decode("utf-8")
                        for a in range(int(ao[i]), int(ao[i + 1]))})
        return out

    def msg_ids(self, batch: Batch) -> np.ndarray:
        """SHA-1(body) of every message: the XML idempotency key (watcher.py:45).  uint8 [n, 20]."""
        return self.replay(batch, _dt.datetime(1, 1, 1), columns=("msg_id",)).cols["msg_id"]

    # -- device-pointer stages (torch tensors) -------------------------------
    def device_columns(self, n: int, total: int, want: Optional[Sequence[str]] = None) -> dict:
        import torch
        want = tuple(want) if want is not None else self.ALL_COLUMNS
        dev = torch.device("cuda", self.device)
        cols = {}
        for name, dt, width in L.COLUMNS:
            if name not in want:
                continue
            tdt = {"u1": torch.uint8, "i4": torch.int32, "i8": torch.int64}[dt]
            shape = (max(total, 1),) if name == "fixed_bytes" else ((n, width) if width > 1 else (n,))
            cols[name] = torch.empty(shape, dtype=tdt, device=dev)
        return cols

    def _dev_outputs(self, cols: dict) -> L.Outputs:
        """struct smsgate_outputs over a set of device columns; built once per column set (a step of the
        multi-GPU pipeline makes three C calls on the same columns and is launch-bound on the host).  The
        cache holds no reference to the tensors: an entry is valid while the dict still has the same size and
        its first and last columns still sit at the same addresses."""
        key = id(cols)
        vals = list(cols.values())
        sig = (len(vals), vals[0].data_ptr(), vals[-1].data_ptr()) if vals else (0, 0, 0)
        hit = self._outputs_cache.get(key)
        if hit is not None and hit[0] == sig:
            return hit[1]
        out = L.Outputs(**{k: v.data_ptr() for k, v in cols.items()})
        if len(self._outputs_cache) > 16:
            self._outputs_cache.clear()
        self._outputs_cache[key] = (sig, out)
        return out

    def scan_keys(self, d_bytes, d_offsets, cols: dict) -> None:
        n = d_offsets.numel() - 1
        out = self._dev_outputs(cols)
        L.check(self._ctx, self._lib.smsgate_scan_keys(self._ctx, d_bytes.data_ptr(), d_offsets.data_ptr(), n,
                                                       C.byref(out)))

    def device_keys(self, d_bytes, d_offsets):
        """MD5 of every body (device tensors in, uint8 [n, 16] device tensor out): the msg_id
        `get_md5_hash` gives a device-sourced message (libs/models.py:97-109, api_gateway/main.py:113)."""
        import torch
        n = d_offsets.numel() - 1
        out = torch.empty((max(n, 1), 16), dtype=torch.uint8, device=d_bytes.device)
        L.check(self._ctx, self._lib.smsgate_device_keys(self._ctx, d_bytes.data_ptr(), d_offsets.data_ptr(), n,
                                                         out.data_ptr()))
        return out[:n]

    def msg_ids_device(self, bodies: Sequence[str]) -> list:
        """`get_md5_hash` over a list of bodies (hex strings), computed on the GPU."""
        import torch
        b = Batch.from_strings(bodies, None, None)
        dev = torch.device("cuda", self.device)
        d_bytes = torch.zeros(b.body_bytes.size + 64, dtype=torch.uint8, device=dev)
        d_bytes[:b.body_bytes.size] = torch.from_numpy(np.array(b.body_bytes, copy=True)).to(dev)
        d_off = torch.from_numpy(np.ascontiguousarray(b.body_offsets, dtype=np.int64)).to(dev)
        out = self.device_keys(d_bytes, d_off).cpu().numpy()
        return [row.tobytes().hex() for row in out]

    def cross_source(self, d_xml_sha1, d_dev_sha1):
        """For every device-sourced message (rows of d_dev_sha1, the SHA-1 of its body) the index of the
        first XML message with the same body (rows of d_xml_sha1), or -1: the bodies the reference stores
        twice because it keys the two sources differently."""
        import torch
        n_xml, n_dev = d_xml_sha1.shape[0], d_dev_sha1.shape[0]
        self.reserve(max(n_xml, 1), 1)
        match = torch.empty(max(n_dev, 1), dtype=torch.int32, device=d_dev_sha1.device)
        L.check(self._ctx, self._lib.smsgate_cross_source(self._ctx, n_xml, d_xml_sha1.data_ptr(), n_dev,
                                                          d_dev_sha1.data_ptr(), match.data_ptr()))
        return match[:n_dev]

    def join_route(self, n: int, now: _dt.datetime, cols: dict) -> None:
        out = self._dev_outputs(cols)
        L.check(self._ctx, self._lib.smsgate_join_route(self._ctx, n, pack_datetime(now), C.byref(out)))

    def join_route_dedup(self, n: int, now: _dt.datetime, base_index: int, cols: dict) -> None:
        """Stages 2 and 3 in one call (the join kernel inserts into the upsert table as it routes)."""
        out = self._dev_outputs(cols)
        L.check(self._ctx, self._lib.smsgate_join_route_dedup(self._ctx, n, pack_datetime(now), int(base_index),
                                                              C.byref(out)))

    def dedup(self, n: int, base_index: int, cols: dict) -> None:
        out = self._dev_outputs(cols)
        L.check(self._ctx, self._lib.smsgate_dedup(self._ctx, n, int(base_index), C.byref(out)))

    def replay_device(self, d_bytes, d_offsets, now: _dt.datetime, cols: dict, base_index: int = 0) -> None:
        n = d_offsets.numel() - 1
        out = self._dev_outputs(cols)
        L.check(self._ctx, self._lib.smsgate_replay_device(self._ctx, d_bytes.data_ptr(), d_offsets.data_ptr(), n,
                                                           pack_datetime(now), int(base_index), C.byref(out)))
