"""Bulk re-parse of the dead-letter queue.

The reference re-parses `sms.failed` one message at a time:
`services/parser_worker/dlq_worker.py:48-78` decodes the payload, wants a
`raw` entry in it and hands the message to the worker's `_process_one`, which
(`services/parser_worker/worker.py:90-99`) decodes it again, unwraps `raw` and
validates it as a `RawSMS` before the usual filter -> parse_sms_llm -> gate
path.  Here a whole queue dump goes through that path at once: the unwrapping
and the pydantic validation are host work per payload (they decide WHICH
messages are replayed), the replay itself is the GPU path, with one
difference to an XML replay: a DLQ message brings its own `msg_id` (the SHA-1
of the body for XML-sourced messages, the MD5 for device-sourced ones), and
upserts are keyed by THAT id, not by a digest computed here.
"""
from __future__ import annotations

import datetime as _dt
import hashlib
import json
from dataclasses import dataclass, field
from typing import Any, List, Optional, Sequence, Union

import numpy as np

from . import _lib as L
from .batch import Batch
from .models import RawSMS
from .replay import ReplayEngine, ReplayResult

# what happened to a payload before the replay proper
NOT_JSON = "not_json"            # dlq_worker.py:48-53: logged, acked, nothing else
NOT_AN_OBJECT = "not_an_object"  # payload.get(...) raises AttributeError in the reference (dlq_worker.py:60)
NO_RAW = "no_raw"                # dlq_worker.py:61-63: "nothing to re-parse"
INVALID_RAW = "invalid_raw"      # worker.py:99-109: RawSMS(**raw) failed -> published to sms.failed again
REPLAYED = "replayed"


def unwrap(payload: Union[bytes, str]) -> tuple:
    """One DLQ payload -> (outcome, RawSMS | None, error text | None), following
    dlq_worker.py:48-66 and worker.py:90-99 step by step."""
    try:
        obj = json.loads(payload)
    except Exception:
        return NOT_JSON, None, None
    if not hasattr(obj, "get"):
        return NOT_AN_OBJECT, None, None
    if obj.get("raw", None) is None:
        return NO_RAW, None, None
    # worker.py:91-95: bytes are decoded (strict UTF-8) and parsed again; `raw` replaces the payload
    try:
        data = payload.decode() if isinstance(payload, (bytes, bytearray)) else payload
        raw_data = json.loads(data)
        if "raw" in raw_data:
            raw_data = raw_data["raw"]
        return REPLAYED, RawSMS(**raw_data), None
    except Exception as err:                       # worker.py:101-109
        return INVALID_RAW, None, str(err)


def key_bytes(msg_id: str) -> bytes:
    """The 20 bytes a message's own id occupies in the dedup key column: the digest itself for a
    40-hex-digit id (what `_iter_sms` assigns), SHA-1 of the id text for anything else (MD5 ids of
    device-sourced messages, hand-made ids).  Equal ids give equal keys; different ids give different
    keys unless SHA-1 collides."""
    if len(msg_id) == 40:
        try:
            b = bytes.fromhex(msg_id)
            if msg_id == b.hex():                  # lower-case hex only: "AB.." and "ab.." are different ids
                return b
        except ValueError:
            pass
    return hashlib.sha1(b"id:" + msg_id.encode("utf-8", "surrogatepass")).digest()


@dataclass
class DlqResult:
    outcomes: List[str]                  # per payload
    errors: List[Optional[str]]          # per payload: the validation error text, for INVALID_RAW
    row: np.ndarray                      # per payload: row in `result`, -1 if it was not replayed
    raws: List[RawSMS]                   # per replayed row
    result: Optional[ReplayResult]       # columns of the replayed rows (None when nothing was replayed)
    needs_phone_timestamp: List[int] = field(default_factory=list)

    def parsed(self, r: int):
        """What `parse_sms_llm(raw)` returns / raises for replayed row r, carrying the message's own ids."""
        raw = self.raws[r]
        return self.result.parsed(r, sender=raw.sender, device_id=raw.device_id, msg_id=raw.msg_id)

    def upserts(self) -> dict:
        """msg_id -> ParsedSMS left behind by re-parsing the whole dump (last occurrence of an id wins)."""
        return {self.raws[int(r)].msg_id: self.parsed(int(r)) for r in self.result.survivor_indices()} \
            if self.result is not None else {}


def reparse(engine: ReplayEngine, payloads: Sequence[Union[bytes, str]], now: _dt.datetime,
            base_index: int = 0) -> DlqResult:
    """Re-parse a dump of `sms.failed` in one pass.  The response cache must be loaded
    (`engine.load_cache`).  Stage calls on device memory: keys + scans of the bodies, then the message's
    own ids take the place of the computed key column, then join + routing and the upsert dedup."""
    import torch
    outcomes, errors, raws = [], [], []
    row = np.full(len(payloads), -1, dtype=np.int64)
    for i, p in enumerate(payloads):
        o, raw, err = unwrap(p)
        outcomes.append(o)
        errors.append(err)
        if raw is not None:
            row[i] = len(raws)
            raws.append(raw)
    if not raws:
        return DlqResult(outcomes, errors, row, raws, None)
    batch = Batch.from_strings([r.body for r in raws], [r.sender for r in raws], None)
    n = batch.n
    dev = torch.device("cuda", engine.device)
    d_bytes = torch.zeros(batch.body_bytes.size + 64, dtype=torch.uint8, device=dev)
    d_bytes[:batch.body_bytes.size] = torch.from_numpy(np.array(batch.body_bytes, copy=True)).to(dev)
    d_off = torch.from_numpy(np.ascontiguousarray(batch.body_offsets, dtype=np.int64)).to(dev)
    cols = engine.device_columns(n, batch.body_bytes.size)
    engine.reserve(n, max(batch.body_bytes.size, 1))
    engine.scan_keys(d_bytes, d_off, cols)
    ids = np.frombuffer(b"".join(key_bytes(r.msg_id) for r in raws), dtype=np.uint8).reshape(n, 20)
    cols["msg_id"].copy_(torch.from_numpy(ids.copy()).to(dev))
    engine.join_route(n, now, cols)
    engine.dedup(n, base_index, cols)
    torch.cuda.synchronize(dev)
    host = {k: v.cpu().numpy() for k, v in cols.items()}
    res = ReplayResult(batch, host, engine.table, base_index)
    # libs/gemini_parser.py:228-229: a cached date without any date text falls back to the phone's timestamp,
    # int(raw.date).  For XML-sourced messages that raises (ISO text); for a numeric date it would not:
    # those rows are listed, not evaluated.
    needs = []
    if engine.table is not None:
        for r in range(n):
            e = int(host["slot"][r])
            # (an XML-sourced message with such an entry ends as UNMATCHED: int() of its ISO date raises
            # inside the try of gemini_parser.py:224-241)
            if e >= 0 and int(host["status"][r]) == L.ST_UNMATCHED and not (int(host["flags"][r]) & L.F_PARSER_OTP):
                f = int(engine.table.eflags[e])
                no_date = (f & L.E_DATE_KEY) and not (f & (L.E_DATE_STRICT | L.E_DATE_HOST | L.E_DATE_NOBODY))
                if no_date and _intable(raws[r].date):
                    needs.append(r)
    return DlqResult(outcomes, errors, row, raws, res, needs)


def _intable(s: Any) -> bool:
    try:
        int(s)
        return True
    except Exception:
        return False
