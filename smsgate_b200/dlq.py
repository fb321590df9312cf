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
INVALID_RAW = "invalid_raw"      # worker.py:99-109: RawSMS(**raw) failed.  The handler means to publish it to sms.failed
                                 # again, but msg.data is already a str by then and `.decode` raises (:105): run as
                                 # written, nothing is published or acked (tests/golden/worker_flow.json)
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

    def parsed(self, r: int):
        """What `parse_sms_llm(raw)` returns / raises for replayed row r, carrying the message's own ids."""
        raw = self.raws[r]
        return self.result.parsed(r, sender=raw.sender, device_id=raw.device_id, msg_id=raw.msg_id)

    def upserts(self) -> dict:
        """msg_id -> ParsedSMS left behind by re-parsing the whole dump (last occurrence of an id wins)."""
        return {self.raws[int(r)].msg_id: self.parsed(int(r)) for r in self.result.survivor_indices()} \
            if self.result is not None else {}


def phone_datetime(raw_date: Any) -> int:
    """`parse_unix_timestamp(int(raw.date), tz="Asia/Yerevan", aware=False)` (libs/gemini_parser.py:229,
    :139-188) as a packed naive datetime, 0 where the reference raises (ISO text, negative, >= 1e14, out of
    range).  Evaluated with Python's own int / float / datetime / zoneinfo, like the reference."""
    import zoneinfo

    from .gemini_cache import pack_datetime
    try:
        ts = float(int(raw_date))
        if ts < 0 or ts >= 1e14:
            return 0
        secs = ts if ts < 1e11 else ts / 1_000
        d = _dt.datetime.fromtimestamp(secs, tz=_dt.timezone.utc).astimezone(zoneinfo.ZoneInfo("Asia/Yerevan"))
        return pack_datetime(d.replace(tzinfo=None))
    except Exception:
        return 0


def replay_raws(engine: ReplayEngine, raws: Sequence[RawSMS], now: _dt.datetime, base_index: int = 0,
                worker_filter: bool = True) -> ReplayResult:
    """Replay messages that bring their own identity (`RawSMS` objects: DLQ entries, device-sourced
    messages): keys + scans of the bodies on the GPU, then the messages' own ids take the place of the
    computed key column, the phone timestamps of numeric `raw.date`s ride along for the cached dates that
    hold no date text (libs/gemini_parser.py:227-229), then join + routing and the upsert dedup.
    `worker_filter=False` leaves the worker's skip filter (worker.py:112-126) out of the routing: what
    `parse_sms_llm` does when it is called directly."""
    import torch
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
    if not worker_filter:
        cols["flags"].bitwise_and_(0xFF & ~L.F_WORKER_SKIP)
    phone = torch.from_numpy(np.array([phone_datetime(r.date) for r in raws], dtype=np.int64)).to(dev)
    engine.set_phone_dates(phone)
    try:
        engine.join_route(n, now, cols)
    finally:
        engine.set_phone_dates(None)
    engine.dedup(n, base_index, cols)
    torch.cuda.synchronize(dev)
    host = {k: v.cpu().numpy() for k, v in cols.items()}
    res = ReplayResult(batch, host, engine.table, base_index)
    res.phone_dates = phone.cpu().numpy()
    return res


def reparse(engine: ReplayEngine, payloads: Sequence[Union[bytes, str]], now: _dt.datetime,
            base_index: int = 0) -> DlqResult:
    """Re-parse a dump of `sms.failed` in one pass.  The response cache must be loaded
    (`engine.load_cache`).  What the reference does with each payload before the replay proper is decided
    per payload on the host (`unwrap`); the replayed rows go through `replay_raws`."""
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
    return DlqResult(outcomes, errors, row, raws, replay_raws(engine, raws, now, base_index))
