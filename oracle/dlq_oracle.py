"""TEST INFRASTRUCTURE ONLY (see oracle/README or DESIGN.md section 2): CPU restatement of the reference's
dead-letter re-parse, one payload at a time.  Nothing under smsgate_b200/ imports this.

Follows services/parser_worker/dlq_worker.py:48-78 (`_handle_dlq_msg` with reparse=True) and the worker's
`_process_one` it calls, services/parser_worker/worker.py:88-109 (decode, unwrap `raw`, RawSMS(**raw)),
then the replay restatement in replay_oracle.py.  `RawSMS` is a field-for-field copy of
libs/models.py:44-57; oracle/gen_golden_dlq.py checks it against the reference's own class.
"""
from __future__ import annotations

import datetime as _dt
import json
from typing import Literal, Optional

from pydantic import BaseModel, Field

from . import replay_oracle as O


class RawSMS(BaseModel):                       # libs/models.py:44-57
    msg_id: str = Field(...)
    sender: str = Field(..., min_length=1)
    body: str = Field(..., min_length=1)
    date: str = Field(...)
    device_id: Optional[str] = Field(None)
    source: Literal["device", "xml"] = Field("device")


def handle(payload, raw_cls=RawSMS):
    """-> (outcome, raw | None).  Outcomes: not_json, not_an_object, no_raw, invalid_raw, replayed."""
    try:
        obj = json.loads(payload)                                # dlq_worker.py:48-49
    except Exception:
        return "not_json", None                                  # :50-53
    try:
        raw_entry = obj.get("raw", None)                         # :60
    except AttributeError:
        return "not_an_object", None                             # escapes the handler in the reference
    if raw_entry is None:
        return "no_raw", None                                    # :61-63
    try:                                                         # worker.py:88-99 via dlq_worker.py:72
        data = payload
        if isinstance(data, bytes):
            data = data.decode()
            d = json.loads(data)
            if "raw" in d:
                d = d["raw"]
        else:
            # a str payload: worker.py:96-98 would not unwrap; the DLQ only ever delivers bytes, and the
            # drop-in treats text like the bytes it encodes to
            d = json.loads(data)
            if "raw" in d:
                d = d["raw"]
        return "replayed", raw_cls(**d)
    except Exception:
        return "invalid_raw", None                               # worker.py:101-109


def reparse(payloads, cache, now: _dt.datetime):
    """-> (outcomes, rows, message outcomes of the replayed rows, survivor, first_seen, last_idx)."""
    outcomes, rows, raws = [], [], []
    for p in payloads:
        o, raw = handle(p)
        outcomes.append(o)
        rows.append(len(raws) if raw is not None else -1)
        if raw is not None:
            raws.append(raw)
    outs = [O.replay_one(r.body, r.sender, r.date, cache, now, device_id=r.device_id) for r in raws]
    surv, first, last, _ = O.dedup((r.msg_id for r in raws), (o.status == O.ST_UPSERT for o in outs))
    return outcomes, rows, raws, outs, surv, first, last
