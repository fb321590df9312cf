"""The oracle against a RUN of the reference's own message handlers (tests/golden/worker_flow.json):
statuses, sms.parsed / sms.failed payload bytes, and the DLQ handler's decisions per payload."""
import datetime as dt
import json

from oracle import dlq_oracle as DO
from oracle import replay_oracle as O
from smsgate_b200.models import ParsedSMS
from tests import _flow as F


def _dump_json(parsed, msg_id):
    """sms.parsed payload of an oracle outcome: the reference's model, the message's own id."""
    import dataclasses
    return ParsedSMS(**dict(dataclasses.asdict(parsed), msg_id=msg_id)).model_dump_json()


def test_worker_flow_matches_reference_run():
    g, bodies, senders = F.load()
    now = dt.datetime.fromisoformat(g["now"])
    seen = set()
    for i, row in enumerate(g["flows"]["worker"]):
        raw = DO.RawSMS(**json.loads(bytes.fromhex(row["payload_hex"])))
        assert raw.body == bodies[i]
        want = F.status_of(row, g["offline_error"])
        out = O.replay_one(raw.body, raw.sender, raw.date, g["cache"], now, device_id=raw.device_id)
        assert out.status == want, (i, raw.body[:50])
        seen.add(want)
        pj = None
        if out.parsed is not None and want in (F.UPSERT, F.NO_MERCHANT):
            pj = _dump_json(out.parsed, raw.msg_id)
        F.check_published(row, want, pj, raw.model_dump())
    assert seen == set(range(9))                 # every branch of the handler occurs in the fixture


def test_dlq_flow_matches_reference_run():
    g, bodies, senders = F.load()
    now = dt.datetime.fromisoformat(g["now"])
    payloads = [bytes.fromhex(r["payload_hex"]) for r in g["flows"]["dlq"]]
    outcomes, rows, raws, outs, surv, first, last = DO.reparse(payloads, g["cache"], now)
    kinds = set()
    for i, row in enumerate(g["flows"]["dlq"]):
        kinds.add(outcomes[i])
        out, raw = (outs[rows[i]], raws[rows[i]]) if rows[i] >= 0 else (None, None)
        try:
            F.check_dlq_row(row, outcomes[i], out, raw, g["offline_error"], _dump_json)
        except AssertionError as e:
            raise AssertionError(f"DLQ payload {i}: {e}") from e
    assert kinds == {"not_json", "no_raw", "not_an_object", "invalid_raw", "replayed"}
