"""Reading tests/golden/worker_flow.json: what the reference's own handlers (worker._process_one,
dlq_worker._handle_dlq_msg, run unmodified by oracle/gen_golden_worker.py against a fake NATS) did with each
message, turned into the status / outcome the replay must report for it."""
import json
import os

from smsgate_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "worker_flow.json")

UPSERT, NO_MERCHANT, SKIPPED, MISS, UNMATCHED, BROKEN, FUTURE, FAILED_EXC, ESCAPED = range(9)


def load():
    with open(GOLDEN, encoding="utf-8") as f:
        g = json.load(f)
    backup = synth.generate(g["n"], seed=g["seed"], **g["kwargs"])
    bodies = backup.bodies + g["extra_bodies"]
    senders = backup.senders + ["BANK"] * len(g["extra_bodies"])
    return g, bodies, senders


def status_of(row, offline_error):
    """The status a replayed message must get, from what the reference's handler did with it."""
    pubs, esc, cnt = row["published"], row["escaped"], row["counters"]
    if esc == "TypeError":
        return ESCAPED                       # aware date: worker.py:174 raises
    if esc == "AttributeError":
        return FUTURE                        # worker.py:177: .decode on a str inside the future-date branch
    subjects = [s for s, _ in pubs]
    if "parsed" in subjects:
        assert subjects == ["parsed", "processing"] and pubs[0][1] == pubs[1][1] and cnt == [1, 0, 0] and row["acks"] == 1
        return UPSERT if json.loads(pubs[0][1])["merchant"] else NO_MERCHANT     # writer.py:70
    if subjects == ["failed"]:
        assert cnt == [0, 0, 1] and row["acks"] == 1
        p = json.loads(pubs[0][1])
        if p.get("reason") == "unmatched":
            return UNMATCHED
        return MISS if p.get("err") == offline_error else FAILED_EXC
    assert not pubs and row["acks"] == 1
    if cnt == [1, 0, 0]:
        return SKIPPED
    assert cnt == [0, 1, 0]
    return BROKEN


def check_published(row, status, parsed_json, raw_dump):
    """Payload bytes the reference put on the wire against ours: sms.parsed carries
    ParsedSMS.model_dump_json() (worker.py:183), an unmatched message {"reason", "raw"} (worker.py:153)."""
    if status in (UPSERT, NO_MERCHANT):
        assert parsed_json == row["published"][0][1]
    elif status == UNMATCHED:
        assert json.dumps({"reason": "unmatched", "raw": raw_dump}) == row["published"][0][1]


def check_dlq_row(row, outcome, out, raw, offline_error, dump_json):
    """One DLQ payload: what the reference's `_handle_dlq_msg(reparse=True)` did (`row`) against the oracle's
    reading of the same payload (`outcome` of dlq_oracle.handle; `out` / `raw` the replayed message, if any).
    `dump_json(parsed, msg_id)` renders the sms.parsed payload."""
    nothing = not row["published"] and row["counters"] == [0, 0, 0]
    if outcome in ("not_json", "no_raw"):
        assert nothing and row["acks"] == 1 and row["escaped"] is None
    elif outcome == "not_an_object":
        assert nothing and row["acks"] == 0 and row["escaped"] == "AttributeError"      # payload.get on a list
    elif outcome == "invalid_raw":
        # worker.py:101-109 means to publish the failure, but `.decode` on the already decoded str raises
        # inside the handler and dlq_worker.py:76 swallows it: nothing happens
        assert nothing and row["acks"] == 0 and row["escaped"] is None
    else:
        st = out.status
        if st in (FUTURE, ESCAPED):
            assert nothing and row["acks"] == 0 and row["escaped"] is None              # swallowed by dlq_worker.py:76
        else:
            assert st == status_of(row, offline_error)
            pj = dump_json(out.parsed, raw.msg_id) if st in (UPSERT, NO_MERCHANT) else None
            check_published(row, st, pj, raw.model_dump())
