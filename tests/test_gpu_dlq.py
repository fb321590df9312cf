"""DLQ re-parse on the GPU against the oracle restatement (oracle/dlq_oracle.py)."""
import datetime as dt
import json

import numpy as np
import pytest

from oracle import dlq_oracle as DO
from oracle import replay_oracle as O
from smsgate_b200 import _lib as L
from smsgate_b200 import dlq, synth
from smsgate_b200.synth import reference_cache_key
from tests._dlq_cases import payloads

pytestmark = pytest.mark.gpu
NOW = dt.datetime(2026, 1, 1)


@pytest.fixture(scope="module")
def engine():
    from smsgate_b200.replay import ReplayEngine
    eng = ReplayEngine(0)
    yield eng
    eng.close()


def test_reparse_matches_oracle(engine):
    backup = synth.generate(3000, seed=73, mixed_locale=True, miss_frac=0.1)
    cache = dict(backup.cache)
    # one entry without any date text: the reference falls back to the phone's timestamp (gemini_parser.py:228)
    victim = next(i for i, b in enumerate(backup.bodies) if i % 10 == 0 and reference_cache_key(b) in cache)
    cache[reference_cache_key(backup.bodies[victim])] = dict(cache[reference_cache_key(backup.bodies[victim])], date="")
    engine.load_cache(cache)
    ps = payloads(backup.bodies, backup.senders)
    got = dlq.reparse(engine, ps, NOW)
    outcomes, rows, raws, outs, surv, first, last = DO.reparse(ps, cache, NOW)
    assert got.outcomes == outcomes
    assert list(got.row) == rows
    assert [r.model_dump() for r in got.raws] == [r.model_dump() for r in raws]
    res = got.result
    # the row with a numeric date and a dateless cache entry takes the phone's timestamp (gemini_parser.py:227-229)
    assert outs[rows[victim]].status == O.ST_UPSERT and int(res.phone_dates[rows[victim]]) > 0
    for r, o in enumerate(outs):
        assert int(res.status[r]) == o.status, r
        assert res.cache_key_hex(r) == o.cache_key.hex()
        if o.parsed is not None and o.status not in (O.ST_FAILED_EXC, O.ST_ESCAPED):
            p = got.parsed(r)
            assert p.msg_id == raws[r].msg_id and p.device_id == raws[r].device_id and p.sender == raws[r].sender
            assert p.date == o.parsed.date and p.amount == o.parsed.amount and p.raw_body == o.parsed.raw_body
    # dedup by the messages' own ids
    for r in range(len(raws)):
        assert bool(res.survivor[r]) == surv[r], r
        assert bool(res.first_seen[r]) == first[r], r
        assert int(res.last_idx[r]) == last[r], r


def test_flows_match_a_run_of_the_reference_handlers(engine):
    """tests/golden/worker_flow.json holds what the reference's own `_process_one` and `_handle_dlq_msg`
    did with every message (acks, publishes with their payload bytes, counters, escaped exceptions).  The
    GPU replay must report the matching status for each, and hand back byte-identical sms.parsed /
    sms.failed payloads."""
    from smsgate_b200.models import RawSMS
    from tests import _flow as F
    g, bodies, senders = F.load()
    now = dt.datetime.fromisoformat(g["now"])
    engine.load_cache(g["cache"])
    # the live path: one RawSMS per message
    rows_w = g["flows"]["worker"]
    raws = [RawSMS(**json.loads(bytes.fromhex(r["payload_hex"]))) for r in rows_w]
    res = dlq.replay_raws(engine, raws, now)
    seen = set()
    for i, row in enumerate(rows_w):
        want = F.status_of(row, g["offline_error"])
        assert int(res.status[i]) == want, (i, raws[i].body[:50])
        seen.add(want)
        pj = None
        if want in (F.UPSERT, F.NO_MERCHANT):
            pj = res.parsed(i, sender=raws[i].sender, device_id=raws[i].device_id, msg_id=raws[i].msg_id).model_dump_json()
        F.check_published(row, want, pj, raws[i].model_dump())
    assert seen == set(range(9))
    # the dead-letter path
    rows_d = g["flows"]["dlq"]
    got = dlq.reparse(engine, [bytes.fromhex(r["payload_hex"]) for r in rows_d], now)
    for i, row in enumerate(rows_d):
        o = got.outcomes[i]
        nothing = not row["published"] and row["counters"] == [0, 0, 0]
        if o in (dlq.NOT_JSON, dlq.NO_RAW):
            assert nothing and row["acks"] == 1 and row["escaped"] is None, i
        elif o == dlq.NOT_AN_OBJECT:
            assert nothing and row["acks"] == 0 and row["escaped"] == "AttributeError", i
        elif o == dlq.INVALID_RAW:
            assert nothing and row["acks"] == 0 and row["escaped"] is None, i
        else:
            r = int(got.row[i])
            st = int(got.result.status[r])
            if st in L.HANDLER_RAISES:
                assert nothing and row["acks"] == 0 and row["escaped"] is None, i
            else:
                assert st == F.status_of(row, g["offline_error"]), i
                pj = got.parsed(r).model_dump_json() if st in (F.UPSERT, F.NO_MERCHANT) else None
                F.check_published(row, st, pj, got.raws[r].model_dump())
