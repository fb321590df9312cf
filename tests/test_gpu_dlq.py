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
    skip = set(got.needs_phone_timestamp)
    assert skip == {rows[victim]}                    # numeric date + dateless cache entry: listed, not evaluated
    for r, o in enumerate(outs):
        if r in skip:
            assert o.status == O.ST_UPSERT and int(res.status[r]) == L.ST_UNMATCHED
            continue
        assert int(res.status[r]) == o.status, r
        assert res.cache_key_hex(r) == o.cache_key.hex()
        if o.parsed is not None and o.status != O.ST_FAILED_EXC:
            p = got.parsed(r)
            assert p.msg_id == raws[r].msg_id and p.device_id == raws[r].device_id and p.sender == raws[r].sender
            assert p.date == o.parsed.date and p.amount == o.parsed.amount and p.raw_body == o.parsed.raw_body
    # dedup by the messages' own ids: the verdicts of every row not touched by the skipped one
    ids = [r.msg_id for r in raws]
    touched = {ids[r] for r in skip}
    for r in range(len(raws)):
        if ids[r] in touched:
            continue
        assert bool(res.survivor[r]) == surv[r], r
        assert bool(res.first_seen[r]) == first[r], r
        assert int(res.last_idx[r]) == last[r], r
    assert sum(surv) > 50 and len(set(ids)) < len(ids)
    ups = got.upserts()
    assert all(k in set(ids) for k in ups)


def test_reparse_nothing_to_do(engine):
    got = dlq.reparse(engine, [b"junk", json.dumps({"err": "x", "entry": "y"}).encode()], NOW)
    assert got.outcomes == [dlq.NOT_JSON, dlq.NO_RAW] and got.result is None and got.upserts() == {}
