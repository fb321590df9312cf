"""The CPU oracle against golden vectors produced by executing the reference
(oracle/gen_golden.py).  No GPU, no CUDA library."""
import datetime as dt

import pytest

from oracle import replay_oracle as O

STATUS = {"UPSERT": O.ST_UPSERT, "PARSED_NO_MERCHANT": O.ST_PARSED_NO_MERCHANT,
          "SKIPPED_FILTER": O.ST_SKIPPED_FILTER, "CACHE_MISS": O.ST_CACHE_MISS,
          "UNMATCHED": O.ST_UNMATCHED, "BROKEN": O.ST_BROKEN, "FUTURE_DATE": O.ST_FUTURE_DATE,
          "FAILED_EXC": O.ST_FAILED_EXC, "ESCAPED": O.ST_ESCAPED}


def _dec(d):
    if d is None:
        return None
    s, digits, e = d.as_tuple()
    return [int(s), "".join(map(str, digits)), int(e)]


def _dump(p):
    if p is None:
        return None
    return {"msg_id": p.msg_id, "device_id": p.device_id, "sender": p.sender, "date": p.date.isoformat(),
            "tz": p.date.tzinfo is not None, "raw_body": p.raw_body, "txn_type": p.txn_type,
            "amount": _dec(p.amount), "currency": p.currency, "card": p.card, "merchant": p.merchant,
            "city": p.city, "address": p.address, "balance": _dec(p.balance),
            "parser_version": p.parser_version}


def test_decimals(golden):
    for row in golden("decimals.json"):
        text, kind = row[0], row[1]
        if kind == "ValueError":
            with pytest.raises(ValueError):
                O.ambiguous_decimal(text)
            assert O.decimal_triplet(text)[0] == O.DEC_ERR_SYNTAX
        else:
            d = O.ambiguous_decimal(text)
            assert _dec(d) == row[2:], text
            err, sign, coef, exp = O.decimal_triplet(text)
            if int(row[3]) > O.INT64_MAX:
                assert err == O.DEC_ERR_OVERFLOW
            else:
                assert (err, sign, coef, exp) == (O.DEC_OK, row[2], int(row[3]), row[4]), text


def test_keys_and_scans(golden):
    for c in golden("keys.json"):
        b = c["body"]
        assert O.msg_id_xml(b).hex() == c["msg_id"]
        assert O.msg_id_device(b).hex() == c["md5"]
        assert O.fixed_body(b) == c["fixed"], b
        assert O.cache_key(O.fixed_body(b)).hex() == c["cache_key"], b
        assert O.worker_skip(b) == c["worker_skip"], b
        assert O.parser_otp(b) == c["parser_otp"], b
        bd = O.body_date(b)
        assert (bd.isoformat() if bd else None) == c["body_date"], b


def test_parse_matrix(golden):
    now = dt.datetime(2026, 1, 1)
    for c in golden("parse_matrix.json"):
        key = O.cache_key(O.fixed_body(c["body"])).hex()
        out = O.replay_one(c["body"], "BANK", c["raw_date"], {key: c["resp"]}, now)
        assert out.status == STATUS[c["status"]], c["name"]
        want = c["parsed"]
        if want is not None:
            want = dict(want, msg_id=out.msg_id.hex())
        assert _dump(out.parsed) == want, c["name"]


@pytest.mark.parametrize("name", ["replay_plain.json", "replay_mixed.json"])
def test_replay(golden, name):
    from smsgate_b200 import synth
    g = golden(name)
    backup = synth.generate(g["n"], seed=g["seed"], **g["kwargs"])
    now = dt.datetime.fromisoformat(g["now"])
    raw_dates = [O.iso_utc_from_ms(ms) for ms in backup.date_ms]
    for lean in (False, True):        # lean = the form bench.py times as the CPU baseline: same results, less side work
        res = O.replay(backup.bodies, backup.senders, raw_dates, backup.cache, now, lean=lean)
        assert len(res.outcomes) == len(g["rows"])
        for i, (o, r) in enumerate(zip(res.outcomes, g["rows"])):
            assert o.msg_id.hex() == r["msg_id"]
            assert raw_dates[i] == r["date"]
            assert o.status == STATUS[r["status"]], i
            if r["cache_key"] is not None and (o.cache_key or not lean):
                assert o.cache_key.hex() == r["cache_key"]
            assert _dump(o.parsed) == r["parsed"], i
        assert {k.hex(): v for k, v in res.upserts.items()} == g["upserts"]
        first = {o.msg_id.hex(): i for i, o in enumerate(res.outcomes) if res.first_seen[i]}
        assert first == g["first"]
        assert sum(res.survivor) == len(g["upserts"])
