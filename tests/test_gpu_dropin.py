"""The per-message drop-ins (smsgate_b200/dropin.py) driven ONLY through the reference's own names and
signatures, against the vectors produced by executing the reference: parse_matrix.json (parse_sms_llm),
decimals.json (parse_ambiguous_decimal), legacy.json (parse_transaction_message / parse_credit_message),
iter_sms.json (_iter_sms)."""
import datetime as dt
import struct
from decimal import Decimal

import pytest

from oracle import replay_oracle as O
from smsgate_b200 import dropin
from smsgate_b200.models import BrokenMessage, RawSMS

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _engine():
    from smsgate_b200.replay import ReplayEngine
    eng = ReplayEngine(0)
    dropin.use_engine(eng)
    yield eng
    dropin.use_engine(None)
    eng.close()


def test_parse_sms_llm(golden):
    from smsgate_b200.replay import CacheMiss
    seen = set()
    for c in golden("parse_matrix.json"):
        if c["status"] == "SKIPPED_FILTER":
            continue                                  # the worker never calls parse_sms_llm for these
        key = O.cache_key(O.fixed_body(c["body"])).hex()
        dropin.load_cache({key: c["resp"]} if c["status"] != "CACHE_MISS" else {})
        raw = RawSMS(msg_id="m-1", sender="BANK", body=c["body"], date=c["raw_date"], device_id="dev-7", source="device")
        seen.add(c["status"])
        if c["status"] == "BROKEN":
            with pytest.raises(BrokenMessage):
                dropin.parse_sms_llm(raw)
        elif c["status"] == "CACHE_MISS":
            with pytest.raises(CacheMiss):
                dropin.parse_sms_llm(raw)
        elif c["status"] == "FAILED_EXC":
            with pytest.raises(Exception):
                dropin.parse_sms_llm(raw)
        elif c["status"] == "UNMATCHED":
            assert dropin.parse_sms_llm(raw) is None, c["name"]
        else:
            p = c["parsed"]
            big = any(int(p[w][1]) > O.INT64_MAX for w in ("amount", "balance"))
            if big:
                with pytest.raises(OverflowError):
                    dropin.parse_sms_llm(raw)
                continue
            got = dropin.parse_sms_llm(raw)
            assert (got.msg_id, got.device_id, got.sender) == ("m-1", "dev-7", "BANK")
            if c["status"] != "ESCAPED":              # an aware datetime: isoformat carries the offset
                assert got.date.isoformat() == p["date"], c["name"]
            assert got.raw_body == p["raw_body"] and got.txn_type.value == p["txn_type"]
            for w in ("amount", "balance"):
                s, digits, e = getattr(got, w).as_tuple()
                assert [s, "".join(map(str, digits)), e] == p[w], c["name"]
            for f in ("currency", "card", "merchant", "city", "address", "parser_version"):
                assert getattr(got, f) == p[f], (c["name"], f)
    assert {"UPSERT", "PARSED_NO_MERCHANT", "UNMATCHED", "BROKEN", "FAILED_EXC", "FUTURE_DATE"} <= seen


def test_parse_ambiguous_decimal(golden):
    rows = golden("decimals.json")
    for r in rows[::3]:
        if r[1] == "ValueError":
            with pytest.raises(ValueError):
                dropin.parse_ambiguous_decimal(r[0])
        elif int(r[3]) > O.INT64_MAX:
            with pytest.raises(OverflowError):
                dropin.parse_ambiguous_decimal(r[0])
        else:
            s, digits, e = dropin.parse_ambiguous_decimal(r[0]).as_tuple()
            assert [s, "".join(map(str, digits)), e] == [r[2], r[3], r[4]], r[0]
    assert dropin.parse_ambiguous_decimal(12.5) == Decimal(12.5)          # non-str: Decimal(x), decimal_utils.py:10-12
    assert dropin.parse_ambiguous_decimal(7) == Decimal(7)
    with pytest.raises(TypeError):
        dropin.parse_ambiguous_decimal(None)


def _bits(d):
    if d is None:
        return None
    out = dict(d)
    for k in ("amount", "balance"):
        out[k] = "%016x" % struct.unpack("<Q", struct.pack("<d", d[k]))[0]
    return out


def test_parse_transaction_and_credit_message(golden):
    cases = golden("legacy.json")["cases"]
    checked = 0
    for c in cases[::2]:
        for fn, want in ((dropin.parse_transaction_message, c["debit"]), (dropin.parse_credit_message, c["credit"])):
            try:
                got = fn(c["body"])
            except OverflowError:
                continue                               # beyond the kernel's exact float range: flagged, not guessed
            assert _bits(got) == want, c["body"]
            checked += 1
    assert checked > 0.9 * len(cases)
    assert dropin.parse_transaction_message("") is None and dropin.parse_credit_message(None) is None


def test_iter_sms(golden, tmp_path):
    for c in golden("iter_sms.json"):
        p = tmp_path / (c["name"] + ".xml")
        p.write_text(c["xml"], encoding="utf-8", newline="")
        rows, error = [], None                     # a generator: what it yields before it raises counts too
        try:
            for r in dropin._iter_sms(p):
                rows.append(r.model_dump())
        except Exception as e:                     # noqa: BLE001 - the type is compared with the reference's
            error = type(e).__name__
        assert rows == c["rows"] and error == c.get("error"), c["name"]
