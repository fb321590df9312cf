"""The kernels' per-message cores (same source, compiled for the CPU by
tests/hostcheck) against the golden vectors.  This pins the device LOGIC in a
GPU-less container; tests/test_gpu_parity.py runs the real kernels."""
import datetime as dt

import numpy as np
import pytest

from oracle import replay_oracle as O
from smsgate_b200 import _lib as L
from smsgate_b200 import gemini_cache, synth
from tests import _hostcheck as H

STATUS = {n: i for i, n in enumerate(L.STATUS_NAMES)}


def test_scan_matches_reference_vectors(golden):
    cases = golden("keys.json")
    bodies = [c["body"] for c in cases]
    cols = H.scan(bodies)
    off = cols["offsets"]
    for i, c in enumerate(cases):
        assert cols["msg_id"][i].tobytes().hex() == c["msg_id"], c["body"]
        fl = int(cols["flags"][i])
        assert bool(fl & L.F_WORKER_SKIP) == c["worker_skip"], c["body"]
        assert bool(fl & L.F_PARSER_OTP) == c["parser_otp"], c["body"]
        a = int(off[i])
        fixed = cols["fixed_bytes"][a:a + int(cols["fixed_len"][i])].tobytes().decode()
        assert fixed == c["fixed"], c["body"]
        assert bool(fl & L.F_MODIFIED) == (c["fixed"] != c["body"]), c["body"]
        assert cols["cache_key"][i].tobytes().hex() == c["cache_key"], c["body"]
        want = 0 if c["body_date"] is None else O.pack_date(dt.date.fromisoformat(c["body_date"]))
        assert int(cols["body_date"][i]) == want, c["body"]


def test_device_key_md5_matches_reference_vectors(golden):
    """get_md5_hash (libs/models.py:97-109): the reference's own outputs, then every padding boundary."""
    import hashlib
    cases = golden("keys.json")
    got = H.md5([c["body"] for c in cases])
    for i, c in enumerate(cases):
        assert got[i].tobytes().hex() == c["md5"], c["body"]
    edge = ["", "a", "x" * 55, "y" * 56, "z" * 57, "q" * 63, "r" * 64, "s" * 65, "t" * 119, "u" * 120, "v" * 121,
            "\u043f\u0440\u0438\u0432\u0435\u0442 " * 40, "w" * 1000]
    got = H.md5(edge)
    for i, b in enumerate(edge):
        assert got[i].tobytes() == hashlib.md5(b.encode()).digest() == O.msg_id_device(b), len(b)


def test_decimals_match_reference_vectors(golden):
    rows = golden("decimals.json")
    coef, exp, flags = H.decimal([r[0] for r in rows])
    for i, r in enumerate(rows):
        err = int(flags[i]) & 3
        if r[1] == "ValueError":
            assert err == L.DEC_SYNTAX, r[0]
        elif int(r[3]) > O.INT64_MAX:
            assert err == L.DEC_OVERFLOW, r[0]
        else:
            assert (err, (int(flags[i]) >> 4) & 1, int(coef[i]), int(exp[i])) == (0, r[2], int(r[3]), r[4]), r[0]


def test_strict_date_is_a_sound_subset_of_strptime():
    import random
    rnd = random.Random(3)
    alphabet = "0123456789" * 4 + "..::  \t-/a"
    cases = ["06.05.25 14:23", "6.5.25 4:23", "06.05.25  14:23", " 6.05.25 14:23", "06.05.25 14:23 ", "32.05.25 14:23",
             "31.02.25 10:00", "29.02.24 23:59", "29.02.23 23:59", "01.01.69 00:00", "01.01.68 00:00", "00.05.25 10:00",
             "06.00.25 10:00", "06.13.25 1:1", "06.05.25 24:00", "06.05.25 14:60", "06.05.25 7:5", "06.05.2025 14:23",
             "31.12.99 23:59", "1.1.00 0:0", "31.04.25 10:00", "30.04.25 10:00", "06.05.25\t14:23", "06.05.25\n14:23",
             "", "1", "06.05.25", "06.05.25 ", "06.05.25 1", "06.05.25 1:", "06.05.25 :1"]
    for _ in range(3000):
        d, m, y = rnd.randint(0, 40), rnd.randint(0, 15), rnd.randint(0, 99)
        h, mi = rnd.randint(0, 30), rnd.randint(0, 70)
        fmt = rnd.choice(["{:02d}.{:02d}.{:02d} {:02d}:{:02d}", "{}.{}.{:02d} {}:{}", "{:02d}.{}.{:02d}  {}:{:02d}"])
        cases.append(fmt.format(d, m, y, h, mi))
    for _ in range(2000):
        cases.append("".join(rnd.choice(alphabet) for _ in range(rnd.randint(0, 16))))
    got = H.strict_date(cases)
    accepted = 0
    for s, p in zip(cases, got):
        try:
            want = dt.datetime.strptime(s, "%d.%m.%y %H:%M")
        except ValueError:
            want = None
        if p > 0:                         # whenever the GPU form accepts, it must agree with strptime
            assert want is not None and gemini_cache.pack_datetime(want) == int(p), s
            accepted += 1
        elif want is not None:            # it may only decline strings outside the plain ASCII form
            assert not s.isascii() or s != s.strip() or any(ch in s for ch in "\x0b\x0c\x1c\x1d\x1e\x1f"), s
    assert accepted > 500


def _table(mapping):
    return gemini_cache.build_table(mapping, H.strict_date)


def test_parse_matrix_status_and_fields(golden):
    now = dt.datetime(2026, 1, 1)
    for c in golden("parse_matrix.json"):
        if c["raw_date"].lstrip("-").isdigit():
            continue                      # device-source raw.date (unix seconds): not an XML-replay case
        sc = H.scan([c["body"]])
        key = sc["cache_key"][0].tobytes().hex()
        table = _table({key: c["resp"]})
        jr = H.join_route(sc, table, gemini_cache.pack_datetime(now))
        assert int(jr["status"][0]) == STATUS[c["status"]], c["name"]
        p = c["parsed"]
        if p is None or c["status"] in ("FAILED_EXC", "ESCAPED"):
            continue
        got_date = gemini_cache.unpack_datetime(jr["date"][0])
        assert got_date.isoformat() == p["date"], c["name"]
        assert jr["card"][0, :int(jr["card_len"][0])].tobytes().decode() == p["card"], c["name"]
        fl = int(jr["dec_flags"][0])
        for which, sh, sg in (("amount", 0, 4), ("balance", 2, 5)):
            sign, digits, exp = p[which]
            if int(digits) > O.INT64_MAX:
                assert (fl >> sh) & 3 == L.DEC_OVERFLOW
            else:
                assert ((fl >> sh) & 3, (fl >> sg) & 1, int(jr[which + "_coef"][0]), int(jr[which + "_exp"][0])) == \
                    (0, sign, int(digits), exp), c["name"]


@pytest.mark.parametrize("name", ["replay_plain.json", "replay_mixed.json"])
def test_replay_against_reference_run(golden, name):
    g = golden(name)
    backup = synth.generate(g["n"], seed=g["seed"], **g["kwargs"])
    now = dt.datetime.fromisoformat(g["now"])
    sc = H.scan(backup.bodies)
    table = _table(backup.cache)
    jr = H.join_route(sc, table, gemini_cache.pack_datetime(now))
    for i, r in enumerate(g["rows"]):
        assert sc["msg_id"][i].tobytes().hex() == r["msg_id"]
        assert int(jr["status"][i]) == STATUS[r["status"]], i
        if r["cache_key"] is not None:
            assert sc["cache_key"][i].tobytes().hex() == r["cache_key"]
        p = r["parsed"]
        if p is not None:
            assert gemini_cache.unpack_datetime(jr["date"][i]).isoformat() == p["date"]
            assert int(jr["amount_coef"][i]) == int(p["amount"][1]) and int(jr["amount_exp"][i]) == p["amount"][2]
            assert int(jr["balance_coef"][i]) == int(p["balance"][1]) and int(jr["balance_exp"][i]) == p["balance"][2]
            assert jr["card"][i, :int(jr["card_len"][i])].tobytes().decode() == p["card"]


def test_numeric_14_2_storage_rounding():
    texts = ["0", "0.005", "0.0049999", "0.015", "1.005", "2.675", "52.00", "1,234.565", "999999999999.994",
             "999999999999.995", "1000000000000", "123456789012345678", "12345678901234567890", "0.00000000000000000001",
             "5", "5.5", "-2.675", "79.825,895", "0.994999999999999999", "abc", ""]
    import random
    rnd = random.Random(9)
    for _ in range(2000):
        texts.append("".join(rnd.choice("0123456789") for _ in range(rnd.randint(1, 16))) + "." +
                     "".join(rnd.choice("0123456789") for _ in range(rnd.randint(0, 6))))
    bodies = [f"body {i}" for i in range(len(texts))]
    sc = H.scan(bodies)
    cache = {sc["cache_key"][i].tobytes().hex(): dict(txn_type="debit", date="06.05.25 14:23", amount=t, currency="usd",
                                                      card="1234", merchant="M", city=None, address=None, balance=t)
             for i, t in enumerate(texts)}
    jr = H.join_route(sc, _table(cache), gemini_cache.pack_datetime(dt.datetime(2026, 1, 1)))
    for i, t in enumerate(texts):
        ov, cents = O.numeric_14_2(t)
        fl = int(jr["dec_flags"][i])
        assert bool(fl & 0x40) == ov and bool(fl & 0x80) == ov, t
        assert int(jr["amount_cents"][i]) == cents and int(jr["balance_cents"][i]) == cents, t


def test_numeric_14_2_against_postgres_documented_cases(golden):
    """The Numeric(14,2) columns (db/models.py:21,27) against PostgreSQL's documented numeric behaviour
    (tests/golden/postgres_numeric.json, sources cited there): rounding to the column's scale, ties away from
    zero, overflow when the rounded value needs more than precision - scale integer digits.  Both the oracle
    and the device code (compiled for the CPU) must agree with every documented case."""
    g = golden("postgres_numeric.json")
    rows = g["ties_away_from_zero"] + g["round_to_scale"] + g["precision"]
    texts = [r[0] for r in rows]
    bodies = [f"pg body {i}" for i in range(len(texts))]
    sc = H.scan(bodies)
    cache = {sc["cache_key"][i].tobytes().hex(): dict(txn_type="debit", date="06.05.25 14:23", amount="1", currency="usd",
                                                      card="1234", merchant="M", city=None, address=None, balance=t)
             for i, t in enumerate(texts)}
    jr = H.join_route(sc, _table(cache), gemini_cache.pack_datetime(dt.datetime(2026, 1, 1)))
    for i, (t, sign, want) in enumerate(rows):
        ov, cents = O.numeric_14_2(t)
        fl = int(jr["dec_flags"][i])
        if want == "overflow":
            assert ov and bool(fl & 0x80), t
        else:
            assert (ov, cents) == (False, want), t
            assert not (fl & 0x80) and int(jr["balance_cents"][i]) == want and ((fl >> 5) & 1) == sign, t


def test_scan_fuzz_against_oracle():
    """Adversarial bodies (tests/_fuzz.py): every scan output of the device code equals the oracle's."""
    from tests._fuzz import bodies
    bs = [b for b in bodies(6000, 81)]
    cols = H.scan(bs)
    off = cols["offsets"]
    for i, b in enumerate(bs):
        fl = int(cols["flags"][i])
        assert cols["msg_id"][i].tobytes() == O.msg_id_xml(b), repr(b)
        assert bool(fl & L.F_WORKER_SKIP) == O.worker_skip(b), repr(b)
        assert bool(fl & L.F_PARSER_OTP) == O.parser_otp(b), repr(b)
        fixed = O.fixed_body(b)
        a = int(off[i])
        got = cols["fixed_bytes"][a:a + int(cols["fixed_len"][i])].tobytes().decode() if fl & L.F_MODIFIED else b
        assert got == fixed and bool(fl & L.F_MODIFIED) == (fixed != b), repr(b)
        assert cols["cache_key"][i].tobytes() == O.cache_key(fixed), repr(b)
        assert int(cols["body_date"][i]) == O.pack_date(O.body_date(b)), repr(b)


def test_decimal_fuzz_against_oracle():
    """Random amount-like strings (both separators, every kind of white space, signs, letters, digits of
    other scripts, 0 to 24 characters): parse_ambiguous_decimal's result, exponent and sign included, and
    the Numeric(14,2) rounding derived from it."""
    import random
    rng = random.Random(17)
    alphabet = "0123456789" * 3 + "..,,  -\u00a0\u202f\u2009\t\nE+eAMD$\u0665\uff11"
    strings = ["".join(rng.choice(alphabet) for _ in range(rng.choice([0, 1, 2, 3, 5, 8, 12, 18, 24])))
               for _ in range(30000)]
    coef, exp, flags = H.decimal(strings)
    for i, t in enumerate(strings):
        want = O.decimal_triplet(t)
        err = int(flags[i]) & 3
        assert err == want[0], repr(t)
        if err == 0:
            assert ((int(flags[i]) >> 4) & 1, int(coef[i]), int(exp[i])) == want[1:], repr(t)


def test_decimal_ascii_fuzz_against_oracle():
    """The one-pass ASCII form of sg_parse_decimal on separator- and sign-heavy strings."""
    import random
    rng = random.Random(23)
    alphabet = "0123456789" * 2 + "....,,,,---  \t+eA"
    strings = ["".join(rng.choice(alphabet) for _ in range(rng.choice([1, 2, 3, 4, 6, 9, 14, 22])))
               for _ in range(80000)]
    strings += ["-", "-.", ".-", "-.5", ".-.5", "1.-2", "-1,2.3", "1,-2.3", "-,1.2", ",-1.2", "1,2,3", "1.2.3", "1,2.3,4",
                "1.2,3.4", "--1", "-1-", "1-", " - 1", "\t-\t1 ", "18446744073709551615", "18446744073709551616",
                "9223372036854775807", "9223372036854775808", "-9223372036854775808", "00000000000000000000001.5"]
    coef, exp, flags = H.decimal(strings)
    for i, t in enumerate(strings):
        want = O.decimal_triplet(t)
        err = int(flags[i]) & 3
        assert err == want[0], repr(t)
        if err == 0:
            assert ((int(flags[i]) >> 4) & 1, int(coef[i]), int(exp[i])) == want[1:], repr(t)


def test_route_fuzz_against_oracle():
    """Cache entries mutated at random (keys dropped, values of the wrong type, cards / dates / amounts of every
    odd shape the reference's parse_sms_llm has an opinion on) against bodies with and without a body date:
    status, date, card and amounts of the device logic equal the oracle's restatement of the hit branch."""
    import random
    rng = random.Random(23)
    now = dt.datetime(2026, 1, 1)
    pool = {
        "txn_type": ["debit", "credit", "otp", "unknown", "DEBIT", "refund", None, 5],
        "date": ["06.05.25 14:23", "10.06.2025 20:51", "31.02.25 10:00", "6.5.25 4:3", "06.05.25  14:23", "yesterday", "",
                 "2025-06-10T20:51:00", "2025-06-10T20:51:00+04:00", "01.01.27 00:00", "12/31/2030", None, 1717000000,
                 "06.05.25 14:23 ", "06.05.25 24:00", "June 10, 2025 8pm"],
        "amount": ["52.00", "1,234.56", "1.234,56", "79 825,89", "-5.00", "abc", "", None, 12.5, 7, "1e5", "5-",
                   "99999999999999999999", "0.005", "1,5", -3, "١٢"],
        "currency": ["USD", "usd", None, "", 5],
        "card": ["1234", "4083***7538", "**18", "*** 7538", "12 34 56", "", None, 1234, "١٢٣٤٥", "ab", "a\u00e9\u00ef\u00f6\u00fc\u00df"],
        "merchant": ["SHOP", "", None, 0],
        "city": ["YEREVAN", None],
        "address": ["null", "MAIN STR 1", None, ""],
        "balance": ["1842.74", "-10", None, "x", "0", 3.5, "1.000.000"],
    }
    bodies, entries = [], []
    for i in range(4000):
        body = rng.choice(["PAYMENT %d 06.05.25 14:23 done", "PAYMENT %d on 10.06.2025", "PAYMENT %d no date here",
                           "PAYMENT %d 31.02.2025 then 01.03.25", "PAYMENT %d 4083***7538 12.12.2030"]) % i
        v = {k: rng.choice(vals) for k, vals in pool.items()}
        if rng.random() < 0.6:                        # mostly sane entries with one or two odd fields
            v.update(txn_type="debit", date="06.05.25 14:23", amount="52.00", currency="usd", card="4083***7538",
                     merchant="SHOP", city=None, address="null", balance="1842.74")
            for k in rng.sample(list(pool), rng.choice([0, 1, 1, 2])):
                v[k] = rng.choice(pool[k])
        for k in list(v):
            if rng.random() < 0.03:
                del v[k]
        if rng.random() < 0.02:
            v = rng.choice([None, "a string", 5, ["list"]])
        bodies.append(body)
        entries.append(v)
    sc = H.scan(bodies)
    cache = {sc["cache_key"][i].tobytes().hex(): entries[i] for i in range(len(bodies))}
    jr = H.join_route(sc, _table(cache), gemini_cache.pack_datetime(now))
    seen = set()
    for i, b in enumerate(bodies):
        o = O.replay_one(b, "BANK", "2024-02-14T09:00:00+00:00", cache, now)
        assert int(jr["status"][i]) == o.status, (i, entries[i], b)
        seen.add(o.status)
        if o.parsed is not None and o.status not in (O.ST_FAILED_EXC, O.ST_ESCAPED):
            assert int(jr["date"][i]) == O.pack_datetime(o.parsed.date), (i, entries[i], b)
            assert jr["card"][i, :int(jr["card_len"][i])].tobytes().decode() == o.parsed.card, (i, entries[i])
            fl = int(jr["dec_flags"][i])
            for w, sh, sg, want in (("amount", 0, 4, o.amount), ("balance", 2, 5, o.balance)):
                assert (fl >> sh) & 3 == want[0], (i, w, entries[i])
                if want[0] == 0:
                    assert ((fl >> sg) & 1, int(jr[w + "_coef"][i]), int(jr[w + "_exp"][i])) == want[1:], (i, w, entries[i])
    assert {O.ST_UPSERT, O.ST_PARSED_NO_MERCHANT, O.ST_UNMATCHED, O.ST_BROKEN, O.ST_FUTURE_DATE, O.ST_FAILED_EXC} <= seen
