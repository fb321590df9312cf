"""TEST INFRASTRUCTURE - long differential fuzz runs on the CPU, beyond what `pytest -m "not gpu"` has time for.

    python scripts/fuzz_campaigns.py <campaign> [first_seed] [last_seed]

Two families:
  * the device logic compiled for the CPU (tests/hostcheck: the same *_core.cuh the kernels include) against the
    oracle: `scan` (keys, filters, rewrite, body date, decimals), `route` (cache entries mutated at random),
    `legacy` (both patterns against CPython's `re`), `xml` (byte mutations of backups against ElementTree),
    `xml_struct` (documents generated from a grammar of odd names, attributes, references);
  * the oracle against the REFERENCE ITSELF, imported from /root/reference (build container only): `ref_units`
    (worker filter lifted from worker.py, parser filter, clean + mask, fix_broken_datetime,
    parse_ambiguous_decimal), `ref_route` (parse_sms_llm + the worker / writer gates on random cache entries),
    `ref_dlq` (the DLQ handler and `_process_one` behind a fake NATS, damaged payloads), `ref_legacy` (parse_transaction_message / parse_credit_message), `ref_iter_sms` (the `_iter_sms` generator on mutated backups: rows, exception, rows before the exception).
Every mismatch prints a line containing MISMATCH; a clean run prints only the per-seed progress lines.
Round 2's runs and the one finding they produced are listed in DESIGN.md section 2."""
import datetime as dt
import hashlib
import os
import random
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

NOW = dt.datetime(2026, 1, 1)

ENTRY_POOL = {
    "txn_type": ["debit", "credit", "otp", "unknown", "DEBIT", "refund", None, 5, "", " debit"],
    "date": ["06.05.25 14:23", "10.06.2025 20:51", "31.02.25 10:00", "6.5.25 4:3", "06.05.25  14:23", "yesterday", "",
             "2025-06-10T20:51:00", "2025-06-10T20:51:00+04:00", "01.01.27 00:00", "12/31/2030", None, 1717000000,
             "06.05.25 14:23 ", "06.05.25 24:00", "June 10, 2025 8pm", "  ", "13.13.25 10:00", "00.00.00 00:00",
             "29.02.24 23:59", "29.02.25 23:59", "1.1.1", "2025", "10:30", "06.05.25", "06.05.25 14:23:59",
             "06-05-25 14:23", "٠٦.٠٥.٢٥ 14:23", 12.5, True, "2025-06-10 20:51:00Z",
             "Tue 10 Jun 2025", "10.06.25 20:51 +0400", "99.99.99 99:99", "31.12.99 23:59", "01.01.00 00:00",
             "01.01.69 00:00", "01.01.68 00:00"],
    "amount": ["52.00", "1,234.56", "1.234,56", "79 825,89", "-5.00", "abc", "", None, 12.5, 7, "1e5", "5-",
               "99999999999999999", "0.005", "1,5", -3, "١٢", True, "0", "-0", " 5 ", [1], {"a": 1}, 1e20,
               float("nan")],
    "currency": ["USD", "usd", None, "", 5, "руб", "ß"],
    "card": ["1234", "4083***7538", "**18", "*** 7538", "12 34 56", "", None, 1234, "١٢٣٤٥", "ab",
             "aéïöüß", "****", "* * 1 2 3 4", 12, "1234567890123456", " 1234", "12\t34"],
    "merchant": ["SHOP", "", None, 0, " ", "Ж"],
    "city": ["YEREVAN", None, 5],
    "address": ["null", "MAIN STR 1", None, "", "NULL", 7],
    "balance": ["1842.74", "-10", None, "x", "0", 3.5, "1.000.000", "", True],
}
ROUTE_BODIES = ["PAYMENT %d 06.05.25 14:23 done", "PAYMENT %d on 10.06.2025", "PAYMENT %d no date here",
                "PAYMENT %d 31.02.2025 then 01.03.25", "PAYMENT %d 4083***7538 12.12.2030",
                "PAYMENT %d 29.02.2024 x 30.02.24", "PAYMENT %d 99.99.9999 and 01.01.27",
                "PAYMENT %d ٠٦.٠٥.٢٠٢٥", "PAYMENT %d 1.1.25 06.05.25x",
                "PAYMENT %d 06.05.20255"]
SANE = dict(txn_type="debit", date="06.05.25 14:23", amount="52.00", currency="usd", card="4083***7538",
            merchant="SHOP", city=None, address="null", balance="1842.74")


def random_entry(rng):
    v = {k: rng.choice(vals) for k, vals in ENTRY_POOL.items()}
    if rng.random() < 0.6:                               # mostly sane entries with up to three odd fields
        for k in rng.sample(list(SANE), rng.randint(6, 9)):
            v[k] = SANE[k]
    for k in list(v):
        if rng.random() < 0.04:
            del v[k]
    if rng.random() < 0.02:
        v = rng.choice([None, "str", 5, [], {}])
    return v


def progress(seed, t0, **kw):
    print("seed", seed, "done", round(time.time() - t0), "s", " ".join(f"{k}={v}" for k, v in kw.items()), flush=True)


# ------------------------------------------------------------------ device logic (CPU build) against the oracle
def scan(first, last):
    from oracle import replay_oracle as O
    from smsgate_b200 import _lib as L
    from tests import _hostcheck as H
    from tests._fuzz import bodies
    t0, bad = time.time(), 0
    for seed in range(first, last):
        bs = list(bodies(8000, 1000 + seed))
        cols = H.scan(bs)
        off = cols["offsets"]
        for i, b in enumerate(bs):
            fl = int(cols["flags"][i])
            fixed = O.fixed_body(b)
            a = int(off[i])
            got = cols["fixed_bytes"][a:a + int(cols["fixed_len"][i])].tobytes().decode() if fl & L.F_MODIFIED else b
            ok = (cols["msg_id"][i].tobytes() == O.msg_id_xml(b) and bool(fl & L.F_WORKER_SKIP) == O.worker_skip(b)
                  and bool(fl & L.F_PARSER_OTP) == O.parser_otp(b) and got == fixed
                  and bool(fl & L.F_MODIFIED) == (fixed != b) and cols["cache_key"][i].tobytes() == O.cache_key(fixed)
                  and int(cols["body_date"][i]) == O.pack_date(O.body_date(b)))
            if not ok:
                bad += 1
                print("SCAN MISMATCH seed", seed, repr(b)[:200], flush=True)
        rng = random.Random(5000 + seed)
        alphabet = "0123456789" * 3 + "..,,  -   \t\nE+eAMD$٥１"
        strings = ["".join(rng.choice(alphabet) for _ in range(rng.choice([0, 1, 2, 3, 5, 8, 12, 18, 24])))
                   for _ in range(40000)]
        coef, exp, flags = H.decimal(strings)
        for i, t in enumerate(strings):
            want = O.decimal_triplet(t)
            err = int(flags[i]) & 3
            if err != want[0] or (err == 0 and ((int(flags[i]) >> 4) & 1, int(coef[i]), int(exp[i])) != want[1:]):
                bad += 1
                print("DECIMAL MISMATCH", repr(t), flush=True)
        progress(seed, t0, bad=bad)


def route(first, last):
    import tests.test_device_logic_on_cpu as T
    from oracle import replay_oracle as O
    from smsgate_b200 import gemini_cache
    from tests import _hostcheck as H
    t0, bad = time.time(), 0
    for seed in range(first, last):
        rng = random.Random(99000 + seed)
        bodies = [rng.choice(ROUTE_BODIES) % i for i in range(3000)]
        entries = [random_entry(rng) for _ in bodies]
        sc = H.scan(bodies)
        cache = {sc["cache_key"][i].tobytes().hex(): entries[i] for i in range(len(bodies))}
        jr = H.join_route(sc, T._table(cache), gemini_cache.pack_datetime(NOW))
        for i, b in enumerate(bodies):
            o = O.replay_one(b, "BANK", "2024-02-14T09:00:00+00:00", cache, NOW)
            try:
                assert int(jr["status"][i]) == o.status, "status"
                if o.parsed is not None and o.status not in (O.ST_FAILED_EXC, O.ST_ESCAPED):
                    assert int(jr["date"][i]) == O.pack_datetime(o.parsed.date), "date"
                    assert jr["card"][i, :int(jr["card_len"][i])].tobytes().decode() == o.parsed.card, "card"
                    fl = int(jr["dec_flags"][i])
                    for w, sh, sg, want in (("amount", 0, 4, o.amount), ("balance", 2, 5, o.balance)):
                        assert (fl >> sh) & 3 == want[0], w
                        if want[0] == 0:
                            assert ((fl >> sg) & 1, int(jr[w + "_coef"][i]), int(jr[w + "_exp"][i])) == want[1:], w
            except AssertionError as e:
                bad += 1
                print("ROUTE MISMATCH", e, repr(b), entries[i], int(jr["status"][i]), o.status, flush=True)
        progress(seed, t0, bad=bad)


LEGACY_PIECES = ["SALE", "PURCHASE", "purchase", " DB ", "INTERNET", "PURCH.COMPLETION.DB", ":", ": ", ", ", ",", "YEREVAN",
                 "Yerevan", "A", "shop 7", "06.05.25", "06/05/2025", "06-05-202", " 14:23", "14:23", ",card ***0018.",
                 "card ***0018", "CARD  ***9999.", ". ", " Amount:", "Amount: ", "amount:", "1,234.50", "5", "1.2.3", " USD",
                 "AMD", ", Balance:", "Balance:", "balance: ", "2", "\n", " ", "\t", "x", "TYPE", "9", ".", "***", "٥",
                 " ", "ſ", "ı", "İ", "é", "_", "Ж", "  ", "OTP", "otp", "PASS=", "CODE:", "\r\n",
                 "\x0b", "\x1c", "\x85", " ", "　", "K", "ǅ", "ß", "12:5", "31.12.1999", "1e5",
                 "١٢"]
LEGACY_BASES = [
    "APPROVED PURCHASE DB SALE: TEST LLC, YEREVAN, TEST STR. 29, 24 AREA,06.05.25 14:23,card ***0018. "
    "Amount:52.00 USD, Balance:1842.74 USD",
    "ACCOUNT REPLENISHMENT: 06.05.25 14:23, card ***0018. Amount:1,000.00 AMD, Balance:2,000.00 AMD",
    "PURCH.COMPLETION.DB INTERNET: SHOP, ONLINE ADDR,06/05/2025 14:23,card ***0018. Amount:1,234.50 AMD, Balance:9.99",
]


def legacy(first, last):
    import tests.test_legacy_logic_on_cpu as T
    t0, bad = time.time(), 0
    for seed in range(first, last):
        bodies = _legacy_bodies(random.Random(9000 + seed))
        try:
            T.check_against_oracle(T._res(bodies), bodies)
        except AssertionError as e:
            bad += 1
            print("LEGACY MISMATCH seed", seed, repr(e)[:300], flush=True)
        progress(seed, t0, bad=bad)


XML_SEEDS = [
    (b'<?xml version="1.0" encoding="UTF-8"?>\n<smses count="3">\n <sms address="BANK" date="1707900464673" '
     b'body="Pay 5&#10;now &amp; go" read="1"/>\n <sms address="\xd0\x91\xd0\xb0\xd0\xbd\xd0\xba" date="17" '
     b'body="\xd0\xa1\xd1\x83\xd0\xbc\xd0\xbc\xd0\xb0 1&#160;234,56"/>\n <sms date="3" address="C" body="z"></sms>\n</smses>\n'),
    (b"<?xml version='1.0' encoding='UTF-8' standalone='yes' ?>\n<!--File Created-->\n"
     b"<?xml-stylesheet type=\"text/xsl\" href=\"sms.xsl\"?>\n<smses count=\"2\">\n"
     b"  <sms protocol=\"0\" address=\"+37499\" date=\"1707900464673\" type=\"1\" "
     b"body=\"hi &amp; bye &#10;x &#x2022; &lt;&gt; &quot;q&quot; &apos;\" read=\"1\" />\n"
     b"  <mms date=\"5\"><parts><part seq=\"0\" text=\"t\"/></parts><sms address=\"in\" date=\"9\" body=\"nested\"/></mms>\n"
     b"  <sms address='A' date=' 1_000 ' body='q\"q'></sms>\n</smses>\n"),
    (b'<smses><sms address="A" date="1" body="x &#x1F600; \xf0\x9f\x98\x80 &#55357;"/><![CDATA[ x ]]><!-- c -->'
     b'<sms body="y" address="&#65;" date="-0"/></smses>'),
]
XML_BYTES = b'<>/"\'&;= #xa0\n\r\t-!?s[]Dc:1_+\x00\xc3\xa9\xff'


def xml(first, last):
    from tests import _hostcheck as H
    from tests.test_xml_logic_on_cpu import compare
    t0 = time.time()
    unsup = checked = bad = 0
    for seed in range(first, last):
        rnd = random.Random(7000 + seed)
        for _ in range(3000):
            d = bytearray(rnd.choice(XML_SEEDS))
            for _ in range(rnd.randint(1, 4)):
                k = rnd.randrange(len(d))
                op = rnd.random()
                if op < 0.3:
                    del d[k]
                elif op < 0.6:
                    d[k] = rnd.choice(XML_BYTES)
                elif op < 0.9:
                    d.insert(k, rnd.choice(XML_BYTES))
                else:
                    j = rnd.randrange(len(d))
                    d[k:k] = d[j:j + rnd.randint(1, 12)]
            data = bytes(d)
            try:
                r = compare(H.xml_extract(data), data)
                unsup += r == "unsupported"
                checked += r == "checked"
            except AssertionError:
                bad += 1
                print("XML MISMATCH", repr(data)[:400], flush=True)
        progress(seed, t0, checked=checked, out_of_dialect=unsup, bad=bad)


def xml_struct(first, last):
    from tests import _hostcheck as H
    from tests.test_xml_logic_on_cpu import compare
    names = [b"sms", b"sms", b"sms", b"mms", b"smses", b"Sms", b"sms2", b"s:ms", b"_sms", b"sm-s", b"\xd1\x81ms", b"sms.", b"a"]
    attrn = [b"body", b"address", b"date", b"read", b"type", b"Body", b"body2", b"xml:lang", b"xmlns", b"xmlns:a", b"a:b",
             b"_x", b"d-ate", b"\xc3\xa9", b"date", b"body"]
    vals = [b"x", b"", b"1707900464673", b"-5", b"+5", b" 7 ", b"1_0", b"a&amp;b", b"&lt;&gt;&quot;&apos;", b"&#10;", b"&#x41;",
            b"&#65", b"&#xZ;", b"&bogus;", b"&#0;", b"&#xD800;", b"&#x10FFFF;", b"&#x110000;", b"\xd0\x96", b"\xf0\x9f\x98\x80",
            b"\xff", b"\xc3", b"a\tb\nc\rd\r\ne", b"<", b">", b"]]>", b"'", b'"', b"&", b"a  b", b"\xc2\xa0", b"\xe2\x80\xa8",
            b"\xef\xbf\xbe", b"\xef\xbf\xbf", b"\x7f", b"\x01", b"9" * 25, b"&#9;&#13;&#32;"]
    ws = [b" ", b" ", b" ", b"\n", b"\t", b"\r\n"]
    prologs = [b"", b"", b"<?xml version='1.0' encoding='UTF-8' standalone='yes' ?>\n", b'<?xml version="1.0"?>', b"\xef\xbb\xbf",
               b"<!-- hi -->\n", b"<?xml-stylesheet type='text/xsl' href='sms.xsl'?>", b"\n", b"<!DOCTYPE smses>"]

    def attr(r):
        q = r.choice([b'"', b'"', b"'"])
        v = r.choice(vals)
        if r.random() < 0.7:
            v = v.replace(q, b"&quot;" if q == b'"' else b"&apos;")
        return r.choice(attrn) + r.choice([b"=", b"=", b" = ", b"= ", b"\n="]) + q + v + q

    def elem(r, depth=0):
        n = r.choice(names)
        attrs = b"".join(r.choice(ws) + attr(r) for _ in range(r.randint(0, 5)))
        if r.random() < 0.6 or depth > 2:
            return b"<" + n + attrs + r.choice([b"/>", b" />", b"/ >", b"></" + n + b">", b"> </" + n + b" >"])
        inner = b"".join(r.choice([elem(r, depth + 1), b"text", b"&amp;", b"<!-- c -->", b"<?pi x?>", b"\n ", b"<![CDATA[x]]>",
                                   b"&#65;"]) for _ in range(r.randint(0, 4)))
        close = n if r.random() < 0.95 else r.choice(names)
        return b"<" + n + attrs + b">" + inner + b"</" + close + b">"
    t0 = time.time()
    unsup = checked = bad = 0
    for seed in range(first, last):
        r = random.Random(31000 + seed)
        for _ in range(3000):
            root = r.choice([b"smses", b"smses", b"smses", b"allsms", b"a:smses"])
            rattr = b"".join(b" " + attr(r) for _ in range(r.randint(0, 2)))
            body = b"".join(r.choice([b"\n  ", b"", b" "]) + elem(r) for _ in range(r.randint(0, 6)))
            data = (b"".join(r.choice(prologs) for _ in range(r.randint(0, 3))) + b"<" + root + rattr + b">" + body +
                    r.choice([b"\n", b""]) + b"</" + root + b">" + r.choice([b"", b"\n", b"<!-- t -->", b"x"]))
            try:
                res = compare(H.xml_extract(data), data)
                unsup += res == "unsupported"
                checked += res == "checked"
            except AssertionError:
                bad += 1
                print("XML MISMATCH", repr(data)[:500], flush=True)
        progress(seed, t0, checked=checked, out_of_dialect=unsup, bad=bad)


# ------------------------------------------------------------------ the oracle against the reference itself
def _reference():
    from oracle.ref_runner import Reference
    return Reference(tempfile.mkdtemp(prefix="smsgate_fuzz_"))


def ref_units(first, last):
    from oracle import gen_golden as G
    from oracle import replay_oracle as O
    from tests._fuzz import bodies
    ref = _reference()
    worker_filter = G._worker_filter(ref)                 # the `if` test of worker.py:112-121, lifted with ast
    gp = ref.gemini_parser
    t0, bad = time.time(), 0
    cur = dt.datetime(2024, 2, 14, 9, 30)
    for seed in range(first, last):
        for b in bodies(6000, 40000 + seed):
            def outcome(f):
                try:
                    with ref.quiet():
                        return f()
                except Exception as e:                    # noqa: BLE001 - the exception type is the outcome
                    return type(e).__name__
            checks = [
                ("worker filter", worker_filter(b), O.worker_skip(b)),
                ("parser filter", any(k in b for k in ["OTP", "CODE:", "PASS:", "PASS=", "Daily limit exceeded:"]), O.parser_otp(b)),
                ("clean + mask", outcome(lambda: gp.mask_card_number_with_prefix(b.replace(" ", " ").replace("•", "*"))),
                 O.fixed_body(b)),
                ("sha1", hashlib.sha1(b.encode()).digest(), O.msg_id_xml(b)),
                ("fix_broken_datetime", outcome(lambda: gp.fix_broken_datetime(b, cur)), outcome(lambda: O.fix_datetime_from_body(b, cur))),
            ]
            for name, a, c in checks:
                if a != c:
                    bad += 1
                    print("REFERENCE MISMATCH", name, repr(b)[:200], a, c, flush=True)
        rng = random.Random(50000 + seed)
        alphabet = "0123456789" * 3 + "..,,  -   \t\nE+eAMD$٥１_"
        for _ in range(30000):
            t = "".join(rng.choice(alphabet) for _ in range(rng.choice([0, 1, 2, 3, 5, 8, 12, 18, 24])))

            def as_tuple(f):
                try:
                    return f(t).as_tuple()
                except Exception as e:                    # noqa: BLE001
                    return type(e).__name__
            a, c = as_tuple(ref.decimal_utils.parse_ambiguous_decimal), as_tuple(O.ambiguous_decimal)
            if a != c:
                bad += 1
                print("REFERENCE MISMATCH parse_ambiguous_decimal", repr(t), a, c, flush=True)
        progress(seed, t0, bad=bad)


def ref_route(first, last):
    import tests.test_oracle_golden as TG
    from oracle import gen_golden as G
    from oracle import replay_oracle as O
    ref = _reference()
    worker_filter = G._worker_filter(ref)
    gp = ref.gemini_parser
    rec = G._RecordingCache(gp.cache)
    gp.cache = rec
    t0, bad, n = time.time(), 0, 0
    for seed in range(first, last):
        rng = random.Random(88000 + seed)
        for i in range(3000):
            body = rng.choice(ROUTE_BODIES) % i
            v = random_entry(rng)
            raw_date = rng.choice(["2024-02-14T09:30:00+00:00", "1707900464", "1707900464673", "abc", "", "17079004.5"])
            fixed = gp.mask_card_number_with_prefix(body.replace(" ", " ").replace("•", "*"))
            key = hashlib.sha256(fixed.encode()).hexdigest()
            rec.inner.clear()
            rec.inner[key] = v
            try:
                status, parsed = G._route(ref, worker_filter, body, "BANK", raw_date)
                want = G._parsed_dump(parsed)
            except Exception as e:                        # noqa: BLE001
                status, want = "RAISED " + type(e).__name__, None
            try:
                out = O.replay_one(body, "BANK", raw_date, {key: O.copy_resp(v)}, G.NOW)
                got_status, got = out.status, TG._dump(out.parsed)
                if want is not None:
                    want = dict(want, msg_id=out.msg_id.hex())
            except Exception as e:                        # noqa: BLE001
                got_status, got = "RAISED " + type(e).__name__, None
            n += 1
            if TG.STATUS.get(status, status) != got_status or got != want:
                bad += 1
                print("REFERENCE MISMATCH parse_sms_llm", repr(body), raw_date, v, "\n   reference:", status, want,
                      "\n   oracle:   ", got_status, got, flush=True)
        progress(seed, t0, messages=n, bad=bad)


def _legacy_bodies(rnd):
    bodies = ["".join(rnd.choice(LEGACY_PIECES) for _ in range(rnd.randint(1, 26))) for _ in range(3000)]
    for _ in range(3000):
        s = list(rnd.choice(LEGACY_BASES))
        for _ in range(rnd.randint(1, 5)):
            k = rnd.randrange(len(s))
            op = rnd.random()
            if op < 0.4:
                del s[k]
            elif op < 0.7:
                s.insert(k, s[k])
            else:
                s[k] = rnd.choice(" ,.:*\n9aA\t٥é")
        bodies.append("".join(s))
    return bodies


def ref_legacy(first, last):
    """oracle/legacy_oracle.py against process_cached.parse_transaction_message / parse_credit_message themselves."""
    import logging

    from oracle import legacy_oracle as LO
    ref = _reference()
    logging.disable(logging.CRITICAL)                     # the reference logs every amount it cannot convert
    pc = ref.process_cached

    def outcome(f, b):
        try:
            with ref.quiet():
                return f(b)
        except Exception as e:                            # noqa: BLE001
            return type(e).__name__
    t0, bad, n = time.time(), 0, 0
    for seed in range(first, last):
        for b in _legacy_bodies(random.Random(9000 + seed)):
            n += 1
            a = (outcome(pc.parse_transaction_message, b), outcome(pc.parse_credit_message, b))
            c = (outcome(LO.parse_debit, b), outcome(LO.parse_credit, b))
            if a != c:
                bad += 1
                print("REFERENCE MISMATCH legacy", repr(b)[:200], a, c, flush=True)
        progress(seed, t0, bodies=n, bad=bad)


def ref_dlq(first, last):
    """dlq_worker._handle_dlq_msg(reparse=True) and worker._process_one, run unmodified behind a fake NATS
    (oracle/gen_golden_worker.Harness), against oracle/dlq_oracle.py on damaged and restructured DLQ payloads and on
    live RawSMS payloads with random cache entries: acks, escaped exception, counters, published bytes."""
    import json

    import torch  # noqa: F401 - before the reference's import stubs go in

    import tests.test_worker_flow_oracle as TW
    from oracle import dlq_oracle as DO
    from oracle import replay_oracle as O
    from oracle.gen_golden_worker import Harness
    from smsgate_b200 import synth
    from tests import _flow as FL
    h = Harness()
    offline = "offline: Gemini is not reachable"
    t0, bad, n = time.time(), 0, 0
    odd_values = [None, 5, 1.5, True, [], ["x"], {}, {"a": 1}, "", "x", "1707900000000", "2024-02-14T09:00:00+00:00"]
    for seed in range(first, last):
        rng = random.Random(66000 + seed)
        bodies = [rng.choice(ROUTE_BODIES) % i for i in range(300)]
        backup = synth.generate(100, seed=500 + seed, mixed_locale=True, otp_frac=0.1)
        bodies += list(backup.bodies)
        cache = {}
        for b in bodies:
            if rng.random() < 0.85:
                v = backup.cache.get(O.cache_key(O.fixed_body(b)).hex())
                cache[O.cache_key(O.fixed_body(b)).hex()] = v if v is not None and rng.random() < 0.5 else random_entry(rng)
        h.ref.seed_cache(cache)
        for b in bodies:
            raw = {"msg_id": hashlib.sha1(b.encode()).hexdigest(), "sender": rng.choice(["BANK", "Ж", "B 1"]), "body": b,
                   "date": rng.choice(["2024-02-14T09:00:00+00:00", "1707900000000", "1707900000", "abc", ""]),
                   "device_id": rng.choice(["xml_backup", None, "d1"]), "source": rng.choice(["xml", "device"])}
            r = rng.random()
            if r < 0.25:                                         # one field bent
                k = rng.choice(list(raw) + ["extra"])
                raw[k] = rng.choice(odd_values)
            elif r < 0.32:
                raw.pop(rng.choice(list(raw)))
            wrap = rng.random()
            if wrap < 0.55:
                obj = {"reason": "unmatched", "raw": raw}
            elif wrap < 0.62:
                obj = {"err": "boom", "entry": raw}              # no `raw`
            elif wrap < 0.68:
                obj = {"raw": rng.choice([None, [raw], "text", 5, json.dumps(raw)])}
            elif wrap < 0.72:
                obj = rng.choice([[1, 2], "str", 5, None, [raw]])
            else:
                obj = raw                                        # the live shape: the RawSMS itself
            data = json.dumps(obj, ensure_ascii=rng.random() < 0.5).encode()
            if rng.random() < 0.06:
                k = rng.randrange(len(data) + 1)
                data = data[:k] + rng.choice([b"\xff", b"{", b"}", b'"', b"\x00", b""]) + data[k + rng.randint(0, 2):]
            n += 1
            try:
                row = h.run_dlq(data)
                outcomes, rows, raws, outs, _, _, _ = DO.reparse([data], cache, NOW)
                out, rw = (outs[0], raws[0]) if rows[0] >= 0 else (None, None)
                FL.check_dlq_row(row, outcomes[0], out, rw, offline, TW._dump_json)
            except AssertionError as e:
                bad += 1
                print("REFERENCE MISMATCH dlq", data[:300], "\n   reference:", str(row)[:400], "\n   oracle:", outcomes, e, flush=True)
            # the live path: the same message as the watcher / gateway publish it, through worker._process_one
            try:
                live = DO.RawSMS(**raw)
            except Exception:                             # noqa: BLE001 - a bent field: not a message the live path sees
                continue
            row = h.run_worker(live.model_dump_json().encode())
            out = O.replay_one(live.body, live.sender, live.date, cache, NOW, device_id=live.device_id)
            try:
                want = FL.status_of(row, offline)
                assert out.status == want, (out.status, want)
                pj = TW._dump_json(out.parsed, live.msg_id) if want in (FL.UPSERT, FL.NO_MERCHANT) else None
                FL.check_published(row, want, pj, live.model_dump())
            except AssertionError as e:
                bad += 1
                print("REFERENCE MISMATCH worker", raw, "\n   reference:", str(row)[:400], "\n   oracle:", out.status, e, flush=True)
        progress(seed, t0, payloads=n, bad=bad)


XML_DATES = [b"99999999999999999999999", b"253402300800000", b"253402300799999", b"-62135596800001", b"-62135596800000",
             b"9223372036854775808", b"9223372036854775807", b"-9223372036854775808", b"-9223372036854775809", b" 5 ",
             b"1_0", b"+7", b"\xd9\xa1\xd9\xa2", b"", b"1e3", b"0x10", b"12345678901234567890", b"9300000000000000000000",
             b"-9300000000000000000000", b"00000000000000000000000000000000000007", b"\xef\xbc\x91"]


def ref_iter_sms(first, last):
    """xml_ingest.iter_sms above the CPU build of the tokenizer against the reference's _iter_sms generator: rows
    yielded, exception type, and the rows yielded before the exception."""
    import re
    from pathlib import Path

    import torch  # noqa: F401 - before the reference's import stubs go in

    from smsgate_b200 import xml_ingest
    from tests import _hostcheck as H
    ref = _reference()
    eng = H.HostTokenizerEngine()
    path = os.path.join(tempfile.mkdtemp(prefix="smsgate_fuzz_xml_"), "a.xml")

    def outcome(gen):
        rows = []
        try:
            for r in gen():
                rows.append(r.model_dump())
            return rows, None
        except Exception as e:                            # noqa: BLE001
            return rows, type(e).__name__
    t0, bad, n = time.time(), 0, 0
    for seed in range(first, last):
        rnd = random.Random(123000 + seed)
        for _ in range(1500):
            doc = bytearray(rnd.choice(XML_SEEDS))
            for _ in range(rnd.randint(0, 3)):
                k = rnd.randrange(len(doc))
                op = rnd.random()
                if op < 0.3:
                    del doc[k]
                elif op < 0.6:
                    doc[k] = rnd.choice(XML_BYTES)
                else:
                    doc.insert(k, rnd.choice(XML_BYTES))
            doc = bytes(doc)
            dates = list(re.finditer(rb'date="[^"]*"', doc))
            if dates and rnd.random() < 0.5:
                m = rnd.choice(dates)
                doc = doc[:m.start()] + b'date="' + rnd.choice(XML_DATES) + b'"' + doc[m.end():]
            with open(path, "wb") as f:
                f.write(doc)
            with ref.quiet():
                want = outcome(lambda: ref.watcher._iter_sms(Path(path)))
            got = outcome(lambda: xml_ingest.iter_sms(path, eng))
            n += 1
            if want != got:
                bad += 1
                print("REFERENCE MISMATCH _iter_sms", doc[:300], "\n   reference:", str(want)[:300], "\n   drop-in:  ", str(got)[:300],
                      flush=True)
        progress(seed, t0, files=n, bad=bad, through_elementtree=eng.host_fallbacks)


CAMPAIGNS = {"ref_dlq": ref_dlq, "ref_iter_sms": ref_iter_sms, "ref_legacy": ref_legacy, "scan": scan, "route": route, "legacy": legacy, "xml": xml, "xml_struct": xml_struct,
             "ref_units": ref_units, "ref_route": ref_route}

if __name__ == "__main__":
    if len(sys.argv) < 2 or sys.argv[1] not in CAMPAIGNS:
        raise SystemExit(__doc__)
    CAMPAIGNS[sys.argv[1]](int(sys.argv[2]) if len(sys.argv) > 2 else 0, int(sys.argv[3]) if len(sys.argv) > 3 else 10)
