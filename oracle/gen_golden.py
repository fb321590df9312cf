"""TEST INFRASTRUCTURE — generates tests/golden/*.json by EXECUTING the reference.

Run once in the build container (the only place `/root/reference` exists):

    python oracle/gen_golden.py

Everything written is an output of the reference's own, unmodified code:
`_iter_sms`, `mask_card_number_with_prefix`, `parse_sms_llm` (its cache probe is
observed through a recording proxy put in place of the module's `cache`
object), `fix_broken_datetime`, `parse_ambiguous_decimal`, and the worker's
inline skip-filter, whose `if` test is lifted out of
services/parser_worker/worker.py with `ast` at generation time and evaluated
as-is (it lives inside a NATS-bound coroutine and cannot be called).  The two
remaining gates (`parsed.date > now`, `parsed.merchant`; worker.py:174,
writer.py:70-73) and upsert-by-msg_id are three lines restated below.
"""
from __future__ import annotations

import ast
import datetime as dt
import json
import os
import random
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle.ref_runner import REFERENCE_ROOT, Reference  # noqa: E402
from smsgate_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
NOW = dt.datetime(2026, 1, 1, 0, 0, 0)


class _RecordingCache:
    """Stands in for gemini_parser.cache; forwards everything and remembers
    the key of every membership test."""

    def __init__(self, inner):
        self.inner = inner
        self.seen = []

    def __contains__(self, k):
        self.seen.append(k)
        return k in self.inner

    def __getitem__(self, k):
        return self.inner[k]

    def __setitem__(self, k, v):
        self.inner[k] = v

    def __len__(self):
        return len(self.inner)


def _worker_filter(ref):
    """Compile the reference's skip test (worker.py:112-121) from its source."""
    src = open(os.path.join(REFERENCE_ROOT, "services/parser_worker/worker.py"), encoding="utf-8").read()
    tree = ast.parse(src)
    for node in ast.walk(tree):
        if isinstance(node, ast.If) and "NOT ENOUGH FUNDS" in ast.unparse(node.test):
            code = compile(ast.Expression(node.test), "worker.py:112", "eval")

            class _R:
                pass

            def test(body):
                r = _R()
                r.body = body
                return bool(eval(code, {"raw_sms": r}))
            return test
    raise RuntimeError("skip filter not found")


def _dec_tuple(d):
    s, digits, e = d.as_tuple()
    return [int(s), "".join(map(str, digits)), int(e)]


def _parsed_dump(p):
    if p is None:
        return None
    return {
        "msg_id": p.msg_id, "device_id": p.device_id, "sender": p.sender,
        "date": p.date.isoformat(), "tz": p.date.tzinfo is not None, "raw_body": p.raw_body,
        "txn_type": p.txn_type.value, "amount": None if p.amount is None else _dec_tuple(p.amount),
        "currency": p.currency, "card": p.card, "merchant": p.merchant, "city": p.city,
        "address": p.address, "balance": None if p.balance is None else _dec_tuple(p.balance),
        "parser_version": p.parser_version,
    }


# ---------------------------------------------------------------- decimals
DECIMAL_CASES = [
    "79,825.89", "79.825,89", "79 825,89", "79\u00a0825,89", "79\u202f825,89", "1,234,567.89",
    "1.234.567,89", "123456", "123.45", "1,23", "1,000", "999,999", "1.234.567", "1.000.000",
    "1,234,567", "1 234 567", "", "   ", "\u00a0", " \u00a0 ", "\t\n", ".5", "5.", "5,", "+5", "00012.50",
    "-5.00", "1e5", "52.00 USD", "AMD 1,000.00", "1.2.3,4", "1,2,3.4", "5.0.0", "None", "null", "5-",
    "--5", ".", ",", "-", "1,234.56.78", "1.234,56,78", "\u0665\u0662.\u0660\u0660", "\uff11\uff12\uff13",
    "0", "0.0", "0.00", "000", "-0", "-0.00", "0,5", "-,5", "-.5", "1-2", "- 5", "5 -", "12345678901234567",
    "123456789012345678", "9223372036854775807", "9223372036854775808", "92233720368547758.07",
    "99999999999999999999", "0.000000000000000001", "1,,2", "1..2", ",1,", ".1.", "1.,2", "1,.2", "a1b2c3",
    "1\u00a0234,5", "\u20091\u2009234,56\u2009", "1'234.56", "1_000", "$1,000.50", "1,000.50$", "(5.00)",
    "5.00-", "1 .5", "1. 5", "1 , 5", "\u066b5", "1\u066b5", "27,252.00", "391,469.09", "3460.00", "1842.74",
    "52.00", "1,00", "1,0", "1,", ",0", "10,000,000.00", "10.000.000,00", "10 000 000,00", "0,00",
    "1.5e3", "Infinity", "NaN", "1.0.0.0", "1,0,0,0", "1.0,0.0", "\n12.5\n", "\x1c7\x1f", "\x857\x85",
    "\u16807", "7\u3000", "\u200b7", "7\u200b",
]


def gen_decimals(ref):
    rnd = random.Random(0x5EED)
    cases = list(DECIMAL_CASES)
    alphabet = "0123456789" * 3 + ".,.,  -" + "\u00a0\u202f" + "Ae+$"
    for _ in range(600):
        cases.append("".join(rnd.choice(alphabet) for _ in range(rnd.randint(0, 14))))
    out = []
    for c in cases:
        try:
            out.append([c, "ok"] + _dec_tuple(ref.decimal_utils.parse_ambiguous_decimal(c)))
        except ValueError:
            out.append([c, "ValueError"])
    return out


# ---------------------------------------------------------------- key / scan cases
def key_bodies():
    t = [
        "APPROVED PURCHASE DB SALE: TEST LLC, MOSKOW, TEST STR. 29, 24 AREA,06.05.25 14:23,card ***0018. "
        "Amount:52.00 USD, Balance:1842.74 USD",
        "APPROVED PURCHASE DB SALE: TEST, MOSKOW,06.05.25 15:11,card ***0018. Amount:3460.00 USD, "
        "Balance:1800.74 USD",
        "DEBIT ACCOUNT\n27,252.00 AMD\n4083***7538,\nAMERIABANK API GATE, AM\n10.06.2025 20:51\n"
        "BALANCE: 391,469.09 AMD",
        "DEBIT ACCOUNT&#10;27,252.00 AMD&#10;4083***7538,&#10;AMERIABANK API GATE, AM&#10;10.06.2025 20:51",
        "a", "ab", "abc", "x" * 55, "x" * 56, "x" * 57, "x" * 63, "x" * 64, "x" * 65, "x" * 119, "x" * 120,
        "x" * 127, "x" * 128, "x" * 129, "y" * 1000, "z" * 5000,
        # card masking
        "4083***7538", "1234***5678***9012", "12345***67890", "1234****5678", "1234**5678", "1234***567",
        "123***4567", "\u0664\u0660\u0668\u0663***\u0667\u0665\u0663\u0668", "4083\u2022\u2022\u20227538",
        "4083\u2022**7538", "4083*\u2022*7538 and 1111***2222", "11112222***33334444***55556666",
        "1111***2222***3333***4444", "1111***22223333***4444", "\uff11\uff12\uff13\uff14***5678",
        "1234***\U0001d7ce\U0001d7cf\U0001d7d0\U0001d7d1", "12\u00a034***5678", "1234*\u00a0**5678",
        "pay 9999***0000.", "***", "****", "1234***", "***5678", "0000***0000" * 3,
        "\u00a0", "\u2022", "a\u00a0b\u2022c", "\u00a0\u00a0\u00a0", "Сумма 1\u00a0234,56 RUB Баланс 79\u202f825,89",
        # worker filter: upper-casing specials
        "otp", "Otp 1234", "your 0TP", "o t p", "code: 12", "Code:12", "CODE :12", "not enough funds",
        "Not Enough Funds", "NOT  ENOUGH FUNDS", "\u0131nsuff\u0131c\u0131ent funds", "insu\ufb03cient funds",
        "insuf\ufb01cient funds", "insu\ufb00icient funds", "in\u017fufficient funds", "credit payment",
        "CREDIT PAYMENT: 10.06.25 12:00, card ***7538. Amount:100,000.00 AMD, Balance:491,469.09 AMD",
        "c2c received", "C2C  RECEIVED", "pass:", "pa\u00df:", "pa\u00df=", "pas\u017f:", "PASS =", "pass=1",
        "person to person", "per\u017fon to person", "Daily limit exceeded", "daily limit exceeded",
        "Daily limit exceeded:", "DAILY LIMIT EXCEEDED", "xDaily limit exceededx", "Daily  limit exceeded",
        "\u0149ot enough funds", "\u0149OT ENOUGH FUNDS", "not enough fund\u00df", "NOT ENOUGH FUNDS\u00df",
        "\ufb01nsufficient funds", "\ufb03nsufficient funds", "credit paymen\ufb06", "credit paymen\u1e97",
        "c2c rece\u0131ved", "C2C RECE\u0130VED", "\u1e9aOTP", "\u01f0OTP", "O\u0308TP", "COTP", "OOTP", "OTOTP",
        "NOT ENOUGH NOT ENOUGH FUNDS", "PERSON TO PERSON TO PERSON", "PASS", "PAS", "CODE", "C2C RECEIVE",
        "INSUFFICIENT FUND", "CREDIT  PAYMENT", "PERSON TO  PERSON", "Daily limit exceede", "DDaily limit exceeded",
        "Daily limit Daily limit exceeded", "OT", "TP", "PASSS:", "PASS;", "CCODE:", "COCODE:", "CODCODE:",
        "\u00b5OTP", "\u03c2otp", "\u0130", "\u0131", "\u00df", "\ufb00\ufb01\ufb02\ufb03\ufb04\ufb05\ufb06",
        # dates in the body
        "on 10.06.2025 and 11.07.26", "on 31.02.2025 then 28.02.25", "99.99.9999 01.01.24", "32.01.2025",
        "00.01.2025 x", "10.13.2025 09.09.09", "10.06.202", "10.06.20251", "110.06.2025", "10.06.25 10.06.2025",
        "1\u0665.06.2025", "10.0\u0666.2025", "10.06.\u0662\u0660\u0662\u0665", "29.02.2024", "29.02.2023 29.02.24",
        "29.02.00", "29.02.1900", "31.04.25", "30.04.25", "1.1.25", "01.01.0000", "01.01.0001", "01.01.69",
        "01.01.68", "3\u0660.01.2025 15.01.25", "\u0661\u0665.01.2025", "2\u0669.02.2024", "10,06,2025",
        "10-06-2025", "10.06.2025", "x10.06.2025x", "12.34.56", "12.12.12.12", "date 07.08.2025 07.08.25",
    ]
    rnd = random.Random(7)
    pieces = ["OTP", "otp", "CODE:", "Pass=", "pass:", " ", "\u00a0", "\u2022", "*", "***", "1234", "5678", "99",
              ".", "10.06.2025", "06.05.25", "\u0664", "\u00df", "\u0131", "\ufb01", "Сумма", "Баланс", "\n",
              "Daily limit exceeded", ":", "NOT ENOUGH FUNDS", "x", "CARD:", "7", "\U0001d7d8", "\u00e9", "\u65e5\u672c"]
    for _ in range(400):
        t.append("".join(rnd.choice(pieces) for _ in range(rnd.randint(1, 14))))
    return t


def gen_keys(ref):
    gp = ref.gemini_parser
    filt = _worker_filter(ref)
    rec = _RecordingCache(gp.cache)
    gp.cache = rec
    sentinel = dt.datetime(1, 1, 1, 12, 34)
    out = []
    try:
        for body in key_bodies():
            raw = ref.models.RawSMS(msg_id="m", sender="s", body=body, date="2025-01-01T00:00:00", source="xml")
            rec.seen.clear()
            with ref.quiet():
                try:
                    got = gp.parse_sms_llm(raw)
                except Exception:                 # cache miss -> Gemini stub raises
                    got = "raised"
            parser_otp = got is None and not rec.seen
            fixed = gp.mask_card_number_with_prefix(body.replace("\u00a0", " ").replace("\u2022", "*"))
            import hashlib
            ck = rec.seen[0] if rec.seen else hashlib.sha256(fixed.encode()).hexdigest()
            with ref.quiet():     # two sentinels: the body may itself say 01.01.0001
                got_a = gp.fix_broken_datetime(body, sentinel).date()
                got_b = gp.fix_broken_datetime(body, sentinel.replace(year=2)).date()
            bd = got_a.isoformat() if got_a == got_b else None
            out.append({"body": body, "msg_id": hashlib.sha1(body.encode()).hexdigest(),
                        "md5": ref.models.get_md5_hash(body), "fixed": fixed, "cache_key": ck,
                        "cache_key_observed": bool(rec.seen), "worker_skip": filt(body),
                        "parser_otp": parser_otp, "body_date": bd})
    finally:
        gp.cache = rec.inner
    return out


# ---------------------------------------------------------------- parse_sms_llm outcome matrix
GOOD = dict(txn_type="debit", date="06.05.25 14:23", amount="52.00", currency="usd", card="***0018",
            merchant="TEST LLC", city="MOSKOW", address="TEST STR. 29", balance="1842.74")


def matrix_cases():
    # NB: no date string that dateutil completes from *today* (e.g. "14:23"): the
    # reference's answer for those changes with the wall clock.
    body = ("APPROVED PURCHASE DB SALE: TEST LLC, MOSKOW, TEST STR. 29, 24 AREA,06.05.25 14:23,card ***0018. "
            "Amount:52.00 USD, Balance:1842.74 USD")
    nodate = "APPROVED PURCHASE TEST LLC card ***0018 Amount 52.00 USD"
    longcard = "DEBIT ACCOUNT\n27,252.00 AMD\n4083***7538,\nAMERIABANK API GATE, AM\n10.06.2025 20:51"
    cases = []

    def add(name, b, resp, raw_date="2025-05-06T10:23:00+00:00"):
        cases.append({"name": name, "body": b, "raw_date": raw_date, "resp": resp})

    def mod(**kw):
        r = dict(GOOD)
        for k, v in kw.items():
            if v == "__del__":
                r.pop(k)
            else:
                r[k] = v
        return r

    add("good", body, GOOD)
    add("good_nodate_body", nodate, GOOD)
    for k in GOOD:
        add(f"missing_{k}", body, mod(**{k: "__del__"}))
        add(f"missing_{k}_nodate", nodate, mod(**{k: "__del__"}))
        add(f"none_{k}", body, mod(**{k: None}))
        add(f"int_{k}", body, mod(**{k: 18}))
    add("otp_txn_missing_amount", body, mod(txn_type="otp", amount="__del__"))
    add("txn_bogus", body, mod(txn_type="refund"))
    add("txn_upper", body, mod(txn_type="DEBIT"))
    for t in ("credit", "otp", "unknown"):
        add(f"txn_{t}", body, mod(txn_type=t))
    add("neg_amount", body, mod(amount="-5.00"))
    add("neg_zero_amount", body, mod(amount="-0.00"))
    add("neg_balance", body, mod(balance="-10,50"))
    add("amount_float", body, mod(amount=52.5))
    add("amount_int", body, mod(amount=52))
    add("amount_bad", body, mod(amount="abc"))
    add("amount_empty", body, mod(amount=""))
    add("balance_bad", body, mod(balance="1,234.56.78"))
    add("amount_big", body, mod(amount="123456789012345678901234.56"))
    add("amount_eu", body, mod(amount="1.234,56", balance="79\u00a0825,89"))
    for c in ("**18", "0018", "***0018", "4083***7538", "40 83", "1", "", "    ", "****", "\u0664\u0660\u0668\u0663",
              "12\u00e945", "123", "1234", "12345", "* 1 * 2 * 3 * 4 * 5"):
        add(f"card_{c!r}", body, mod(card=c))
    add("address_null", body, mod(address="null"))
    add("address_NULL", body, mod(address="NULL"))
    add("merchant_empty", body, mod(merchant=""))
    add("merchant_none", body, mod(merchant=None))
    add("currency_empty", body, mod(currency=""))
    add("currency_mixed", body, mod(currency="aMd"))
    add("extra_key", body, dict(GOOD, note="x"))
    for d in ("06.05.25 14:23", "6.5.25 4:23", "06.05.25  14:23", " 6.05.25 14:23", "06.05.25 14:23 ",
              "32.05.25 14:23", "31.02.25 10:00", "06.05.2025 14:23", "10.06.2025 20:51", "2025-05-06T14:23:00",
              "2025-05-06T14:23:00+04:00", "2025-05-06 14:23:05.250000", "yesterday", "", "   ",
              "06.05.25", "May 6 2025 2:23PM", "1\u0665.05.25 14:23", "06.05.25\u00a014:23", "06.05.25 24:00",
              "06.05.25 14:60", "1718000000", "29.02.24 23:59", "29.02.23 23:59", "01.01.69 00:00", "01.01.68 00:00",
              "13.13.25 10:00", "12.13.25 10:00", "13.12.25 10:00", "00.05.25 10:00", "06.05.25T14:23"):
        add(f"date_{d!r}", body, mod(date=d))
        add(f"date_{d!r}_nodate", nodate, mod(date=d))
    add("date_empty_unixraw", nodate, mod(date=""), raw_date="1718000000")
    add("date_empty_unixms", nodate, mod(date=""), raw_date="1718000000000")
    add("date_empty_unixbig", nodate, mod(date=""), raw_date="171800000000000")
    add("date_empty_neg", nodate, mod(date=""), raw_date="-5")
    add("future", body.replace("06.05.25", "06.05.27"), mod(date="06.05.27 14:23"))
    add("future_by_body", body.replace("06.05.25", "06.05.2031"), GOOD)
    add("aware_nobody", nodate, mod(date="2025-05-06T14:23:00+04:00"))
    add("longcard", longcard, dict(txn_type="debit", date="10.06.2025 20:51", amount="27,252.00", currency="AMD",
                                   card="4083***7538", merchant="AMERIABANK API GATE", city="AM", address="null",
                                   balance="391,469.09"))
    add("body_bad_date_then_good", body.replace("06.05.25", "31.02.2025 x 07.05.25"), GOOD)
    add("body_unicode_digit_date", nodate + " 1\u0665.06.2025", GOOD)
    add("resp_none", body, None)
    add("resp_list", body, [1, 2])
    add("resp_str", body, "oops")
    add("resp_empty", body, {})
    add("otp_body", "Your OTP is 1234 " + body, GOOD)
    add("daily_limit_colon", "Daily limit exceeded: " + body, GOOD)
    return cases


def _route(ref, filt, body, sender, raw_date, msg_id="m"):
    """worker.py:112-189 + writer.py:65-84 with the reference's own parse_sms_llm."""
    gp = ref.gemini_parser
    gp.cache.seen.clear()
    if filt(body):
        return "SKIPPED_FILTER", None
    raw = ref.models.RawSMS(msg_id=msg_id, sender=sender, body=body, date=raw_date,
                            device_id="xml_backup", source="xml")
    rec = gp.cache
    try:
        with ref.quiet():
            parsed = gp.parse_sms_llm(raw)
    except gp.BrokenMessage:
        return "BROKEN", None
    except Exception:
        if rec.seen and rec.seen[0] not in rec.inner:
            return "CACHE_MISS", None
        return "FAILED_EXC", None
    if parsed is None:
        return "UNMATCHED", None
    try:
        if parsed.date > NOW:                       # worker.py:174 with a frozen clock
            return "FUTURE_DATE", parsed
    except TypeError:
        return "ESCAPED", parsed
    return ("UPSERT" if parsed.merchant else "PARSED_NO_MERCHANT"), parsed   # writer.py:70


def gen_matrix(ref):
    import hashlib
    gp = ref.gemini_parser
    filt = _worker_filter(ref)
    out = []
    rec = _RecordingCache(gp.cache)
    gp.cache = rec
    try:
        for c in matrix_cases():
            fixed = gp.mask_card_number_with_prefix(c["body"].replace("\u00a0", " ").replace("\u2022", "*"))
            key = hashlib.sha256(fixed.encode()).hexdigest()
            rec.inner.clear()
            rec.inner[key] = c["resp"]
            status, parsed = _route(ref, filt, c["body"], "BANK", c["raw_date"])
            out.append(dict(c, status=status, parsed=_parsed_dump(parsed)))
    finally:
        gp.cache = rec.inner
        rec.inner.clear()
    return out


# ---------------------------------------------------------------- _iter_sms
XML_CASES = {
    "basic": '<smses count="2"><sms protocol="0" address="BANK" date="1707900464673" type="1" '
             'body="hello &amp; bye" read="1"/><sms address="B2" date="1707900465000" body="second"/></smses>',
    "entities": '<smses><sms address="A" date="1000" body="l1&#10;l2&#9;t&#160;n&#x2022;b &lt;&gt;&quot;&apos;"/></smses>',
    "literal_ws": '<smses><sms address="A" date="1000" body="a\nb\tc\r\nd"/></smses>',
    "no_date": '<smses><sms address="A" body="x"/></smses>',
    "nested_mms": '<smses><mms><sms address="A" date="1" body="inner"/></mms><sms address="A" date="2" body="outer"/></smses>',
    "missing_body": '<smses><sms address="A" date="1"/></smses>',
    "empty_body": '<smses><sms address="A" date="1" body=""/></smses>',
    "missing_address": '<smses><sms date="1" body="x"/></smses>',
    "bad_date": '<smses><sms address="A" date="12x" body="x"/></smses>',
    "negative_date": '<smses><sms address="A" date="-1000" body="x"/></smses>',
    "surrogate_ref": '<smses><sms address="A" date="1" body="&#55357;&#56832;"/></smses>',
    "astral": '<smses><sms address="A" date="1" body="ok &#128512; \U0001f600"/></smses>',
    "cyrillic": '<smses><sms address="Банк" date="1718000000123" body="Покупка 1&#160;234,56 RUB"/></smses>',
    "not_xml": 'plain text',
    "unclosed": '<smses><sms address="A" date="1" body="x">',
    "single_quotes": "<smses><sms address='A' date='5' body='it\"s'/></smses>",
    "other_root_child": '<smses><call number="1"/><sms address="A" date="7" body="z"/></smses>',
    "dup_attr": '<smses><sms address="A" address="B" date="7" body="z"/></smses>',
    "cdata_text": '<smses><sms address="A" date="7" body="z">text<![CDATA[x]]></sms></smses>',
    # _iter_sms is a generator: what it yields before it raises is part of its behaviour
    "bad_date_after_two": '<smses><sms address="A" date="1000" body="one"/><sms address="B" date="2000" body="two"/>'
                          '<sms address="C" date="12x" body="three"/><sms address="D" date="4000" body="four"/></smses>',
    "empty_body_after_one": '<smses><sms address="A" date="1000" body="one"/><sms address="B" date="2000" body=""/>'
                            '<sms address="C" date="3000" body="three"/></smses>',
    "year_10000_after_one": '<smses><sms address="A" date="253402300799999" body="last ms of 9999"/>'
                            '<sms address="B" date="253402300800000" body="year 10000"/></smses>',
    "beyond_int64_after_one": '<smses><sms address="A" date="-62135596800000" body="first ms of year 1"/>'
                              '<sms address="B" date="9223372036854775808" body="2^63"/></smses>',
    "beyond_time_t_after_one": '<smses><sms address="A" date="1000" body="one"/>'
                               '<sms address="B" date="100000000000000000000000" body="1e23"/></smses>',
    "odd_int_forms": '<smses><sms address="A" date=" +1_000 " body="one"/><sms address="B" date="-0" body="two"/>'
                     '<sms address="C" date="\u0661\u0662" body="arabic-indic twelve"/></smses>',
}


def gen_iter_sms(ref):
    out = []
    tmp = tempfile.mkdtemp()
    for name, text in XML_CASES.items():
        p = os.path.join(tmp, name + ".xml")
        with open(p, "w", encoding="utf-8", newline="") as f:
            f.write(text)
        from pathlib import Path
        rows = []                                  # with "error": what the generator yielded before it raised
        try:
            for r in ref.watcher._iter_sms(Path(p)):
                rows.append(r.model_dump())
            out.append({"name": name, "xml": text, "rows": rows})
        except Exception as e:
            out.append({"name": name, "xml": text, "rows": rows, "error": type(e).__name__})
    return out


# ---------------------------------------------------------------- a whole small replay
def gen_replay(ref, n, seed, **kw):
    from pathlib import Path
    gp = ref.gemini_parser
    filt = _worker_filter(ref)
    backup = synth.generate(n, seed=seed, **kw)
    tmp = tempfile.mkdtemp()
    xml = os.path.join(tmp, "sms.xml")
    synth.write_xml(xml, backup)
    rec = _RecordingCache(gp.cache)
    gp.cache = rec
    rows = []
    upserts = {}
    inserted_first = {}
    try:
        rec.inner.clear()
        for k, v in backup.cache.items():
            rec.inner[k] = v
        for i, raw in enumerate(ref.watcher._iter_sms(Path(xml))):
            status, parsed = _route(ref, filt, raw.body, raw.sender, raw.date, msg_id=raw.msg_id)
            ck = rec.seen[0] if rec.seen else None
            rows.append({"msg_id": raw.msg_id, "sender": raw.sender, "date": raw.date, "status": status,
                         "cache_key": ck, "parsed": _parsed_dump(parsed)})
            if status == "UPSERT":               # pocketbase.py:168-200 / upsert.py:19-31
                inserted_first.setdefault(raw.msg_id, i)
                upserts[raw.msg_id] = i
    finally:
        gp.cache = rec.inner
        rec.inner.clear()
    return {"n": n, "seed": seed, "kwargs": kw, "now": NOW.isoformat(), "rows": rows,
            "upserts": upserts, "first": inserted_first}


# ---------------------------------------------------------------- legacy regex path
def legacy_bodies():
    b1 = ("APPROVED PURCHASE DB SALE: TEST LLC, MOSKOW, TEST STR. 29, 24 AREA,06.05.25 14:23,card ***0018. "
          "Amount:52.00 USD, Balance:1842.74 USD")
    b2 = "APPROVED PURCHASE DB SALE: TEST, MOSKOW,06.05.25 15:11,card ***0018. Amount:3460.00 USD, Balance:1800.74 USD"
    cr = "CREDIT PAYMENT: 10.06.25 12:00, card ***7538. Amount:100,000.00 AMD, Balance:491,469.09 AMD"
    t = [b1, b2, cr,
         "DEBIT ACCOUNT\n27,252.00 AMD\n4083***7538,\nAMERIABANK API GATE, AM\n10.06.2025 20:51\nBALANCE: 391,469.09 AMD",
         "APPROVED PURCHASE DB INTERNET: SHOP 1, YEREVAN, ABOVYAN 12, 06.05.25 14:23,card ***0018. Amount:1,234.50 AMD, Balance:7.000,00 AMD",
         "PURCH.COMPLETION.DB INTERNET: SHOP, YEREVAN, 06/05/2025 14:23, card ***0018. Amount:5 usd, Balance:6",
         "purchase: shop, yerevan, addr, 06-05-25 14:23, CARD ***0018. amount:5 usd, balance:6",
         "SALE: A,B,C,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2",
         "SALE: A, YEREVAN,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2",
         "SALE: A, YEREVAN, ,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2",
         "SALE:  , YEREVAN, x,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2",
         "SALE: ,X, y,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2",
         "SALE:\n A\n,\n B,06.05.25\n14:23,\ncard\n***0018.\nAmount:\n1 USD,\nBalance:\n2",
         "xxSALE: A, B,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2 trailing text",
         "SALE: A, B,06.05.25 14:23,card ***0018. Amount:1.2.3 USD, Balance:2",
         "SALE: A, B,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2.3.4",
         "SALE: A, B,06.05.25 14:23,card ***0018. Amount:, USD, Balance:2",
         "SALE: A, B,06.05.25 14:23,card ***0018. Amount:. USD, Balance:2",
         "SALE: A, B,06.05.25 14:23,card ***0018. Amount:1,000,000.555 USD, Balance:123456789012345.67",
         "SALE: A, B,06.05.25 14:23,card ***0018. Amount:١٢.٥ USD, Balance:１２",
         "SALE: A, B,٠٦.05.25 14:23,card ***٠٠١٨. Amount:1 USÐ, Balance:2",
         "SALE: A, B,06.05.25 14:23,card ***0018. Amount:1 U_1, Balance:2",
         "SALE: A, B,06.05.25 14:23,card ***0018. Amount:1 US, Balance:2",
         "SALE: A, B,06.05.25 14:23,card ***0018. Amount:1 USDX, Balance:2",
         "SALE: A, B,06.05.20251 14:23,card ***0018. Amount:1 USD, Balance:2",
         "SALE: A, B,06.05.2025 14:23,card ***0018. Amount:1 USD, Balance:2",
         "SALE: A, B,06.05.202 14:23,card ***0018. Amount:1 USD, Balance:2",
         "SALE: A, B,06.05.2 14:23,card ***0018. Amount:1 USD, Balance:2",
         "ſALE: A, B,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2",
         "SALE: A, B,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2\n",
         "PURCHAſE DB İNTERNET: A, B,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2",
         "PURCHASE DB ıNTERNET: A, B,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2",
         "PURCHASE DB  INTERNET : A, B,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2",
         "PURCHASE SALE: A, B,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2",
         "SALE: A: B, SALE: C, D,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2",
         "SALE: A, B,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2 SALE: E, F,07.05.25 15:00,card ***0019. Amount:3 EUR, Balance:4",
         "SALE: A, B, 01.01.25 x,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2",
         "SALE: A, B,06.05.25 14:23 ,card ***0018. Amount:1 USD, Balance:2",
         "SALE: A, B,06.05.25  14:23,  card   ***0018.   Amount:   1   USD,   Balance:   2",
         "SALE: A, B,06.05.25 14:23,card **0018. Amount:1 USD, Balance:2",
         "SALE: A, B,06.05.25 14:23,card ****0018. Amount:1 USD, Balance:2",
         "SALE: A, B,06.05.25 14:23,card ***00189. Amount:1 USD, Balance:2",
         "SALE:A,B,06.05.25 14:23,card ***0018.Amount:1 USD,Balance:2",
         "SALE:  A , B ,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2",
         "", " ", "\n", "SALE:", "SALE: ,", "OTP SALE: A, B,06.05.25 14:23,card ***0018. Amount:1 USD, Balance:2",
         "pass=1", "code: 5", "oſp", "ﬆ", "CODE :",
         # credit
         "ACCOUNT REPLENISHMENT: 06.05.25 14:23, card ***0018. Amount:1,000.00 AMD, Balance:2,000.00 AMD",
         "ACCOUNT REPLENISHMENT: 06.05.25 14:23, card ***0018. Amount:1,000.00 AMD, Balance:2,000.00 AMD\n",
         "ACCOUNT REPLENISHMENT: 06.05.25 14:23, card ***0018. Amount:1,000.00 AMD, Balance:2,000.00 AMD\n\n",
         "ACCOUNT REPLENISHMENT: 06.05.25 14:23, card ***0018. Amount:1,000.00 AMD, Balance:2,000.00 AMD x",
         "line one\nACCOUNT REPLENISHMENT: 06.05.25 14:23, card ***0018. Amount:1,000.00 AMD, Balance:2,000.00 AMD",
         "a.b ACCOUNT REPLENISHMENT: 06.05.25 14:23, card ***0018. Amount:1 AMD, Balance:2 AMD",
         "a.b_c 9: 06.05.25 14:23, card ***0018. Amount:1 AMD, Balance:2 AMD",
         ": 06.05.25 14:23, card ***0018. Amount:1 AMD, Balance:2 AMD",
         "x : 06.05.25 14:23, card ***0018. Amount:1 AMD, Balance:2 AMD",
         "Пополнение счета: 06.05.25 14:23, card ***0018. Amount:1 AMD, Balance:2 руб",
         "T1: 06.05.25 14:23, T2: 07.05.25 15:24, card ***0018. Amount:1 AMD, Balance:2 AMD",
         "TYPE: 06.05.2025 14:23, card ***0018. Amount:1.5.5 AMD, Balance:2 AMD",
         "TYPE: 06.05.25 14:23, card ***0018. Amount:1 AMD, Balance:2 AMD AMD",
         "TYPE: 06.05.25 14:23, card ***0018. Amount:1 AMD, Balance:2  AMD  ",
         "TYPE: 06.05.25 14:23, card ***0018. Amount:1 AMD, Balance:2 AM",
         "TYPE: 06.05.25 14:23, card ***0018. Amount:1 AMD, Balance:2AMD",
         "credit payment: 10.06.25 12:00, card ***7538. amount:100,000.00 amd, balance:491,469.09 amd",
         ]
    rnd = random.Random(11)
    pieces = ["SALE", "PURCHASE", " DB ", "INTERNET", ":", ": ", ", ", ",", "YEREVAN", "A", "shop 7", "06.05.25", "06.05.2025",
              " 14:23", "14:23", ",card ***0018.", "card ***0018", ". ", " Amount:", "Amount: ", "1,234.50", "5", " USD", "AMD",
              ", Balance:", "Balance:", "2", "\n", " ", "x", "TYPE", "9", ".", "***", "٥", " "]
    for _ in range(500):
        t.append("".join(rnd.choice(pieces) for _ in range(rnd.randint(2, 22))))
    # well-formed shapes with random fields
    for i in range(120):
        sep = rnd.choice([".", "/", "-"])
        d = f"{rnd.randint(1,28):02d}{sep}{rnd.randint(1,12):02d}{sep}{rnd.choice(['25','2025','202'])}"
        kw = rnd.choice(["PURCHASE", "SALE", "PURCHASE DB INTERNET", "PURCH.COMPLETION.DB INTERNET", "Purchase", "sale"])
        city = rnd.choice(["YEREVAN, ", "Yerevan, ", "GYUMRI, ", ""])
        addr = rnd.choice(["ABOVYAN 12", "", "A, B, C", " "])
        amt = rnd.choice(["1", "52.00", "1,234.56", "1.234,56", "7.000,00", "0.10", "99999999.99", "12345678901234567", "1,2,3"])
        t.append(f"APPROVED {kw}: M{i} LLC, {city}{addr},{d} {rnd.randint(0,23):02d}:{rnd.randint(0,59):02d},card ***{rnd.randint(0,9999):04d}. "
                 f"Amount:{amt} {rnd.choice(['USD','amd','EUR'])}, Balance:{rnd.choice(['1842.74','0','1,000,000.01'])} USD")
        t.append(f"{rnd.choice(['ACCOUNT REPLENISHMENT','C2C','Refund 1'])}: {d} 12:00, card ***{rnd.randint(0,9999):04d}. "
                 f"Amount:{amt} AMD, Balance:491,469.09 AMD")
    return t


def _legacy_dump(d):
    if d is None:
        return None
    import struct
    out = dict(d)
    for k in ("amount", "balance"):
        out[k] = "%016x" % struct.unpack("<Q", struct.pack("<d", d[k]))[0]
    return out


def gen_legacy(ref):
    import logging
    logging.disable(logging.CRITICAL)
    pc = ref.process_cached
    cases = []
    for b in legacy_bodies():
        cases.append({"body": b, "debit": _legacy_dump(pc.parse_transaction_message(b)),
                      "credit": _legacy_dump(pc.parse_credit_message(b))})
    # the whole of process_sms_from_cache on a small source cache, through real diskcache directories
    import diskcache
    tmp = tempfile.mkdtemp()
    src, deb, cre = (os.path.join(tmp, n) for n in ("src", "deb", "cre"))
    bodies = legacy_bodies()[:140]
    with diskcache.Cache(src) as c:
        for i, b in enumerate(bodies):
            c.set(str(1707900000000 + i), {"address": "BANK", "date": str(1707900000000 + i), "body": b})
        c.set("already", {"body": bodies[0], "status": "processed"})
        c.set("nobody", {"address": "X"})
    pc.process_sms_from_cache(src, deb, cre)
    with diskcache.Cache(src) as c:
        statuses = {k: c.get(k).get("status") for k in c}
    def dump_cache(path):
        with diskcache.Cache(path) as c:
            return {k: dict(_legacy_dump({kk: vv for kk, vv in c.get(k).items() if kk not in ("msg_id", "original_body")}),
                            msg_id=c.get(k)["msg_id"], original_body=c.get(k)["original_body"]) for k in c}
    logging.disable(logging.NOTSET)
    return {"cases": cases, "process": {"bodies": bodies, "statuses": statuses, "debit": dump_cache(deb),
                                        "credit": dump_cache(cre)}}


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = Reference(tempfile.mkdtemp())
    try:
        def dump(name, obj):
            with open(os.path.join(OUT, name), "w", encoding="utf-8") as f:
                json.dump(obj, f, ensure_ascii=True, separators=(",", ":"))
            print(name, os.path.getsize(os.path.join(OUT, name)), "bytes")

        dump("decimals.json", gen_decimals(ref))
        dump("keys.json", gen_keys(ref))
        dump("parse_matrix.json", gen_matrix(ref))
        dump("iter_sms.json", gen_iter_sms(ref))
        dump("legacy.json", gen_legacy(ref))
        dump("replay_plain.json", gen_replay(ref, 400, 0x5EED + 1, dup_frac=0.1))
        dump("replay_mixed.json", gen_replay(ref, 500, 0x5EED + 5, mixed_locale=True, miss_frac=0.3,
                                             dup_frac=0.2, hot_key_zipf=1.2))
    finally:
        ref.close()


if __name__ == "__main__":
    main()
