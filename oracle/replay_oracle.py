"""TEST INFRASTRUCTURE — CPU restatement of the reference's bulk-replay path.

This module is the *checker*, never the product: only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` / `--impl reference`
legs may import it.  The shipped path (`smsgate_b200`) never does and fails
loudly when its CUDA library is missing.

Parity pin: the reference's own tests hold no vectors for this path
(SURVEY.md §4), so the oracle is pinned against outputs of the reference
itself, executed unmodified in the build container by `oracle/gen_golden.py`
and committed as `tests/golden/*.json`; `tests/test_oracle_golden.py` replays
them through this module on CPU.

Every function cites the reference lines it restates (paths relative to the
reference root).  Third-party arithmetic the reference leans on is used the
same way here: hashlib (OpenSSL SHA-1/SHA-256/MD5), `decimal`, `re`,
`datetime.strptime`, dateutil 2.9.0.post0, pydantic v2, diskcache 5.6.3.
"""
from __future__ import annotations

import datetime as _dt
import hashlib
import re
from dataclasses import dataclass, field
from decimal import Decimal, InvalidOperation
from enum import Enum
from typing import Any, Iterable, Optional

from dateutil.parser import parse as _dateutil_parse
from pydantic import BaseModel, Field

# --------------------------------------------------------------------------
# Status codes shared with include/smsgate_b200.h (SMSGATE_ST_*)
# --------------------------------------------------------------------------
ST_UPSERT = 0            # reached pb_writer with a truthy merchant -> upserted
ST_PARSED_NO_MERCHANT = 1  # published to sms.parsed, writer acks without upsert
ST_SKIPPED_FILTER = 2    # worker skip-filter hit: acked, nothing published
ST_CACHE_MISS = 3        # would call Gemini; offline replay -> DLQ set
ST_UNMATCHED = 4         # parse_sms_llm returned None -> sms.failed
ST_BROKEN = 5            # BrokenMessage: acked and dropped
ST_FUTURE_DATE = 6       # parsed.date > now: the handler then raises at worker.py:177 (.decode on a str), nothing published
ST_FAILED_EXC = 7        # parse_sms_llm raised something else -> sms.failed
ST_ESCAPED = 8           # aware date: the comparison at worker.py:174 raises TypeError outside every try

DLQ_STATUSES = (ST_CACHE_MISS, ST_UNMATCHED, ST_FAILED_EXC)     # pinned by tests/golden/worker_flow.json


# --------------------------------------------------------------------------
# Keys
# --------------------------------------------------------------------------
def msg_id_xml(body: str) -> bytes:
    """Idempotency key of an XML-sourced message: SHA-1 of the UTF-8 body
    (services/xml_watcher/watcher.py:45)."""
    return hashlib.sha1(body.encode()).digest()


def msg_id_device(body: str) -> bytes:
    """Idempotency key of a device-sourced message: MD5 of the UTF-8 body
    (libs/models.py:97-109 via services/api_gateway/main.py:113)."""
    return hashlib.md5(body.encode("utf-8")).digest()


_WORKER_UPPER_KEYS = (
    "OTP", "CODE:", "NOT ENOUGH FUNDS", "INSUFFICIENT FUNDS", "CREDIT PAYMENT",
    "C2C RECEIVED", "PASS:", "PASS=", "PERSON TO PERSON",
)
_WORKER_RAW_KEY = "Daily limit exceeded"


def worker_skip(body: str) -> bool:
    """The inline skip test of the parser worker: nine needles looked up in
    `body.upper()` plus one case-sensitive needle in the body itself
    (services/parser_worker/worker.py:112-121)."""
    up = body.upper()
    return any(k in up for k in _WORKER_UPPER_KEYS) or _WORKER_RAW_KEY in body


_PARSER_KEYS = ("OTP", "CODE:", "PASS:", "PASS=", "Daily limit exceeded:")


def parser_otp(body: str) -> bool:
    """Case-sensitive OTP test at the top of parse_sms_llm
    (libs/gemini_parser.py:198-200)."""
    return any(k in body for k in _PARSER_KEYS)


_CARD_RE = re.compile(r"\d{4}\*{3}(\d{4})")


def fixed_body(body: str) -> str:
    """NBSP -> space, U+2022 -> '*', then every leftmost non-overlapping
    `dddd***dddd` becomes `CARD:` + its last four digits
    (libs/gemini_parser.py:202-203 and :121-137)."""
    cleaned = body.replace("\u00a0", " ").replace("\u2022", "*")
    return _CARD_RE.sub(lambda m: "CARD:" + m.group(1), cleaned)


def cache_key(fixed: str) -> bytes:
    """Gemini response-cache key: SHA-256 of the rewritten body
    (libs/gemini_parser.py:207)."""
    return hashlib.sha256(fixed.encode()).digest()


# --------------------------------------------------------------------------
# Decimal canonicalisation
# --------------------------------------------------------------------------
def ambiguous_decimal(text: Any) -> Decimal:
    """libs/decimal_utils.py:4-63.  Raises ValueError like the reference."""
    if not isinstance(text, str):
        return Decimal(text)                                     # :10-12
    s = text.strip().replace(" ", "")                            # :15
    if s == "":
        return Decimal("0.0")                                    # :16-17
    dot, comma = s.rfind("."), s.rfind(",")                      # :20-21
    if dot >= 0 and comma >= 0:                                  # :25-33
        s = s.replace(".", "").replace(",", ".") if comma > dot else s.replace(",", "")
    elif comma >= 0:                                             # :35-44
        s = s.replace(",", "") if s.count(",") > 1 else s.replace(",", ".")
    elif dot >= 0 and s.count(".") > 1:                          # :46-53
        head, _, tail = s.rpartition(".")
        s = head.replace(".", "") + "." + tail
    s = re.sub(r"[^0-9.-]", "", s)                               # :60
    try:
        return Decimal(s)                                        # :62
    except InvalidOperation:
        raise ValueError(f"cannot convert {text!r} (cleaned {s!r})")


DEC_OK = 0
DEC_ERR_SYNTAX = 1      # reference raises ValueError
DEC_ERR_OVERFLOW = 2    # coefficient does not fit int64 (GPU representation limit)
INT64_MAX = (1 << 63) - 1


def decimal_triplet(text: str) -> tuple[int, int, int, int]:
    """(err, sign, coefficient, exponent) the way the GPU reports a decimal:
    `Decimal.as_tuple()` of the reference result with the digits folded into
    one integer.  The int64 overflow status is this build's representation
    limit (hard parts), not a reference behaviour."""
    try:
        d = ambiguous_decimal(text)
    except ValueError:
        return (DEC_ERR_SYNTAX, 0, 0, 0)
    sign, digits, exp = d.as_tuple()
    coef = int("".join(map(str, digits))) if digits else 0
    if coef > INT64_MAX:
        return (DEC_ERR_OVERFLOW, 0, 0, 0)
    return (DEC_OK, sign, coef, exp)


# --------------------------------------------------------------------------
# Dates
# --------------------------------------------------------------------------
def custom_datetime(value):
    """strptime('%d.%m.%y %H:%M') else dateutil (libs/gemini_parser.py:106-119)."""
    try:
        return _dt.datetime.strptime(value, "%d.%m.%y %H:%M")
    except Exception:
        return _dateutil_parse(value)


_BODY_DATE_FORMS = ((re.compile(r"\d{2}\.\d{2}\.\d{4}"), "%d.%m.%Y"),
                    (re.compile(r"\d{2}\.\d{2}\.\d{2}"), "%d.%m.%y"))


def body_date(body: str) -> Optional[_dt.date]:
    """The date `fix_broken_datetime` would take from the body: the first hit
    of dd.mm.yyyy if it is a real calendar date, else the first hit of
    dd.mm.yy if that is one, else nothing (libs/gemini_parser.py:73-100)."""
    for rx, fmt in _BODY_DATE_FORMS:
        m = rx.search(body)
        if not m:
            continue
        try:
            return _dt.datetime.strptime(m.group(0), fmt).date()
        except ValueError:
            continue
    return None


def fix_datetime_from_body(body: str, current):
    """libs/gemini_parser.py:67-104: replace the date part with the body's,
    keep the time.  `current.time()` raises AttributeError for a non-datetime,
    exactly as in the reference."""
    d = body_date(body)
    if d is None:
        return current
    return _dt.datetime.combine(d, current.time())


def pack_datetime(d: _dt.datetime) -> int:
    """Order-preserving int64 packing of a naive datetime used across the
    C ABI: y<<46 | m<<42 | d<<37 | H<<32 | M<<26 | S<<20 | microsecond."""
    return ((d.year << 46) | (d.month << 42) | (d.day << 37) | (d.hour << 32)
            | (d.minute << 26) | (d.second << 20) | d.microsecond)


def unpack_datetime(p: int) -> _dt.datetime:
    return _dt.datetime(p >> 46, (p >> 42) & 15, (p >> 37) & 31, (p >> 32) & 31,
                        (p >> 26) & 63, (p >> 20) & 63, p & 0xFFFFF)


def pack_date(d: Optional[_dt.date]) -> int:
    """int32 packing of the body date: y<<9 | m<<5 | d, 0 when absent."""
    return 0 if d is None else (d.year << 9) | (d.month << 5) | d.day


# --------------------------------------------------------------------------
# The cache-hit branch of parse_sms_llm
# --------------------------------------------------------------------------
class _TxnType(str, Enum):          # libs/models.py:35-41
    DEBIT = "debit"
    CREDIT = "credit"
    OTP = "otp"
    UNKNOWN = "unknown"


class _Core(BaseModel):             # libs/llm_core.py:9-19
    txn_type: _TxnType
    date: _dt.datetime
    amount: Optional[Decimal] = Field(None, ge=0)
    currency: Optional[str]
    card: Optional[str]
    merchant: Optional[str]
    city: Optional[str]
    address: Optional[str]
    balance: Optional[Decimal]


class BrokenMessage(Exception):     # libs/gemini_parser.py:20
    pass


@dataclass
class Parsed:
    """The fields of the reference's ParsedSMS (libs/models.py:60-95) that the
    replay path produces; `raw_body` is the rewritten body (:254)."""
    msg_id: str
    device_id: Optional[str]
    sender: str
    date: _dt.datetime
    raw_body: str
    txn_type: str
    amount: Optional[Decimal]
    currency: Optional[str]
    card: Optional[str]
    merchant: Optional[str]
    city: Optional[str]
    address: Optional[str]
    balance: Optional[Decimal]
    parser_version: str = "llm-0.2.0"


def parse_hit(msg_id: str, sender: str, device_id: Optional[str], raw_date: str,
              body: str, fixed: str, resp: Any) -> Optional[Parsed]:
    """Everything parse_sms_llm does after the cache lookup succeeded
    (libs/gemini_parser.py:224-271).  `resp` is a private copy of the cached
    dict.  Returns None / raises BrokenMessage / raises whatever the reference
    would let escape."""
    try:
        try:
            resp["date"] = custom_datetime(resp["date"])                    # :226
        except Exception as e:                                              # :227-229
            if "String does not contain a date" in str(e):
                # parse_unix_timestamp(int(raw.date), "Asia/Yerevan", aware=False)
                # (:139-188); int() of an ISO string raises and escapes to :238
                ts = float(int(raw_date))
                if ts < 0 or ts >= 1e14:
                    raise Exception("not a unix timestamp")
                secs = ts if ts < 1e11 else ts / 1000
                import zoneinfo
                resp["date"] = (_dt.datetime.fromtimestamp(secs, tz=_dt.timezone.utc)
                                .astimezone(zoneinfo.ZoneInfo("Asia/Yerevan"))
                                .replace(tzinfo=None))
        resp["date"] = fix_datetime_from_body(body, resp["date"])           # :230
        resp["card"] = resp["card"].replace("*", "").replace(" ", "")       # :232
        if len(resp["card"]) > 4:
            resp["card"] = resp["card"][:4]                                 # :233-234
        resp["amount"] = ambiguous_decimal(str(resp["amount"]))             # :235
        resp["balance"] = ambiguous_decimal(str(resp["balance"]))           # :236
        core = _Core.model_validate(resp)                                   # :237
    except Exception:
        resp["txn_type"]            # :239 reads it; a missing key escapes as KeyError
        return None                                                         # :241
    address = "" if core.address == "null" else core.address                # :243-244
    if len(str(core.card)) < 4:                                             # :246-247
        raise BrokenMessage("no card number")
    return Parsed(msg_id=msg_id, device_id=device_id, sender=sender, date=core.date,
                  raw_body=fixed, txn_type=core.txn_type.value, amount=core.amount,
                  currency=core.currency.upper() if core.currency else core.currency,  # models.py:93-95
                  card=core.card, merchant=core.merchant, city=core.city,
                  address=address, balance=core.balance)


# --------------------------------------------------------------------------
# Routing and upsert/dedup semantics
# --------------------------------------------------------------------------
@dataclass
class MessageOutcome:
    msg_id: bytes
    cache_key: bytes
    worker_skip: bool
    parser_otp: bool
    fixed: str
    hit: bool
    status: int
    body_date: int                        # pack_date
    parsed: Optional[Parsed] = None
    amount: tuple = (0, 0, 0, 0)          # decimal_triplet of the cached amount string
    balance: tuple = (0, 0, 0, 0)
    amount_cents: tuple = (False, 0)      # numeric_14_2 of the same strings
    balance_cents: tuple = (False, 0)


def copy_resp(v: Any) -> Any:
    return dict(v) if isinstance(v, dict) else v


def replay_one(body: str, sender: str, raw_date: str, cache, now: _dt.datetime,
               device_id: Optional[str] = "xml_backup", lean: bool = False) -> MessageOutcome:
    """One message through worker -> parse_sms_llm (hit branch) -> writer gate
    (services/parser_worker/worker.py:112-189, services/pb_writer/writer.py:65-84).
    `cache` is any mapping keyed by the 64-char hex cache key.

    `lean=True` is the form bench.py TIMES as the CPU baseline: the reference's own order of work and nothing the
    reference does not compute - the filters first, the rewrite and the cache key only for a message that gets
    that far, no decimal triplets / Numeric(14,2) cents / body-date column (those exist for the parity checks of
    the GPU columns).  Status, parsed record and dedup verdict are the same either way
    (tests/test_oracle_golden.py)."""
    mid = msg_id_xml(body)
    if lean:
        out = MessageOutcome(msg_id=mid, cache_key=b"", worker_skip=worker_skip(body), parser_otp=False, fixed="",
                             hit=False, status=ST_SKIPPED_FILTER, body_date=0)
        if out.worker_skip:                                       # worker.py:112-126
            return out
        out.parser_otp = parser_otp(body)                         # gemini_parser.py:198-200
        if out.parser_otp:
            out.status = ST_UNMATCHED
            return out
        out.fixed = fixed = fixed_body(body)                      # :202-203
        out.cache_key = ck = cache_key(fixed)                     # :207
    else:
        fixed = fixed_body(body)
        ck = cache_key(fixed)
        out = MessageOutcome(msg_id=mid, cache_key=ck, worker_skip=worker_skip(body),
                             parser_otp=parser_otp(body), fixed=fixed, hit=False,
                             status=ST_SKIPPED_FILTER, body_date=pack_date(body_date(body)))
    hexkey = ck.hex()
    out.hit = hexkey in cache                                     # gemini_parser.py:208
    resp0 = cache[hexkey] if out.hit else None                    # :209
    if isinstance(resp0, dict) and not lean:
        out.amount = decimal_triplet(str(resp0.get("amount")))
        out.balance = decimal_triplet(str(resp0.get("balance")))
        out.amount_cents = numeric_14_2(str(resp0.get("amount")))
        out.balance_cents = numeric_14_2(str(resp0.get("balance")))
    if out.worker_skip:                                           # worker.py:112-126
        return out
    if out.parser_otp:                                            # gemini_parser.py:198-200
        out.status = ST_UNMATCHED
        return out
    if not out.hit:                                               # :210-222 (network)
        out.status = ST_CACHE_MISS
        return out
    try:
        parsed = parse_hit(mid.hex(), sender, device_id, raw_date, body, fixed, copy_resp(resp0))
    except BrokenMessage:                                         # worker.py:136-140
        out.status = ST_BROKEN
        return out
    except Exception:                                             # worker.py:141-149
        out.status = ST_FAILED_EXC
        return out
    if parsed is None:                                            # worker.py:151-158
        out.status = ST_UNMATCHED
        return out
    out.parsed = parsed
    try:
        future = parsed.date > now                                # worker.py:174
    except TypeError:                                             # aware vs naive: escapes
        out.status = ST_ESCAPED
        return out
    if future:
        out.status = ST_FUTURE_DATE
    elif parsed.merchant:                                         # writer.py:70-74
        out.status = ST_UPSERT
    else:
        out.status = ST_PARSED_NO_MERCHANT
    return out


@dataclass
class ReplayResult:
    outcomes: list
    survivor: list = field(default_factory=list)     # index is the last occurrence of its msg_id among upserts
    first_seen: list = field(default_factory=list)   # index is the first occurrence (insert, not patch)
    last_idx: list = field(default_factory=list)     # index of the surviving occurrence, -1 when not upserted
    upserts: dict = field(default_factory=dict)      # msg_id -> index of the surviving record


def dedup(msg_ids: Iterable[bytes], eligible: Iterable[bool]):
    """Upsert-by-msg_id: first occurrence inserts, every later one overwrites
    all non-key columns, so the survivor is the LAST occurrence in replay order
    (libs/pocketbase.py:168-200, services/pb_writer/upsert.py:19-31,
    db/models.py:16)."""
    msg_ids = list(msg_ids)
    eligible = list(eligible)
    first: dict = {}
    last: dict = {}
    for i, (k, ok) in enumerate(zip(msg_ids, eligible)):
        if not ok:
            continue
        first.setdefault(k, i)
        last[k] = i
    n = len(msg_ids)
    survivor = [False] * n
    first_seen = [False] * n
    last_idx = [-1] * n
    for i, (k, ok) in enumerate(zip(msg_ids, eligible)):
        if ok:
            survivor[i] = last[k] == i
            first_seen[i] = first[k] == i
            last_idx[i] = last[k]
    return survivor, first_seen, last_idx, last


def replay(bodies, senders, raw_dates, cache, now: _dt.datetime, lean: bool = False) -> ReplayResult:
    outs = [replay_one(b, s, d, cache, now, lean=lean) for b, s, d in zip(bodies, senders, raw_dates)]
    res = ReplayResult(outcomes=outs)
    res.survivor, res.first_seen, res.last_idx, res.upserts = dedup(
        (o.msg_id for o in outs), (o.status == ST_UPSERT for o in outs))
    return res


def iso_utc_from_ms(date_ms: int) -> str:
    """RawSMS.date for XML input (services/xml_watcher/watcher.py:42-43,52)."""
    return _dt.datetime.fromtimestamp(date_ms / 1_000, tz=_dt.timezone.utc).isoformat()


def numeric_14_2(text: str):
    """What a Numeric(14,2) column keeps of parse_ambiguous_decimal(text) (db/models.py:21,27):
    (overflow, |cents|).  PostgreSQL rounds numeric halves away from zero and rejects more than
    12 integer digits.  Derived from PostgreSQL's documented behaviour; no reference code runs it."""
    from decimal import ROUND_HALF_UP
    err, sign, coef, exp = decimal_triplet(text)
    if err == DEC_ERR_OVERFLOW:
        return (True, 0)
    if err != DEC_OK:
        return (False, 0)
    cents = int((Decimal((0, tuple(map(int, str(coef))), exp))).quantize(Decimal("0.01"), rounding=ROUND_HALF_UP) * 100)
    return (True, 0) if cents > 99_999_999_999_999 else (False, cents)
