"""Deterministic synthetic SMS backups and Gemini response caches.

Shapes follow SURVEY.md §8(d): the three bodies of the reference's own tests
(tests/test_parsers.py:14-16, :30-32, :45-46), a `PURCHASE DB INTERNET` shape,
a credit shape, OTP/CODE:/PASS= shapes and, for the mixed-locale config,
Cyrillic bodies with NBSP / U+202F thousands separators and comma decimals,
`4083***7538` / `4083•••7538` card forms and multi-part long messages.

The cache side is what the reference itself would have stored: key =
SHA-256 of the rewritten body (libs/gemini_parser.py:202-207), value = the
dict Gemini is asked for (libs/gemini_parser.py:46-61).  Key derivation here
uses hashlib/re directly; it is data generation, independent of both the CUDA
path and `oracle/`.
"""
from __future__ import annotations

import datetime as _dt
import hashlib
import re
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

_CARD_RE = re.compile(r"\d{4}\*{3}(\d{4})")

MERCHANT_WORDS = ("YEREVAN CITY", "SAS", "ZOVQ", "EVRIKA", "ALFA PHARM", "GLOVO", "YANDEX GO",
                  "WILDBERRIES", "OZON", "NATALI PHARM", "CARREFOUR", "PARMA", "KFC", "TASHIR PIZZA",
                  "TEST LLC", "AMERIABANK API GATE", "IDRAM", "EASYPAY", "TELCELL", "UCOM", "VIVA-MTS",
                  "GAZPROM ARMENIA", "VEOLIA JUR", "ENA", "ZARA", "LC WAIKIKI", "APPLE.COM/BILL",
                  "GOOGLE *YOUTUBE", "NETFLIX.COM", "SPOTIFY", "BOLT.EU", "GG TAXI")
CITIES = ("YEREVAN", "MOSKOW", "GYUMRI", "VANADZOR", "DILIJAN", "ABOVYAN", "TBILISI", "AM")
STREETS = ("TEST STR. 29, 24 AREA", "ABOVYAN 12", "KOMITAS AVE 60", "TUMANYAN 8", "BAGHRAMYAN 2/1",
           "MASHTOTS AVE 50", "ARSHAKUNYATS 34", "TIGRAN METS 4")
CURRENCIES = ("AMD", "USD", "EUR", "RUB")
RU_MERCHANTS = ("ПЯТЕРОЧКА", "МАГНИТ", "ЛЕНТА", "ЯНДЕКС ТАКСИ", "ОЗОН", "АПТЕКА 36.6", "ВКУСВИЛЛ",
                "ПЕРЕКРЕСТОК", "ДИКСИ", "СБЕРМАРКЕТ")


def reference_cache_key(body: str) -> str:
    """The key the reference stores a Gemini answer under."""
    cleaned = body.replace("\u00a0", " ").replace("\u2022", "*")
    fixed = _CARD_RE.sub(lambda m: "CARD:" + m.group(1), cleaned)
    return hashlib.sha256(fixed.encode()).hexdigest()


def _fmt_amount(cents: int, thousands: str = ",", point: str = ".") -> str:
    whole, frac = divmod(int(cents), 100)
    s = f"{whole:,}".replace(",", thousands)
    return f"{s}{point}{frac:02d}"


@dataclass
class Backup:
    """One synthetic backup: parallel per-message lists plus the cache that
    makes the requested share of bodies hit."""
    bodies: list
    senders: list
    date_ms: list
    cache: dict                       # hex key -> response dict
    kinds: list = field(default_factory=list)


def generate(n: int, *, seed: int = 0x5EED, otp_frac: float = 0.05, dup_frac: float = 0.0,
             miss_frac: float = 0.0, mixed_locale: bool = False, hot_key_zipf: Optional[float] = None,
             base_date: _dt.datetime = _dt.datetime(2024, 2, 14, 9, 0)) -> Backup:
    """Generate `n` messages.  `dup_frac` of them repeat an earlier body
    (uniformly, or Zipf(`hot_key_zipf`)-distributed over the distinct set);
    `miss_frac` of the distinct non-OTP bodies are left out of the cache."""
    rng = np.random.Generator(np.random.PCG64(seed))
    n_dup = int(round(n * dup_frac))
    n_distinct = n - n_dup
    # vector draws first so the stream is independent of the formatting below
    kind_u = rng.random(n_distinct)
    merch_i = rng.integers(0, len(MERCHANT_WORDS), n_distinct)
    merch_no = rng.integers(0, 1000, n_distinct)
    city_i = rng.integers(0, len(CITIES), n_distinct)
    street_i = rng.integers(0, len(STREETS), n_distinct)
    ccy_i = rng.integers(0, len(CURRENCIES), n_distinct)
    amt = np.exp(rng.uniform(np.log(100.0), np.log(999_999_999.0), n_distinct)).astype(np.int64)
    bal = rng.integers(0, 99_999_999_999, n_distinct)
    card = rng.integers(0, 10000, n_distinct)
    card_hi = rng.integers(0, 10000, n_distinct)
    miss_u = rng.random(n_distinct)
    locale_u = rng.random(n_distinct)
    otp_code = rng.integers(0, 1_000_000, n_distinct)

    bodies, kinds = [], []
    cache: dict = {}
    # one distinct (date, time) per message keeps bodies pairwise distinct; string tables avoid strftime per message
    base_min = base_date.hour * 60 + base_date.minute
    n_days = (base_min + min(n_distinct, 900_000)) // 1440 + n_distinct // 900_000 + 2
    day0 = base_date.date()
    d2_tab = [(day0 + _dt.timedelta(days=k)).strftime("%d.%m.%y") for k in range(n_days)]
    d4_tab = [(day0 + _dt.timedelta(days=k)).strftime("%d.%m.%Y") for k in range(n_days)]
    hm_tab = [f"{k // 60:02d}:{k % 60:02d}" for k in range(1440)]
    for i in range(n_distinct):
        tmin = base_min + i % 900_000
        day = tmin // 1440 + i // 900_000
        d2, d4, hm = d2_tab[day], d4_tab[day], hm_tab[tmin % 1440]
        m = f"{MERCHANT_WORDS[merch_i[i]]} {merch_no[i]}"
        city = CITIES[city_i[i]]
        ccy = CURRENCIES[ccy_i[i]]
        a, b = int(amt[i]), int(bal[i])
        c4 = f"{card[i]:04d}"
        resp = None
        u = kind_u[i]
        if mixed_locale and locale_u[i] < 0.40:
            sep = "\u00a0" if (i & 1) else "\u202f"
            a_s, b_s = _fmt_amount(a, sep, ","), _fmt_amount(b, sep, ",")
            rm = RU_MERCHANTS[merch_i[i] % len(RU_MERCHANTS)]
            body = (f"Покупка {rm} {merch_no[i]}. Карта *{c4}. Сумма {a_s} RUB. "
                    f"Баланс {b_s} RUB. {d2} {hm}")
            kind = "ru"
            resp = dict(txn_type="debit", date=f"{d2} {hm}", amount=a_s, currency="RUB", card=c4,
                        merchant=f"{rm} {merch_no[i]}", city=None, address=None, balance=b_s)
        elif mixed_locale and locale_u[i] < 0.50:
            stars = "***" if (i & 1) else "\u2022\u2022\u2022"
            a_s, b_s = _fmt_amount(a), _fmt_amount(b)
            body = (f"DEBIT ACCOUNT\n{a_s} AMD\n{card_hi[i]:04d}{stars}{c4},\n{m}, AM\n"
                    f"{d4} {hm}\nBALANCE: {b_s} AMD")
            kind = "longcard"
            resp = dict(txn_type="debit", date=f"{d4} {hm}", amount=a_s, currency="AMD",
                        card=f"{card_hi[i]:04d}***{c4}", merchant=m, city="AM", address="null",
                        balance=b_s)
        elif mixed_locale and locale_u[i] < 0.60:
            a_s, b_s = _fmt_amount(a), _fmt_amount(b)
            filler = " ".join(f"{STREETS[(street_i[i] + k) % len(STREETS)]};" for k in range(8 + i % 40))
            body = (f"APPROVED PURCHASE DB SALE: {m}, {city}, {filler},{d2} {hm},card ***{c4}. "
                    f"Amount:{a_s} {ccy}, Balance:{b_s} {ccy}")
            kind = "multipart"
            resp = dict(txn_type="debit", date=f"{d2} {hm}", amount=a_s, currency=ccy, card=c4,
                        merchant=m, city=city, address=filler, balance=b_s)
        elif u < otp_frac:
            j = i % 3
            body = (f"Your OTP is {otp_code[i]:06d}. Do not share it. {d2} {hm}",
                    f"CODE: {otp_code[i]:06d} for payment {_fmt_amount(a)} {ccy} at {m} {d2} {hm}",
                    f"3DS PASS={otp_code[i]:06d} {m} {_fmt_amount(a)} {ccy} {d2} {hm}")[j]
            kind = "otp"
        elif u < otp_frac + 0.19:
            a_s, b_s = _fmt_amount(a), _fmt_amount(b)
            addr = STREETS[street_i[i]]
            body = (f"APPROVED PURCHASE DB SALE: {m}, {city}, {addr},{d2} {hm},card ***{c4}. "
                    f"Amount:{a_s} {ccy}, Balance:{b_s} {ccy}")
            kind = "sale_addr"
            resp = dict(txn_type="debit", date=f"{d2} {hm}", amount=a_s, currency=ccy, card=c4,
                        merchant=m, city=city, address=addr, balance=b_s)
        elif u < otp_frac + 0.38:
            a_s, b_s = _fmt_amount(a), _fmt_amount(b)
            body = (f"APPROVED PURCHASE DB SALE: {m}, {city},{d2} {hm},card ***{c4}. "
                    f"Amount:{a_s} {ccy}, Balance:{b_s} {ccy}")
            kind = "sale"
            resp = dict(txn_type="debit", date=f"{d2} {hm}", amount=a_s, currency=ccy, card=c4,
                        merchant=m, city=city, address="", balance=b_s)
        elif u < otp_frac + 0.57:
            a_s, b_s = _fmt_amount(a), _fmt_amount(b)
            body = (f"DEBIT ACCOUNT\n{a_s} AMD\n{card_hi[i]:04d}***{c4},\n{m}, AM\n"
                    f"{d4} {hm}\nBALANCE: {b_s} AMD")
            kind = "debit_account"
            resp = dict(txn_type="debit", date=f"{d4} {hm}", amount=a_s, currency="AMD",
                        card=f"{card_hi[i]:04d}***{c4}", merchant=m, city="AM", address="null",
                        balance=b_s)
        elif u < otp_frac + 0.76:
            a_s, b_s = _fmt_amount(a), _fmt_amount(b)
            body = (f"APPROVED PURCHASE DB INTERNET: {m}, {city}, {d2} {hm},card ***{c4}. "
                    f"Amount:{a_s} {ccy}, Balance:{b_s} {ccy}")
            kind = "internet"
            resp = dict(txn_type="debit", date=f"{d2} {hm}", amount=a_s, currency=ccy.lower(), card=c4,
                        merchant=m, city=city, address=None, balance=b_s)
        else:
            a_s, b_s = _fmt_amount(a), _fmt_amount(b)
            body = (f"ACCOUNT REPLENISHMENT: {d2} {hm}, card ***{c4}. "
                    f"Amount:{a_s} {ccy}, Balance:{b_s} {ccy}")
            kind = "credit"
            resp = dict(txn_type="credit", date=f"{d2} {hm}", amount=a_s, currency=ccy, card=f"***{c4}",
                        merchant="ACCOUNT REPLENISHMENT" if (i % 7) else None, city=None, address=None,
                        balance=b_s)
        bodies.append(body)
        kinds.append(kind)
        if resp is not None and miss_u[i] >= miss_frac:
            cache[reference_cache_key(body)] = resp

    if n_dup:
        if hot_key_zipf:
            src = np.minimum(rng.zipf(hot_key_zipf, n_dup) - 1, n_distinct - 1)
        else:
            src = rng.integers(0, n_distinct, n_dup)
        order = rng.permutation(n)
        all_src = np.concatenate([np.arange(n_distinct), src])[order]
        bodies = [bodies[j] for j in all_src]
        kinds = [kinds[j] for j in all_src]
    senders = ["BANK" if (i % 5) else "AMERIABANK" for i in range(n)]
    date_ms = [1_707_900_000_000 + 1000 * i for i in range(n)]
    return Backup(bodies=bodies, senders=senders, date_ms=date_ms, cache=cache, kinds=kinds)


def xml_escape_attr(s: str) -> str:
    """Attribute escaping an SMS Backup & Restore file uses: the five XML
    specials plus newlines as character references."""
    return (s.replace("&", "&amp;").replace("<", "&lt;").replace(">", "&gt;")
             .replace('"', "&quot;").replace("\n", "&#10;").replace("\t", "&#9;").replace("\r", "&#13;"))


def write_xml(path: str, backup: Backup) -> None:
    """`<smses count=N><sms protocol address date type body read/>…` — the
    shape read_xml.py:32-37 and watcher.py:40-44 expect."""
    with open(path, "w", encoding="utf-8") as f:
        f.write("<?xml version='1.0' encoding='UTF-8' standalone='yes' ?>\n")
        f.write(f'<smses count="{len(backup.bodies)}">\n')
        for body, sender, ms in zip(backup.bodies, backup.senders, backup.date_ms):
            f.write(f'  <sms protocol="0" address="{xml_escape_attr(sender)}" date="{ms}" type="1" '
                    f'body="{xml_escape_attr(body)}" read="1" />\n')
        f.write("</smses>\n")


# --------------------------------------------------------------------------- multi-GPU workload shapes
def _owner_of_sha1(body: str, world: int) -> int:
    """Rank that owns the body's msg_id in the global dedup: range partition of the SHA-1's top 32 bits
    (same formula as key_owner() in the kernels and dist.owner_of_prefix)."""
    p = int.from_bytes(hashlib.sha1(body.encode("utf-8")).digest()[:4], "big")
    return (p * world) >> 32


def shard_with_cross_duplicates(n: int, rank: int, world: int, *, dup_frac: float = 0.10, seed: int = 0x5EED,
                                pool_size: Optional[int] = None) -> Backup:
    """Rank `rank`'s shard of BASELINE.json configs[2]: n messages, `dup_frac` of them copies of bodies that
    live in a pool COMMON to all ranks (same seed everywhere), placed at random positions - so about
    dup_frac of the whole job are duplicates spread across shards.  The rest is the rank's own stream."""
    n_dup = int(round(n * dup_frac))
    own = generate(n - n_dup, seed=seed + 7919 * rank)
    pool = generate(pool_size or max(n_dup, 1000), seed=seed + 104729)
    rng = np.random.Generator(np.random.PCG64(seed + 31 * rank + 1))
    pick = rng.integers(0, len(pool.bodies), n_dup)
    pos = rng.permutation(n)
    bodies = [None] * n
    kinds = [None] * n
    for j, p in enumerate(pos[:n - n_dup]):
        bodies[p] = own.bodies[j]
        kinds[p] = own.kinds[j]
    for j, p in enumerate(pos[n - n_dup:]):
        bodies[p] = pool.bodies[pick[j]]
        kinds[p] = pool.kinds[pick[j]]
    cache = dict(own.cache)
    cache.update(pool.cache)
    senders = ["BANK" if (i % 5) else "AMERIABANK" for i in range(n)]
    date_ms = [1_707_900_000_000 + 1000 * (rank * n + i) for i in range(n)]
    return Backup(bodies=bodies, senders=senders, date_ms=date_ms, cache=cache, kinds=kinds)


def shard_skewed(n: int, rank: int, world: int, *, dup_frac: float = 0.5, zipf: float = 1.2,
                 hot_owner_frac: float = 0.5, seed: int = 0x5EED) -> Backup:
    """Rank `rank`'s shard of BASELINE.json configs[3]: `dup_frac` of the messages are
    duplicates drawn Zipf(`zipf`) from a distinct set COMMON to all ranks (hot keys repeat on every rank),
    and `hot_owner_frac` of the rank's own distinct bodies are re-drawn until their msg_id prefix belongs to
    owner rank 0, so rank 0's inbox receives several times what the others do."""
    n_dup = int(round(n * dup_frac))
    n_own = n - n_dup
    own = generate(n_own, seed=seed + 7919 * rank + 3)
    common = generate(max(n_dup // 4, 1000), seed=seed + 15485863)
    rng = np.random.Generator(np.random.PCG64(seed + 57 * rank + 2))
    cache = dict(common.cache)
    bodies_own = list(own.bodies)
    n_hot = int(round(n_own * hot_owner_frac)) if world > 1 else 0
    for j in range(n_hot):
        body = bodies_own[j]
        resp = own.cache.get(reference_cache_key(body))
        k = 0
        cand = body
        while _owner_of_sha1(cand, world) != 0:
            k += 1
            cand = f"{body} ref{k}"
        bodies_own[j] = cand
        if resp is not None:
            cache[reference_cache_key(cand)] = resp
    for j in range(n_hot, n_own):
        key = reference_cache_key(bodies_own[j])
        if key in own.cache:
            cache[key] = own.cache[key]
    src = np.minimum(rng.zipf(zipf, n_dup) - 1, len(common.bodies) - 1)
    pos = rng.permutation(n)
    bodies = [None] * n
    for j, p in enumerate(pos[:n_own]):
        bodies[p] = bodies_own[j]
    for j, p in enumerate(pos[n_own:]):
        bodies[p] = common.bodies[src[j]]
    senders = ["BANK" if (i % 5) else "AMERIABANK" for i in range(n)]
    date_ms = [1_707_900_000_000 + 1000 * (rank * n + i) for i in range(n)]
    return Backup(bodies=bodies, senders=senders, date_ms=date_ms, cache=cache, kinds=[])
