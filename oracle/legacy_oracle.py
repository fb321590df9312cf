"""TEST INFRASTRUCTURE — CPU restatement of the reference's legacy offline replay
(read_xml.py -> process_cached.py).  Same rules as replay_oracle.py: a checker
for tests / smoke / the CPU bench legs only, never imported by the product.

Pinned by tests/golden/legacy.json, which oracle/gen_golden.py produces by running
the reference's own `parse_transaction_message`, `parse_credit_message` and
`process_sms_from_cache` (process_cached.py:150-268) unmodified.
"""
from __future__ import annotations

import hashlib
import re
import struct
from decimal import Decimal, InvalidOperation
from typing import Any, Dict, Optional

# process_cached.py:98-120, the debit pattern (VERBOSE there, written out here)
DEBIT_RE = re.compile(
    r"^.*?(?:PURCHASE\s+DB\s+INTERNET|PURCH\.COMPLETION\.DB\s+INTERNET|PURCHASE|SALE):\s*"
    r"(?P<merchant>.+?),\s*(?:(?P<city>YEREVAN),\s*)?"
    r"(?P<address>.*?)(?=,\s*\d{2}[./-]\d{2}[./-]\d{2,4}\s),\s*(?P<date>\d{2}[./-]\d{2}[./-]\d{2,4})\s+"
    r"(?P<time>\d{2}:\d{2}),\s*card\s+\*{3}(?P<card>\d{4})\.\s*Amount:\s*(?P<amount>[\d,.]+)\s+"
    r"(?P<currency>\w{3}),\s*Balance:\s*(?P<balance>[\d,.]+)",
    re.IGNORECASE | re.DOTALL)

# process_cached.py:122-135, the credit pattern (IGNORECASE only: '.' stops at newlines)
CREDIT_RE = re.compile(
    r"^.*?(?P<type>[\w\s]+?):\s*(?P<date>\d{2}[./-]\d{2}[./-]\d{2,4})\s+(?P<time>\d{2}:\d{2}),\s*"
    r"card\s+\*{3}(?P<card>\d{4})\.\s*Amount:\s*(?P<amount>[\d.,]+)\s+(?P<currency>\w{3}),\s*"
    r"Balance:\s*(?P<balance>[\d.,]+)\s+(?:\w{3})\s*$",
    re.IGNORECASE)


def to_float(value: str) -> float:
    """process_cached.py:141-143 then float(): only commas are dropped."""
    return float(Decimal(value.replace(",", "")))


def parse_debit(message: str) -> Optional[Dict[str, Any]]:
    """process_cached.py:150-177."""
    if not message:
        return None
    m = DEBIT_RE.search(message)
    if not m:
        return None
    g = m.groupdict()
    try:
        return {"direction": "debit", "merchant": g["merchant"].strip(),
                "city": (g.get("city") or "ONLINE").upper(), "address": g["address"].strip() or "N/A",
                "date": g["date"].strip(), "time": g["time"].strip(), "card": "***" + g["card"],
                "amount": to_float(g["amount"]), "currency": g["currency"].strip(),
                "balance": to_float(g["balance"])}
    except (InvalidOperation, TypeError):
        return None


def parse_credit(message: str) -> Optional[Dict[str, Any]]:
    """process_cached.py:180-204."""
    if not message:
        return None
    m = CREDIT_RE.search(message)
    if not m:
        return None
    g = m.groupdict()
    try:
        return {"direction": "credit", "type": g["type"].strip().upper(), "date": g["date"].strip(),
                "time": g["time"].strip(), "card": "***" + g["card"], "amount": to_float(g["amount"]),
                "currency": g["currency"].strip(), "balance": to_float(g["balance"])}
    except (InvalidOperation, TypeError):
        return None


LEG_SKIPPED_OTP, LEG_DEBIT, LEG_CREDIT, LEG_FAILED = 0, 1, 2, 3


def process_body(body: Optional[str]):
    """One source-cache entry through process_cached.py:233-263.
    Returns (status, stripped body, sha256 digest of it, parsed dict | None)."""
    text = (body or "").strip()                                           # :233
    up = text.upper()
    digest = hashlib.sha256(text.encode("utf-8")).digest()                # :254
    if "OTP" in up or "PASS=" in up or "CODE:" in up:                     # :236
        return LEG_SKIPPED_OTP, text, digest, None
    parsed = parse_debit(text)                                            # :243
    if parsed:
        return LEG_DEBIT, text, digest, parsed
    parsed = parse_credit(text)                                           # :248
    if parsed:
        return LEG_CREDIT, text, digest, parsed
    return LEG_FAILED, text, digest, None


def f64_bits(x: float) -> int:
    return struct.unpack("<Q", struct.pack("<d", x))[0]


# ---------------------------------------------------------------------------
# save_to_pocketbase.py:34-78 — the PocketBase record builders of the legacy caches.  (The module cannot be
# imported: its line 17 `from pocketbase import ` is a syntax error in the reference itself.  The golden
# vectors are produced by oracle/gen_golden_pb.py from the text of lines 34-78 executed as they stand.)
# ---------------------------------------------------------------------------
def dt_str(date, time_):                                             # :34-43
    from datetime import datetime
    for fmt in ("%d.%m.%Y", "%d/%m/%Y", "%d-%m-%Y", "%d.%m.%y", "%d/%m/%y", "%d-%m-%y"):
        try:
            dt = datetime.strptime(f"{date} {time_}", f"{fmt} %H:%M")
            return dt.strftime("%Y-%m-%d %H:%M:%S")
        except ValueError:
            continue
    return None


def build_sms_data(record):                                          # :46-62
    pb_dt = dt_str(record.get("date", ""), record.get("time", ""))
    if not pb_dt:
        return None
    return {"merchant": record.get("merchant"), "city": record.get("city"), "address": record.get("address"),
            "datetime": pb_dt, "card": record.get("card"), "amount": str(record.get("amount", 0.0)),
            "currency": record.get("currency"), "balance": str(record.get("balance", 0.0)),
            "msg_id": record.get("msg_id"), "original_body": record.get("original_body")}


def build_transactions(record):                                      # :65-78
    pb_dt = dt_str(record.get("date", ""), record.get("time", ""))
    if not pb_dt:
        return None
    return {"transaction_id": record.get("msg_id"), "transaction_type": record.get("type", record.get("direction")),
            "amount": record.get("amount"), "currency": record.get("currency"), "balance_after": record.get("balance"),
            "timestamp": pb_dt, "status": "parsed"}
