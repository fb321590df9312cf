"""The reference's domain types, field for field (libs/models.py:35-109).

Inside a checkout of the reference, `libs.models` is importable and these
names resolve to the reference's own classes, so the drop-in hands back the
very objects the rest of the pipeline expects.  Elsewhere the mirrors below
are used; they validate and serialise identically (pydantic v2)."""
from __future__ import annotations

import datetime as _dt
import hashlib
from decimal import Decimal
from enum import Enum
from typing import Literal, Optional

try:  # the real thing, when running inside the reference tree
    from libs.models import ParsedSMS, RawSMS, TxnType, get_md5_hash  # type: ignore  # noqa: F401
except Exception:  # pragma: no cover - exercised wherever the reference is absent
    from pydantic import BaseModel, Field, field_validator

    class TxnType(str, Enum):                      # libs/models.py:35-41
        DEBIT = "debit"
        CREDIT = "credit"
        OTP = "otp"
        UNKNOWN = "unknown"

    class RawSMS(BaseModel):                       # libs/models.py:44-57
        msg_id: str
        sender: str = Field(..., min_length=1)
        body: str = Field(..., min_length=1)
        date: str
        device_id: Optional[str] = None
        source: Literal["device", "xml"] = "device"

    class ParsedSMS(BaseModel):                    # libs/models.py:60-95
        msg_id: str
        device_id: Optional[str]
        sender: str
        date: _dt.datetime
        raw_body: str
        txn_type: TxnType
        amount: Optional[Decimal] = None
        currency: Optional[str] = None
        card: Optional[str] = Field(None, min_length=4, max_length=4)
        merchant: Optional[str] = None
        city: Optional[str] = None
        address: Optional[str] = None
        balance: Optional[Decimal] = None
        parser_version: str = "0.1.0"

        @field_validator("currency")
        def _currency_upper(cls, v):               # noqa: N805  libs/models.py:93-95
            return v.upper() if v else v

    def get_md5_hash(input_string: str) -> str:    # libs/models.py:97-109
        return hashlib.md5(input_string.encode("utf-8")).hexdigest()


class BrokenMessage(Exception):
    """libs/gemini_parser.py:20 — raised when the cached card has fewer than four characters."""


__all__ = ["RawSMS", "ParsedSMS", "TxnType", "BrokenMessage", "get_md5_hash"]
