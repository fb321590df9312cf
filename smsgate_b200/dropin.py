"""Per-message drop-ins with the reference's exact names and signatures.

The batch API (`ReplayEngine.replay`, `legacy.parse_bodies`, `ReplayEngine.decimal_batch`) is how the GPU is
used well; these shims are for call sites of the reference that handle ONE message and should not have to
change: same name, same arguments, same return value, same exceptions.  Each call is a one-message batch
on the module-level engine (`use_engine` / `get_engine`), so it costs a kernel launch and a round trip -
correct, not fast.

    libs.gemini_parser.parse_sms_llm(raw)              libs/gemini_parser.py:193
    libs.decimal_utils.parse_ambiguous_decimal(s)      libs/decimal_utils.py:4
    services.xml_watcher.watcher._iter_sms(xml_path)   services/xml_watcher/watcher.py:35
    process_cached.parse_transaction_message(message)  process_cached.py:150
    process_cached.parse_credit_message(message)       process_cached.py:180
    process_cached.process_sms_from_cache(src, p, c)   process_cached.py:210
    read_xml.parse_and_cache_sms(xml_file, cache_dir)  read_xml.py:9
"""
from __future__ import annotations

import datetime as _dt
from decimal import Decimal
from pathlib import Path
from typing import Any, Dict, Iterable, Optional, Union

from . import _lib as L
from .models import BrokenMessage, ParsedSMS, RawSMS  # noqa: F401  (re-exported: the reference's names)

_engine = None


def use_engine(engine) -> None:
    """Make `engine` (a ReplayEngine with the response cache loaded) the one these functions run on."""
    global _engine
    _engine = engine


def get_engine():
    """The module-level engine; created on GPU 0 at first use.  There is no CPU path: without a B200 this
    raises."""
    global _engine
    if _engine is None:
        from .replay import ReplayEngine
        _engine = ReplayEngine(0)
    return _engine


def load_cache(cache) -> None:
    """`cache = diskcache.Cache(".gemini_cache")` of libs/gemini_parser.py:33: a directory, a diskcache.Cache
    or any mapping hex key -> response dict."""
    get_engine().load_cache(cache)


# ------------------------------------------------------------------ libs/gemini_parser.py
def parse_sms_llm(raw: RawSMS) -> Optional[ParsedSMS]:
    """libs/gemini_parser.py:193-271 for one message against the loaded response cache: the parser's OTP
    filter, clean + card masking, cache key, cache join, date / card / amount canonicalisation on the GPU.
    Returns the ParsedSMS or None exactly where the reference does; raises BrokenMessage for a card shorter
    than four characters, and `replay.CacheMiss` where the reference would call Gemini (no network here)."""
    from .dlq import replay_raws
    res = replay_raws(get_engine(), [raw], _dt.datetime.max, worker_filter=False)
    return res.parsed(0, sender=raw.sender, device_id=raw.device_id, msg_id=raw.msg_id)


# ------------------------------------------------------------------ libs/decimal_utils.py
def parse_ambiguous_decimal(num_str: Any) -> Decimal:
    """libs/decimal_utils.py:4-63.  A non-str goes straight to Decimal() as in the reference (:10-12); a
    str is canonicalised on the GPU (smsgate_decimal_batch) into sign / coefficient / exponent, from which
    the exact Decimal is built.  ValueError where the reference raises it."""
    if not isinstance(num_str, str):
        return Decimal(num_str)
    coef, exp, flags = get_engine().decimal_batch([num_str])
    err = int(flags[0]) & 3
    if err == L.DEC_SYNTAX:
        raise ValueError(f"Не удалось преобразовать '{num_str}' в Decimal")
    if err == L.DEC_OVERFLOW:
        raise OverflowError(f"{num_str!r}: more than 18 significant digits (representation limit of the GPU path)")
    return Decimal(((int(flags[0]) >> 4) & 1, tuple(int(c) for c in str(int(coef[0]))), int(exp[0])))


# ------------------------------------------------------------------ services/xml_watcher/watcher.py
def _iter_sms(xml_path: Union[str, Path]) -> Iterable[RawSMS]:
    """services/xml_watcher/watcher.py:35-54: the RawSMS of every top-level <sms>, tokenized and keyed
    (SHA-1) on the GPU.  Raises what the reference raises (ET.ParseError, ValueError, ValidationError)."""
    from .xml_ingest import iter_sms
    return iter_sms(xml_path, get_engine())


# ------------------------------------------------------------------ process_cached.py / read_xml.py
def parse_transaction_message(message: str) -> Optional[Dict[str, Any]]:
    """process_cached.parse_transaction_message (process_cached.py:150-177): TRANSACTION_RE applied to the
    message as given."""
    from .legacy import parse_one
    return parse_one(get_engine(), message, credit=False)


def parse_credit_message(message: str) -> Optional[Dict[str, Any]]:
    """process_cached.parse_credit_message (process_cached.py:180-204): CREDIT_PAYMENT_RE applied to the
    message as given."""
    from .legacy import parse_one
    return parse_one(get_engine(), message, credit=True)


def process_sms_from_cache(source_dir: str, purchase_dest_dir: str, credit_dest_dir: str) -> None:
    """process_cached.process_sms_from_cache (process_cached.py:210-268)."""
    from .legacy import process_sms_from_cache as run
    run(get_engine(), source_dir, purchase_dest_dir, credit_dest_dir)


def parse_and_cache_sms(xml_file: str, cache_dir: str) -> None:
    """read_xml.parse_and_cache_sms (read_xml.py:9-54)."""
    from .xml_ingest import parse_and_cache_sms as run
    run(xml_file, cache_dir, get_engine())
