"""XML backup ingestion: the host side of `_iter_sms` (services/xml_watcher/watcher.py:35-54)
and of `parse_and_cache_sms` (read_xml.py:9-54).

The file's bytes go to the GPU tokenizer (smsgate_xml_scan / _emit), which hands back
body / address / date of every top-level <sms> already decoded - entity and character references, attribute
white-space normalisation - and checked against ElementTree on every document the tests and the fuzz runs
throw at it.  Two kinds of file still go through ElementTree, and both are counted in `engine.host_fallbacks`:
a file in an XML dialect the tokenizer leaves to a general parser (DOCTYPE, CDATA, namespaces, other
encodings), and a file with a `date` attribute the reference raises on, so that the exception and the
messages yielded before it are exactly the reference's.
"""
from __future__ import annotations

import datetime as _dt
import xml.etree.ElementTree as ET
from pathlib import Path
from typing import Iterable, Optional, Union

import numpy as np

from .batch import Batch
from .models import RawSMS


def read_backup(xml_path: Union[str, Path]) -> Batch:
    """Top-level <sms> elements of a backup as one ragged batch, in document order.
    Raises what the reference raises: ET.ParseError for malformed XML, ValueError
    for a non-integer `date` (watcher.py:37-44)."""
    root = ET.parse(xml_path).getroot()                       # watcher.py:37-38
    bodies, senders, dates = [], [], []
    for elem in root.findall("sms"):                          # :40  (nested <sms> under <mms> are not direct children)
        bodies.append(elem.get("body", ""))                   # :41
        dates.append(int(elem.get("date", "0")))              # :42
        senders.append(elem.get("address", ""))               # :44
    return Batch.from_strings(bodies, senders, dates)


def read_backup_gpu(xml_path: Union[str, Path], engine) -> Batch:
    """The same batch as read_backup, tokenized on the GPU (smsgate_xml_scan / _emit): the file
    bytes go to the device, body / address / date come back already decoded.  Raises the
    reference's exceptions (ET.ParseError, ValueError for a non-integer date); a file that
    uses XML outside the GPU dialect is parsed by read_backup instead."""
    from .replay import UnsupportedXML
    from . import _lib as L
    data = np.fromfile(str(xml_path), dtype=np.uint8)
    try:
        d = engine.xml_extract(data)
    except UnsupportedXML:
        engine.host_fallbacks += 1               # a dialect the tokenizer does not decide: ElementTree, and it is counted
        return read_backup(xml_path)
    n = d["n"]
    off = d["body_offsets"].cpu().numpy()
    soff = d["sender_offsets"].cpu().numpy()
    body = d["body_bytes"].cpu().numpy()[:int(off[-1])] if n else np.zeros(0, np.uint8)
    sb = d["sender_bytes"].cpu().numpy()[:int(soff[-1])] if n else np.zeros(0, np.uint8)
    mflags = d["mflags"].cpu().numpy()[:n]
    date_ms = d["date_ms"].cpu().numpy()[:n]
    bad = np.nonzero(mflags & (L.XML_DATE_BAD | L.XML_DATE_BIG))[0]
    if len(bad):
        if mflags[bad[0]] & L.XML_DATE_BAD:
            raise ValueError(f"invalid literal for int() in the date attribute of <sms> #{int(bad[0])}")
        raise OverflowError(f"date attribute of <sms> #{int(bad[0])} is out of range")
    senders = [sb[int(soff[i]):int(soff[i + 1])].tobytes().decode("utf-8") for i in range(n)]
    return Batch(body, off.astype(np.int64), senders, date_ms.astype(np.int64))


def iso_utc(date_ms: int) -> str:
    """RawSMS.date of an XML message (watcher.py:43, :52)."""
    return _dt.datetime.fromtimestamp(date_ms / 1_000, tz=_dt.timezone.utc).isoformat()


def iter_sms(xml_path: Union[str, Path], engine) -> Iterable[RawSMS]:
    """Drop-in for `_iter_sms`: same RawSMS objects in the same order, `msg_id` =
    SHA-1(body) computed on the GPU for the whole file at once.  A message that
    fails RawSMS validation (empty body or sender) raises at that message, after
    the earlier ones were yielded, like the reference's generator."""
    try:
        batch = read_backup_gpu(xml_path, engine)
    except (ValueError, OverflowError):
        # a `date` attribute int() refuses, or one beyond int64: the reference's generator has yielded the messages
        # in front of it by the time it raises (watcher.py:42-43 run per element)
        engine.host_fallbacks += 1
        yield from _iter_sms_until_bad_date(xml_path, engine)
        return
    if batch.n == 0:
        return
    ids = engine.msg_ids(batch)
    for i in range(batch.n):
        yield RawSMS(source="xml", device_id="xml_backup", msg_id=ids[i].tobytes().hex(), sender=batch.senders[i],
                     date=iso_utc(int(batch.date_ms[i])), body=batch.body(i))


def _iter_sms_until_bad_date(xml_path: Union[str, Path], engine) -> Iterable[RawSMS]:
    """`_iter_sms` for a file whose dates make it raise: ElementTree element by element as the reference does
    (watcher.py:40-44), the messages in front of the failing one yielded first (msg_ids still from the GPU), then
    the reference's own exception from `int()` / `datetime.fromtimestamp()`."""
    root = ET.parse(xml_path).getroot()
    bodies, senders, dates, failure = [], [], [], None
    for elem in root.findall("sms"):
        try:
            ts_ms = int(elem.get("date", "0"))                                    # :42
            date = _dt.datetime.fromtimestamp(ts_ms / 1_000, tz=_dt.timezone.utc).isoformat()   # :43, :52
        except Exception as e:                                # noqa: BLE001 - re-raised below, after the yields
            failure = e
            break
        bodies.append(elem.get("body", ""))
        senders.append(elem.get("address", ""))
        dates.append(date)
    if bodies:
        ids = engine.msg_ids(Batch.from_strings(bodies, senders, [0] * len(bodies)))
        for i, body in enumerate(bodies):
            yield RawSMS(source="xml", device_id="xml_backup", msg_id=ids[i].tobytes().hex(), sender=senders[i],
                         date=dates[i], body=body)
    if failure is not None:
        raise failure


def parse_and_cache_sms(xml_file: str, cache_dir: str, engine=None) -> Optional[int]:
    """read_xml.parse_and_cache_sms (read_xml.py:9-54): every top-level <sms> goes into the DiskCache under
    its `date` attribute with its WHOLE attribute dict as the value (:34-41; later duplicates overwrite);
    elements without a `date` (or with an empty one) are skipped (:43-44); a missing or malformed file is
    reported, not raised (:48-51).  The file is tokenized and its attributes decoded on the GPU
    (ReplayEngine.xml_attribs); the DiskCache writes are host work - storage.  A file in an XML dialect the
    tokenizer leaves to a general parser goes through ElementTree and is counted in engine.host_fallbacks.
    Returns the number of stored messages (the reference only prints it), None when nothing was parsed."""
    from diskcache import Cache
    from .replay import UnsupportedXML
    cache = Cache(cache_dir, fanout_cache=True)               # :19 (the kwarg is stored as a setting, as in the reference)
    stored = None
    try:
        if engine is None:
            from .dropin import get_engine
            engine = get_engine()
        with open(xml_file, "rb") as f:                        # FileNotFoundError as ET.parse raises it (:25)
            data = f.read()
        try:
            attribs = engine.xml_attribs(np.frombuffer(data, dtype=np.uint8))
        except UnsupportedXML:
            engine.host_fallbacks += 1
            attribs = [elem.attrib for elem in ET.fromstring(data).findall("sms")]
        stored = 0
        for data_dict in attribs:                              # :32
            key = data_dict.get("date")                        # :37
            if key:
                cache.set(key, data_dict)                      # :41
                stored += 1
    except FileNotFoundError:
        print(f"XML file not found: {xml_file}")
    except ET.ParseError:
        print("could not parse the XML file")
    finally:
        cache.close()
    return stored
