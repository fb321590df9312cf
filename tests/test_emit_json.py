"""smsgate_emit_json (host C++: payload bytes assembled from the replay's columns and the cache texts) against
what pydantic produces for the same records: `ParsedSMS.model_dump_json()` (libs/nats_utils.py:127,
worker.py:183) byte for byte, and the PocketBase record dict (libs/pocketbase.py:296-309) as JSON.
Runs without a GPU: the columns come from the device code compiled for the CPU (tests/hostcheck)."""
import datetime as dt
import json

import numpy as np

from smsgate_b200 import _lib as L
from smsgate_b200 import gemini_cache, synth
from smsgate_b200.batch import Batch
from smsgate_b200.replay import ReplayResult
from tests import _hostcheck as H

NOW = dt.datetime(2026, 1, 1)


def _result(bodies, senders, cache):
    sc = H.scan(bodies)
    table = gemini_cache.build_table(cache, H.strict_date)
    jr = H.join_route(sc, table, gemini_cache.pack_datetime(NOW))
    cols = dict(sc)
    cols.pop("offsets")
    cols.update(jr)
    batch = Batch.from_strings(bodies, senders, range(len(bodies)))
    return ReplayResult(batch, cols, table, 0), table


import pytest


@pytest.mark.parametrize("threads", ["1", "5"])
def test_payload_bytes_equal_pydantic(monkeypatch, threads):
    monkeypatch.setenv("SMSGATE_EMIT_THREADS", threads)      # one thread / rows split over five (large batches get that by default)
    backup = synth.generate(4000, seed=91, mixed_locale=True, miss_frac=0.1)
    cache = dict(backup.cache)
    keys = list(cache)
    # strings that need JSON escaping, non-ASCII text, odd amounts, a non-ASCII currency (left to pydantic)
    cache[keys[0]] = dict(cache[keys[0]], merchant='A "quoted" \\\\ back\\slash', city="line\\nbreak\\ttab\\x01\\x1f", address="null")
    cache[keys[1]] = dict(cache[keys[1]], merchant="Ереван ☕ 中文", city="", address="ул. Абовяна, 12")
    cache[keys[2]] = dict(cache[keys[2]], amount="0.0000001", balance="-12,5")
    cache[keys[3]] = dict(cache[keys[3]], amount="0", balance="000123.4500")
    cache[keys[4]] = dict(cache[keys[4]], currency="руб")
    cache[keys[5]] = dict(cache[keys[5]], currency=None, merchant="x")
    cache[keys[6]] = dict(cache[keys[6]], currency="", city=None)
    cache[keys[7]] = dict(cache[keys[7]], date="2024-02-14 09:00:00.250000")
    cache[keys[0]]["city"] = "line\\nbreak\\ttab\\x01\\x1f".encode().decode("unicode_escape")
    cache[keys[0]]["merchant"] = 'A "quoted" \\ back\\slash'
    senders = [s if i % 9 else 'S"e\\nder ß' for i, s in enumerate(backup.senders)]
    bodies = list(backup.bodies)
    res, table = _result(bodies, senders, cache)
    idx = np.nonzero(np.isin(res.cols["status"], (L.ST_UPSERT, L.ST_PARSED_NO_MERCHANT)))[0]
    assert len(idx) > 3000
    off, data, fb = res.emit_json(idx, 0)
    raw = data.tobytes()
    n_fb = 0
    for k, i in enumerate(idx):
        want = res.parsed(int(i)).model_dump_json().encode("utf-8")
        if fb[k]:
            n_fb += 1
            assert off[k] == off[k + 1]
            continue
        assert raw[off[k]:off[k + 1]] == want, (int(i), raw[off[k]:off[k + 1]], want)
    assert 1 <= n_fb <= 3                                  # the non-ASCII currency (str.upper() is Python's)
    # the PocketBase record of the same messages
    off, data, fb = res.emit_json(idx, 1)
    raw = data.tobytes()
    for k, i in enumerate(idx):
        if fb[k]:
            continue
        p = res.parsed(int(i))
        want = {"msg_id": p.msg_id, "original_body": p.raw_body, "sender": p.sender, "datetime": p.date.isoformat(),
                "card": p.card, "amount": str(p.amount), "currency": p.currency,
                "balance": str(p.balance) if p.balance is not None else None, "merchant": p.merchant,
                "address": p.address, "city": p.city, "txn_type": p.txn_type.value}
        got = json.loads(raw[off[k]:off[k + 1]])
        assert got == want and list(got) == list(want), int(i)


def test_survivor_payloads_use_the_native_emitter():
    backup = synth.generate(1500, seed=92, dup_frac=0.3)
    res, table = _result(backup.bodies, backup.senders, backup.cache)
    from oracle import replay_oracle as O
    surv, first, last, _ = O.dedup([bytes(r) for r in res.cols["msg_id"]], (int(s) == 0 for s in res.cols["status"]))
    res.cols["survivor"] = np.array(surv, np.uint8)
    got = res.parsed_payloads()
    want = [res.parsed(int(i)).model_dump_json().encode() for i in res.survivor_indices()]
    assert got == want and len(got) > 500


_FUZZ_ALPHABET = ['a', 'Z', '0', ' ', '"', '\\', '/', '\n', '\t', '\r', '\x00', '\x01', '\x1f', '\x7f', '\x80', '\xa0',
                  'é', 'Я', '中', '☕', '😀', ' ', ' ', '﻿', '퟿', '', '￿', '<', '>', '&',
                  "'", '\x08', '\x0c', '\x0b']


def test_emitter_fuzz_against_pydantic(monkeypatch):
    """Random cache entries - text from an alphabet of everything JSON escapes or pydantic treats specially,
    amounts in every shape parse_ambiguous_decimal accepts, upper-casable and non-ASCII currencies - rendered by
    the native emitter and by pydantic.  The emitter may decline a row (non-ASCII currency: str.upper() is
    Python's; a coefficient beyond int64), it may never differ."""
    import random
    monkeypatch.setenv("SMSGATE_EMIT_THREADS", "3")
    rng = random.Random(0xE417)

    def text(maxlen=12):
        return ''.join(rng.choice(_FUZZ_ALPHABET) for _ in range(rng.randint(0, maxlen)))

    def amount():
        k = rng.random()
        if k < 0.3:
            return f"{rng.randint(0, 10 ** rng.randint(1, 12))}.{rng.randint(0, 99):02d}"
        if k < 0.4:
            return f"{rng.randint(0, 999)},{rng.randint(0, 999):03d}.{rng.randint(0, 9)}"
        if k < 0.5:
            return f"{rng.randint(0, 999)}.{rng.randint(0, 999):03d},{rng.randint(0, 99):02d}"
        if k < 0.6:
            return rng.choice(["0", "0.0", "-0", "0.00", ".5", "5.", "+5", "00012.50", "1e5", "", "   ", "000",
                               "0.0000000", "1.0000000000", "999999999999999999", "9999999999999999999",
                               "0.000000000000000001", "-0.00", "1 234,5", "1\xa0234,56"])
        if k < 0.75:
            return "0." + "0" * rng.randint(0, 10) + str(rng.randint(1, 999))
        if k < 0.85:
            return str(rng.randint(0, 10 ** 17)) + "0" * rng.randint(0, 3)
        return f"{rng.randint(0, 99999)}.{rng.randint(0, 9999999):07d}"

    backup = synth.generate(2500, seed=93)
    cache = dict(backup.cache)
    for key in list(cache):
        if rng.random() < 0.8:
            cache[key] = dict(cache[key], merchant=text() or "m", city=rng.choice([None, text(), ""]),
                              address=rng.choice([None, text(20), "null", ""]), amount=amount(),
                              balance=rng.choice([amount(), "-" + amount(), None]),
                              currency=rng.choice(["amd", "USD", "eUr", "", None, "ru", "Ab1", "ß", "ı", "a b"]),
                              txn_type=rng.choice(["debit", "credit", "otp", "unknown"]))
    senders = [text(8) for _ in backup.senders]
    res, table = _result(list(backup.bodies), senders, cache)
    idx = np.nonzero(np.isin(res.cols["status"], (L.ST_UPSERT, L.ST_PARSED_NO_MERCHANT)))[0]
    assert len(idx) > 1200
    rendered = declined_for_a_reason = 0
    for mode in (0, 1):
        off, data, fb = res.emit_json(idx, mode)
        raw = data.tobytes()
        for k, i in enumerate(idx):
            if fb[k]:
                cur = res.table.values[int(res.cols["slot"][i])].get("currency")
                over = (int(res.cols["dec_flags"][i]) & 3) == L.DEC_OVERFLOW or ((int(res.cols["dec_flags"][i]) >> 2) & 3) == L.DEC_OVERFLOW
                assert over or (cur is not None and not cur.isascii()), (int(i), cur)
                declined_for_a_reason += 1
                continue
            p = res.parsed(int(i))
            got = raw[off[k]:off[k + 1]]
            if mode == 0:
                assert got == p.model_dump_json().encode("utf-8"), (int(i), got)
            else:
                want = {"msg_id": p.msg_id, "original_body": p.raw_body, "sender": p.sender,
                        "datetime": p.date.isoformat(), "card": p.card, "amount": str(p.amount),
                        "currency": p.currency, "balance": str(p.balance) if p.balance is not None else None,
                        "merchant": p.merchant, "address": p.address, "city": p.city, "txn_type": p.txn_type.value}
                assert json.loads(got) == want, (int(i), got)
            rendered += 1
    assert rendered > 1500 and declined_for_a_reason > 50


def test_packed_records_give_the_same_bytes(monkeypatch):
    """The emitter reading the 64-byte records of smsgate_replay_host_records8 (what the end-to-end call returns)
    against the emitter reading separate columns."""
    from smsgate_b200.replay import _RecordColumns
    backup = synth.generate(3000, seed=94, mixed_locale=True, dup_frac=0.2)
    res, table = _result(list(backup.bodies), list(backup.senders), backup.cache)
    dtype = np.dtype(L.RECORD_FIELDS)
    assert dtype.itemsize == L.RECORD_BYTES_OUT
    records = np.zeros(len(backup.bodies), dtype=dtype)
    for f in dtype.names:
        if f in res.cols:
            records[f] = res.cols[f]
    packed = _RecordColumns(records, np.zeros(len(backup.bodies), np.uint8), engine=None)
    packed._fetched.update({k: res.cols[k] for k in _RecordColumns._ON_DEVICE})       # as if fetched from the device
    res_packed = ReplayResult(res.batch, packed, table, 0)
    idx = np.nonzero(np.isin(res.cols["status"], (L.ST_UPSERT, L.ST_PARSED_NO_MERCHANT)))[0]
    assert len(idx) > 2000
    for threads in ("1", "4"):
        monkeypatch.setenv("SMSGATE_EMIT_THREADS", threads)
        for mode in (0, 1):
            a, b = res.emit_json(idx, mode), res_packed.emit_json(idx, mode)
            assert all(np.array_equal(x, y) for x, y in zip(a, b)), (threads, mode)
