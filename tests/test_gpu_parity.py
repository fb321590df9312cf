"""Parity of the CUDA path, called through the C ABI, against (a) the golden
vectors produced by executing the reference, (b) the CPU oracle on seeded
synthetic backups, (c) size-independent properties at BASELINE.json's 1 M size.
Bit-exact everywhere: this path is integer / byte work."""
import datetime as dt

import numpy as np
import pytest

from oracle import replay_oracle as O
from smsgate_b200 import _lib as L
from smsgate_b200 import gemini_cache, synth
from smsgate_b200.batch import Batch

pytestmark = pytest.mark.gpu
STATUS = {n: i for i, n in enumerate(L.STATUS_NAMES)}
NOW = dt.datetime(2026, 1, 1)


@pytest.fixture(scope="module")
def engine():
    from smsgate_b200.replay import ReplayEngine
    eng = ReplayEngine(0)
    yield eng
    eng.close()


def _run(engine, bodies, cache, now=NOW, senders=None):
    engine.load_cache(cache)
    batch = Batch.from_strings(bodies, senders or ["BANK"] * len(bodies), range(len(bodies)))
    return engine.replay(batch, now)


def test_keys_scans_rewrite_golden(engine, golden):
    cases = golden("keys.json")
    res = _run(engine, [c["body"] for c in cases], {})
    for i, c in enumerate(cases):
        assert res.msg_id_hex(i) == c["msg_id"], c["body"]
        assert res.cache_key_hex(i) == c["cache_key"], c["body"]
        fl = int(res.flags[i])
        assert bool(fl & L.F_WORKER_SKIP) == c["worker_skip"], c["body"]
        assert bool(fl & L.F_PARSER_OTP) == c["parser_otp"], c["body"]
        assert res.fixed_body(i) == c["fixed"], c["body"]
        assert bool(fl & L.F_MODIFIED) == (c["fixed"] != c["body"])
        want = 0 if c["body_date"] is None else O.pack_date(dt.date.fromisoformat(c["body_date"]))
        assert int(res.body_date[i]) == want, c["body"]


def test_decimal_batch_golden(engine, golden):
    rows = golden("decimals.json")
    coef, exp, flags = engine.decimal_batch([r[0] for r in rows])
    for i, r in enumerate(rows):
        err = int(flags[i]) & 3
        if r[1] == "ValueError":
            assert err == L.DEC_SYNTAX, r[0]
        elif int(r[3]) > O.INT64_MAX:
            assert err == L.DEC_OVERFLOW, r[0]
        else:
            assert (err, (int(flags[i]) >> 4) & 1, int(coef[i]), int(exp[i])) == (0, r[2], int(r[3]), r[4]), r[0]


def test_parse_matrix_golden(engine, golden):
    for c in golden("parse_matrix.json"):
        if c["raw_date"].lstrip("-").isdigit():
            continue    # device-source raw.date; XML replay always carries an ISO string (watcher.py:52)
        key = O.cache_key(O.fixed_body(c["body"])).hex()
        res = _run(engine, [c["body"]], {key: c["resp"]})
        assert int(res.status[0]) == STATUS[c["status"]], c["name"]
        p = c["parsed"]
        if p is None:
            continue
        if c["status"] in ("FAILED_EXC", "ESCAPED"):
            continue
        big = any(int(p[w][1]) > O.INT64_MAX for w in ("amount", "balance"))
        if big:
            with pytest.raises(OverflowError):
                res.parsed(0)
            continue
        got = res.parsed(0, sender="BANK")
        assert got.date.isoformat() == p["date"], c["name"]
        assert got.raw_body == p["raw_body"]
        assert got.txn_type.value == p["txn_type"]
        for w in ("amount", "balance"):
            s, digits, e = getattr(got, w).as_tuple()
            assert [s, "".join(map(str, digits)), e] == p[w], c["name"]
        for f in ("currency", "card", "merchant", "city", "address", "parser_version"):
            assert getattr(got, f) == p[f], (c["name"], f)


@pytest.mark.parametrize("name", ["replay_plain.json", "replay_mixed.json"])
def test_replay_golden(engine, golden, name):
    g = golden(name)
    backup = synth.generate(g["n"], seed=g["seed"], **g["kwargs"])
    now = dt.datetime.fromisoformat(g["now"])
    engine.load_cache(backup.cache)
    res = engine.replay(Batch.from_strings(backup.bodies, backup.senders, backup.date_ms), now)
    for i, r in enumerate(g["rows"]):
        assert res.msg_id_hex(i) == r["msg_id"]
        assert int(res.status[i]) == STATUS[r["status"]], i
        if r["cache_key"] is not None:
            assert res.cache_key_hex(i) == r["cache_key"]
        p = r["parsed"]
        if p is not None:
            got = res.parsed(i)
            assert got.model_dump(mode="json")["date"] == p["date"]
            assert got.raw_body == p["raw_body"] and got.card == p["card"] and got.sender == p["sender"]
            assert str(got.amount) == str(O.Decimal((p["amount"][0], tuple(map(int, p["amount"][1])), p["amount"][2])))
            assert got.currency == p["currency"] and got.merchant == p["merchant"] and got.address == p["address"]
    ups = {res.msg_id_hex(int(i)): int(i) for i in res.survivor_indices()}
    assert ups == g["upserts"]
    first = {res.msg_id_hex(int(i)): int(i) for i in np.nonzero(res.first_seen)[0]}
    assert first == g["first"]
    for i in range(res.n):
        if int(res.status[i]) == L.ST_UPSERT:
            assert int(res.last_idx[i]) == g["upserts"][res.msg_id_hex(i)]
        else:
            assert int(res.last_idx[i]) == -1


def _compare_with_oracle(res, backup, now):
    raw_dates = [O.iso_utc_from_ms(ms) for ms in backup.date_ms]
    ora = O.replay(backup.bodies, backup.senders, raw_dates, backup.cache, now)
    n = len(backup.bodies)
    assert res.n == n
    want_ids = np.frombuffer(b"".join(o.msg_id for o in ora.outcomes), np.uint8).reshape(n, 20)
    want_keys = np.frombuffer(b"".join(o.cache_key for o in ora.outcomes), np.uint8).reshape(n, 32)
    assert np.array_equal(res.msg_id, want_ids)
    assert np.array_equal(res.cache_key, want_keys)
    assert np.array_equal(res.status, np.array([o.status for o in ora.outcomes], np.uint8))
    assert np.array_equal((res.flags & L.F_WORKER_SKIP) != 0, np.array([o.worker_skip for o in ora.outcomes]))
    assert np.array_equal((res.flags & L.F_PARSER_OTP) != 0, np.array([o.parser_otp for o in ora.outcomes]))
    assert np.array_equal(res.body_date, np.array([o.body_date for o in ora.outcomes], np.int32))
    assert np.array_equal(res.slot >= 0, np.array([o.hit for o in ora.outcomes]))
    assert np.array_equal(res.survivor != 0, np.array(ora.survivor))
    assert np.array_equal(res.first_seen != 0, np.array(ora.first_seen))
    assert np.array_equal(res.last_idx, np.array(ora.last_idx, np.int64))
    for i, o in enumerate(ora.outcomes):
        if o.hit:
            fl = int(res.dec_flags[i])
            for w, sh, sg, want in (("amount", 0, 4, o.amount), ("balance", 2, 5, o.balance)):
                assert (fl >> sh) & 3 == want[0], (i, w)
                if want[0] == 0:
                    assert ((fl >> sg) & 1, int(res.cols[w + "_coef"][i]), int(res.cols[w + "_exp"][i])) == want[1:]
            assert (bool(fl & 0x40), int(res.amount_cents[i])) == o.amount_cents, i
            assert (bool(fl & 0x80), int(res.balance_cents[i])) == o.balance_cents, i
        if o.parsed is not None and o.status not in (O.ST_FAILED_EXC, O.ST_ESCAPED):
            assert int(res.date[i]) == O.pack_datetime(o.parsed.date), i
            assert res.fixed_body(i) == o.fixed
            assert res.cols["card"][i, :int(res.card_len[i])].tobytes().decode() == o.parsed.card


@pytest.mark.parametrize("kw", [
    dict(n=20000, seed=11),
    dict(n=20000, seed=12, dup_frac=0.5, hot_key_zipf=1.2),
    dict(n=12000, seed=13, mixed_locale=True, miss_frac=0.3, dup_frac=0.1),
])
def test_synthetic_vs_oracle(engine, kw):
    kw = dict(kw)
    n = kw.pop("n")
    backup = synth.generate(n, **kw)
    engine.load_cache(backup.cache)
    res = engine.replay(Batch.from_strings(backup.bodies, backup.senders, backup.date_ms), NOW)
    _compare_with_oracle(res, backup, NOW)


def test_edges_empty_ragged_giant(engine):
    # empty batch
    engine.load_cache({})
    res = engine.replay(Batch.from_strings([], [], []), NOW)
    assert res.n == 0
    # bodies of every length around the SHA block and staging boundaries, empty bodies, one giant body
    rng = np.random.Generator(np.random.PCG64(5))
    alphabet = np.array(list("ABCDEFGHIJKLMNOPQRSTUVWXYZ 0123456789.:*,"))
    bodies = ["", "a", ""]
    stage = 176 * 1024                                   # K1_STAGE: the scan kernel's staging buffer
    for ln in list(range(0, 200)) + [255, 256, 257, 1000, 4095, 4096, 4097, 44 * 1024, stage - 17, stage - 16,
                                     stage - 1, stage, stage + 1, stage + 16]:
        bodies.append("".join(rng.choice(alphabet, ln)))
    giant = "4083***7538 " + "".join(rng.choice(alphabet, 400_000)) + " 1111\u2022\u2022\u20222222 OTP\u00a0"
    bodies += [giant, "tail after giant", giant[:50_000], "x" * 200_000 + "Daily limit exceeded: 10.06.2025"]
    # 700 x 4 KB bodies: a CTA's share of these (>= 64 messages) overflows the 176 KB stage -> sub-batches,
    # some of them rewritten (card form / bullet) so that the in-place rewrite runs inside a sub-batch too
    for i in range(700):
        b = "".join(rng.choice(alphabet, 4000)) + str(i)
        if i % 5 == 0:
            b = b[:1000] + " 4083***7538 " + b[1000:]
        if i % 7 == 0:
            b = b[:3000] + "\u2022\u00a0" + b[3000:]
        bodies.append(b)
    cache = {O.cache_key(O.fixed_body(b)).hex(): dict(synth_resp) for b, synth_resp in
             ((b, dict(txn_type="debit", date="06.05.25 14:23", amount="1,00", currency="usd", card="1234",
                       merchant="M", city=None, address=None, balance="2")) for b in bodies[::3])}
    res = _run(engine, bodies, cache)
    backup = synth.Backup(bodies=bodies, senders=["BANK"] * len(bodies), date_ms=list(range(len(bodies))), cache=cache)
    _compare_with_oracle(res, backup, NOW)


def test_long_cache_value_strings(engine):
    """Cached amount / balance / date strings far longer than real ones (padding, 90-digit numbers): same
    results as the oracle, also mixed with ordinary entries in one batch."""
    bodies = ["PURCHASE: SHOP %d, CITY, 06.05.25 14:23, card ***%04d. Amount:%d.00 USD, Balance:5.00 USD" % (i, i, i)
              for i in range(600)]
    cache = {}
    for i, b in enumerate(bodies):
        v = dict(txn_type="debit", date="06.05.25 14:23", amount="%d,50" % i, currency="usd", card="1234***%04d" % i,
                 merchant="M", city=None, address=None, balance="2")
        if i % 3 == 0:
            v["amount"] = " " * 70 + "1 234,5"                    # 77 bytes of amount
        if i % 3 == 1:
            v["balance"] = "9" * 19 + "." + "0" * 70              # overflows int64
        if i % 5 == 0:
            v["date"] = "  10.06.2025 20:51" + " " * 60           # host-resolved date, long text
        cache[O.cache_key(O.fixed_body(b)).hex()] = v
    res = _run(engine, bodies, cache)
    backup = synth.Backup(bodies=bodies, senders=["BANK"] * len(bodies), date_ms=list(range(len(bodies))), cache=cache)
    _compare_with_oracle(res, backup, NOW)
    assert (res.status == L.ST_UPSERT).sum() > 100


def test_stage_calls_match_fused_replay(engine):
    """scan_keys / join_route / dedup called one by one on device tensors give
    the same columns as replay_host."""
    import torch
    backup = synth.generate(5000, seed=21, dup_frac=0.2, mixed_locale=True, miss_frac=0.1)
    engine.load_cache(backup.cache)
    batch = Batch.from_strings(backup.bodies, backup.senders, backup.date_ms)
    ref = engine.replay(batch, NOW, base_index=1000)
    d_bytes = torch.from_numpy(batch.body_bytes.copy()).cuda()
    d_off = torch.from_numpy(batch.body_offsets.copy()).cuda()
    cols = engine.device_columns(batch.n, d_bytes.numel())
    engine.reserve(batch.n, d_bytes.numel())
    engine.scan_keys(d_bytes, d_off, cols)
    engine.join_route(batch.n, NOW, cols)
    engine.dedup(batch.n, 1000, cols)
    torch.cuda.synchronize()
    for name in ("msg_id", "cache_key", "flags", "body_date", "fixed_len", "slot", "status", "amount_coef",
                 "balance_coef", "amount_exp", "balance_exp", "dec_flags", "date", "card", "card_len", "survivor",
                 "first_seen", "last_idx", "amount_cents", "balance_cents"):
        assert np.array_equal(cols[name].cpu().numpy(), ref.cols[name]), name
    assert int(ref.last_idx.max()) >= 1000
    # stages 2 + 3 fused, and the whole path in one call: the same columns again
    for mode in ("fused", "whole"):
        cols2 = engine.device_columns(batch.n, d_bytes.numel())
        if mode == "fused":
            engine.scan_keys(d_bytes, d_off, cols2)
            engine.join_route_dedup(batch.n, NOW, 1000, cols2)
        else:
            engine.replay_device(d_bytes, d_off, NOW, cols2, base_index=1000)
        torch.cuda.synchronize()
        for name in ("msg_id", "slot", "status", "amount_coef", "date", "survivor", "first_seen", "last_idx", "amount_cents"):
            assert np.array_equal(cols2[name].cpu().numpy(), ref.cols[name]), (mode, name)


def test_pinned_buffers_match_pageable(engine):
    """Pinned host buffers take the store-to-host kernel, pageable ones the copy engine: same columns,
    for a batch large enough to be split into chunks and with odd chunk boundaries."""
    backup = synth.generate(150_001, seed=31, dup_frac=0.2, mixed_locale=True, miss_frac=0.1)
    engine.load_cache(backup.cache)
    batch = Batch.from_strings(backup.bodies, backup.senders, backup.date_ms)
    a = engine.replay(batch, NOW, base_index=7)
    b = engine.replay(batch, NOW, base_index=7, pinned=True)
    for name in a.cols:
        assert np.array_equal(np.asarray(a.cols[name]), np.asarray(b.cols[name])), name
    for i in (0, 1, 77_777, 150_000):
        assert a.fixed_body(i) == b.fixed_body(i) == O.fixed_body(backup.bodies[i])


def test_packed_records_match_columns(engine):
    """smsgate_replay_host_records returns the values of smsgate_replay_host, record by record."""
    backup = synth.generate(150_001, seed=32, dup_frac=0.3, mixed_locale=True, miss_frac=0.1)
    engine.load_cache(backup.cache)
    batch = Batch.from_strings(backup.bodies, backup.senders, backup.date_ms)
    a = engine.replay(batch, NOW, base_index=5)
    for pinned, compact in ((True, True), (False, True), (True, False), (False, False)):
        b = engine.replay_records(batch, NOW, base_index=5, pinned=pinned, compact_verdict=compact)
        assert b.cols.nbytes == batch.n * (64 + (1 if compact else 8))
        for name in b.cols.keys():
            assert np.array_equal(np.asarray(a.cols[name]), np.asarray(b.cols[name])), name
        assert b.msg_id_hex(150_000) == a.msg_id_hex(150_000)
        assert np.array_equal(b.survivor_indices(), a.survivor_indices())
    empty = engine.replay_records(Batch.from_strings([], [], []), NOW)
    assert empty.n == 0 and len(empty.cols["status"]) == 0


def test_record_result_materialises_like_column_result(engine):
    """A result that came back as packed records hands out the same ParsedSMS objects, payloads and
    exceptions as the column result: the columns the record leaves out are fetched from the device on
    first use, a stale result says so instead of reading another batch's data, detach() keeps it."""
    from smsgate_b200.replay import CacheMiss
    backup = synth.generate(6_000, seed=77, dup_frac=0.3, mixed_locale=True, miss_frac=0.2)
    engine.load_cache(backup.cache)
    batch = Batch.from_strings(backup.bodies, backup.senders, backup.date_ms)
    a = engine.replay(batch, NOW)
    ua, pa, ra = a.upserts(), a.parsed_payloads(), a.pocketbase_records()
    b = engine.replay_records(batch, NOW)
    assert b.upserts() == ua and b.parsed_payloads() == pa and b.pocketbase_records() == ra
    assert any("CARD:" in p.raw_body for p in ua.values())            # masked cards are in the batch
    miss = int(np.nonzero(a.status == L.ST_CACHE_MISS)[0][0])
    with pytest.raises(CacheMiss) as e1:
        a.parsed(miss)
    with pytest.raises(CacheMiss) as e2:
        b.parsed(miss)
    assert str(e1.value) == str(e2.value)
    # same-size batch again: the earlier result's pinned arrays are reused and its device columns are gone
    c = engine.replay_records(batch, NOW).detach()
    d = engine.replay_records(Batch.from_strings(backup.bodies[::-1], backup.senders, backup.date_ms), NOW)
    d2 = engine.replay_records(batch, NOW)
    with pytest.raises(RuntimeError, match="later replay"):
        d.cols["card"]
    assert c.upserts() == ua                                           # the detached one is unaffected
    assert d2.upserts() == ua


def test_native_cache_loader_equals_python_loader(engine, tmp_path):
    """engine.load_cache(<directory>) - SQLite + pickles decoded in C++, dates classified on the GPU - builds
    the same table, array for array, as the per-entry Python path over diskcache's own view of the cache,
    and replays identically; the response dicts behind it are unpickled only on demand."""
    import diskcache
    from tests.test_cache_loader import _extra_entries
    backup = synth.generate(8000, seed=52, mixed_locale=True, miss_frac=0.1)
    c = diskcache.Cache(str(tmp_path))
    with c.transact():
        for k, v in backup.cache.items():
            c[k] = v
    _extra_entries(c)
    odd_dates = ["2024-02-14 09:00:00", "14:23 06.05.2025", "", "yesterday", "2024-02-14T09:00:00+04:00", "06.05.25 14:23"]
    some = [k for k in list(backup.cache)[:len(odd_dates)]]
    for k, d in zip(some, odd_dates):
        c[k] = dict(backup.cache[k], date=d)
    c["f" * 64] = dict(backup.cache[some[0]], merchant="X" * 40_000)          # side file
    via_diskcache = {k: c[k] for k in c if isinstance(k, str) and k in c}
    c.close()
    want = gemini_cache.build_table(via_diskcache, engine.strict_date_batch)
    got = engine.load_cache(str(tmp_path))
    for f in ("keys", "val_offsets", "val_bytes", "eflags", "date_packed"):
        assert np.array_equal(getattr(got, f), getattr(want, f)), f
    assert got.skipped == want.skipped and got.host_dates == want.host_dates
    assert 0 < got.host_decoded < 10 and got.host_dates_resolved < want.m // 4   # dates outside the strict form: the host chain, per entry
    assert all(got.values[e] == want.values[e] for e in range(0, want.m, 97))
    batch = Batch.from_strings(backup.bodies, backup.senders, backup.date_ms)
    a = engine.replay(batch, NOW)
    engine.load_cache(via_diskcache)
    b = engine.replay(batch, NOW)
    for name in ("status", "slot", "amount_coef", "date", "card", "survivor"):
        assert np.array_equal(a.cols[name], b.cols[name]), name
    engine.load_cache(str(tmp_path))
    assert engine.replay(batch, NOW).upserts() == b.upserts()                   # materialised through LazyValues


def test_device_keys_and_cross_source(engine, golden):
    """MD5 device keys against the reference's vectors and the oracle, and the report of bodies that the
    reference stores twice (same body seen from the XML backup and from a device)."""
    import torch
    cases = golden("keys.json")
    assert engine.msg_ids_device([c["body"] for c in cases]) == [c["md5"] for c in cases]
    backup = synth.generate(30_000, seed=41, mixed_locale=True)
    got = engine.msg_ids_device(backup.bodies)
    assert got == [O.msg_id_device(b).hex() for b in backup.bodies]
    # device batch: every third XML body again (in another order) plus bodies the backup does not hold
    rng = np.random.Generator(np.random.PCG64(7))
    again = [int(i) for i in rng.permutation(len(backup.bodies))[:10_000]]
    other = synth.generate(5_000, seed=42).bodies
    dev_bodies = [backup.bodies[i] for i in again] + other
    xml_set = {}
    for i, b in enumerate(backup.bodies):
        xml_set.setdefault(b, i)
    want = np.array([xml_set.get(b, -1) for b in dev_bodies], np.int32)

    def sha1_rows(bodies):
        return torch.from_numpy(np.frombuffer(b"".join(O.msg_id_xml(b) for b in bodies), np.uint8)
                                .reshape(len(bodies), 20).copy()).cuda()
    match = engine.cross_source(sha1_rows(backup.bodies), sha1_rows(dev_bodies)).cpu().numpy()
    assert np.array_equal(match, want)
    assert (match >= 0).sum() >= 10_000
    empty = engine.cross_source(sha1_rows(backup.bodies)[:0], sha1_rows(dev_bodies)).cpu().numpy()
    assert np.all(empty == -1)


def test_scan_fuzz_against_oracle(engine):
    """Adversarial bodies (tests/_fuzz.py) through the scan kernel: keys, filters, rewrite, body date."""
    from tests._fuzz import bodies
    bs = bodies(40_000, 82)
    engine.load_cache({})
    res = engine.replay(Batch.from_strings(bs, ["B"] * len(bs), range(len(bs))), NOW)
    want_ids = np.frombuffer(b"".join(O.msg_id_xml(b) for b in bs), np.uint8).reshape(len(bs), 20)
    assert np.array_equal(res.msg_id, want_ids)
    fixed = [O.fixed_body(b) for b in bs]
    assert np.array_equal(res.cache_key, np.frombuffer(b"".join(O.cache_key(f) for f in fixed), np.uint8).reshape(len(bs), 32))
    assert np.array_equal((res.flags & L.F_WORKER_SKIP) != 0, np.array([O.worker_skip(b) for b in bs]))
    assert np.array_equal((res.flags & L.F_PARSER_OTP) != 0, np.array([O.parser_otp(b) for b in bs]))
    assert np.array_equal((res.flags & L.F_MODIFIED) != 0, np.array([f != b for f, b in zip(fixed, bs)]))
    assert np.array_equal(res.body_date, np.array([O.pack_date(O.body_date(b)) for b in bs], np.int32))
    for i in range(0, len(bs), 7):
        assert res.fixed_body(i) == fixed[i], repr(bs[i])


def test_wire_formats_of_the_survivors(engine):
    """The `sms.parsed` JSON payloads and the PocketBase record dicts of the surviving records equal what
    the oracle's parsed objects serialise to (libs/nats_utils.py:127, libs/pocketbase.py:296-309)."""
    import json
    backup = synth.generate(3000, seed=91, dup_frac=0.3, mixed_locale=True, miss_frac=0.1)
    engine.load_cache(backup.cache)
    res = engine.replay(Batch.from_strings(backup.bodies, backup.senders, backup.date_ms), NOW)
    ora = O.replay(backup.bodies, backup.senders, [O.iso_utc_from_ms(m) for m in backup.date_ms], backup.cache, NOW)
    surv = [i for i, s in enumerate(ora.survivor) if s]
    payloads, records = res.parsed_payloads(), res.pocketbase_records()
    assert len(payloads) == len(records) == len(surv) > 500
    for k, i in enumerate(surv):
        p = ora.outcomes[i].parsed
        got = json.loads(payloads[k])
        assert got["msg_id"] == p.msg_id and got["raw_body"] == p.raw_body and got["sender"] == p.sender
        assert got["date"] == p.date.isoformat() and got["amount"] == str(p.amount) and got["balance"] == str(p.balance)
        assert got["txn_type"] == p.txn_type and got["card"] == p.card and got["merchant"] == p.merchant
        assert got["currency"] == p.currency and got["address"] == p.address and got["city"] == p.city
        assert got["device_id"] == "xml_backup" and got["parser_version"] == "llm-0.2.0"
        r = records[k]
        assert r == {"msg_id": p.msg_id, "original_body": p.raw_body, "sender": p.sender, "datetime": p.date.isoformat(),
                     "card": p.card, "amount": str(p.amount), "currency": p.currency, "balance": str(p.balance),
                     "merchant": p.merchant, "address": p.address, "city": p.city, "txn_type": r["txn_type"]}
        assert str(getattr(r["txn_type"], "value", r["txn_type"])) == p.txn_type


def test_full_size_properties(engine):
    """BASELINE.json config 2 size (1 M messages, 100 % hit for non-OTP bodies):
    properties that do not need the oracle at full size, plus a 2 000-message
    sample checked against it."""
    n = 1_000_000
    backup = synth.generate(n, seed=0x5EED + 2)
    engine.load_cache(backup.cache)
    batch = Batch.from_strings(backup.bodies, backup.senders, backup.date_ms)
    res = engine.replay(batch, NOW, columns=[c for c in engine.ALL_COLUMNS if c != "fixed_bytes"])
    kinds = np.array(backup.kinds)
    otp = kinds == "otp"
    # every OTP shape is skipped by the worker filter, nothing else is
    assert np.array_equal((res.flags & L.F_WORKER_SKIP) != 0, otp)
    # every other body hits the cache
    assert np.all(res.slot[~otp] >= 0)
    # the join is a bijection onto the loaded entries
    hit_slots = res.slot[res.slot >= 0]
    assert len(np.unique(hit_slots)) == len(hit_slots) == len(backup.cache)
    # keys are pairwise distinct (bodies are), so dedup keeps every upsert as both first and last
    ups = res.status == L.ST_UPSERT
    assert np.array_equal(res.survivor != 0, ups) and np.array_equal(res.first_seen != 0, ups)
    assert np.array_equal(res.last_idx[ups], np.nonzero(ups)[0])
    assert np.all(res.last_idx[~ups] == -1)
    # idempotence: replaying the same batch gives the same columns
    res2 = engine.replay(batch, NOW, columns=["msg_id", "cache_key", "status", "survivor"])
    for c in ("msg_id", "cache_key", "status", "survivor"):
        assert np.array_equal(res2.cols[c], res.cols[c])
    # doubling the batch (every message twice): survivors are exactly the second copies
    half = 100_000
    bodies2 = backup.bodies[:half] * 2
    res3 = engine.replay(Batch.from_strings(bodies2, ["B"] * (2 * half), range(2 * half)), NOW,
                         columns=["msg_id", "status", "survivor", "first_seen", "last_idx"])
    u = res3.status == L.ST_UPSERT
    assert np.array_equal(res3.survivor[:half] != 0, np.zeros(half, bool))
    assert np.array_equal(res3.survivor[half:] != 0, u[half:])
    assert np.array_equal(res3.first_seen[:half] != 0, u[:half])
    assert np.array_equal(res3.last_idx[:half][u[:half]], np.nonzero(u[:half])[0] + half)
    # sample against the oracle
    idx = np.random.Generator(np.random.PCG64(1)).choice(n, 2000, replace=False)
    for i in idx:
        i = int(i)
        o = O.replay_one(backup.bodies[i], backup.senders[i], O.iso_utc_from_ms(backup.date_ms[i]), backup.cache, NOW)
        assert res.msg_id_hex(i) == o.msg_id.hex() and res.cache_key_hex(i) == o.cache_key.hex()
        assert int(res.status[i]) == o.status
        if o.parsed is not None:
            assert int(res.date[i]) == O.pack_datetime(o.parsed.date)
            assert int(res.amount_coef[i]) == o.amount[2] and int(res.amount_exp[i]) == o.amount[3]
            assert int(res.balance_coef[i]) == o.balance[2] and int(res.balance_exp[i]) == o.balance[3]
