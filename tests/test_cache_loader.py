"""the response cache read straight from diskcache's SQLite file equals what the
reference sees through diskcache itself (`key in cache`, `cache[key]`, libs/gemini_parser.py:208-209)."""
import time

import diskcache
import numpy as np

from smsgate_b200 import gemini_cache, synth


def _fill(path):
    c = diskcache.Cache(path)
    backup = synth.generate(3000, seed=51, mixed_locale=True)
    for k, v in backup.cache.items():
        c[k] = v
    c["f" * 64] = {"txn_type": "debit", "merchant": "X" * 40_000}      # > 32 KiB: diskcache moves it to a file
    c["e" * 64] = "a plain string"                                      # MODE_RAW
    c["d" * 64] = b"raw bytes"
    c["c" * 64] = 12.5
    c["b" * 64] = None
    c.set("a" * 64, {"txn_type": "credit"}, expire=0.01)                # expired by the time we read
    c.set("9" * 64, {"txn_type": "credit"}, expire=3600)
    c[b"bytes key"] = {"txn_type": "debit"}                             # never equal to a str key
    c[12345] = {"txn_type": "debit"}
    c[("tuple", 1)] = {"txn_type": "debit"}
    first = next(iter(backup.cache))
    c[first] = {"txn_type": "unknown", "overwritten": True}             # UPDATE in place keeps the rowid
    del c[list(backup.cache)[5]]
    time.sleep(0.05)
    return c, backup


def test_sqlite_reader_matches_diskcache(tmp_path):
    c, backup = _fill(str(tmp_path))
    want = {}
    for k in c:                       # diskcache's own iteration (rowid order) and lookup
        if isinstance(k, str) and k in c:
            want[k] = c[k]
    c.close()
    got = gemini_cache.read_diskcache_sqlite(str(tmp_path))
    assert list(got) == list(want)
    assert got == want
    assert "a" * 64 not in got and "9" * 64 in got
    assert got["f" * 64]["merchant"] == "X" * 40_000


def test_table_from_sqlite_equals_table_from_mapping(tmp_path):
    """The flat arrays smsgate_cache_load receives are the same whichever way the cache was read."""
    c, backup = _fill(str(tmp_path))
    via_diskcache = {k: c[k] for k in c if isinstance(k, str) and k in c}
    c.close()

    def no_gpu(strings):               # host stand-in for ReplayEngine.strict_date_batch: "not strict" for all
        return np.full(len(strings), -1, np.int64)
    a = gemini_cache.build_table(via_diskcache, no_gpu)
    b = gemini_cache.build_table(gemini_cache.read_diskcache_sqlite(str(tmp_path)), no_gpu)
    for f in ("keys", "val_offsets", "val_bytes", "eflags", "date_packed"):
        assert np.array_equal(getattr(a, f), getattr(b, f)), f
    assert a.values == b.values and a.skipped == b.skipped


def _extra_entries(c):
    """Values at the edges of the native decoder's pickle subset."""
    base = dict(txn_type="debit", date="06.05.25 14:23", amount="52.00", currency="usd", card="***0018",
                merchant="TEST LLC", city=None, address="null", balance="1,842.74")
    odd = {
        "int_amount": dict(base, amount=52, balance=-7),
        "big_int": dict(base, amount=2 ** 40, balance=2 ** 70),            # LONG1 beyond int64: Python
        "float_amount": dict(base, amount=52.5),                           # float repr: Python
        "bool_none": dict(base, amount=True, balance=None),
        "shared_string": dict(base, amount="77.00", balance="77.00"[:]),   # the same object twice: BINGET
        "no_card": {k: v for k, v in base.items() if k != "card"},
        "card_none": dict(base, card=None),
        "merchant_empty": dict(base, merchant=""),
        "currency_int": dict(base, currency=5),
        "txn_bogus": dict(base, txn_type="bogus"),
        "extra_keys": dict(base, note="x", nested_ok=None),
        "nested_list": dict(base, tags=["a", "b"]),                        # a container: Python
        "nested_dict": dict(base, meta={"k": 1}),
        "date_unicode": dict(base, date="٠٦.٠٥.٢٥ 14:23"),                 # non-ASCII date: host chain
        "date_long": dict(base, date="x" * 40),
        "date_int": dict(base, date=20250506),
        "date_missing": {k: v for k, v in base.items() if k != "date"},
        "long_merchant": dict(base, merchant="Ж" * 300),                   # BINUNICODE (4-byte length)
        "not_a_dict": ["a", "list"],
        "empty_dict": {},
    }
    amount = "99.00"
    odd["shared_string"] = dict(base, amount=amount, balance=amount)
    import hashlib
    names = {}
    for name, v in odd.items():
        k = hashlib.sha256(name.encode()).hexdigest()
        c[k] = v
        names[k] = name
    return names


import pytest


@pytest.mark.parametrize("connections", ["1", "3", "16"])
def test_native_reader_matches_python_reader(tmp_path, monkeypatch, connections):
    """smsgate_dcache_* (C++: SQLite + pickle subset, no Python object per entry) against the per-entry
    Python path, array by array, for every entry it decodes itself; what it declines is flagged, not guessed.
    Read on one connection and as 3 / 16 rowid ranges at once (what a large cache gets by default)."""
    monkeypatch.setenv("SMSGATE_DCACHE_THREADS", connections)
    import ctypes as C

    from smsgate_b200 import _lib as L
    c, backup = _fill(str(tmp_path))
    names = _extra_entries(c)
    c.close()
    by_name, host, keys, m = _native_vs_python(tmp_path, names)
    assert {n for n, e in by_name.items() if n and host[e] == 1} == {"big_int", "float_amount", "nested_list", "nested_dict", "not_a_dict"}
    assert {n for n, e in by_name.items() if n and host[e] == 2} == {"date_unicode", "date_long", "date_int"}
    assert host[[e for e in range(m) if keys[e].tobytes().hex() == "f" * 64][0]] == 1      # value in a side file
    assert (host == 0).sum() > 2500


def _native_vs_python(tmp_path, names):
    """Every array the native reader fills against the per-entry Python path on the same directory."""
    import ctypes as C

    from smsgate_b200 import _lib as L

    def no_gpu(strings):
        return np.full(len(strings), -1, np.int64)
    want = gemini_cache.build_table(gemini_cache.read_diskcache_sqlite(str(tmp_path)), no_gpu)
    lib = L.lib()
    h = C.c_void_p()
    err = C.create_string_buffer(256)
    assert lib.smsgate_dcache_open(str(tmp_path).encode(), time.time(), C.byref(h), err, 256) == 0, err.value
    m, vb, tb, sk, nh = (C.c_int64(0) for _ in range(5))
    lib.smsgate_dcache_sizes(h, C.byref(m), C.byref(vb), C.byref(tb), C.byref(sk), C.byref(nh))
    assert m.value == want.m and sk.value == want.skipped
    keys = np.zeros((m.value, 32), np.uint8)
    off = np.zeros(4 * m.value + 1, np.int32)
    vbytes = np.zeros(vb.value + 1, np.uint8)
    eflags = np.zeros(m.value, np.uint16)
    rowids = np.zeros(m.value, np.int64)
    host = np.zeros(m.value, np.uint8)
    toff = np.zeros(5 * m.value + 1, np.int64)
    tbytes = np.zeros(tb.value + 1, np.uint8)
    tkind = np.zeros(5 * m.value, np.uint8)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    assert lib.smsgate_dcache_fill(h, p(keys), p(off), p(vbytes), p(eflags), p(rowids), p(host), p(toff), p(tbytes), p(tkind)) == 0
    lib.smsgate_dcache_close(h)
    assert np.array_equal(keys, want.keys)                       # same entries, same (rowid) order
    assert nh.value == int((host == 1).sum())
    date_bits = L.E_DATE_STRICT | L.E_DATE_HOST | L.E_DATE_NOBODY | L.E_DATE_TZ | L.E_DATE_PHONE
    by_name = {}
    for e in range(m.value):
        name = names.get(keys[e].tobytes().hex())
        by_name[name] = e
        if host[e] == 1:
            continue
        for k in (0, 1, 3):                                       # str(amount), str(balance), card
            a = vbytes[off[4 * e + k]:off[4 * e + k + 1]].tobytes()
            b = want.val_bytes[want.val_offsets[4 * e + k]:want.val_offsets[4 * e + k + 1]].tobytes()
            assert a == b, (name, k)
        assert int(eflags[e]) == int(want.eflags[e]) & ~date_bits, name
        got_date = vbytes[off[4 * e + 2]:off[4 * e + 3]].tobytes()
        if host[e] == 0:
            assert got_date == want.val_bytes[want.val_offsets[4 * e + 2]:want.val_offsets[4 * e + 3]].tobytes(), name
        else:
            assert got_date == b""                                # the host chain decides this date
        v = want.values[e]
        for j, field in enumerate(("txn_type", "currency", "merchant", "city", "address")):
            kind = int(tkind[5 * e + j])
            text = tbytes[toff[5 * e + j]:toff[5 * e + j + 1]].tobytes().decode()
            if not isinstance(v, dict) or field not in v:
                assert kind == 0, (name, field)
            elif v[field] is None:
                assert kind == 1
            elif isinstance(v[field], str):
                assert kind == 2 and text == v[field], (name, field)
            else:
                assert kind not in (0, 1, 2)
    return by_name, host, keys, m.value


def test_native_reader_fuzz_over_pickle_protocols(tmp_path, monkeypatch):
    """Random response dicts - every value type the decoder's subset has an opinion on, shared and repeated
    strings, long strings, extra keys - stored by diskcache with pickle protocols 2 to 5 (a cache directory may
    have been written by an older interpreter).  What the native reader decodes itself must equal the Python
    path; declining an entry is always allowed."""
    import hashlib
    import random
    monkeypatch.setenv("SMSGATE_DCACHE_THREADS", "4")
    rng = random.Random(0xCAC4E)
    words = ["", "a", "52.00", "1,842.74", "TEST LLC", "Ереван", "☕", "x" * 255, "y" * 256, "z" * 70000, "null",
             "06.05.25 14:23", "06.05.2025 14:23", "2025-05-06T14:23:00+04:00", "***0018", "4083***7538", "usd"]

    def value():
        k = rng.random()
        if k < 0.55:
            return rng.choice(words)
        if k < 0.65:
            return None
        if k < 0.75:
            return rng.choice([0, 1, -1, 255, 256, 65535, 65536, 2 ** 31 - 1, 2 ** 31, -2 ** 31, 2 ** 63 - 1, 2 ** 63, -2 ** 63, 10 ** 30])
        if k < 0.8:
            return rng.choice([True, False])
        if k < 0.85:
            return rng.choice([0.0, 12.5, -1e300, float("inf")])
        if k < 0.9:
            return rng.choice([b"bytes", ("t", 1), ["l"], {"d": 1}, {1, 2}, frozenset(), bytearray(b"x")])
        return "".join(rng.choice("aЖ☕ \"\\\n0,.-") for _ in range(rng.randint(0, 30)))
    fields = ["txn_type", "date", "amount", "currency", "card", "merchant", "city", "address", "balance"]
    total_native = 0
    for proto in (2, 3, 4, 5):
        d = tmp_path / f"p{proto}"
        c = diskcache.Cache(str(d), disk_pickle_protocol=proto)
        names = {}
        for j in range(400):
            keep = [f for f in fields if rng.random() < 0.9]
            rng.shuffle(keep)
            shared = value()
            v = {f: (shared if rng.random() < 0.2 else value()) for f in keep}
            if rng.random() < 0.5:
                v["txn_type"] = rng.choice(["debit", "credit", "otp", "unknown", "bogus"])
            if rng.random() < 0.2:
                v[rng.choice(["note", "x" * 300, "Ж"])] = value()
            if rng.random() < 0.03:
                v = rng.choice([[], "str", 5, None, (1, 2), {}])
            name = f"p{proto}_{j}"
            k = hashlib.sha256(name.encode()).hexdigest()
            c[k] = v
            names[k] = name
        c.close()
        by_name, host, keys, m = _native_vs_python(d, names)
        total_native += int((host != 1).sum())
    assert total_native > 600
