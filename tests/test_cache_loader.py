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
