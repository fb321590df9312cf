"""Host-side timing of the response-cache loader, CPU only, three ways over the same
`.gemini_cache` directory:
  1. what the reference does per message: diskcache `key in cache` + `cache[key]` (libs/gemini_parser.py:208-209)
  2. round 1's loader: one SQL query, pickle.loads per row, then the per-entry Python loop of build_table
  3. the native reader: SQLite + pickle subset decoded in C++ into the flat arrays (smsgate_dcache_*)
(the date classification that follows 2 and 3 runs on the GPU and is not part of this host-side figure).
usage: python scripts/bench_cache_load.py [entries] [--skip-diskcache]"""
import ctypes as C
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import diskcache  # noqa: E402
import numpy as np  # noqa: E402

from smsgate_b200 import _lib as L  # noqa: E402
from smsgate_b200 import gemini_cache, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 100_000
d = tempfile.mkdtemp(prefix="gemini_cache_")
c = diskcache.Cache(d, size_limit=1 << 38)
m = 0
keys_sample = []
for part in range(0, n, 500_000):                     # generated and stored in parts: 10 M entries do not fit a dict comfortably
    backup = synth.generate(min(500_000, n - part), seed=61 + part)
    with c.transact():
        for k, v in backup.cache.items():
            c[k] = v
    m += len(backup.cache)
    keys_sample += list(backup.cache)[:20_000]
    del backup
line = f"{m} entries ({os.path.getsize(os.path.join(d, 'cache.db')) / 1e6:.0f} MB cache.db), host cores {os.cpu_count()}: "
if "--skip-diskcache" not in sys.argv:
    t0 = time.perf_counter()
    a = {k: c[k] for k in keys_sample if k in c}
    t1 = time.perf_counter()
    line += f"diskcache per key {1e6 * (t1 - t0) / len(keys_sample):.1f} us/entry (sample of {len(keys_sample)}); "
c.close()


def no_gpu(strings):
    return np.full(len(strings), -1, np.int64)


t0 = time.perf_counter()
py0 = t0
b = gemini_cache.read_diskcache_sqlite(d)
t1 = time.perf_counter()
tb = gemini_cache.build_table(b, no_gpu)
t2 = time.perf_counter()
py_total = t2 - py0
line += (f"python loader {t2 - t0:.2f} s = {1e6 * (t2 - t0) / m:.2f} us/entry (SQL + unpickle {t1 - t0:.2f} s, "
         f"build_table {t2 - t1:.2f} s); ")
del b
lib = L.lib()
h = C.c_void_p()
err = C.create_string_buffer(256)


def native_open_seconds(connections):
    """smsgate_dcache_open alone, best of three, with the number of read connections forced."""
    os.environ["SMSGATE_DCACHE_THREADS"] = str(connections)
    best = None
    for _ in range(3):
        hh = C.c_void_p()
        a = time.perf_counter()
        assert lib.smsgate_dcache_open(d.encode(), time.time(), C.byref(hh), err, 256) == 0, err.value
        b_ = time.perf_counter()
        lib.smsgate_dcache_close(hh)
        best = b_ - a if best is None else min(best, b_ - a)
    os.environ.pop("SMSGATE_DCACHE_THREADS")
    return best


one, eight = native_open_seconds(1), native_open_seconds(8)
t0 = time.perf_counter()
assert lib.smsgate_dcache_open(d.encode(), time.time(), C.byref(h), err, 256) == 0, err.value
mm, vb, tbs, sk, nh = (C.c_int64(0) for _ in range(5))
lib.smsgate_dcache_sizes(h, C.byref(mm), C.byref(vb), C.byref(tbs), C.byref(sk), C.byref(nh))
keys = np.zeros((mm.value, 32), np.uint8)
off = np.zeros(4 * mm.value + 1, np.int32)
vbytes = np.zeros(vb.value + 1, np.uint8)
eflags = np.zeros(mm.value, np.uint16)
p = lambda x: x.ctypes.data_as(C.c_void_p)  # noqa: E731
assert lib.smsgate_dcache_fill(h, p(keys), p(off), p(vbytes), p(eflags), None, None, None, None, None) == 0
lib.smsgate_dcache_close(h)
t1 = time.perf_counter()
assert mm.value == m and np.array_equal(keys, tb.keys) and nh.value == 0
assert np.array_equal(off[1::4] - off[0:-1:4], tb.val_offsets[1::4] - tb.val_offsets[0:-1:4])      # same amount strings
line += (f"native reader {t1 - t0:.2f} s = {1e6 * (t1 - t0) / m:.2f} us/entry ({py_total / (t1 - t0):.0f}x faster than the python "
         f"loader; reading alone, best of 3: one connection {1e6 * one / m:.2f} us/entry, eight rowid ranges at once "
         f"{1e6 * eight / m:.2f} us/entry)")
print(line)
import shutil  # noqa: E402
shutil.rmtree(d, ignore_errors=True)
