"""Host-side timing of the response-cache loader: diskcache's own per-key access
(what the reference does per message, libs/gemini_parser.py:208-209) against one SQL query over the same
file.  CPU only.  usage: python scripts/bench_cache_load.py [entries]"""
import os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import diskcache
from smsgate_b200 import gemini_cache, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
backup = synth.generate(n, seed=61)
d = tempfile.mkdtemp(prefix="gemini_cache_")
c = diskcache.Cache(d, size_limit=1 << 36)
with c.transact():
    for k, v in backup.cache.items():
        c[k] = v
m = len(backup.cache)
t0 = time.perf_counter()
a = {k: c[k] for k in backup.cache if k in c}
t1 = time.perf_counter()
c.close()
b = gemini_cache.read_diskcache_sqlite(d)
t2 = time.perf_counter()
assert a == b
print(f"{m} entries: diskcache `key in cache` + `cache[key]` {1e6 * (t1 - t0) / m:.1f} us/entry, "
      f"one SQL query + unpickle {1e6 * (t2 - t1) / m:.2f} us/entry ({(t1 - t0) / (t2 - t1):.0f}x)")
