"""Times the per-message work of the legacy offline replay (process_cached.py:228-256: strip, OTP filter,
the two regular expressions, float conversion, SHA-256 of the body): the oracle port on one host core
(which runs CPython's own `re`, i.e. what the reference runs) against smsgate_legacy_parse on the GPU,
host arrays in, host arrays out.  The DiskCache reads / writes around it are not part of either number."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import legacy_oracle as LO          # noqa: E402  (CPU baseline leg of a bench script)
from smsgate_b200 import legacy, synth          # noqa: E402
from smsgate_b200.replay import ReplayEngine    # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
n_cpu = min(n, 50_000)
backup = synth.generate(n, seed=0x5EED + 2)
t0 = time.perf_counter()
cpu = [LO.process_body(b) for b in backup.bodies[:n_cpu]]
t_cpu = time.perf_counter() - t0
eng = ReplayEngine(0)
best = 1e9
for _ in range(3):
    t0 = time.perf_counter()
    res = legacy.parse_bodies(eng, backup.bodies)
    best = min(best, time.perf_counter() - t0)
# the C call alone (host arrays in and out), without building the arrays from the Python list
import ctypes as C
import numpy as np
from smsgate_b200 import _lib as L
from smsgate_b200.batch import ragged_from_strings
t0 = time.perf_counter()
data, offsets = ragged_from_strings(backup.bodies)
t_enc = time.perf_counter() - t0
cols = dict(status=np.zeros(n, np.uint8), lflags=np.zeros(n, np.uint8), strip=np.zeros((n, 2), np.int32),
            caps=np.full((n, 2 * L.LEG_GROUPS), -1, np.int32), amount=np.zeros(n, np.float64),
            balance=np.zeros(n, np.float64), body_hash=np.zeros((n, 32), np.uint8))
out = L.LegacyOutputs(**{k: v.ctypes.data for k, v in cols.items()})
call = 1e9
for _ in range(3):
    t0 = time.perf_counter()
    L.check(eng._ctx, L.lib().smsgate_legacy_parse(eng._ctx, n, offsets.ctypes.data_as(C.c_void_p),
                                                   data.ctypes.data_as(C.c_void_p), C.byref(out)))
    call = min(call, time.perf_counter() - t0)
import collections
kinds = collections.Counter(int(s) for s in res.cols["status"])
for i in range(0, n_cpu, 997):
    assert int(res.cols["status"][i]) == cpu[i][0], i
print(f"{n} bodies; statuses (0 otp, 1 debit, 2 credit, 3 failed): {dict(kinds)}")
print(f"oracle port (CPython re), 1 core, {n_cpu} bodies: {1e6 * t_cpu / n_cpu:.1f} us/body = {n_cpu / t_cpu / 1e3:.1f} k bodies/s")
print(f"smsgate_legacy_parse incl. UTF-8 encoding of the list and both copies: {best * 1e3:.0f} ms = {n / best / 1e6:.2f} M bodies/s")
print(f"  of which: UTF-8 encoding of the Python list {t_enc * 1e3:.0f} ms; smsgate_legacy_parse alone (pageable host arrays, "
      f"copies + kernel) {call * 1e3:.1f} ms = {n / call / 1e6:.1f} M bodies/s")
