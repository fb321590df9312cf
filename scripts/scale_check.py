"""BASELINE.json configs[2] per-GPU size (12.5 M messages, ~1.45 GB of bodies) on one GPU: the 1 M bench
batch repeated 12.5 times, so every upsert key occurs 12 or 13 times.  Checks that the large batch gives,
copy by copy, the statuses / keys of the 1 M batch, that exactly the last copy of every upsert survives
and the first is first_seen, and times the device-resident path and the host call."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                      # noqa: E402
from smsgate_b200 import _lib as L, synth         # noqa: E402
from smsgate_b200.batch import Batch              # noqa: E402
from smsgate_b200.replay import ReplayEngine      # noqa: E402

base_n = 1_000_000
reps = float(sys.argv[1]) if len(sys.argv) > 1 else 12.5
backup = synth.generate(base_n, seed=0x5EED + 2)
b1 = Batch.from_strings(backup.bodies, backup.senders, backup.date_ms)
n = int(base_n * reps)
full, part = n // base_n, n % base_n
tb = int(b1.body_offsets[-1])
bytes_ = np.concatenate([np.tile(b1.body_bytes, full), b1.body_bytes[:int(b1.body_offsets[part])]])
offs = np.concatenate([(b1.body_offsets[:-1][None, :] + (np.arange(full, dtype=np.int64) * tb)[:, None]).ravel(),
                       b1.body_offsets[:part + 1] + full * tb])
assert len(offs) == n + 1 and offs[-1] == len(bytes_)
eng = ReplayEngine(0)
eng.load_cache(backup.cache)
ref = eng.replay_records(b1, bench.NOW, pinned=False)
hb, ho = torch.from_numpy(bytes_).pin_memory(), torch.from_numpy(offs).pin_memory()     # pinned, like bench.py
big = Batch(hb.numpy(), ho.numpy(), None, None)
t0 = time.perf_counter()
res = eng.replay_records(big, bench.NOW, pinned=True)
t_first = time.perf_counter() - t0
t0 = time.perf_counter()
res = eng.replay_records(big, bench.NOW, pinned=True)
t_host = time.perf_counter() - t0
idx = np.arange(n) % base_n
assert np.array_equal(res.status, ref.status[idx])
assert np.array_equal(res.msg_id, ref.msg_id[idx])
assert np.array_equal(res.amount_coef, ref.amount_coef[idx]) and np.array_equal(res.date, ref.date[idx])
ups = res.status == L.ST_UPSERT
copies = np.where(np.arange(base_n) < part, full + 1, full)
last_copy = (copies[idx] - 1) * base_n + idx
assert np.array_equal(res.survivor != 0, ups & (np.arange(n) == last_copy))
assert np.array_equal(res.first_seen != 0, ups & (np.arange(n) < base_n))
assert np.array_equal(res.last_idx[ups], last_copy[ups]) and np.all(res.last_idx[~ups] == -1)
# device-resident
d_bytes = torch.from_numpy(bytes_).cuda()
d_offs = torch.from_numpy(offs).cuda()
cols = eng.device_columns(n, len(bytes_), [c for c in eng.ALL_COLUMNS if c != "fixed_bytes"])
eng.reserve(n, len(bytes_))
ts = []
for _ in range(4):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); eng.replay_device(d_bytes, d_offs, bench.NOW, cols); e1.record(); e1.synchronize()
    ts.append(e0.elapsed_time(e1))
assert np.array_equal(cols["survivor"].cpu().numpy(), res.survivor)
print(f"{n} messages, {len(bytes_) / 1e9:.2f} GB of bodies, every key {full}-{full + 1} times: all checks passed")
print(f"device-resident replay: {min(ts):.2f} ms = {n / min(ts) / 1e6:.2f} G msgs/s; host call (pinned in, records out): "
      f"{t_host * 1e3:.0f} ms = {n / t_host / 1e6:.0f} M msgs/s (first call incl. allocations {t_first * 1e3:.0f} ms)")
