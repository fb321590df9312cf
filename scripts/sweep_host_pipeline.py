"""End-to-end host call (smsgate_replay_host_records8, pinned buffers) under different chunkings of the batch:
divisor / min_chunk / offsets-first (smsgate_set_host_pipeline), with the per-chunk timeline of the best one.
usage: python scripts/sweep_host_pipeline.py > gpurun_out/host_pipeline.json"""
import ctypes as C
import datetime as dt
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from smsgate_b200 import _lib as L  # noqa: E402
from smsgate_b200 import synth  # noqa: E402
from smsgate_b200.batch import Batch  # noqa: E402
from smsgate_b200.replay import ReplayEngine  # noqa: E402

NOW = dt.datetime(2026, 1, 1)
n = 1_000_000
backup = synth.generate(n, seed=0x5EED + 2)
batch = Batch.from_strings(backup.bodies, backup.senders, backup.date_ms)
eng = ReplayEngine(0)
eng.load_cache(backup.cache)
h_bytes = torch.from_numpy(batch.body_bytes.copy()).pin_memory()
h_offs = torch.from_numpy(batch.body_offsets.copy()).pin_memory()
hb = Batch(h_bytes.numpy(), h_offs.numpy(), backup.senders, None)
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
out = []
for div, mn, first in ((5, 65536, 0), (3, 131072, 1), (3, 131072, 0), (3, 100000, 1), (3, 163840, 1), (3, 196608, 1),
                       (4, 131072, 1), (4, 100000, 1), (2, 131072, 1), (2, 196608, 1), (5, 131072, 1), (3, 262144, 1)):
    eng.set_host_pipeline(div, mn, bool(first))
    ts = []
    for it in range(9):
        flush.zero_()
        torch.cuda.synchronize()
        e0, e1 = ev(), ev()
        e0.record()
        eng.replay_records(hb, NOW, pinned=True)
        e1.record()
        e1.synchronize()
        if it >= 3:
            ts.append(e0.elapsed_time(e1))
    tl = (C.c_float * 40)()
    nc = L.lib().smsgate_replay_host_timeline(eng._ctx, tl, 40)
    row = {"divisor": div, "min_chunk": mn, "offsets_first": first, "e2e_ms": sum(ts) / len(ts), "min_ms": min(ts), "chunks": nc,
           "copy_in_done_ms": round(tl[nc - 1], 3), "last_kernel_ms": round(tl[2 * nc - 1], 3), "dedup_ms": round(tl[2 * nc], 3),
           "all_back_ms": round(tl[2 * nc + 1], 3)}
    out.append(row)
    print(row, file=sys.stderr, flush=True)
print(json.dumps(out, indent=1))
