"""Timeline of smsgate_replay_host on the bench workload (GPU box)."""
import ctypes as C, sys, datetime as dt
import numpy as np, torch
sys.path.insert(0, ".")
from smsgate_b200 import synth, _lib as L
from smsgate_b200.batch import Batch
from smsgate_b200.replay import ReplayEngine
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
backup = synth.generate(n, seed=0x5EED + 2)
eng = ReplayEngine(0)
eng.load_cache(backup.cache)
b = Batch.from_strings(backup.bodies, backup.senders, backup.date_ms)
hb_bytes = torch.from_numpy(b.body_bytes.copy()).pin_memory(); hb_offs = torch.from_numpy(b.body_offsets.copy()).pin_memory()
hb = Batch(hb_bytes.numpy(), hb_offs.numpy(), backup.senders, None)
for it in range(5):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    res = eng.replay_records(hb, bench.NOW, pinned=True)
    e1.record(); e1.synchronize()
    ms = (C.c_float * 40)()
    nc = eng._lib.smsgate_replay_host_timeline(eng._ctx, ms, 40)
    print("total %.3f ms; chunks %d" % (e0.elapsed_time(e1), nc))
    print("  h2d done :", " ".join("%.2f" % ms[c] for c in range(nc)))
    print("  kern done:", " ".join("%.2f" % ms[nc + c] for c in range(nc)))
    print("  dedup %.2f  all stored %.2f" % (ms[2 * nc], ms[2 * nc + 1]))
