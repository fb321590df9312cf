"""Times builds of the library against each other on one GPU: smsgate_b200/variants/lib*.so (made by
smsgate_b200.build.build_variant) on the hit100 and mixed workloads (VARIANT_WORKLOADS=hit100,skew,... picks others), keys kernel and whole step, with the L2
flushed between steps.  Every variant's key columns are compared with the first one's.
Usage: python scripts/time_variants.py [name ...] > gpurun_out/variants.json"""
import glob
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import datetime as dt  # noqa: E402

import torch  # noqa: E402

from smsgate_b200 import _lib as L  # noqa: E402
from smsgate_b200 import synth  # noqa: E402
from smsgate_b200.batch import Batch  # noqa: E402

NOW = dt.datetime(2026, 1, 1)
N = int(os.environ.get("VARIANT_MSGS", "1000000"))


def engine_for(path):
    from smsgate_b200.replay import ReplayEngine
    L._lib = None
    L.LIB_PATH = path
    return ReplayEngine(0)


def main():
    names = sys.argv[1:]
    libs = sorted(glob.glob(os.path.join(ROOT, "smsgate_b200", "variants", "lib*.so")))
    if names:
        libs = [p for p in libs if os.path.basename(p)[3:-3] in names]
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    out = {}
    known = {"hit100": dict(seed=0x5EED + 2), "mixed": dict(seed=0x5EED + 5, mixed_locale=True, miss_frac=0.3),
             "dup10": dict(seed=0x5EED + 3, dup_frac=0.10), "skew": dict(seed=0x5EED + 4, dup_frac=0.5, hot_key_zipf=1.2)}
    for wl in os.environ.get("VARIANT_WORKLOADS", "hit100,mixed").split(","):
        kw = known[wl]
        backup = synth.generate(N, **kw)
        batch = Batch.from_strings(backup.bodies, backup.senders, backup.date_ms)
        total = int(batch.body_offsets[-1])
        d_bytes = torch.from_numpy(batch.body_bytes.copy()).cuda()
        d_offs = torch.from_numpy(batch.body_offsets.copy()).cuda()
        table, ref_sum = None, None
        for path in libs:
            name = os.path.basename(path)[3:-3]
            eng = engine_for(path)
            table = eng.load_cache(backup.cache if table is None else table)
            eng.reserve(N, total)
            cols = eng.device_columns(N, total)
            k1, step = [], []
            for it in range(13):
                flush.zero_()
                e0, e1, e2 = ev(), ev(), ev()
                e0.record()
                eng.scan_keys(d_bytes, d_offs, cols)
                e1.record()
                eng.join_route_dedup(N, NOW, 0, cols)
                e2.record()
                e2.synchronize()
                if it >= 3:
                    k1.append(e0.elapsed_time(e1))
                    step.append(e0.elapsed_time(e2))
            phases = None
            if hasattr(L.lib(), "smsgate_debug_scan_phases"):          # -DK1_PHASE_TIMERS build: clocks per phase, summed over CTAs
                import ctypes as C
                buf = (C.c_ulonglong * 8)()
                L.lib().smsgate_debug_scan_phases(None, 1)
                eng.scan_keys(d_bytes, d_offs, cols)
                torch.cuda.synchronize()
                L.lib().smsgate_debug_scan_phases(buf, 0)
                tot = float(sum(buf[:4])) or 1.0
                phases = {k: round(buf[i] / tot, 4) for i, k in enumerate(("stage_sort", "scan", "partition", "hash"))}
                phases["clocks_per_cta"] = tot / 148
            chk = [int(cols[c].to(torch.int64).sum().item()) for c in ("msg_id", "cache_key", "flags", "status", "survivor", "first_seen", "last_idx")]
            if ref_sum is None:
                ref_sum = chk
            out.setdefault(wl, {})[name] = {"k_scan_keys_ms": sum(k1) / len(k1), "k_scan_keys_min_ms": min(k1),
                                            "step_ms": sum(step) / len(step), "join_dedup_ms": (sum(step) - sum(k1)) / len(step), "same_results_as_first": chk == ref_sum,
                                            "phase_share": phases}
            print(wl, name, out[wl][name], file=sys.stderr, flush=True)
            eng.close()
            del cols
            torch.cuda.empty_cache()
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
