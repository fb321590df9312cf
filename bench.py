#!/usr/bin/env python
"""bench.py — bulk replay throughput (BASELINE.json metric) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W]            our CUDA path
  python bench.py --impl reference [--gpus N] [--steps K] ...   the CPU path of the reference
                                                                (oracle port, all host cores)

One "step" = one pass of the whole hot path (keys + scans, cache join + amounts
+ routing, dedup) over one batch of synthetic messages.  Workload at every N is
BASELINE.json configs[1] per GPU: 1 M synthetic bank-notification SMS, every
non-OTP body present in the loaded Gemini response cache (weak scaling: each
rank replays its own 1 M shard; msg_ids are globally deduplicated through the
peer-memory exchange by key-hash prefix when N > 1, smsgate_b200/dist.py).

Prints ONE JSON line (rank 0).  `value` = messages/s with the batch resident in
HBM; `e2e` = the same through the host-buffer C-ABI call
(smsgate_replay_host_records: pinned host memory in, one 64-byte result record
and one dedup verdict per message back out, copies inside the timed region).
"""
from __future__ import annotations

import argparse
import datetime as dt
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "msgs_per_sec_bulk_replay"
UNIT = "msgs/s"
NOW = dt.datetime(2026, 1, 1)
N_PER_GPU = 1_000_000
RESULT_COLUMNS = ("msg_id", "status", "slot", "amount_coef", "balance_coef", "amount_exp", "balance_exp",
                  "dec_flags", "date", "survivor", "first_seen", "last_idx")


def workload_name(n):
    return WORKLOADS["hit100"].format(n=n)


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.path = tempfile.mktemp(suffix=".csv")

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        for line in open(self.path):
            p = [x.strip() for x in line.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                mx.append(float(p[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak_hbm():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy bandwidth)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md, 6.65 TB/s)"


# --------------------------------------------------------------------------- CPU arms
def _cpu_replay_shard(args):
    """One worker of the CPU arm: the oracle port over a shard, against an
    on-disk diskcache like the reference's `.gemini_cache` (2 SQLite queries
    + unpickle per hit, libs/gemini_parser.py:208-209)."""
    bodies, senders, date_ms, cache_dir = args
    import contextlib
    import io

    import diskcache

    from oracle import replay_oracle as O
    cache = diskcache.Cache(cache_dir)
    raw_dates = [O.iso_utc_from_ms(m) for m in date_ms]
    with contextlib.redirect_stdout(io.StringIO()):
        res = O.replay(bodies, senders, raw_dates, cache, NOW, lean=True)     # what the reference computes, nothing more
    cache.close()
    return [(o.msg_id, o.status) for o in res.outcomes]


def _seed_diskcache(entries: dict) -> str:
    import diskcache
    d = tempfile.mkdtemp(prefix="gemini_cache_")
    c = diskcache.Cache(d, size_limit=1 << 36)       # default 1 GiB would evict silently
    with c.transact():
        for k, v in entries.items():
            c[k] = v
    c.close()
    return d


def cpu_arm(backup, n_sample: int, workers: int, steps: int, warmup: int):
    """Times the oracle port on `n_sample` messages of the workload per step.
    Returns (msgs_per_s, ms_per_step)."""
    from oracle import replay_oracle as O
    from smsgate_b200.synth import reference_cache_key
    n_sample = min(n_sample, len(backup.bodies))
    bodies, senders, date_ms = backup.bodies[:n_sample], backup.senders[:n_sample], backup.date_ms[:n_sample]
    need = {}
    for b in bodies:
        k = reference_cache_key(b)
        if k in backup.cache:
            need[k] = backup.cache[k]
    cache_dir = _seed_diskcache(need)
    shards = []
    per = (n_sample + workers - 1) // workers
    for w in range(workers):
        lo, hi = w * per, min(n_sample, (w + 1) * per)
        if lo < hi:
            shards.append((bodies[lo:hi], senders[lo:hi], date_ms[lo:hi], cache_dir))
    pool = None
    if workers > 1:
        import multiprocessing as mp
        pool = mp.get_context("fork").Pool(workers)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        parts = pool.map(_cpu_replay_shard, shards) if pool else [_cpu_replay_shard(s) for s in shards]
        flat = [x for p in parts for x in p]             # shard order = replay order: last writer wins
        O.dedup((m for m, _ in flat), (s == O.ST_UPSERT for _, s in flat))
        t1 = time.perf_counter()
        if it >= warmup:
            times.append(t1 - t0)
    if pool:
        pool.close()
    import shutil
    shutil.rmtree(cache_dir, ignore_errors=True)
    mean = sum(times) / len(times)
    return n_sample / mean, mean * 1e3, n_sample


def run_reference(args, rank, world):
    if rank != 0:
        return
    from smsgate_b200 import synth
    cores = os.cpu_count() or 1
    n_sample = min(N_PER_GPU, 3000 * cores)
    backup = synth.generate(max(n_sample, 1000), seed=0x5EED + 2)
    value, ms, n_used = cpu_arm(backup, n_sample, cores, args.steps, args.warmup)
    sample = (f"{n_used} messages of the workload per step, bodies already in memory (no XML parse), "
              f"oracle port of worker filter + parse_sms_llm hit branch + writer gate + dict upsert, "
              f"diskcache-backed response cache, {cores} processes; the port's lean form (the reference's order of work, "
              f"nothing the reference does not compute) runs within a tenth of the reference's own functions on the same "
              f"input (profiles/r2_cpu_units_reference_vs_port.txt: 11.9 k vs 11.0 k msgs/s on one core)")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8/u32/int64", "data": "synthetic",
            "config": {"workload": workload_name(N_PER_GPU), "sample_msgs_per_step": n_used},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------- our arm
WORKLOADS = {
    "hit100": "configs[1]: {n} synthetic bank-notification SMS per GPU, 100% cache hit for non-OTP bodies",
    "dup10": "configs[2] per-GPU shape: {n} messages per GPU, 10% of the job duplicates spread across shards "
             "(common pool), every rank its own stream otherwise",
    "skew": "configs[3]: {n} messages per GPU, 50% duplicates Zipf(1.2) from a set common to all ranks, 50% of each "
            "rank's distinct bodies re-drawn until msg_id prefix owner = rank 0",
    "mixed": "configs[4]: {n} mixed-locale bodies (40% Cyrillic with NBSP / U+202F thousands and comma decimals, "
             "10% card forms, 10% multi-part 400-1200 B), 30% cache misses routed to the DLQ set",
}


def make_backup(workload: str, n: int, rank: int, world: int):
    from smsgate_b200 import synth
    if workload == "hit100":
        return synth.generate(n, seed=0x5EED + 2 + 7919 * rank)
    if workload == "dup10":
        return synth.shard_with_cross_duplicates(n, rank, world, dup_frac=0.10, seed=0x5EED + 3)
    if workload == "skew":
        return synth.shard_skewed(n, rank, world, seed=0x5EED + 4)
    if workload == "mixed":
        return synth.generate(n, seed=0x5EED + 5 + 7919 * rank, mixed_locale=True, miss_frac=0.3)
    raise ValueError(workload)


class Resident:
    """One workload resident on this rank's GPU: batch in HBM, cache loaded, output columns allocated."""

    def __init__(self, eng, backup, rank, ov=None):
        import torch

        from smsgate_b200.batch import Batch
        self.backup = backup
        self.batch = Batch.from_strings(backup.bodies, backup.senders, backup.date_ms)
        self.n = self.batch.n
        self.total = int(self.batch.body_offsets[-1])
        self.base_index = rank * self.n
        eng.load_cache(backup.cache)
        eng.reserve(self.n, self.total)
        if ov is not None:                              # the other lane probes the same table
            ov.refresh_cache()
            ov.reserve(self.n, self.total)
        self.h_bytes = torch.from_numpy(self.batch.body_bytes.copy()).pin_memory()
        self.h_offs = torch.from_numpy(self.batch.body_offsets.copy()).pin_memory()
        self.d_bytes = self.h_bytes.cuda()
        self.d_offs = self.h_offs.cuda()
        self.cols = eng.device_columns(self.n, self.total)


def time_resident(eng, R, exchange, steps, warmup, flush, barrier):
    """`steps` device-resident passes of the whole path over R, one batch at a time (nothing overlapped).
    Returns (mean step ms, mean k_scan_keys ms)."""
    import torch
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    k1_ms, step_ms = [], []
    barrier()                                           # ranks build their workloads at different speeds: line up before the first exchange
    for it in range(warmup + steps):
        if it == warmup:
            barrier()
        flush.zero_()                                   # evict the previous step's lines from L2 (not timed)
        e0, e1, e2 = ev(), ev(), ev()
        e0.record()
        eng.scan_keys(R.d_bytes, R.d_offs, R.cols)
        e1.record()
        eng.join_route_dedup(R.n, NOW, R.base_index, R.cols)  # stages 2 + 3 as smsgate_replay_device runs them
        if exchange is not None:
            exchange.run(R.cols, R.base_index)
        e2.record()
        e2.synchronize()
        if it >= warmup:
            k1_ms.append(e0.elapsed_time(e1))
            step_ms.append(e0.elapsed_time(e2))
    barrier()
    return sum(step_ms) / len(step_ms), sum(k1_ms) / len(k1_ms)


def time_overlapped(ov, R, steps, warmup, barrier, copies=3):
    """`steps` batches through dist.OverlappedReplay, timed as ONE bracket on the device: the exchange of
    batch k runs on the side stream under the keys kernel of batch k+1.  Inputs rotate through `copies`
    copies (each ~ the size of L2) and outputs through the lanes' column sets, so no step finds its lines
    in L2 and no flush sits inside the bracket.  Returns (ms per step, mean k_scan_keys ms)."""
    import torch
    ins = [(R.d_bytes, R.d_offs)] + [(R.d_bytes.clone(), R.d_offs.clone()) for _ in range(copies - 1)]
    colsets = [R.cols] + [ov.engines[l].device_columns(R.n, R.total) for l in range(1, len(ov.engines))]
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    pairs = []

    def go(k, timed):
        b, o = ins[k % copies]
        pe = (ev(), ev(), ev()) if timed else None
        ov.submit(b, o, colsets[ov.k % len(colsets)], R.n, NOW, R.base_index, scan_events=pe)
        if timed:
            pairs.append(pe)

    barrier()
    for k in range(warmup):
        go(k, False)
    ov.drain()
    barrier()
    l0 = ov.launch_count
    start, end = ev(), ev()
    start.record(ov.main)
    h0 = time.perf_counter()
    for k in range(steps):
        go(warmup + k, True)
    time_overlapped.host_submit_ms = (time.perf_counter() - h0) * 1e3 / steps      # host time to queue one step
    for d in ov.done:
        ov.main.wait_event(d)
    end.record(ov.main)
    ov.drain()
    barrier()
    time_overlapped.launches = ov.launch_count - l0
    k1 = sum(p[0].elapsed_time(p[1]) for p in pairs) / len(pairs)
    # where a step of the main stream goes: keys kernel, join + upsert kernels, and the wait before the next keys
    # kernel starts (lane reuse behind its exchange, exchange kernels taking the SMs at the boundary)
    time_overlapped.main_stream_ms = {
        "k_scan_keys": k1, "join_route_dedup": sum(p[1].elapsed_time(p[2]) for p in pairs) / len(pairs),
        "between_steps": (sum(pairs[k][2].elapsed_time(pairs[k + 1][0]) for k in range(len(pairs) - 1)) / (len(pairs) - 1)
                          if len(pairs) > 1 else 0.0)}
    return start.elapsed_time(end) / steps, k1


def exchange_parity(eng, dist, rank, world, ex, shape: dict, n_per: int = 20_000):
    """The multi-GPU data plane against the oracle: one GLOBAL generator call sliced by rank, replayed and
    exchanged; msg_id / status / verdict columns gathered to rank 0 and compared with
    oracle.replay_oracle.dedup over hashlib's SHA-1 of the same bodies.  Returns (ok, records received by
    every rank) on rank 0, (None, None) elsewhere; exits non-zero on a mismatch."""
    import hashlib

    import numpy as np
    import torch

    from smsgate_b200 import synth
    from smsgate_b200.batch import Batch
    backup = synth.generate(world * n_per, seed=0x5EED + 77, **shape)
    lo, hi = rank * n_per, (rank + 1) * n_per
    b = Batch.from_strings(backup.bodies[lo:hi], backup.senders[lo:hi], backup.date_ms[lo:hi])
    eng.load_cache(backup.cache)
    total = int(b.body_offsets[-1])
    eng.reserve(n_per, total)
    d_bytes = torch.zeros(total + 64, dtype=torch.uint8, device="cuda")
    d_bytes[:total] = torch.from_numpy(b.body_bytes.copy()).cuda()
    d_offs = torch.from_numpy(b.body_offsets.copy()).cuda()
    cols = eng.device_columns(n_per, total + 64)
    torch.cuda.synchronize()
    dist.barrier()
    for _ in range(2):                                   # twice: the arrival counters alternate
        eng.replay_device(d_bytes, d_offs, NOW, cols, base_index=lo)
        ex.run(cols, lo)
    torch.cuda.synchronize()
    received = torch.tensor([ex.last_received()], dtype=torch.int64, device="cuda")
    gathered = {}
    for name in ("msg_id", "status", "survivor", "first_seen", "last_idx"):
        t = cols[name][:n_per].contiguous()
        out = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
        dist.gather(t, out, dst=0)
        if rank == 0:
            gathered[name] = torch.cat(out).cpu().numpy()
    recv_all = [torch.empty_like(received) for _ in range(world)] if rank == 0 else None
    dist.gather(received, recv_all, dst=0)
    if rank != 0:
        return None, None
    from oracle import replay_oracle as O
    ids = [hashlib.sha1(x.encode("utf-8")).digest() for x in backup.bodies]
    if gathered["msg_id"].tobytes() != b"".join(ids):
        raise SystemExit("exchange parity FAILED: msg_id differs from hashlib.sha1")
    surv, first, last, _ = O.dedup(ids, (int(s) == O.ST_UPSERT for s in gathered["status"]))
    if not (np.array_equal(gathered["survivor"] != 0, np.array(surv)) and
            np.array_equal(gathered["first_seen"] != 0, np.array(first)) and
            np.array_equal(gathered["last_idx"], np.array(last, np.int64))):
        raise SystemExit("exchange parity FAILED: global dedup verdicts differ from oracle.replay_oracle.dedup")
    return True, [int(x.item()) for x in recv_all]


def run_ours(args, rank, world, local_rank):
    import torch

    from smsgate_b200.batch import Batch
    from smsgate_b200.replay import ReplayEngine

    # pinned staging buffers next to the GPU: bind before anything is allocated (smsgate_b200/numa.py)
    from smsgate_b200.numa import bind_to_gpu
    bound = bind_to_gpu(local_rank)
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    n = args.msgs
    eng = ReplayEngine(local_rank)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")     # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- multi-GPU only: the exchange against the oracle, before anything is timed
    parity = None
    if world > 1:
        from smsgate_b200.dist import make_global_dedup
        parity = {}
        ex_small = make_global_dedup(eng, dist, 20_000, prefer_p2p=not args.nccl_exchange)
        for name, shape in (("dup50_zipf", dict(dup_frac=0.5, hot_key_zipf=1.2)), ("dup10", dict(dup_frac=0.1))):
            ok, received = exchange_parity(eng, dist, rank, world, ex_small, shape)
            parity[name] = {"ok": ok, "records_received_per_rank": received}
        del ex_small

    # two batches in flight per rank: the exchange of one under the keys kernel of the next (dist.OverlappedReplay)
    from smsgate_b200.dist import OverlappedReplay
    ov = OverlappedReplay(eng, dist if world > 1 else None, n, prefer_p2p=not args.nccl_exchange,
                          reserve_sms=(args.reserve_sms if world > 1 else 0), lanes=args.lanes,
                          exchange_first=not args.main_first, prefilter_bits=args.prefilter_bits)
    torch.cuda.synchronize()
    torch.cuda.set_stream(ov.main)                       # events, flushes and barriers below follow the engines' stream
    exchange = ov.exchanges[0]

    def allmax(vals):
        t = torch.tensor(vals, dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(x) for x in t.cpu()]

    def allgather_int(v):
        t = torch.tensor([v], dtype=torch.int64, device="cuda")
        if dist is None:
            return [int(v)]
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [int(x.item()) for x in out]

    # ---- the other configurations of BASELINE.json that this N can run: device-resident step only
    extra = {}
    extra_names = [w for w in (("dup10", "skew") if world > 1 else ("mixed",)) if w != args.workload]
    if args.no_extra:
        extra_names = []
    for name in extra_names:
        R = Resident(eng, make_backup(name, n, rank, world), rank, ov)
        sync_step, _ = time_resident(eng, R, exchange, 3, 3, flush, barrier)
        step, k1 = time_overlapped(ov, R, max(3, args.steps // 2), 3, barrier)
        step, k1, sync_step = allmax([step, k1, sync_step])
        upserts = int((R.cols["status"] == 0).sum().item())
        survivors = int(R.cols["survivor"].sum().item())
        firsts = int(R.cols["first_seen"].sum().item())
        info = {"workload": WORKLOADS[name].format(n=n), "ms_per_step": step, "k_scan_keys_ms": k1,
                "one_batch_at_a_time_ms": sync_step,
                "value": world * n / (step * 1e-3), "unit": UNIT, "mean_body_bytes": round(R.total / R.n, 1),
                "body_mb_per_s_per_gpu": R.total / (step * 1e-3) / 1e6,
                "upserts_per_rank": allgather_int(upserts), "survivors_per_rank": allgather_int(survivors)}
        if exchange is not None:
            info["records_received_per_rank"] = allgather_int(exchange.last_received())
            # every distinct upserted key has exactly one first and one last occurrence in the whole job
            info["survivors_equal_firsts"] = sum(allgather_int(firsts)) == sum(info["survivors_per_rank"])
            if not info["survivors_equal_firsts"]:
                raise SystemExit(f"{name}: global first/last counts disagree")
        extra[name] = info
        del R
        torch.cuda.empty_cache()

    # ---- the headline workload
    R = Resident(eng, make_backup(args.workload, n, rank, world), rank, ov)
    backup, total, base_index = R.backup, R.total, R.base_index
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    sync_step, sync_k1 = time_resident(eng, R, exchange, max(3, args.steps // 2), 3, flush, barrier)
    step, k1 = time_overlapped(ov, R, args.steps, args.warmup, barrier)
    launches = time_overlapped.launches
    received = allgather_int(exchange.last_received()) if exchange is not None else None

    # ---- end to end through the host-buffer C-ABI call (pinned in, result records out)
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    hb = Batch(R.h_bytes.numpy(), R.h_offs.numpy(), backup.senders, None)
    e2e_ms = []
    for it in range(args.warmup + args.steps):
        flush.zero_()
        torch.cuda.synchronize()
        e0, e1 = ev(), ev()
        e0.record()
        res = eng.replay_records(hb, NOW, base_index=base_index, pinned=True, exchange=exchange)
        e1.record()
        e1.synchronize()
        if it >= args.warmup:
            e2e_ms.append(e0.elapsed_time(e1))
    clocks = sampler.stop() if rank == 0 else None
    h2d = total + 8 * (n + 1)
    d2h = int(res.cols.nbytes)          # result bytes copied back per step (records + verdicts)
    upserts_local = int((res.status == 0).sum())

    step, e2e, k1, sync_step, sync_k1 = allmax([step, sum(e2e_ms) / len(e2e_ms), k1, sync_step, sync_k1])   # max over ranks
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    mean_len = total / n
    k1_bytes = total + n * (8 + 20 + 32 + 1 + 4 + 4)          # algorithmic bytes of k_scan_keys
    peak, peak_src = measured_peak_hbm()
    achieved = k1_bytes / (k1 * 1e-3) / 1e9
    traffic = None
    prof = {}
    try:
        with open(os.path.join(ROOT, "profiles", "k_scan_keys_traffic.json")) as f:
            prof = json.load(f)
            traffic = prof.get("dram_bytes_per_launch")
    except Exception:
        pass
    cpu_cores = 1
    cpu_val = cpu_n = None
    if world == 1:                     # the CPU baseline is reported at N=1 only
        cpu_val, cpu_ms, cpu_n = cpu_arm(backup, args.cpu_sample, cpu_cores, 1, 0)
    line = {
        "metric": METRIC, "value": world * n / (step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8/u32/int64", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload].format(n=n), "msgs_per_gpu": n, "mean_body_bytes": round(mean_len, 1),
                   "cache_entries": len(backup.cache),
                   "l2": "inputs larger than L2: 3 copies of the batch (~124 MB each) and 2 output sets rotate, so no "
                         "step finds its lines in the 126 MB L2; no flush inside the timed bracket (the one-batch and "
                         "e2e figures flush a 512 MiB buffer between steps instead)",
                   "timing": "one CUDA-event bracket around all K steps on the engines' stream, max over ranks; "
                             "batch k's dedup exchange overlaps batch k+1's keys kernel (two lanes)",
                   "scan_sms_reserved_for_exchange": (args.reserve_sms if world > 1 else 0), "lanes": args.lanes, "prefilter_bits_per_key": args.prefilter_bits, "exchange_stream_priority": ("low" if args.main_first else "high"),
                   "host_cpus_bound": (len(bound) if bound else None),
                   "parallelism": f"shard-by-index x{world}" + (f", {type(exchange).__name__} exchange by msg_id prefix" if world > 1 else "")},
        "clocks": clocks,
        "e2e": {"value": world * n / (e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e},
        "gpu_launches": int(launches),
        "host_submit_ms_per_step": time_overlapped.host_submit_ms,
        "main_stream_ms_per_step": time_overlapped.main_stream_ms,
        "roofline": {"bound": "hbm", "kernel": "k_scan_keys", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "kernel_ms": k1, "kernel_share_of_step": k1 / step,
                     "note": "SHA-1 + SHA-256 rounds make this kernel integer-ALU bound, not HBM bound (DESIGN.md); "
                             "alu_pipe_busy_pct / fma_pipe_busy_pct / issue_slots_busy_pct are the ncu figures of "
                             "the same kernel (profiles/, one capture per change), the ceiling it actually runs against",
                     "alu_pipe_busy_pct": prof.get("alu_pipe_busy_pct"),
                     "fma_pipe_busy_pct": prof.get("fma_pipe_busy_pct"),
                     "issue_slots_busy_pct": prof.get("issue_slots_busy_pct"),
                     "path_algorithmic_gbs": n * (mean_len + 229) / (step * 1e-3) / 1e9},
        "cpu_baseline": ({"value": cpu_val, "unit": UNIT, "cores": cpu_cores, "kind": "port",
                          "sample": f"{cpu_n} messages of the same workload, bodies in memory, oracle port with a "
                                    f"diskcache-backed response cache, single process.  Port vs the reference's own functions on "
                                    f"configs[0] in the build container (profiles/r2_cpu_units_reference_vs_port.txt): same "
                                    f"upsert set, the reference runs this unit at ~11.0 k msgs/s, the port (its lean form: "
                                    f"the reference's order of work and nothing the reference does not compute) at ~11.9 k"}
                         if world == 1 else None),
        "upserts_per_gpu": upserts_local,
        "one_batch_at_a_time": {"ms_per_step": sync_step, "k_scan_keys_ms": sync_k1, "value": world * n / (sync_step * 1e-3),
                                "note": "per-step CUDA events, L2 flushed between steps, nothing overlapped: the latency "
                                        "of a single batch (round 1's `value` was measured this way)"},
        "configs": extra,
    }
    if world > 1:
        line["exchange_parity"] = "ok" if all(v["ok"] for v in parity.values()) else "FAILED"
        line["exchange_parity_cases"] = parity
        line["exchange_records_received_per_rank"] = received
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--msgs", type=int, default=N_PER_GPU, help="messages per GPU (default: configs[1])")
    ap.add_argument("--cpu-sample", type=int, default=60_000)
    ap.add_argument("--nccl-exchange", action="store_true", help="use the NCCL all-to-all exchange instead of peer memory")
    ap.add_argument("--workload", default="hit100", choices=sorted(WORKLOADS), help="headline workload (default: configs[1])")
    ap.add_argument("--reserve-sms", type=int, default=0, help="SMs the keys kernel leaves to the exchange stream (N>1)")
    ap.add_argument("--lanes", type=int, default=3, help="batches in flight per rank (dist.OverlappedReplay)")
    ap.add_argument("--main-first", action="store_true", help="give the keys / join stream the higher priority (default: the exchange stream)")
    ap.add_argument("--prefilter-bits", type=int, default=16, help="bits per key of the whole job in the exchange's bitmaps")
    ap.add_argument("--no-extra", action="store_true", help="skip the `configs` object (the other BASELINE.json shapes)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
