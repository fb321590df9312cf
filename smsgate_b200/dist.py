"""Global upsert-by-msg_id across the GPUs of one node.

The reference gets global idempotency from a unique key on `msg_id`
(db/models.py:16; services/pb_writer/upsert.py:19-31; libs/pocketbase.py:168-200).
With the backup sharded by index over `world` ranks that becomes exactly one
exchange step: records of locally distinct keys travel to the rank that owns the
key's hash prefix, the owner decides first / last occurrence globally, the
verdicts travel back.  torch.distributed is the plumbing
(NCCL over NVLink on the GPUs; gloo in the CPU tests of this module's logic);
the record building, merging and applying are CUDA kernels behind the C ABI
(smsgate_dedup_export / _merge / _apply).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Sequence

import numpy as np

from . import _lib as L


def owner_of_prefix(prefix32: np.ndarray, world: int) -> np.ndarray:
    """Owner rank of a key whose first four digest bytes, big-endian, are `prefix32`:
    a range partition of the 32-bit prefix space (same formula as key_owner() in the kernels)."""
    return ((prefix32.astype(np.uint64) * np.uint64(world)) >> np.uint64(32)).astype(np.int64)


class ExchangePlan:
    """Variable-size all-to-all of fixed-width rows plus its mirror image."""

    def __init__(self, dist, device=None, group=None):
        self.dist = dist
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.device = device

    def exchange_counts(self, send_counts: Sequence[int]) -> List[int]:
        import torch
        s = torch.tensor(list(send_counts), dtype=torch.int64, device=self.device)
        r = torch.empty_like(s)
        self.dist.all_to_all_single(r, s, group=self.group)
        return [int(x) for x in r.cpu().tolist()]

    def rows(self, send, send_counts: Sequence[int], recv_counts: Sequence[int], out=None):
        """send: tensor [sum(send_counts), width] grouped by destination rank.
        Returns the tensor [sum(recv_counts), width] grouped by source rank."""
        import torch
        n_recv = int(sum(recv_counts))
        if out is None:
            out = torch.empty((n_recv,) + tuple(send.shape[1:]), dtype=send.dtype, device=send.device)
        else:
            out = out[:n_recv]
        self.dist.all_to_all_single(out, send[:int(sum(send_counts))], output_split_sizes=list(recv_counts),
                                    input_split_sizes=list(send_counts), group=self.group)
        return out


class GlobalDedup:
    """One rank's side of the global dedup.  `run(cols, base_index)` expects
    smsgate_dedup to have been called on the local batch and rewrites
    cols['survivor'], cols['first_seen'], cols['last_idx'] with the global verdicts."""

    def __init__(self, engine, dist, n_max: int, group=None):
        import torch
        self.engine = engine
        self.torch = torch
        dev = torch.device("cuda", engine.device)
        self.plan = ExchangePlan(dist, device=dev, group=group)
        self.n_max = max(int(n_max), 1)
        self.records = torch.empty((max(n_max, 1), L.RECORD_BYTES), dtype=torch.uint8, device=dev)
        self.rec_slot = torch.empty(max(n_max, 1), dtype=torch.int32, device=dev)
        self.h_counts = np.zeros(self.plan.world, dtype=np.int64)
        self._recv = None
        self._verdicts = None
        self._back = None
        self.launches = 0
        self.bytes_sent = 0

    def _buf(self, name: str, rows: int, width: int, dtype):
        t = getattr(self, name)
        if t is None or t.shape[0] < rows:
            t = self.torch.empty((max(rows, 1), width), dtype=dtype, device=self.records.device)
            setattr(self, name, t)
        return t

    def run(self, cols: dict, base_index: int) -> None:
        self.run_outputs(self.engine._dev_outputs(cols), cols["status"].numel(), base_index)

    def run_outputs(self, out: "L.Outputs", n: int, base_index: int) -> None:
        """Same, on a raw `smsgate_outputs` of device pointers (e.g. smsgate_device_view)."""
        eng, lib, torch = self.engine, L.lib(), self.torch
        if n > self.n_max:
            raise ValueError(f"shard of {n} messages exceeds the n_max={self.n_max} this exchange was sized for")
        l0 = eng.launch_count
        L.check(eng._ctx, lib.smsgate_dedup_export(eng._ctx, n, int(base_index), self.plan.world, C.byref(out),
                                                   self.records.data_ptr(), self.rec_slot.data_ptr(),
                                                   self.h_counts.ctypes.data_as(C.c_void_p)))
        send_counts = [int(x) for x in self.h_counts]
        n_rec = sum(send_counts)
        recv_counts = self.plan.exchange_counts(send_counts)
        n_recv = sum(recv_counts)
        recv = self.plan.rows(self.records, send_counts, recv_counts,
                              out=self._buf("_recv", n_recv, L.RECORD_BYTES, torch.uint8))
        L.check(eng._ctx, lib.smsgate_reserve_merge(eng._ctx, n_recv))
        verdicts = self._buf("_verdicts", n_recv, 2, torch.int64)
        L.check(eng._ctx, lib.smsgate_dedup_merge(eng._ctx, n_recv, recv.data_ptr(), verdicts.data_ptr()))
        back = self.plan.rows(verdicts, recv_counts, send_counts, out=self._buf("_back", n_rec, 2, torch.int64))
        L.check(eng._ctx, lib.smsgate_dedup_apply(eng._ctx, n, int(base_index), n_rec, self.rec_slot.data_ptr(),
                                                  back.data_ptr(), C.byref(out)))
        self.launches += 0          # the context counts its own launches
        self.bytes_sent += n_rec * L.RECORD_BYTES + n_recv * 16
        self.last_counts = (send_counts, recv_counts)
        self._last_received = n_recv
        self.last_launches = eng.launch_count - l0


    def last_received(self) -> int:
        """Records this rank received as owner during the last run."""
        return int(getattr(self, "_last_received", 0))


class GlobalDedupP2P:
    """The same global dedup with the exchange fused into the kernels: records and
    verdicts are stored straight into the peers' memory over NVLink (buffers mapped
    with torch's symmetric memory), with four cross-GPU barriers (two for the
    bitmap prefilter, two for the records and their verdicts) and no host
    synchronisation per batch.  Every rank must use the same `n_max`."""

    def __init__(self, engine, dist, n_max: int, group=None, prefilter: bool = True, prefilter_bits: int = 16):
        import torch
        import torch.distributed._symmetric_memory as symm_mem
        self.engine = engine
        self.torch = torch
        group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        dev = torch.device("cuda", engine.device)
        n_max = max(int(n_max), 1)
        self.capacity = self.world * n_max
        cap_slots = 16
        while cap_slots < 2 * max(n_max, 8):
            cap_slots <<= 1
        # one symmetric allocation per buffer kind; rendezvous gives every rank's mapping
        self._inbox = symm_mem.empty(self.capacity * L.P2P_RECORD_BYTES, dtype=torch.uint8, device=dev)
        self._count = symm_mem.empty(16, dtype=torch.int32, device=dev)
        self._gfl = symm_mem.empty(2 * cap_slots, dtype=torch.int64, device=dev)
        self._count.zero_()
        self._flags = symm_mem.empty(L.MAX_WORLD, dtype=torch.int32, device=dev)     # smsgate_p2p_barrier arrival flags
        self._flags.zero_()
        self.ovr_words = cap_slots // 32
        self._ovr = symm_mem.empty(2 * self.ovr_words, dtype=torch.int32, device=dev)   # override bits, one half per parity
        self._ovr.zero_()
        bufs = [self._inbox, self._count, self._gfl]
        # prefilter bitmaps: `prefilter_bits` bits per key of the whole job (a false positive only costs one record;
        # the bitmaps cross NVLink twice per batch, so fewer bits trade records for bitmap traffic)
        self.bitmap_words = 0
        if prefilter and self.world > 1:
            words = 64
            while words * 32 < prefilter_bits * n_max * self.world and words < (1 << 26):
                words <<= 1
            self.bitmap_words = words
            self._stage = symm_mem.empty(self.world * (words // self.world + 12), dtype=torch.int32, device=dev)
            self._collide = symm_mem.empty(words, dtype=torch.int32, device=dev)
            bufs += [self._stage, self._collide]
        bufs.append(self._ovr)
        bufs.append(self._flags)
        torch.cuda.synchronize()
        self._h = [symm_mem.rendezvous(t, group) for t in bufs]
        self.desc = L.P2P()
        self.desc.world, self.desc.rank, self.desc.capacity = self.world, self.rank, self.capacity
        self.desc.bitmap_words = self.bitmap_words
        self.desc.gfl_slots = cap_slots
        self.desc.ovr_words = cap_slots // 32
        self.n_max = n_max
        for w in range(self.world):
            self.desc.inbox[w] = self._h[0].buffer_ptrs[w]
            self.desc.in_count[w] = self._h[1].buffer_ptrs[w]
            self.desc.gfl[w] = self._h[2].buffer_ptrs[w]
            self.desc.flags[w] = self._h[-1].buffer_ptrs[w]
            self.desc.ovr[w] = self._h[-2].buffer_ptrs[w]
            if self.bitmap_words:
                self.desc.stage[w] = self._h[3].buffer_ptrs[w]
                self.desc.collide[w] = self._h[4].buffer_ptrs[w]
        L.check(engine._ctx, L.lib().smsgate_p2p_reserve(engine._ctx, self.capacity))
        if self.bitmap_words:
            L.check(engine._ctx, L.lib().smsgate_p2p_reserve_filter(engine._ctx, self.bitmap_words))
            # the join kernel marks the keys in the bitmap while it inserts them into the upsert table
            L.check(engine._ctx, L.lib().smsgate_p2p_fuse_filter(engine._ctx, self.bitmap_words))
        self._h[1].barrier(channel=0)          # every rank's counters and flags are zero before anyone scatters
        self.step = 0
        self.epoch = 0
        self.launches = 0
        self.bytes_sent = 0

    def run(self, cols: dict, base_index: int) -> None:
        self.run_outputs(self.engine._dev_outputs(cols), cols["status"].numel(), base_index)

    def run_outputs(self, out: "L.Outputs", n: int, base_index: int) -> None:
        eng, lib = self.engine, L.lib()
        if n > self.n_max:
            raise ValueError(f"shard of {n} messages exceeds the n_max={self.n_max} this exchange was sized for")
        parity = self.step & 1
        # filter_push, barrier, filter_reduce, barrier, scatter, barrier, merge, barrier, resolve: queued by one C
        # call (smsgate_p2p_exchange); the barriers are in-stream kernels (smsgate_p2p_barrier), no host round trip
        used = lib.smsgate_p2p_exchange(eng._ctx, n, int(base_index), parity, self.epoch, C.byref(out), C.byref(self.desc))
        if used < 0:
            L.check(eng._ctx, used)
        self.epoch += used
        self.step += 1

    def last_received(self) -> int:
        """Records that arrived in this rank's inbox during the last run (synchronises)."""
        if self.step == 0:
            return 0
        self.torch.cuda.synchronize()
        return int(self._count[(self.step - 1) & 1].item())

    def overflowed(self) -> bool:
        """True if an inbox ever overflowed (some rank exceeded n_max).  Synchronises."""
        flag = C.c_int(0)
        out = L.Outputs()
        L.check(self.engine._ctx, L.lib().smsgate_p2p_resolve(self.engine._ctx, 0, 0, 0, C.byref(out), C.byref(self.desc),
                                                              C.byref(flag)))
        if flag.value == 2:
            raise RuntimeError("a cross-GPU barrier of the exchange timed out (a rank fell behind by more than ~10 s)")
        return bool(flag.value)


def make_global_dedup(engine, dist, n_max: int, prefer_p2p: bool = True, prefilter_bits: int = 16):
    """GlobalDedupP2P when peer mappings are available, else the NCCL all-to-all path."""
    if prefer_p2p:
        try:
            return GlobalDedupP2P(engine, dist, n_max, prefilter_bits=prefilter_bits)
        except Exception as e:       # no symmetric memory in this build / topology: fall back to NCCL
            import warnings
            warnings.warn(f"peer-memory exchange unavailable ({e!r}); using the NCCL all-to-all path")
    return GlobalDedup(engine, dist, n_max)


class OverlappedReplay:
    """Device-resident replay of a STREAM of batches on one rank with the global dedup exchange of batch k
    running under the keys kernel of batch k+1.

    The exchange is a chain of small latency-bound kernels and cross-GPU barriers that depends on the whole
    batch's keys, so inside one batch nothing can hide it; across batches everything can.  A few contexts
    ("lanes", three by default) take turns: each owns its dedup scratch, its output columns and its peer-memory buffers and
    shares the one response cache (smsgate_cache_share).  Lane work is queued on `main` (smsgate_scan_keys,
    smsgate_join_route_dedup), its exchange on `side` behind an event; a
    lane is reused only after its previous exchange has finished.  The exchange kernels and the in-stream barrier
    are built to fit on an SM next to the keys kernel's CTA (128 threads, 32 registers), so they run on every SM
    while it has them all; `reserve_sms` > 0 additionally keeps SMs free of it (smsgate_set_scan_ctas).

    The reference has no counterpart: its upsert is one synchronous HTTP/SQL round trip per message
    (libs/pocketbase.py:168-200, services/pb_writer/upsert.py:19-31)."""

    def __init__(self, engine, dist, n_max: int, prefer_p2p: bool = True, reserve_sms: int = 0, lanes: int = 3,
                 exchange_first: bool = True, prefilter_bits: int = 16):
        import torch

        from .replay import ReplayEngine
        self.torch = torch
        dev = engine.device
        lo, hi = torch.cuda.Stream.priority_range()          # (lowest priority, highest priority)
        # `exchange_first`: the exchange stream has the HIGHER priority.  Its kernels are tiny next to the keys kernel
        # (tens of microseconds each at full speed); while that kernel holds every SM they trickle through the one
        # co-resident block per SM, and at every kernel boundary of the main stream they take the freed SMs first
        # and are done in microseconds - instead of starving behind 0.16 ms of join blocks and making the lane wait.
        self.main = torch.cuda.Stream(device=dev, priority=lo if exchange_first else hi)
        self.side = torch.cuda.Stream(device=dev, priority=hi if exchange_first else lo)
        self.engines = [engine]
        for _ in range(lanes - 1):
            e = ReplayEngine(dev)
            e.share_cache_of(engine)
            self.engines.append(e)
        sm = torch.cuda.get_device_properties(dev).multi_processor_count
        self.exchanges = []
        for e in self.engines:
            e.set_scan_ctas(max(1, sm - reserve_sms) if reserve_sms > 0 else 0)
            self.exchanges.append(make_global_dedup(e, dist, n_max, prefer_p2p=prefer_p2p, prefilter_bits=prefilter_bits)
                                  if dist is not None else None)
        torch.cuda.synchronize(dev)
        for e in self.engines:
            e.set_stream(self.main)
        self.joined = [torch.cuda.Event() for _ in self.engines]
        self.done = [torch.cuda.Event() for _ in self.engines]
        self.k = 0

    def refresh_cache(self) -> None:
        """After the first engine reloaded its cache: the other lanes follow."""
        for e in self.engines[1:]:
            e.share_cache_of(self.engines[0])

    def reserve(self, n: int, total: int) -> None:
        for e in self.engines:
            e.reserve(n, total)

    def submit(self, d_bytes, d_offs, cols, n: int, now, base_index: int, scan_events=None) -> int:
        """Queue one batch on the next lane (`cols` must be a column set used by that lane only).
        Returns the lane."""
        torch = self.torch
        lane = self.k % len(self.engines)
        eng, ex = self.engines[lane], self.exchanges[lane]
        with torch.cuda.stream(self.main):
            self.main.wait_event(self.done[lane])            # the lane's previous exchange read these columns
            if scan_events is not None:
                scan_events[0].record(self.main)
            eng.scan_keys(d_bytes, d_offs, cols)
            if scan_events is not None:
                scan_events[1].record(self.main)
            eng.join_route_dedup(n, now, base_index, cols)
            if scan_events is not None and len(scan_events) > 2:
                scan_events[2].record(self.main)
            self.joined[lane].record(self.main)
        if ex is not None:
            with torch.cuda.stream(self.side):
                self.side.wait_event(self.joined[lane])
                eng.set_stream(self.side)
                try:
                    ex.run(cols, base_index)
                finally:
                    eng.set_stream(self.main)
                self.done[lane].record(self.side)
        else:
            self.done[lane].record(self.main)
        self.k += 1
        return lane

    def drain(self) -> None:
        self.main.synchronize()
        self.side.synchronize()

    @property
    def launch_count(self) -> int:
        return sum(e.launch_count for e in self.engines)

    def close(self) -> None:
        self.drain()
        for e in self.engines[1:]:
            e.close()
