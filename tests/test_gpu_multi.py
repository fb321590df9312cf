"""Global dedup kernels (export / merge / apply).

1. One GPU playing R ranks: R contexts on cuda:0, each with its own shard; the
   all-to-all is done by slicing tensors in this process.  Checks the three
   kernels and the record layout against the oracle's global upsert.
2. Two real ranks over NCCL when the box has two GPUs (skipped otherwise)."""
import ctypes as C
import datetime as dt
import os
import socket

import numpy as np
import pytest

from oracle import replay_oracle as O
from smsgate_b200 import _lib as L
from smsgate_b200 import synth
from smsgate_b200.batch import Batch

pytestmark = pytest.mark.gpu
NOW = dt.datetime(2026, 1, 1)


def _global_oracle(backup):
    raw = [O.iso_utc_from_ms(m) for m in backup.date_ms]
    return O.replay(backup.bodies, backup.senders, raw, backup.cache, NOW)


@pytest.mark.parametrize("world", [2, 3, 8])
def test_ranks_emulated_on_one_gpu(world):
    import torch

    from smsgate_b200.replay import ReplayEngine
    lib = L.lib()
    n_per = 4000
    backup = synth.generate(n_per * world, seed=300 + world, dup_frac=0.5, hot_key_zipf=1.2, miss_frac=0.05)
    ora = _global_oracle(backup)
    engines, cols, recs, slots, counts = [], [], [], [], []
    base0 = (1 << 34)                                             # indices beyond 32 bits
    for r in range(world):
        eng = ReplayEngine(0)
        eng.load_cache(backup.cache)
        lo, hi = r * n_per, (r + 1) * n_per
        b = Batch.from_strings(backup.bodies[lo:hi], backup.senders[lo:hi], backup.date_ms[lo:hi])
        d_bytes = torch.from_numpy(b.body_bytes.copy()).cuda()
        d_off = torch.from_numpy(b.body_offsets.copy()).cuda()
        c = eng.device_columns(b.n, d_bytes.numel())
        eng.reserve(b.n, d_bytes.numel())
        eng.replay_device(d_bytes, d_off, NOW, c, base_index=base0 + lo)
        rec = torch.empty((n_per, L.RECORD_BYTES), dtype=torch.uint8, device="cuda")
        slot = torch.empty(n_per, dtype=torch.int32, device="cuda")
        h_counts = np.zeros(world, np.int64)
        out = eng._dev_outputs(c)
        L.check(eng._ctx, lib.smsgate_dedup_export(eng._ctx, b.n, base0 + lo, world, C.byref(out), rec.data_ptr(),
                                                   slot.data_ptr(), h_counts.ctypes.data_as(C.c_void_p)))
        engines.append(eng); cols.append(c); recs.append(rec); slots.append(slot); counts.append(h_counts.copy())
    torch.cuda.synchronize()
    # all-to-all by slicing: owner o receives, from every rank r in order, r's bucket o
    starts = [np.concatenate([[0], np.cumsum(c)]) for c in counts]
    verdict_back = [torch.empty((int(counts[r].sum()), 2), dtype=torch.int64, device="cuda") for r in range(world)]
    for o in range(world):
        parts = [recs[r][int(starts[r][o]):int(starts[r][o + 1])] for r in range(world)]
        recv = torch.cat(parts) if parts else torch.empty((0, 32), dtype=torch.uint8, device="cuda")
        n_recv = recv.shape[0]
        eng = engines[o]
        L.check(eng._ctx, lib.smsgate_reserve_merge(eng._ctx, n_recv))
        ver = torch.empty((max(n_recv, 1), 2), dtype=torch.int64, device="cuda")
        L.check(eng._ctx, lib.smsgate_dedup_merge(eng._ctx, n_recv, recv.contiguous().data_ptr(), ver.data_ptr()))
        torch.cuda.synchronize()
        pos = 0
        for r in range(world):
            k = int(counts[r][o])
            verdict_back[r][int(starts[r][o]):int(starts[r][o + 1])] = ver[pos:pos + k]
            pos += k
    got_s, got_f, got_l = [], [], []
    for r in range(world):
        eng = engines[r]
        out = eng._dev_outputs(cols[r])
        n_rec = int(counts[r].sum())
        L.check(eng._ctx, lib.smsgate_dedup_apply(eng._ctx, n_per, base0 + r * n_per, n_rec, slots[r].data_ptr(),
                                                  verdict_back[r].data_ptr(), C.byref(out)))
        torch.cuda.synchronize()
        got_s.append(cols[r]["survivor"].cpu().numpy())
        got_f.append(cols[r]["first_seen"].cpu().numpy())
        got_l.append(cols[r]["last_idx"].cpu().numpy())
        eng.close()
    assert np.array_equal(np.concatenate(got_s) != 0, np.array(ora.survivor))
    assert np.array_equal(np.concatenate(got_f) != 0, np.array(ora.first_seen))
    want_l = np.array([(-1 if x < 0 else base0 + x) for x in ora.last_idx], np.int64)
    assert np.array_equal(np.concatenate(got_l), want_l)


def _nccl_worker(rank, world, port, n_per, q, p2p=False, prefilter=True):
    import torch
    import torch.distributed as dist

    from smsgate_b200.dist import GlobalDedup, GlobalDedupP2P
    from smsgate_b200.replay import ReplayEngine
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        backup = synth.generate(n_per * world, seed=500, dup_frac=0.5, hot_key_zipf=1.2)
        lo, hi = rank * n_per, (rank + 1) * n_per
        b = Batch.from_strings(backup.bodies[lo:hi], backup.senders[lo:hi], backup.date_ms[lo:hi])
        eng = ReplayEngine(rank)
        eng.load_cache(backup.cache)
        d_bytes = torch.from_numpy(b.body_bytes.copy()).cuda()
        d_off = torch.from_numpy(b.body_offsets.copy()).cuda()
        c = eng.device_columns(b.n, d_bytes.numel())
        eng.reserve(b.n, d_bytes.numel())
        ex = GlobalDedupP2P(eng, dist, b.n, prefilter=prefilter) if p2p else GlobalDedup(eng, dist, b.n)
        for _ in range(3 if p2p else 1):                 # several batches: counters alternate, inboxes are reused
            eng.replay_device(d_bytes, d_off, NOW, c, base_index=lo)
            ex.run(c, lo)
        torch.cuda.synchronize()
        if p2p:
            assert not ex.overflowed()
        q.put((rank, c["survivor"].cpu().numpy(), c["first_seen"].cpu().numpy(), c["last_idx"].cpu().numpy()))
        eng.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
@pytest.mark.parametrize("p2p,world,prefilter", [(False, 2, False), (True, 2, True), (True, 2, False), (True, 1, True)])
def test_ranks_over_nccl_or_peer_memory(p2p, world, prefilter):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    n_per = 20000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = [ctx.Process(target=_nccl_worker, args=(r, world, port, n_per, q, p2p, prefilter)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=500) for _ in range(world)), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    backup = synth.generate(n_per * world, seed=500, dup_frac=0.5, hot_key_zipf=1.2)
    ora = _global_oracle(backup)
    assert np.array_equal(np.concatenate([r[1] for r in res]) != 0, np.array(ora.survivor))
    assert np.array_equal(np.concatenate([r[2] for r in res]) != 0, np.array(ora.first_seen))
    assert np.array_equal(np.concatenate([r[3] for r in res]), np.array(ora.last_idx, np.int64))


class _EmulatedPeers:
    """`world` ranks of the peer-memory exchange on ONE GPU: every rank is its own context on cuda:0,
    the "peer" pointers of the smsgate_p2p descriptor are plain device tensors of the other contexts,
    and since all contexts share one stream a cross-GPU barrier is simply the order of the calls.
    Drives k_bits_build/_push/_reduce and k_p2p_scatter/_clear/_insert/_verdict/_resolve exactly as
    dist.GlobalDedupP2P.run_outputs does."""

    def __init__(self, engines, n_max, prefilter):
        import torch
        self.engines, self.world = engines, len(engines)
        world = self.world
        self.capacity = world * n_max
        cap_slots = 16
        while cap_slots < 2 * max(n_max, 8):
            cap_slots <<= 1
        dev = "cuda"
        self.inbox = [torch.zeros(self.capacity * L.P2P_RECORD_BYTES, dtype=torch.uint8, device=dev) for _ in range(world)]
        self.count = [torch.zeros(16, dtype=torch.int32, device=dev) for _ in range(world)]
        self.gfl = [torch.zeros(2 * cap_slots, dtype=torch.int64, device=dev) for _ in range(world)]
        self.ovr = [torch.zeros(2 * (cap_slots // 32), dtype=torch.int32, device=dev) for _ in range(world)]
        self.words = 0
        if prefilter:
            words = 64
            while words * 32 < 16 * n_max * world:
                words <<= 1
            self.words = words
            self.stage = [torch.zeros(world * (words // world + 12), dtype=torch.int32, device=dev) for _ in range(world)]
            self.collide = [torch.zeros(words, dtype=torch.int32, device=dev) for _ in range(world)]
        self.desc = []
        lib = L.lib()
        for r, eng in enumerate(engines):
            d = L.P2P()
            d.world, d.rank, d.capacity, d.bitmap_words, d.gfl_slots = world, r, self.capacity, self.words, cap_slots
            d.ovr_words = cap_slots // 32
            for w in range(world):
                d.inbox[w] = self.inbox[w].data_ptr()
                d.in_count[w] = self.count[w].data_ptr()
                d.gfl[w] = self.gfl[w].data_ptr()
                d.ovr[w] = self.ovr[w].data_ptr()
                if self.words:
                    d.stage[w] = self.stage[w].data_ptr()
                    d.collide[w] = self.collide[w].data_ptr()
            L.check(eng._ctx, lib.smsgate_p2p_reserve(eng._ctx, self.capacity))
            if self.words:
                L.check(eng._ctx, lib.smsgate_p2p_reserve_filter(eng._ctx, self.words))
                if world != 2:           # the join kernel marks the bitmap itself (as dist.GlobalDedupP2P sets it up)
                    L.check(eng._ctx, lib.smsgate_p2p_fuse_filter(eng._ctx, self.words))
            self.desc.append(d)
        self.step = 0

    def run(self, outs, ns, bases):
        lib, parity = L.lib(), self.step & 1
        E, D = self.engines, self.desc
        rng = range(self.world)
        if self.words:
            for r in rng:
                L.check(E[r]._ctx, lib.smsgate_p2p_filter_push(E[r]._ctx, ns[r], C.byref(outs[r]), C.byref(D[r])))
            for r in rng:
                L.check(E[r]._ctx, lib.smsgate_p2p_filter_reduce(E[r]._ctx, C.byref(D[r])))
        for r in rng:
            L.check(E[r]._ctx, lib.smsgate_p2p_scatter(E[r]._ctx, ns[r], bases[r], parity, C.byref(outs[r]), C.byref(D[r])))
        for r in rng:
            L.check(E[r]._ctx, lib.smsgate_p2p_merge(E[r]._ctx, parity, C.byref(D[r])))
        for r in rng:
            L.check(E[r]._ctx, lib.smsgate_p2p_resolve(E[r]._ctx, ns[r], bases[r], parity, C.byref(outs[r]), C.byref(D[r]), None))
        self.step += 1

    def overflowed(self):
        flag = C.c_int(0)
        out = L.Outputs()
        any_over = False
        for r, eng in enumerate(self.engines):
            L.check(eng._ctx, L.lib().smsgate_p2p_resolve(eng._ctx, 0, 0, 0, C.byref(out), C.byref(self.desc[r]), C.byref(flag)))
            any_over |= bool(flag.value)
        return any_over


@pytest.mark.parametrize("world,prefilter,ragged", [(2, True, False), (3, True, True), (8, True, False),
                                                    (2, False, False), (5, False, True)])
def test_peer_memory_exchange_emulated_on_one_gpu(world, prefilter, ragged):
    """The default multi-GPU data plane (bitmap prefilter + peer stores) against the oracle's global
    upsert, several batches in a row (the arrival counters alternate, inboxes and bitmaps are reused),
    with shards of unequal size when `ragged` (one of them empty)."""
    import torch

    from smsgate_b200.replay import ReplayEngine
    n_per = 3000
    sizes = [n_per] * world
    if ragged:
        sizes = [max(0, n_per - 700 * r) for r in range(world)]
        sizes[-1] = 0
    engines = [ReplayEngine(0) for _ in range(world)]
    try:
        peers = _EmulatedPeers(engines, n_per, prefilter)
        base0 = (1 << 33) + 5
        for batch_no, seed in enumerate((700 + world, 900 + world, 1100 + world)):
            backup = synth.generate(sum(sizes), seed=seed, dup_frac=(0.5, 0.1, 0.0)[batch_no],
                                    hot_key_zipf=(1.2, None, None)[batch_no], miss_frac=0.05)
            ora = _global_oracle(backup)
            cols, outs, bases = [], [], []
            lo = 0
            for r, eng in enumerate(engines):
                hi = lo + sizes[r]
                eng.load_cache(backup.cache)
                b = Batch.from_strings(backup.bodies[lo:hi], backup.senders[lo:hi], backup.date_ms[lo:hi])
                d_bytes = torch.zeros(b.body_bytes.size + 64, dtype=torch.uint8, device="cuda")
                d_bytes[:b.body_bytes.size] = torch.from_numpy(b.body_bytes.copy()).cuda()
                d_off = torch.from_numpy(b.body_offsets.copy()).cuda()
                c = eng.device_columns(max(b.n, 1), d_bytes.numel())
                eng.reserve(n_per, d_bytes.numel())
                eng.replay_device(d_bytes, d_off, NOW, c, base_index=base0 + lo)
                cols.append(c); outs.append(eng._dev_outputs(c)); bases.append(base0 + lo)
                lo = hi
            peers.run(outs, sizes, bases)
            torch.cuda.synchronize()
            got_s = np.concatenate([cols[r]["survivor"][:sizes[r]].cpu().numpy() for r in range(world)])
            got_f = np.concatenate([cols[r]["first_seen"][:sizes[r]].cpu().numpy() for r in range(world)])
            got_l = np.concatenate([cols[r]["last_idx"][:sizes[r]].cpu().numpy() for r in range(world)])
            assert np.array_equal(got_s != 0, np.array(ora.survivor)), batch_no
            assert np.array_equal(got_f != 0, np.array(ora.first_seen)), batch_no
            want_l = np.array([(-1 if x < 0 else base0 + x) for x in ora.last_idx], np.int64)
            assert np.array_equal(got_l, want_l), batch_no
        assert not peers.overflowed()
        # a shard beyond what the verdict arrays were sized for is refused, not written out of bounds
        small = L.P2P.from_buffer_copy(peers.desc[0])
        small.gfl_slots = 16
        rc = L.lib().smsgate_p2p_scatter(engines[0]._ctx, sizes[0], 0, 0, C.byref(outs[0]), C.byref(small))
        assert rc == L.E_LIMIT
    finally:
        for eng in engines:
            eng.close()


def test_in_stream_barrier_single_rank_and_timeout():
    """smsgate_p2p_barrier: with one rank it signals and finds its own flag; a rank whose peer never arrives
    gives up after about ten seconds and reports it instead of hanging the GPU."""
    import torch

    from smsgate_b200.replay import ReplayEngine
    with ReplayEngine(0) as eng:
        lib = L.lib()
        peers = _EmulatedPeers([eng], 64, False)
        flags = torch.zeros(L.MAX_WORLD, dtype=torch.int32, device="cuda")
        d = peers.desc[0]
        d.flags[0] = flags.data_ptr()
        for epoch in (1, 2, 3):
            L.check(eng._ctx, lib.smsgate_p2p_barrier(eng._ctx, C.byref(d), epoch))
        torch.cuda.synchronize()
        assert int(flags[0].item()) == 3 and not peers.overflowed()
        # two ranks, the second never calls: rank 0's wait times out
        other = torch.zeros(L.MAX_WORLD, dtype=torch.int32, device="cuda")
        d.world = 2
        d.inbox[1], d.in_count[1], d.gfl[1], d.ovr[1], d.flags[1] = d.inbox[0], d.in_count[0], d.gfl[0], d.ovr[0], other.data_ptr()
        L.check(eng._ctx, lib.smsgate_p2p_barrier(eng._ctx, C.byref(d), 4))
        torch.cuda.synchronize()
        flag = C.c_int(0)
        out = L.Outputs()
        L.check(eng._ctx, lib.smsgate_p2p_resolve(eng._ctx, 0, 0, 0, C.byref(out), C.byref(d), C.byref(flag)))
        assert flag.value == 2 and int(other[0].item()) == 4
