"""Host-side logic of the global dedup exchange (smsgate_b200/dist.py) on CPU:
two gloo ranks, the three kernels replaced by numpy stand-ins that use the
kernels' record layout and owner formula, the result checked against the
oracle's upsert-by-msg_id over the concatenated backup."""
import hashlib
import os
import socket

import numpy as np
import pytest

from oracle import replay_oracle as O
from smsgate_b200 import _lib as L
from smsgate_b200 import synth
from smsgate_b200.dist import ExchangePlan, owner_of_prefix


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _export(ids, eligible, base, world):
    """numpy stand-in for smsgate_dedup (local) + smsgate_dedup_export."""
    first, last = {}, {}
    for i, (k, ok) in enumerate(zip(ids, eligible)):
        if ok:
            first.setdefault(k, i)
            last[k] = i
    keys = list(first)
    if keys:
        prefix = np.array([int.from_bytes(k[:4], "big") for k in keys], dtype=np.uint64)
        owner = owner_of_prefix(prefix, world)
    else:
        owner = np.zeros(0, np.int64)
    order = np.argsort(owner, kind="stable")
    recs = np.zeros((len(keys), L.RECORD_BYTES), np.uint8)
    for r, j in enumerate(order):
        k = keys[j]
        f, l = base + first[k], base + last[k]
        recs[r, :20] = np.frombuffer(k, np.uint8)
        recs[r, 20:24] = np.frombuffer(int(f & 0xFFFFFFFF).to_bytes(4, "little"), np.uint8)
        recs[r, 24:28] = np.frombuffer(int(l & 0xFFFFFFFF).to_bytes(4, "little"), np.uint8)
        recs[r, 28:30] = np.frombuffer(int(f >> 32).to_bytes(2, "little"), np.uint8)
        recs[r, 30:32] = np.frombuffer(int(l >> 32).to_bytes(2, "little"), np.uint8)
    counts = np.bincount(owner, minlength=world).tolist()
    return recs, [keys[j] for j in order], counts


def _merge(recs):
    """numpy stand-in for smsgate_dedup_merge."""
    first, last = {}, {}
    rows = []
    for r in recs:
        k = r[:20].tobytes()
        f = int.from_bytes(r[20:24].tobytes(), "little") | int.from_bytes(r[28:30].tobytes(), "little") << 32
        l = int.from_bytes(r[24:28].tobytes(), "little") | int.from_bytes(r[30:32].tobytes(), "little") << 32
        first[k] = min(first.get(k, f), f)
        last[k] = max(last.get(k, l), l)
        rows.append(k)
    return np.array([[first[k], last[k]] for k in rows], dtype=np.int64).reshape(len(rows), 2)


def _worker(rank, world, port, n_per, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        backup = synth.generate(n_per * world, seed=77, dup_frac=0.6, hot_key_zipf=1.1)
        lo, hi = rank * n_per, (rank + 1) * n_per
        bodies = backup.bodies[lo:hi]
        ids = [hashlib.sha1(b.encode()).digest() for b in bodies]
        eligible = [not O.worker_skip(b) for b in bodies]          # any per-message predicate will do here
        base = (1 << 33) + lo                                      # exercises the 48-bit index split
        recs, keys, send_counts = _export(ids, eligible, base, world)
        plan = ExchangePlan(dist)
        recv_counts = plan.exchange_counts(send_counts)
        got = plan.rows(torch.from_numpy(recs), send_counts, recv_counts).numpy()
        # every record that arrived here is owned by this rank
        if len(got):
            pre = np.array([int.from_bytes(r[:4].tobytes(), "big") for r in got], dtype=np.uint64)
            assert np.all(owner_of_prefix(pre, world) == rank)
        verdicts = _merge(got)
        back = plan.rows(torch.from_numpy(verdicts), recv_counts, send_counts).numpy()
        assert back.shape == (len(keys), 2)
        glob = {k: (int(back[r, 0]), int(back[r, 1])) for r, k in enumerate(keys)}
        survivor = [ok and glob[k][1] == base + i for i, (k, ok) in enumerate(zip(ids, eligible))]
        first_seen = [ok and glob[k][0] == base + i for i, (k, ok) in enumerate(zip(ids, eligible))]
        last_idx = [glob[k][1] if ok else -1 for k, ok in zip(ids, eligible)]
        q.put((rank, survivor, first_seen, last_idx, send_counts, recv_counts))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_exchange_matches_global_upsert():
    import torch.multiprocessing as mp
    world, n_per = 2, 1500
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_per, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted(q.get(timeout=240) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    backup = synth.generate(n_per * world, seed=77, dup_frac=0.6, hot_key_zipf=1.1)
    ids = [hashlib.sha1(b.encode()).digest() for b in backup.bodies]
    eligible = [not O.worker_skip(b) for b in backup.bodies]
    survivor, first_seen, last_idx, _ = O.dedup(ids, eligible)
    base = 1 << 33
    got_s = sum((r[1] for r in results), [])
    got_f = sum((r[2] for r in results), [])
    got_l = sum((r[3] for r in results), [])
    assert got_s == survivor
    assert got_f == first_seen
    assert got_l == [(-1 if x < 0 else base + x) for x in last_idx]
    # the mirrored exchange really is the mirror: what rank a sent to b is what b received from a
    assert results[0][4][1] == results[1][5][0] and results[1][4][0] == results[0][5][1]


def test_owner_formula_is_a_range_partition():
    pre = np.array([0, 1, 0x1FFFFFFF, 0x20000000, 0x7FFFFFFF, 0x80000000, 0xFFFFFFFF], dtype=np.uint64)
    assert owner_of_prefix(pre, 8).tolist() == [0, 0, 0, 1, 3, 4, 7]
    assert owner_of_prefix(pre, 1).tolist() == [0] * 7
    assert owner_of_prefix(pre, 3).tolist() == [0, 0, 0, 0, 1, 1, 2]
