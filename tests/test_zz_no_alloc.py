"""smsgate_dedup_static: the allocation-free variant of the upsert-by-msg_id
stage, for callers that are not allowed to touch the CUDA allocator (a loop
whose device memory was handed out at start-up, a process where another
component owns the allocator).  It must give exactly the verdicts of
smsgate_dedup, also on a second call that finds the first call's table still
in the built-in array, and must refuse a batch beyond its fixed capacity.

Collected last on purpose (file name), and self-contained so it also passes
alone in a fresh process."""
import ctypes as C
import datetime as dt

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
NOW = dt.datetime(2026, 1, 1)


def _dedup_both_ways(eng, backup, base_index):
    import torch

    from smsgate_b200 import _lib as L
    from smsgate_b200.batch import Batch
    lib = L.lib()
    batch = Batch.from_strings(backup.bodies, backup.senders, backup.date_ms)
    n = batch.n
    d_bytes = torch.from_numpy(batch.body_bytes.copy()).cuda()
    d_off = torch.from_numpy(batch.body_offsets.copy()).cuda()
    cols = eng.device_columns(n, d_bytes.numel())
    eng.reserve(n, d_bytes.numel())
    eng.scan_keys(d_bytes, d_off, cols)
    eng.join_route(n, NOW, cols)
    eng.dedup(n, base_index, cols)                      # the existing entry point: the expected answer
    torch.cuda.synchronize()
    want = {k: cols[k].cpu().numpy().copy() for k in ("survivor", "first_seen", "last_idx")}
    got = {"survivor": torch.full((n,), 7, dtype=torch.uint8, device="cuda"),
           "first_seen": torch.full((n,), 7, dtype=torch.uint8, device="cuda"),
           "last_idx": torch.full((n,), -7, dtype=torch.int64, device="cuda")}
    err = C.create_string_buffer(256)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.smsgate_dedup_static(0, stream, n, base_index, cols["msg_id"].data_ptr(), cols["status"].data_ptr(),
                                  got["survivor"].data_ptr(), got["first_seen"].data_ptr(),
                                  got["last_idx"].data_ptr(), err, len(err))
    assert rc == 0, err.value
    torch.cuda.synchronize()
    for k in want:
        assert np.array_equal(got[k].cpu().numpy(), want[k]), k       # integer verdicts: bit-exact
    # and directly against the CPU oracle's upsert-by-msg_id over hashlib's SHA-1 of the same bodies,
    # so that the two entry points cannot be wrong together
    import hashlib

    from oracle import replay_oracle as O
    ids = [hashlib.sha1(b.encode("utf-8")).digest() for b in backup.bodies]
    status = cols["status"].cpu().numpy()
    surv, first, last, _ = O.dedup(ids, (int(s) == O.ST_UPSERT for s in status))
    assert np.array_equal(got["survivor"].cpu().numpy() != 0, np.array(surv))
    assert np.array_equal(got["first_seen"].cpu().numpy() != 0, np.array(first))
    assert np.array_equal(got["last_idx"].cpu().numpy(),
                          np.array([(-1 if x < 0 else base_index + x) for x in last], np.int64))
    return int(want["survivor"].sum())


def test_static_dedup_matches_context_dedup_twice_and_refuses_oversize():
    import torch

    from smsgate_b200 import _lib as L
    from smsgate_b200 import synth
    from smsgate_b200.replay import ReplayEngine
    with ReplayEngine(0) as eng:
        # first call: duplicate-heavy batch; second call: a different, smaller batch that would be
        # answered wrongly if the table left by the first call were not reset
        first = synth.generate(30000, seed=41, dup_frac=0.5, hot_key_zipf=1.2)
        eng.load_cache(first.cache)
        s1 = _dedup_both_ways(eng, first, base_index=0)
        second = synth.generate(9000, seed=42, dup_frac=0.3)
        eng.load_cache(second.cache)
        s2 = _dedup_both_ways(eng, second, base_index=5_000_000)
        assert 0 < s2 < s1
        # the same first batch again: the array now holds the second call's table
        eng.load_cache(first.cache)
        assert _dedup_both_ways(eng, first, base_index=0) == s1
    # a batch beyond the built-in capacity is refused, with a message, before anything is launched
    err = C.create_string_buffer(256)
    dummy = torch.zeros(64, dtype=torch.uint8, device="cuda")
    rc = L.lib().smsgate_dedup_static(0, None, L.DEDUP_STATIC_MAX + 1, 0, dummy.data_ptr(), dummy.data_ptr(),
                                      None, None, None, err, len(err))
    assert rc == L.E_LIMIT
    assert b"at most" in err.value and str(L.DEDUP_STATIC_MAX).encode() in err.value
