"""The upsert slot protocol of `dedup_insert_one` (smsgate_b200/csrc/smsgate_b200.cu) as a step machine, run under
random interleavings on the CPU.  The kernel decides from values it read a moment ago whether an atomicMin /
atomicMax is still needed; this model makes every memory access of a message one step, lets a scheduler pick which
message moves next, and checks that whatever the order, every key ends with {first, last} = {lowest, highest index
carrying it} and every message knows its key's slot - the reference's upsert-by-msg_id semantics
(libs/pocketbase.py:168-200: first occurrence inserts, the last one's payload survives).

It is a model of the protocol, not the kernel: the kernel itself is checked against the oracle in
tests/test_gpu_parity.py and tests/test_zz_no_alloc.py on the GPU."""
import random

EMPTY = None


def _message(i, key, home, tab, keys, slot_of):
    """One message's walk over the table; yields after every access to shared memory."""
    cap = len(tab)
    pos = home
    while True:
        cur = tab[pos]                                   # the 8-byte load of {first, last}
        yield
        if cur is EMPTY:
            cur_now = tab[pos]                           # 64-bit CAS(slot, EMPTY, {i, i})
            if cur_now is EMPTY:
                tab[pos] = (i, i)
                yield
                break
            cur = cur_now                                # claimed in between: the CAS hands back what is there
            yield
        first_then, last_then = cur
        other = keys[first_then]                         # the key of the message the slot points at
        yield
        if other == key:
            if i < first_then:                           # atomicMin(first)
                f, l = tab[pos]
                tab[pos] = (min(f, i), l)
                yield
            if i > last_then:                            # atomicMax(last)
                f, l = tab[pos]
                tab[pos] = (f, max(l, i))
                yield
            break
        pos = (pos + 1) % cap
    slot_of[i] = pos


def _run(keys, homes, cap, rng):
    tab = [EMPTY] * cap
    slot_of = [None] * len(keys)
    live = [_message(i, k, homes[k], tab, keys, slot_of) for i, k in enumerate(keys)]
    while live:
        g = rng.choice(live)
        try:
            next(g)
        except StopIteration:
            live.remove(g)
    return tab, slot_of


def test_any_interleaving_gives_lowest_and_highest_index_per_key():
    rng = random.Random(0x5EED)
    for trial in range(3000):
        n_keys = rng.randint(1, 4)
        n = rng.randint(1, 9)
        cap = 8
        keys = [rng.randrange(n_keys) for _ in range(n)]
        homes = {k: rng.randrange(cap) if rng.random() < 0.5 else 3 for k in range(n_keys)}   # forced collisions
        tab, slot_of = _run(keys, homes, cap, rng)
        want = {}
        for i, k in enumerate(keys):
            f, l = want.get(k, (i, i))
            want[k] = (min(f, i), max(l, i))
        seen = {}
        for i, k in enumerate(keys):
            assert slot_of[i] is not None
            assert tab[slot_of[i]] == want[k], (trial, keys, homes, tab, slot_of)
            assert seen.setdefault(k, slot_of[i]) == slot_of[i]          # one slot per key
        assert sum(s is not EMPTY for s in tab) == len(want)


def test_sequential_order_is_one_of_them():
    keys = [0, 1, 0, 2, 1, 0]

    class First:
        @staticmethod
        def choice(seq):
            return seq[0]
    tab, slot_of = _run(keys, {0: 1, 1: 1, 2: 2}, 4, First)
    assert tab[slot_of[0]] == (0, 5) and tab[slot_of[1]] == (1, 4) and tab[slot_of[3]] == (3, 3)
