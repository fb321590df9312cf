"""Host half of the DLQ re-parse (which payloads are replayed, with which ids) against the oracle
restatement of dlq_worker.py / worker.py and against the golden outcomes produced with the reference's
own RawSMS class (oracle/gen_golden_dlq.py)."""
import json

from oracle import dlq_oracle as DO
from smsgate_b200 import dlq, synth
from tests._dlq_cases import payloads


def test_unwrap_matches_oracle():
    backup = synth.generate(400, seed=71, mixed_locale=True)
    for p in payloads(backup.bodies, backup.senders):
        want, wraw = DO.handle(p)
        got, raw, err = dlq.unwrap(p)
        assert got == want, p[:80]
        if wraw is not None:
            assert raw.model_dump() == wraw.model_dump()
        assert (err is not None) == (got == dlq.INVALID_RAW)


def test_unwrap_matches_reference_vectors(golden):
    for case in golden("dlq.json"):
        p = bytes.fromhex(case["payload_hex"])
        got, raw, _ = dlq.unwrap(p)
        assert got == case["outcome"], p[:80]
        assert (raw.model_dump() if raw is not None else None) == case["raw"]
        assert DO.handle(p)[0] == case["outcome"]


def test_key_bytes():
    a = "f0f5b4237706abd1f15a59e2f1b3ff80c180af31"
    assert dlq.key_bytes(a) == bytes.fromhex(a)
    assert dlq.key_bytes(a.upper()) != dlq.key_bytes(a)
    assert dlq.key_bytes("5508ae317ec9270448afb568667be895") != dlq.key_bytes("5508ae317ec9270448afb568667be896")
    assert len({dlq.key_bytes(s) for s in ("", "a", "dup-1", "dup-2", a, a.upper(), a[:39] + "g")}) == 7
