"""Build container only: writes tests/golden/dlq.json — for every test payload, what the reference's
dead-letter re-parse does with it before the replay proper.  The decisions are made by the reference's own
code where it is callable: `RawSMS` is libs.models.RawSMS (imported from /root/reference through
oracle/ref_runner.py); the surrounding statements (dlq_worker.py:48-66, worker.py:88-99) are inline in
async NATS handlers and are followed by oracle/dlq_oracle.py:handle, which is run here WITH the reference's
class.  Usage: python -m oracle.gen_golden_dlq"""
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import dlq_oracle as DO          # noqa: E402
from oracle.ref_runner import Reference      # noqa: E402
from smsgate_b200 import synth               # noqa: E402
from tests._dlq_cases import payloads        # noqa: E402


def main():
    ref = Reference(tempfile.mkdtemp(prefix="ref_dlq_"))
    backup = synth.generate(200, seed=72, mixed_locale=True)
    cases = []
    for p in payloads(backup.bodies, backup.senders):
        outcome, raw = DO.handle(p, raw_cls=ref.models.RawSMS)
        own, own_raw = DO.handle(p)
        assert own == outcome and (raw is None) == (own_raw is None)
        if raw is not None:
            assert raw.model_dump() == own_raw.model_dump()
        cases.append({"payload_hex": p.hex(), "outcome": outcome, "raw": raw.model_dump() if raw is not None else None})
    ref.close()
    with open(os.path.join(ROOT, "tests", "golden", "dlq.json"), "w") as f:
        json.dump(cases, f, indent=0)
    print(len(cases), "cases", {o: sum(c["outcome"] == o for c in cases) for o in {c["outcome"] for c in cases}})


if __name__ == "__main__":
    main()
