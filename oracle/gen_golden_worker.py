"""Build container only: writes tests/golden/worker_flow.json by RUNNING the reference's own handlers —
`services.parser_worker.worker._process_one` (worker.py:74-189) and
`services.parser_worker.dlq_worker._handle_dlq_msg` (dlq_worker.py:39-78) — unmodified, against a fake NATS
connection that records what each message causes: acks, publishes (subject + payload bytes), the three
Prometheus counters the worker bumps, and any exception that escapes the handler.  Only the transport is
faked (nc.jetstream().publish, msg.ack, ensure_stream) and `datetime.now()` inside worker.py is frozen;
the filter, RawSMS validation, parse_sms_llm against a seeded `.gemini_cache`, the date gate and the payload
formats are the reference's code.  Gemini is unreachable here, so a cache miss surfaces as the reference's
own "exception in parse_sms_llm" branch (worker.py:141-149).
Usage: python -m oracle.gen_golden_worker"""
import asyncio
import datetime as dt
import json
import logging
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.ref_runner import REFERENCE_ROOT, Reference      # noqa: E402
from smsgate_b200 import synth                               # noqa: E402
from smsgate_b200.synth import reference_cache_key           # noqa: E402
from tests._dlq_cases import payloads                        # noqa: E402

NOW = dt.datetime(2026, 1, 1)


class _Meta:
    sequence = 1


class _Msg:
    def __init__(self, data):
        self.data = data
        self.metadata = _Meta()
        self.acks = 0

    async def ack(self):
        self.acks += 1

    def __repr__(self):
        return "<msg>"


class _JS:
    def __init__(self, log):
        self.log = log

    async def publish(self, subject, payload, **kw):
        self.log.append((subject, payload))


class _NC:
    def __init__(self):
        self.log = []

    def jetstream(self):
        return _JS(self.log)


def special_cache(backup):
    """Cache of the synthetic backup with a handful of entries bent into the corners of parse_sms_llm."""
    cache = dict(backup.cache)
    hits = [i for i, b in enumerate(backup.bodies) if reference_cache_key(b) in cache]
    bend = {}

    def mod(k, **kw):
        i = hits[k]
        key = reference_cache_key(backup.bodies[i])
        cache[key] = dict(cache[key], **kw)
        bend[i] = sorted(kw)
    mod(0, date="")                                   # no date text: the phone's timestamp stands in
    mod(10, date="")                                  # the same on a message whose payload carries a numeric date
    mod(20, date="   ")
    mod(3, date="2024-02-14T09:00:00+04:00")          # timezone-aware: the date gate's comparison raises
    mod(5, merchant=None)
    mod(6, date="01.01.2031 10:00")                   # future
    mod(7, amount="-5.00")
    mod(8, card="**18")                               # BrokenMessage
    mod(9, card=None)
    mod(11, txn_type="bogus")
    mod(12, amount="1.234,56", balance="79 825,89")
    mod(13, date="yesterday")
    mod(30, date="")
    mod(40, date="")
    # the except-handler of parse_sms_llm itself raises (resp_data['txn_type'], gemini_parser.py:239): the
    # exception reaches the worker's `except Exception` (worker.py:141-149) -> {"err", "entry"} on sms.failed
    k14 = reference_cache_key(backup.bodies[hits[14]])
    cache[k14] = {k: v for k, v in cache[k14].items() if k != "txn_type"}
    cache[k14]["card"] = None
    bend[hits[14]] = ["-txn_type", "card"]
    cache[reference_cache_key(backup.bodies[hits[17]])] = "not a dict"
    bend[hits[17]] = ["value"]
    return cache, bend


class Harness:
    """The reference's two handlers, imported unmodified, behind a fake NATS: `run(handler, payload bytes)` reports
    acks, the exception that escaped (if any), the three Prometheus counters' deltas and what was published."""

    def __init__(self):
        logging.disable(logging.CRITICAL)
        self.ref = ref = Reference(tempfile.mkdtemp(prefix="ref_worker_"))
        sys.path.insert(0, os.path.join(REFERENCE_ROOT, "services", "parser_worker"))     # `from metrics import ...`
        if "redis" not in sys.modules:                       # metrics.py only names the type (absent package: stub)
            import types
            redis, ra = types.ModuleType("redis"), types.ModuleType("redis.asyncio")
            ra.Redis = type("Redis", (), {})
            redis.asyncio = ra
            sys.modules.update({"redis": redis, "redis.asyncio": ra})
        with ref.quiet():
            import services.parser_worker.dlq_worker as D
            import services.parser_worker.worker as W
        from libs import nats_utils

        class _Frozen(dt.datetime):
            @classmethod
            def now(cls, tz=None):
                return NOW

        W.datetime = _Frozen

        async def _no_stream(nc):
            return None
        W.ensure_stream = _no_stream
        D.ensure_stream = _no_stream
        self.W, self.D = W, D
        self.subjects = {nats_utils.SUBJECT_FAILED: "failed", nats_utils.SUBJECT_PARSED: "parsed",
                         getattr(nats_utils, "SUBJECT_PROCESSING", "?"): "processing"}

    def counters(self):
        W = self.W
        return [c._value.get() for c in (W.PARSED_OK, W.PARSED_SKIP, W.PARSED_FAIL)]

    def run(self, coro_fn, data):
        nc, msg = _NC(), _Msg(data)
        before = self.counters()
        escaped = None
        try:
            with self.ref.quiet():
                asyncio.run(coro_fn(nc, msg))
        except Exception as e:                                   # what the worker loop would die of
            escaped = type(e).__name__
        after = self.counters()
        return {"acks": msg.acks, "escaped": escaped,
                "counters": [int(a - b) for a, b in zip(after, before)],          # PARSED_OK, PARSED_SKIP, PARSED_FAIL
                "published": [[self.subjects.get(s, s), p.decode("utf-8")] for s, p in nc.log]}

    def run_dlq(self, data):
        return self.run(lambda nc, m: self.D._handle_dlq_msg(nc, m, reparse=True), data)

    def run_worker(self, data):
        return self.run(self.W._process_one, data)


def main():
    h = Harness()
    ref, W, D = h.ref, h.W, h.D
    run = h.run

    backup = synth.generate(160, seed=74, mixed_locale=True, miss_frac=0.1, otp_frac=0.08)
    cache, bend = special_cache(backup)
    # bodies WITHOUT date text (every synthetic body carries one, and fix_broken_datetime then replaces the
    # cached date's day): the cached date decides alone - aware / future / missing / phone timestamp
    base = dict(txn_type="debit", amount="52.00", currency="usd", card="***0018", merchant="TEST LLC", city="AM",
                address="null", balance="1,842.74")
    for k, date in enumerate(["2024-02-14T09:00:00+04:00", "01.01.31 10:00", "14.02.24 09:30", "", "2031-05-06 07:08:09",
                              "2024-02-14 09:00:00Z", "yesterday", None]):
        body = f"PAYMENT {k} TEST LLC 52.00 USD card 0018"
        backup.bodies.append(body)
        backup.senders.append("BANK")
        backup.date_ms.append(1_707_900_000_000 + 1000 * (160 + k) * (10 if k % 2 else 1))
        cache[reference_cache_key(body)] = dict(base, date=date)
    ref.seed_cache(cache)

    flows = {"dlq": [], "worker": []}
    for p in payloads(backup.bodies, backup.senders):
        flows["dlq"].append(dict(payload_hex=p.hex(), **run(lambda nc, m: D._handle_dlq_msg(nc, m, reparse=True), p)))
    # the live path: one RawSMS JSON per message, as the watcher / gateway publish it (libs/nats_utils.py:118-121)
    for i, (b, s, ms) in enumerate(zip(backup.bodies, backup.senders, backup.date_ms)):
        import hashlib
        numeric = (i % 10 == 0) or i in (163, 165)
        raw = ref.models.RawSMS(msg_id=hashlib.sha1(b.encode()).hexdigest(), sender=s, body=b,
                                date=str(ms) if numeric else dt.datetime.fromtimestamp(ms / 1000, tz=dt.timezone.utc).isoformat(),
                                device_id=None if numeric else "xml_backup", source="device" if numeric else "xml")
        data = raw.model_dump_json().encode()
        flows["worker"].append(dict(payload_hex=data.hex(), **run(W._process_one, data)))
    ref.close()
    out = {"now": NOW.isoformat(), "seed": 74, "n": 160, "extra_bodies": backup.bodies[160:], "extra_date_ms": backup.date_ms[160:], "kwargs": dict(mixed_locale=True, miss_frac=0.1, otp_frac=0.08),
           "cache": cache, "bent": {str(k): v for k, v in bend.items()}, "flows": flows,
           "offline_error": "offline: Gemini is not reachable"}
    with open(os.path.join(ROOT, "tests", "golden", "worker_flow.json"), "w", encoding="utf-8") as f:
        json.dump(out, f, ensure_ascii=False, indent=0)
    for name, rows in flows.items():
        kinds = {}
        for r in rows:
            k = (r["acks"], r["escaped"], tuple(r["counters"]), tuple(s for s, _ in r["published"]))
            kinds[k] = kinds.get(k, 0) + 1
        print(name, len(rows))
        for k, v in sorted(kinds.items(), key=str):
            print("   ", v, k)


if __name__ == "__main__":
    main()
