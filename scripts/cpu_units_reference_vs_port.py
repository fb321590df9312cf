"""BUILD CONTAINER ONLY (needs /root/reference): BASELINE.md section 4's timed units A, B, C on configs[0]
(a 10 000-message synthetic XML backup, 100 % cache hit) with the REFERENCE'S OWN functions (imported
unmodified through oracle/ref_runner.py) next to the oracle port that bench.py's CPU arm has to use on the GPU
box, where the reference is absent.  Same input, same seeded `.gemini_cache`, one process, best of 3, and the
two sides' results are compared before any time is printed.
Usage: python scripts/cpu_units_reference_vs_port.py > profiles/r2_cpu_units_reference_vs_port.txt"""
import contextlib
import datetime as dt
import io
import os
import shutil
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import legacy_oracle as LO          # noqa: E402
from oracle import replay_oracle as O           # noqa: E402
from oracle.gen_golden import _worker_filter    # noqa: E402
from oracle.ref_runner import Reference         # noqa: E402
from smsgate_b200 import synth                  # noqa: E402

N = 10_000
NOW = dt.datetime(2026, 1, 1)


def best_of(fn, k=3):
    ts, out = [], None
    for _ in range(k):
        t0 = time.perf_counter()
        out = fn()
        ts.append(time.perf_counter() - t0)
    return min(ts), out


def main():
    import diskcache
    work = tempfile.mkdtemp(prefix="cpu_units_")
    backup = synth.generate(N, seed=0x5EED + 1)
    xml = os.path.join(work, "sms-backup.xml")
    synth.write_xml(xml, backup)
    ref = Reference(os.path.join(work, "ref"))
    ref.seed_cache(backup.cache)
    filt = _worker_filter(ref)
    quiet = lambda: contextlib.redirect_stdout(io.StringIO())  # noqa: E731
    print(f"configs[0]: {N} messages, XML {os.path.getsize(xml) / 1e6:.1f} MB, {len(backup.cache)} cache entries, "
          f"host cores {os.cpu_count()}, single process, best of 3\n")

    # ---- unit A: _iter_sms
    ta_ref, raws = best_of(lambda: list(ref.watcher._iter_sms(xml)))
    from smsgate_b200.xml_ingest import read_backup

    def port_a():
        b = read_backup(xml)
        return [(O.msg_id_xml(b.body(i)).hex(), b.senders[i], O.iso_utc_from_ms(int(b.date_ms[i]))) for i in range(b.n)]
    ta_port, port_rows = best_of(port_a)
    assert [(r.msg_id, r.sender, r.date) for r in raws] == port_rows
    print(f"A  _iter_sms (XML -> RawSMS, SHA-1 msg_id)              reference {N / ta_ref:10.0f} msgs/s   port {N / ta_port:10.0f} msgs/s")

    # ---- unit B: worker filter -> parse_sms_llm -> gates -> dict upsert
    def ref_b():
        ups = {}
        gp = ref.gemini_parser
        with quiet():
            for raw in raws:
                if filt(raw.body):
                    continue
                try:
                    parsed = gp.parse_sms_llm(raw)
                except Exception:
                    continue
                if parsed is None or parsed.date > NOW or not parsed.merchant:
                    continue
                ups[raw.msg_id] = parsed                                     # last writer wins
        return ups
    tb_ref, ups_ref = best_of(ref_b)
    cache_dir = os.path.join(work, "port_cache")
    c = diskcache.Cache(cache_dir, size_limit=1 << 36)
    with c.transact():
        for k, v in backup.cache.items():
            c[k] = v

    def port_b():
        with quiet():
            res = O.replay(backup.bodies, backup.senders, [r.date for r in raws], c, NOW, lean=True)
        return {o.msg_id.hex(): o for o, s in zip(res.outcomes, res.survivor) if s}
    tb_port, ups_port = best_of(port_b)
    c.close()
    assert set(ups_ref) == set(ups_port)
    for k, p in ups_ref.items():
        q = ups_port[k].parsed
        assert (p.date, p.amount, p.balance, p.card, p.merchant, p.raw_body) == (q.date, q.amount, q.balance, q.card, q.merchant, q.raw_body)
    print(f"B  filter -> parse_sms_llm -> gates -> dict upsert       reference {N / tb_ref:10.0f} msgs/s   port {N / tb_port:10.0f} msgs/s"
          f"   ({len(ups_ref)} upserts, identical)")

    # ---- unit C: parse_and_cache_sms + process_sms_from_cache
    def ref_c():
        d = tempfile.mkdtemp(dir=work)
        with quiet():
            ref.read_xml.parse_and_cache_sms(xml, os.path.join(d, "src"))
            ref.process_cached.process_sms_from_cache(os.path.join(d, "src"), os.path.join(d, "deb"), os.path.join(d, "cre"))
        with diskcache.Cache(os.path.join(d, "deb")) as a, diskcache.Cache(os.path.join(d, "cre")) as b:
            return len(a), len(b)
    tc_ref, counts_ref = best_of(ref_c, 2)

    def port_c():
        deb = cre = 0
        for b in backup.bodies:
            st = LO.process_body(b)[0]
            deb += st == 1
            cre += st == 2
        return deb, cre
    tc_port, counts_port = best_of(port_c, 2)
    assert counts_ref == counts_port, (counts_ref, counts_port)
    print(f"C  parse_and_cache_sms + process_sms_from_cache          reference {N / tc_ref:10.0f} msgs/s   "
          f"port (per-message work only, no DiskCache writes) {N / tc_port:10.0f} msgs/s   ({counts_ref[0]} debit, {counts_ref[1]} credit)")
    print("\nUnit B is the figure set beside the GPU msgs/s (bench.py cpu_baseline / --impl reference run the PORT, because"
          "\n/root/reference does not exist on the GPU box); this table is the cross-check that the port costs what the"
          "\nreference costs on the same input and produces the same upsert set.")
    ref.close()
    shutil.rmtree(work, ignore_errors=True)


if __name__ == "__main__":
    main()
