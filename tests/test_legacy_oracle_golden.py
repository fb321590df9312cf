"""The legacy-path oracle (oracle/legacy_oracle.py) against vectors produced by
running the reference's process_cached.py (oracle/gen_golden.py).  CPU only."""
from oracle import legacy_oracle as LO


def _dump(d):
    if d is None:
        return None
    out = dict(d)
    for k in ("amount", "balance"):
        out[k] = "%016x" % LO.f64_bits(d[k])
    return out


def test_regex_parsers(golden):
    for c in golden("legacy.json")["cases"]:
        assert _dump(LO.parse_debit(c["body"])) == c["debit"], c["body"]
        assert _dump(LO.parse_credit(c["body"])) == c["credit"], c["body"]


def test_process_flow(golden):
    g = golden("legacy.json")["process"]
    names = {LO.LEG_SKIPPED_OTP: "skipped_otp", LO.LEG_DEBIT: "processed", LO.LEG_CREDIT: "processed",
             LO.LEG_FAILED: "failed_to_parse"}
    debit, credit = {}, {}
    for i, b in enumerate(g["bodies"]):
        key = str(1707900000000 + i)
        st, text, digest, parsed = LO.process_body(b)
        assert names[st] == g["statuses"][key], b
        if parsed:
            rec = dict(_dump(parsed), msg_id=key, original_body=text)
            (debit if st == LO.LEG_DEBIT else credit)[digest.hex()] = rec
    assert debit == g["debit"] and credit == g["credit"]
    assert g["statuses"]["already"] == "processed"
    assert LO.process_body(None)[0] == LO.LEG_FAILED and g["statuses"]["nobody"] == "failed_to_parse"
