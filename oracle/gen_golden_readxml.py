"""Build container only: writes tests/golden/read_xml.json by running the reference's own
`read_xml.parse_and_cache_sms` (read_xml.py:9-54, imported unmodified through oracle/ref_runner.py) on a
handful of backup files and dumping the DiskCache it leaves behind (keys in iteration order, values as stored).
Usage: python -m oracle.gen_golden_readxml"""
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.ref_runner import Reference      # noqa: E402
from smsgate_b200 import synth               # noqa: E402


def documents():
    b = synth.generate(60, seed=81, mixed_locale=True)
    rows = []
    for i, (body, sender, ms) in enumerate(zip(b.bodies, b.senders, b.date_ms)):
        date = str(ms)
        if i % 11 == 3:
            date = str(b.date_ms[i - 1])                       # duplicate key: the later message overwrites
        attrs = [("protocol", "0"), ("address", sender), ("date", date), ("type", "1"), ("subject", "null"),
                 ("body", body), ("toa", "null"), ("sc_toa", "null"), ("service_center", "+37491000000"), ("read", "1"),
                 ("status", "-1"), ("locked", "0"), ("date_sent", str(ms - 5)), ("sub_id", "1"),
                 ("readable_date", "14 Feb 2024 09:00:00"), ("contact_name", "(Unknown) & <Co>")]
        if i % 13 == 5:
            attrs = [a for a in attrs if a[0] != "date"]       # no date: skipped with a warning
        if i % 17 == 6:
            attrs = [(k, "" if k == "date" else v) for k, v in attrs]      # empty date: falsy key, skipped too
        if i % 7 == 2:
            attrs = attrs[::-1]                                # attribute order is the dict's order
        rows.append("  <sms " + " ".join(f'{k}="{synth.xml_escape_attr(v)}"' for k, v in attrs) + " />")
    head = ("<?xml version='1.0' encoding='UTF-8' standalone='yes' ?>\n<!--File Created By SMS Backup & Restore-->\n"
            f'<smses count="{len(rows)}" backup_set="x" backup_date="1718244809" type="full">\n')
    full = (head + "\n".join(rows) + '\n  <mms date="5"><parts><part seq="0" text="t"/></parts><sms date="9" body="nested"/></mms>\n'
            "</smses>\n").encode("utf-8")
    return [
        ("backup", full),
        ("entities", b'<smses><sms date="1" body="a &amp; b &#10;c&#x2022; &lt;&gt; &quot;q&apos;" address="A&#160;B" x="tab\there\r\nnl"/>'
                     b"<sms date=' 2 ' body='single \"quotes\"'></sms><sms body='no date'/><sms date='1' body='again'/></smses>"),
        ("malformed", b'<smses><sms date="1" body="x"/>'),
        ("empty_root", b"<smses/>"),
        ("missing_file", None),
        ("doctype", b'<!DOCTYPE smses><smses><sms date="7" body="dialect the GPU tokenizer leaves to ElementTree"/></smses>'),
    ]


def main():
    import diskcache
    work = tempfile.mkdtemp(prefix="ref_readxml_")
    ref = Reference(work)
    cases = []
    for name, data in documents():
        xml = os.path.join(work, name + ".xml")
        cdir = os.path.join(work, name + "_cache")
        if data is not None:
            with open(xml, "wb") as f:
                f.write(data)
        with ref.quiet():
            ref.read_xml.parse_and_cache_sms(xml, cdir)
        with diskcache.Cache(cdir) as c:
            entries = [[k, c[k]] for k in c]
        cases.append({"name": name, "xml_hex": None if data is None else data.hex(), "entries": entries})
        print(name, len(entries))
    ref.close()
    with open(os.path.join(ROOT, "tests", "golden", "read_xml.json"), "w", encoding="utf-8") as f:
        json.dump(cases, f, ensure_ascii=False, indent=0)


if __name__ == "__main__":
    main()
