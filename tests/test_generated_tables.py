"""The committed table headers are what their generators produce now: a change to the needles or the
alphabet in gen_tables.py that is not followed by re-running BOTH generators (gen_regex.py shares the
upper-case class ids) must fail here, not as a wrong flag on the GPU."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "smsgate_b200", "csrc")


def test_table_headers_are_up_to_date(tmp_path):
    env = dict(os.environ, SMSGATE_GEN_OUT=str(tmp_path))
    for gen, header in (("gen_tables.py", "scan_tables.h"), ("gen_regex.py", "regex_tables.h")):
        subprocess.run([sys.executable, os.path.join(CSRC, gen)], check=True, cwd=CSRC, env=env,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        fresh = open(os.path.join(str(tmp_path), header)).read()
        committed = open(os.path.join(CSRC, header)).read()
        assert fresh == committed, f"{header} is stale: run `python {gen}` in smsgate_b200/csrc"
