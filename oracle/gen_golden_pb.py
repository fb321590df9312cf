"""Build container only: tests/golden/pb_builders.json from the reference's own text.  save_to_pocketbase.py
cannot be imported (its line 17 is a syntax error), so lines 11-15 (imports) and 34-78 (`_dt_str`,
`_build_sms_data`, `_build_transactions`) are executed exactly as they stand in /root/reference.
Usage: python -m oracle.gen_golden_pb"""
import json
import logging
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests._pb_cases import pairs, records   # noqa: E402

SRC = "/root/reference/save_to_pocketbase.py"


def main():
    lines = open(SRC, encoding="utf-8").read().split("\n")
    text = "\n".join(lines[10:15] + lines[33:78])
    ns = {}
    logging.disable(logging.CRITICAL)
    exec(compile(text, SRC, "exec"), ns)
    out = {"dt_str": [[d, t, ns["_dt_str"](d, t)] for d, t in pairs()],
           "sms_data": [[r, ns["_build_sms_data"](r)] for r in records()],
           "transactions": [[r, ns["_build_transactions"](r)] for r in records()]}
    with open(os.path.join(ROOT, "tests", "golden", "pb_builders.json"), "w") as f:
        json.dump(out, f, ensure_ascii=True, indent=0)
    ok = sum(1 for _, _, v in out["dt_str"] if v)
    print(len(out["dt_str"]), "pairs,", ok, "parse;", len(out["sms_data"]), "records")


if __name__ == "__main__":
    main()
