"""Generates regex_tables.h: the two legacy patterns of process_cached.py:98-135
compiled for the backtracking matcher in legacy_core.cuh, plus the Unicode tables
the matcher needs.

The patterns are parsed with CPython's own regex parser (`re._parser`), so group
numbering, repeat bounds and alternation order are exactly what `re` uses; the
parse tree is then lowered to a small instruction set whose execution order is
the order `re` tries alternatives in (first alternative first, greedy loops longest
first, lazy loops shortest first, a failed path pops the most recent choice).

Instruction word: op in bits 0-7, a in bits 8-19, b in bits 20-31.
"""
from __future__ import annotations

import os
import re
import re._constants as C
import re._parser as P
import sys

HERE = os.path.dirname(os.path.abspath(__file__))

DEBIT = (r"^.*?(?:PURCHASE\s+DB\s+INTERNET|PURCH\.COMPLETION\.DB\s+INTERNET|PURCHASE|SALE):\s*"
         r"(?P<merchant>.+?),\s*(?:(?P<city>YEREVAN),\s*)?"
         r"(?P<address>.*?)(?=,\s*\d{2}[./-]\d{2}[./-]\d{2,4}\s),\s*(?P<date>\d{2}[./-]\d{2}[./-]\d{2,4})\s+"
         r"(?P<time>\d{2}:\d{2}),\s*card\s+\*{3}(?P<card>\d{4})\.\s*Amount:\s*(?P<amount>[\d,.]+)\s+"
         r"(?P<currency>\w{3}),\s*Balance:\s*(?P<balance>[\d,.]+)")
CREDIT = (r"^.*?(?P<type>[\w\s]+?):\s*(?P<date>\d{2}[./-]\d{2}[./-]\d{2,4})\s+(?P<time>\d{2}:\d{2}),\s*"
          r"card\s+\*{3}(?P<card>\d{4})\.\s*Amount:\s*(?P<amount>[\d.,]+)\s+(?P<currency>\w{3}),\s*"
          r"Balance:\s*(?P<balance>[\d.,]+)\s+(?:\w{3})\s*$")

OPS = dict(MATCH=0, CHAR=1, CHAR_CI=2, ANY=3, ANY_NONL=4, CLASS=5, SPLIT=6, JMP=7, SAVE=8, BOL=9, EOL=10,
           ASSERT_BEGIN=11, ASSERT_END=12)
# character classes: bit mask of categories / literal sets
CLS = dict(DIGIT=1, SPACE=2, WORD=4, SEP=8, COMMA_DOT=16)


def ins(op, a=0, b=0):
    assert 0 <= a < 4096 and 0 <= b < 4096
    return OPS[op] | a << 8 | b << 20


def class_mask(items):
    mask = 0
    lits = set()
    for kind, val in items:
        if kind is C.CATEGORY:
            mask |= {C.CATEGORY_DIGIT: CLS["DIGIT"], C.CATEGORY_SPACE: CLS["SPACE"], C.CATEGORY_WORD: CLS["WORD"]}[val]
        elif kind is C.LITERAL:
            lits.add(chr(val))
        else:
            raise ValueError(kind)
    if lits == set("./-"):
        mask |= CLS["SEP"]
    elif lits == set(",."):
        mask |= CLS["COMMA_DOT"]
    elif lits:
        raise ValueError(lits)
    return mask


def lower_seq(seq, prog, dotall):
    for op, av in seq:
        if op is C.AT:
            prog.append(ins("BOL" if av is C.AT_BEGINNING else "EOL"))
            assert av in (C.AT_BEGINNING, C.AT_END)
        elif op is C.LITERAL:
            ch = chr(av)
            if ch.isalpha():
                assert ch.isascii()
                prog.append(ins("CHAR_CI", ord(ch.lower())))
            else:
                prog.append(ins("CHAR", av))
        elif op is C.ANY:
            prog.append(ins("ANY" if dotall else "ANY_NONL"))
        elif op is C.IN:
            prog.append(ins("CLASS", class_mask(av)))
        elif op is C.SUBPATTERN:
            group, add_flags, del_flags, body = av
            assert not add_flags and not del_flags
            if group is not None:
                prog.append(ins("SAVE", 2 * group))
            lower_seq(body, prog, dotall)
            if group is not None:
                prog.append(ins("SAVE", 2 * group + 1))
        elif op is C.BRANCH:
            _, alts = av
            jumps = []
            for k, alt in enumerate(alts):
                if k < len(alts) - 1:
                    split_at = len(prog)
                    prog.append(None)
                    lower_seq(alt, prog, dotall)
                    jumps.append(len(prog))
                    prog.append(None)
                    prog[split_at] = ins("SPLIT", split_at + 1, len(prog))      # try this alternative, else the next
                else:
                    lower_seq(alt, prog, dotall)
            for j in jumps:
                prog[j] = ins("JMP", len(prog))
        elif op in (C.MAX_REPEAT, C.MIN_REPEAT):
            lo, hi, body = av
            greedy = op is C.MAX_REPEAT
            for _ in range(lo):
                lower_seq(body, prog, dotall)
            if hi is C.MAXREPEAT:
                top = len(prog)
                prog.append(None)
                lower_seq(body, prog, dotall)
                prog.append(ins("JMP", top))
                out = len(prog)
                prog[top] = ins("SPLIT", top + 1, out) if greedy else ins("SPLIT", out, top + 1)
            else:
                pending = []
                for _ in range(hi - lo):
                    at = len(prog)
                    prog.append(None)
                    lower_seq(body, prog, dotall)
                    pending.append(at)
                out = len(prog)
                for at in pending:
                    prog[at] = ins("SPLIT", at + 1, out) if greedy else ins("SPLIT", out, at + 1)
        elif op is C.ASSERT:
            direction, body = av
            assert direction == 1
            prog.append(ins("ASSERT_BEGIN"))
            lower_seq(body, prog, dotall)
            prog.append(ins("ASSERT_END"))
        else:
            raise ValueError(op)


def compile_pattern(pattern, flags):
    tree = P.parse(pattern, flags)
    prog = []
    lower_seq(tree, prog, bool(flags & re.S))
    prog.append(ins("MATCH"))
    assert all(p is not None for p in prog) and len(prog) < 4096
    groups = dict(tree.state.groupdict)
    return prog, groups


def ranges_where(pred):
    runs = []
    for cp in range(0x80, 0x110000):
        if 0xD800 <= cp <= 0xDFFF:
            continue
        if pred(chr(cp)):
            if runs and cp == runs[-1][1] + 1:
                runs[-1][1] = cp
            else:
                runs.append([cp, cp])
    return runs


def main():
    debit, dg = compile_pattern(DEBIT, re.I | re.S)
    credit, cg = compile_pattern(CREDIT, re.I)
    word = ranges_where(lambda ch: re.match(r"\w", ch) is not None)
    # case-insensitive literals: which non-ASCII code points match an ASCII letter under re.IGNORECASE
    exc = []
    for l in sorted(set(ch.lower() for ch in DEBIT + CREDIT if ch.isalpha())):
        rx = re.compile(l, re.I)
        for cp in range(0x80, 0x110000):
            if not (0xD800 <= cp <= 0xDFFF) and rx.fullmatch(chr(cp)):
                exc.append((ord(l), cp))
    # OTP filter of process_cached.py:236 on body.upper(): same alphabet / class ids as the worker filter
    sys.path.insert(0, HERE)
    import gen_tables as G
    alphabet = G.ci_alphabet()
    cls = {ch: i + 2 for i, ch in enumerate(alphabet)}
    dense, out = G.aho(["OTP", "PASS=", "CODE:"], alphabet)
    n = len(dense)
    otp = [[0] * G.STRIDE for _ in range(n + 1)]
    for s in range(n):
        otp[s][1] = s
        for ch, c in cls.items():
            t = dense[s][ch]
            otp[s][c] = n if out[t] else t
    otp[n] = [n] * G.STRIDE
    row = G.STRIDE * 2

    with open(os.path.join(os.environ.get("SMSGATE_GEN_OUT") or HERE, "regex_tables.h"), "w") as f:
        w = f.write
        w("// GENERATED by gen_regex.py - do not edit.\n#pragma once\n#include <stdint.h>\n\n")
        for name, v in OPS.items():
            w(f"#define RX_{name} {v}\n")
        for name, v in CLS.items():
            w(f"#define RX_CLS_{name} {v}\n")
        w(f"#define RX_DEBIT_LEN {len(debit)}\n#define RX_CREDIT_LEN {len(credit)}\n")
        for name, g in dg.items():
            w(f"#define RX_DEBIT_G_{name.upper()} {g}\n")
        for name, g in cg.items():
            w(f"#define RX_CREDIT_G_{name.upper()} {g}\n")
        w(f"#define RX_MAX_GROUPS {max(max(dg.values()), max(cg.values())) + 1}\n")
        w(f"#define RX_N_WORD_RANGES {len(word)}\n#define RX_N_CI_EXC {len(exc)}\n")
        w(f"#define LEG_OTP_STATES {n + 1}\n#define LEG_OTP_ACCEPT_OFF {n * row}\n\n")

        def arr(name, ctype, vals):
            w(f"static const {ctype} {name}[{len(vals)}] = {{\n")
            for i in range(0, len(vals), 16):
                w("  " + ",".join(str(v) for v in vals[i:i + 16]) + ",\n")
            w("};\n\n")
        arr("H_RX_DEBIT", "uint32_t", debit)
        arr("H_RX_CREDIT", "uint32_t", credit)
        arr("H_RX_WORD_LO", "uint32_t", [a for a, _ in word])
        arr("H_RX_WORD_HI", "uint32_t", [b for _, b in word])
        arr("H_RX_CI_EXC", "uint32_t", [l << 24 | cp for l, cp in exc])
        arr("H_LEG_OTP_NEXT", "uint16_t", [v * row for r in otp for v in r])
    print("debit", len(debit), "credit", len(credit), "word ranges", len(word), "ci exceptions",
          [(chr(l), hex(cp)) for l, cp in exc], "otp states", n + 1)


if __name__ == "__main__":
    main()
