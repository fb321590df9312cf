"""Generates scan_tables.h: the byte-class table and the three dense automata
the scan kernel walks once per body byte.

  CI  Aho-Corasick automaton over the UPPER-CASED symbol stream for the nine
      needles of the worker skip-filter (services/parser_worker/worker.py:112-121).
      One absorbing accept state.
  CS  automaton over RAW bytes for the case-sensitive needles: the parser's OTP
      list (libs/gemini_parser.py:198) and the worker's `Daily limit exceeded`
      (worker.py:120).  One sticky fact rides in the state number (second block
      of states = "Daily limit exceeded" seen); one absorbing state = parser
      needle seen.
  DM  product of two KMP automata over code-point classes {other, digit, '.',
      '*'}: the body-date forms dd.mm.yy / dd.mm.yyyy (libs/gemini_parser.py:73-76)
      and the card form dddd***dddd (libs/gemini_parser.py:131).  Rows are
      sorted by date progress (nothing / short form seen / long form seen) so
      the kernel spots the two events with one compare per byte; whether the
      card form was seen is read off the final row.

All three tables hold, per (row, class), the BYTE OFFSET of the next row inside
its own table (uint16), and the byte-class word holds each automaton's class
already multiplied by the entry size, so one step is `next = tab[state + cls]`.

Run `python smsgate_b200/csrc/gen_tables.py` after changing a needle; the
header is committed so the build needs no Python.
"""
from __future__ import annotations

import os
import re
import sys
from collections import deque

CI_NEEDLES = ["OTP", "CODE:", "NOT ENOUGH FUNDS", "INSUFFICIENT FUNDS", "CREDIT PAYMENT",
              "C2C RECEIVED", "PASS:", "PASS=", "PERSON TO PERSON"]
CS_PARSER = ["OTP", "CODE:", "PASS:", "PASS=", "Daily limit exceeded:"]
CS_WORKER = "Daily limit exceeded"
NBSP = b"\xc2\xa0"

STRIDE = 32            # classes per CI / CS row
DM_STRIDE = 8

# upper()-expansions that contain ASCII (probed over all of Unicode 15.0)
UPPER_SPECIALS = {
    0x00DF: "SS", 0x0131: "I", 0x0149: "\x00N", 0x017F: "S", 0x01F0: "J\x00",
    0x1E96: "H\x00", 0x1E97: "T\x00", 0x1E98: "W\x00", 0x1E99: "Y\x00", 0x1E9A: "A\x00",
    0xFB00: "FF", 0xFB01: "FI", 0xFB02: "FL", 0xFB03: "FFI", 0xFB04: "FFL", 0xFB05: "ST", 0xFB06: "ST",
}  # "\x00" stands for a non-ASCII output character


def digit_runs():
    cps = [cp for cp in range(0x110000) if re.match(r"\d", chr(cp))]
    runs = []
    for cp in cps:
        if runs and cp == runs[-1][1] + 1:
            runs[-1][1] = cp
        else:
            runs.append([cp, cp])
    return [(a, b - a + 1) for a, b in runs]


def aho(needles, alphabet):
    """Dense goto table of the Aho-Corasick automaton; returns (next, accepting-set)."""
    goto = [{}]
    out = [set()]
    for n in needles:
        s = 0
        for ch in n:
            if ch not in goto[s]:
                goto.append({})
                out.append(set())
                goto[s][ch] = len(goto) - 1
            s = goto[s][ch]
        out[s].add(n)
    fail = [0] * len(goto)
    q = deque()
    for ch, t in goto[0].items():
        q.append(t)
    dense = [dict() for _ in goto]
    order = [0]
    while q:
        r = q.popleft()
        order.append(r)
        for ch, t in goto[r].items():
            q.append(t)
            f = fail[r]
            while f and ch not in goto[f]:
                f = fail[f]
            fail[t] = goto[f].get(ch, 0) if goto[f].get(ch, 0) != t else 0
            out[t] |= out[fail[t]]
    for s in order:
        for ch in alphabet:
            if ch in goto[s]:
                dense[s][ch] = goto[s][ch]
            else:
                dense[s][ch] = dense[fail[s]][ch] if s else 0
    return dense, out


# The scan kernel's first pass runs ONE upper-case automaton whose needles are the worker's keywords plus the
# upper-cased form of every case-sensitive needle that is not among them: a body in which no such needle
# occurs after upper-casing cannot contain a case-sensitive needle either, so the case-sensitive automaton
# only has to run over the few bodies this automaton flags.
CI2_EXTRA = [n.upper() for n in CS_PARSER + [CS_WORKER] if n.upper() not in CI_NEEDLES]


def ci_alphabet():
    return sorted(set("".join(CI_NEEDLES + CI2_EXTRA)))


def build_ci(needles=None):
    needles = CI_NEEDLES if needles is None else needles
    alphabet = ci_alphabet()
    cls = {ch: i + 2 for i, ch in enumerate(alphabet)}      # 0 = OTHER, 1 = KEEP (continuation byte)
    assert len(cls) + 2 <= STRIDE
    dense, out = aho(needles, alphabet)
    n = len(dense)
    ACC = n                                                # absorbing
    table = [[0] * STRIDE for _ in range(n + 1)]
    for s in range(n):
        for c in range(STRIDE):
            table[s][c] = 0
        table[s][1] = s
        for ch, c in cls.items():
            t = dense[s][ch]
            table[s][c] = ACC if out[t] else t
    for c in range(STRIDE):
        table[ACC][c] = ACC
    return table, cls, ACC


def build_cs():
    needles = CS_PARSER + [CS_WORKER]
    alphabet = sorted(set("".join(needles)))
    cls = {ch: i + 1 for i, ch in enumerate(alphabet)}     # 0 = OTHER
    assert len(cls) + 1 <= STRIDE
    dense, out = aho(needles, alphabet)
    n = len(dense)
    # state id = W * n + automaton state (W = "Daily limit exceeded" seen), then the parser-hit state
    ACC = 2 * n
    sym_of = {k: ch for ch, k in cls.items()}
    table = [[0] * STRIDE for _ in range(2 * n + 1)]
    for W in range(2):
        for s in range(n):
            for c in range(STRIDE):
                sym = sym_of.get(c)
                t = dense[s][sym] if sym is not None else 0
                w2 = 1 if (W or CS_WORKER in out[t]) else 0
                table[W * n + s][c] = ACC if any(x in CS_PARSER for x in out[t]) else w2 * n + t
    for c in range(STRIDE):
        table[ACC][c] = ACC
    return table, cls, n, ACC


def kmp(pattern, classes):
    """Dense KMP automaton; state = matched prefix length, len(pattern) is reported
    by the caller.  After a full match the search restarts from scratch (regex
    `sub`/`search` semantics: non-overlapping)."""
    m = len(pattern)
    nxt = [[0] * len(classes) for _ in range(m + 1)]
    for s in range(m + 1):
        for ci, c in enumerate(classes):
            if s == m:
                text = c                      # restart after a complete match
            else:
                text = pattern[:s] + c
            k = min(len(text), m)
            while k > 0 and text[-k:] != pattern[:k]:
                k -= 1
            nxt[s][ci] = k
    return nxt


def build_dm():
    # class ids: 0 other, 1 digit, 2 '.', 3 '*', 4 keep
    classes = ["o", "d", ".", "*"]
    date = kmp("dd.dd.dddd", classes)        # progress 8 = short form complete, 10 = long form complete
    card = kmp("dddd***dddd", classes)       # 11 = card form complete
    start = (0, 0, 0, 0, 0)                  # (date progress, B found, A found, card progress, card found)
    seen = {start}
    work = [start]
    trans = {}
    while work:
        key = work.pop()
        dp, fb, fa, cp, fc = key
        row = []
        for ci in range(4):
            ndp, nfb, nfa = dp, fb, fa
            if not fa:
                ndp = date[dp][ci]
                if ndp >= 8:
                    nfb = 1
                if ndp == 10:
                    nfa = 1
                    ndp = 0
            ncp, nfc = cp, fc
            if not fc:
                ncp = card[cp][ci]
                if ncp == 11:
                    nfc = 1
                    ncp = 0
            nk = (ndp, nfb, nfa, ncp, nfc)
            if nk not in seen:
                seen.add(nk)
                work.append(nk)
            row.append(nk)
        trans[key] = row
    order = sorted(seen, key=lambda k: (k[1] + k[2], k[4], k[0], k[3]))
    assert order[0] == start
    idx = {k: i for i, k in enumerate(order)}
    first_b = min(i for i, k in enumerate(order) if k[1])
    first_a = min(i for i, k in enumerate(order) if k[2])
    table = [[0] * DM_STRIDE for _ in order]
    for k, i in idx.items():
        for ci in range(4):
            table[i][ci] = idx[trans[k][ci]]
        for ci in range(4, DM_STRIDE):
            table[i][ci] = i                                 # keep
    has_card = [1 if k[4] else 0 for k in order]
    return table, len(order), first_b, first_a, has_card


def lead_len(b):
    if b < 0x80:
        return 1
    if 0xC2 <= b <= 0xDF:
        return 2
    if 0xE0 <= b <= 0xEF:
        return 3
    if 0xF0 <= b <= 0xF4:
        return 4
    return 0


def main():
    ci, ci_cls, CI_ACC = build_ci()
    ci2, ci2_cls, CI2_ACC = build_ci(CI_NEEDLES + CI2_EXTRA)
    assert ci2_cls == ci_cls
    cs, cs_cls, CS_N, CS_ACC = build_cs()
    dm, dm_n, dm_first_b, dm_first_a, dm_has_card = build_dm()
    runs = digit_runs()

    special_leads = {0xC2, 0xE2}                            # U+00A0 and U+2022, rewritten by `clean`
    for cp in UPPER_SPECIALS:
        special_leads.add(chr(cp).encode()[0])
    for a, n in runs:
        if a >= 0x80:
            special_leads.add(chr(a).encode()[0])
            assert chr(a).encode()[0] == chr(a + n - 1).encode()[0]

    CI_ROW = STRIDE * 2          # bytes per row (uint16 entries)
    CS_ROW = STRIDE * 2
    DM_ROW = DM_STRIDE * 2
    # byte class word: byte0 = CI class * 2, byte1 = CS class * 2, byte2 = DM class * 2, bit 31 = special lead
    cls_word = []
    for b in range(256):
        ch = chr(b)
        if b < 0x80:
            cic = ci_cls.get(ch.upper(), 0)
            csc = cs_cls.get(ch, 0)
            dmc = 1 if ch.isdigit() else 2 if ch == "." else 3 if ch == "*" else 0
            sp = 0
        elif b < 0xC0:
            cic, csc, dmc, sp = 1, 0, 4, 0                  # continuation byte: CI / DM keep their state
        else:
            cic, csc, dmc = 0, 0, 0
            sp = 1 if b in special_leads else 0
        cls_word.append((cic * 2) | (csc * 2) << 8 | (dmc * 2) << 16 | sp << 31)

    here = os.environ.get("SMSGATE_GEN_OUT") or os.path.dirname(os.path.abspath(__file__))
    with open(os.path.join(here, "scan_tables.h"), "w") as f:
        w = f.write
        w("// GENERATED by gen_tables.py - do not edit.\n#pragma once\n#include <stdint.h>\n\n")
        w(f"#define SCAN_STRIDE {STRIDE}\n#define DM_STRIDE {DM_STRIDE}\n")
        w(f"#define CI_STATES {len(ci)}\n#define CI_ROW_BYTES {CI_ROW}\n#define CI_ACCEPT_OFF {CI_ACC * CI_ROW}\n")
        w(f"#define CI2_STATES {len(ci2)}\n#define CI2_ACCEPT_OFF {CI2_ACC * CI_ROW}\n")
        w(f"#define CS_STATES {len(cs)}\n#define CS_ROW_BYTES {CS_ROW}\n#define CS_DAILY_OFF {CS_N * CS_ROW}\n"
          f"#define CS_ACCEPT_OFF {CS_ACC * CS_ROW}\n")
        w(f"#define DM_STATES {dm_n}\n#define DM_ROW_BYTES {DM_ROW}\n#define DM_FIRST_B_OFF {dm_first_b * DM_ROW}\n"
          f"#define DM_FIRST_A_OFF {dm_first_a * DM_ROW}\n")
        for ch, c in sorted(ci_cls.items(), key=lambda kv: kv[1]):
            name = {" ": "SP", ":": "COLON", "=": "EQ"}.get(ch, ch)
            w(f"#define CI_CLS_{name} {c * 2}\n")
        w("#define CI_CLS_OTHER 0\n#define CI_CLS_KEEP 2\n")
        w("#define DM_CLS_OTHER 0\n#define DM_CLS_DIGIT 2\n#define DM_CLS_DOT 4\n#define DM_CLS_STAR 6\n#define DM_CLS_KEEP 8\n")
        w("// class ids for the card-form KMP table used by the rewrite (not pre-multiplied)\n")
        w("#define KMP_CLS_OTHER 0\n#define KMP_CLS_DIGIT 1\n#define KMP_CLS_STAR 3\n\n")

        def arr(name, ctype, rows):
            flat = [v for r in rows for v in r]
            w(f"static const {ctype} {name}[{len(flat)}] = {{\n")
            for i in range(0, len(flat), 24):
                w("  " + ",".join(str(v) for v in flat[i:i + 24]) + ",\n")
            w("};\n\n")

        arr("H_BYTE_CLASS", "uint32_t", [cls_word])
        arr("H_CI_NEXT", "uint16_t", [[v * CI_ROW for v in r] for r in ci])
        arr("H_CI2_NEXT", "uint16_t", [[v * CI_ROW for v in r] for r in ci2])
        arr("H_CS_NEXT", "uint16_t", [[v * CS_ROW for v in r] for r in cs])
        arr("H_DM_NEXT", "uint16_t", [[v * DM_ROW for v in r] for r in dm])
        arr("H_DM_HAS_CARD", "uint8_t", [dm_has_card])
        w(f"#define N_DIGIT_RUNS {len(runs)}\n")
        arr("H_DIGIT_RUN_START", "uint32_t", [[a for a, _ in runs]])
        arr("H_DIGIT_RUN_LEN", "uint8_t", [[n for _, n in runs]])
        arr("H_CARD_KMP", "uint8_t", kmp("dddd***dddd", ["o", "d", ".", "*"]))
    print("CI2 states", len(ci2), "extra needles", CI2_EXTRA)
    print("CI states", len(ci), "CS states", len(cs), "DM states", dm_n, "first_b", dm_first_b, "first_a", dm_first_a,
          "special leads", sorted(hex(b) for b in special_leads))


if __name__ == "__main__":
    sys.exit(main())
