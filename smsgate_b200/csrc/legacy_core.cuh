// Per-message logic of the legacy offline replay (process_cached.py:210-268):
// strip, OTP filter on body.upper(), the debit and the credit regular
// expression (process_cached.py:98-135) run by a backtracking matcher that tries
// alternatives in the order Python's `re` does, `_to_decimal` -> float, and the
// SHA-256 of the stripped body.  __host__ __device__ like the other cores, so
// tests/hostcheck can run it on the CPU against the golden vectors.
#pragma once
#include "regex_tables.h"
#include "scan_core.cuh"

struct LegacyTables {
    const uint32_t* debit;      // H_RX_DEBIT
    const uint32_t* credit;     // H_RX_CREDIT
    const uint32_t* word_lo;    // \w ranges above ASCII
    const uint32_t* word_hi;
    const uint32_t* ci_exc;     // letter << 24 | code point that matches it under IGNORECASE
    const uint8_t* otp_next;    // uint16 [LEG_OTP_STATES][32]: OTP / PASS= / CODE: on the upper-cased stream
    ScanTables scan;            // byte classes, digit runs
};

#define LEG_SKIPPED_OTP 0
#define LEG_DEBIT 1
#define LEG_CREDIT 2
#define LEG_FAILED 3
#define LEG_F_BUDGET 0x01       // the matcher gave up (step or stack budget): status is LEG_FAILED, not trustworthy
#define LEG_F_INEXACT 0x02      // amount/balance has more digits than the exact float path handles

#define RX_STACK 128
#define RX_MAX_STEPS 4000000

SG_HD bool sg_rx_is_word(uint32_t cp, const LegacyTables& L) {
    if (cp < 0x80) return (cp >= '0' && cp <= '9') || ((cp | 0x20) >= 'a' && (cp | 0x20) <= 'z') || cp == '_';
    if (cp == 0xFFFFFFFFu) return false;
    int lo = 0, hi = RX_N_WORD_RANGES - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (L.word_lo[mid] <= cp) lo = mid; else hi = mid - 1;
    }
    return cp >= L.word_lo[lo] && cp <= L.word_hi[lo];
}

SG_HD bool sg_rx_class(uint32_t mask, uint32_t cp, const LegacyTables& L) {
    if ((mask & RX_CLS_DIGIT) && sg_digit_value(cp, L.scan.t->digit_run_start, L.scan.t->digit_run_len) >= 0) return true;
    if ((mask & RX_CLS_SPACE) && sg_is_space(cp)) return true;
    if ((mask & RX_CLS_WORD) && sg_rx_is_word(cp, L)) return true;
    if ((mask & RX_CLS_SEP) && (cp == '.' || cp == '/' || cp == '-')) return true;
    if ((mask & RX_CLS_COMMA_DOT) && (cp == ',' || cp == '.')) return true;
    return false;
}

// Backtracking matcher anchored at offset 0 (both patterns start with ^, and
// re.search without MULTILINE can only succeed there).  caps[2g], caps[2g+1] =
// byte span of group g, -1 when unset.  Returns 1 on a match, 0 on no match,
// -1 when a budget ran out.
SG_HD int sg_rx_exec(const uint32_t* prog, const uint8_t* s, int len, int* caps, const LegacyTables& L) {
    uint32_t ka[RX_STACK];      // kind << 30 | payload   (0 alternative: pc, 1 capture undo: slot, 2 lookahead mark)
    int32_t kb[RX_STACK];       // pos / old capture value / pos
    int sp = 0, pc = 0, pos = 0;
    for (int k = 0; k < 2 * RX_MAX_GROUPS; ++k) caps[k] = -1;
    for (int steps = 0; steps < RX_MAX_STEPS; ++steps) {
        const uint32_t w = prog[pc];
        const uint32_t op = w & 255, a = (w >> 8) & 4095, b = w >> 20;
        bool ok = true;
        switch (op) {
            case RX_MATCH: return 1;
            case RX_CHAR: case RX_CHAR_CI: case RX_ANY: case RX_ANY_NONL: case RX_CLASS: {
                if (pos >= len) { ok = false; break; }
                uint32_t cp;
                const int n = sg_utf8_decode(s + pos, len - pos, &cp);
                if (op == RX_CHAR) ok = cp == a;
                else if (op == RX_CHAR_CI) {
                    if (cp < 0x80) ok = (cp | 0x20) == a;
                    else {
                        ok = false;
                        for (int k = 0; k < RX_N_CI_EXC; ++k) ok = ok || L.ci_exc[k] == (a << 24 | cp);
                    }
                } else if (op == RX_ANY_NONL) ok = cp != '\n';
                else if (op == RX_CLASS) ok = sg_rx_class(a, cp, L);
                if (ok) { pos += n; ++pc; }
                break;
            }
            case RX_SPLIT:
                if (sp >= RX_STACK) return -1;
                ka[sp] = b; kb[sp] = pos; ++sp;
                pc = (int)a;
                break;
            case RX_JMP: pc = (int)a; break;
            case RX_SAVE:
                if (sp >= RX_STACK) return -1;
                ka[sp] = (1u << 30) | a; kb[sp] = caps[a]; ++sp;
                caps[a] = pos;
                ++pc;
                break;
            case RX_BOL: ok = pos == 0; if (ok) ++pc; break;
            case RX_EOL: ok = pos == len || (pos == len - 1 && s[pos] == '\n'); if (ok) ++pc; break;
            case RX_ASSERT_BEGIN:
                if (sp >= RX_STACK) return -1;
                ka[sp] = 2u << 30; kb[sp] = pos; ++sp;
                ++pc;
                break;
            case RX_ASSERT_END:
                // the lookahead succeeded: drop the choices made inside it, go back to where it started
                while (sp > 0 && (ka[sp - 1] >> 30) != 2) --sp;
                pos = kb[sp - 1];
                --sp;
                ++pc;
                break;
            default: return -1;
        }
        if (ok) continue;
        // failure: undo to the most recent alternative
        for (;;) {
            if (sp == 0) return 0;
            --sp;
            const uint32_t kind = ka[sp] >> 30;
            if (kind == 0) { pc = (int)ka[sp]; pos = kb[sp]; break; }
            if (kind == 1) caps[ka[sp] & 0xFFF] = kb[sp];
            // kind 2: a lookahead that failed as a whole: keep unwinding
        }
    }
    return -1;
}

// str.strip() on a byte span: trims the 29 Unicode whitespace code points.
SG_HD void sg_strip_span(const uint8_t* s, int* lo, int* hi) {
    int a = *lo, b = *hi;
    while (a < b) {
        uint32_t cp;
        const int n = sg_utf8_decode(s + a, b - a, &cp);
        if (!sg_is_space(cp)) break;
        a += n;
    }
    while (b > a) {
        int p = b - 1;
        while (p > a && (s[p] & 0xC0) == 0x80) --p;
        uint32_t cp;
        sg_utf8_decode(s + p, b - p, &cp);
        if (!sg_is_space(cp)) break;
        b = p;
    }
    *lo = a; *hi = b;
}

// Correctly rounded (nearest, ties to even) coef / 10^frac for coef < 2^64, frac <= 22, in
// 128-bit integer arithmetic: scale so that the quotient has 54..55 bits, divide, round on
// the remainder.  Only used when the one-division path below is not exact.
SG_HD double sg_div_pow10_exact(uint64_t coef, int frac) {
    if (coef == 0) return 0.0;
    unsigned __int128 den = 1;
    for (int k = 0; k < frac; ++k) den *= 10;                 // < 2^74
    unsigned __int128 num = coef;
    int sh = 0;                                               // value = num / den * 2^-sh
    while ((num >> 53) < den) { num <<= 1; ++sh; }            // now num/den >= 2^53 (num < den * 2^54 < 2^128)
    while ((num >> 55) >= den) { den <<= 1; --sh; }           // and < 2^55
    unsigned __int128 q = num / den, rem = num % den;
    // q has 54 or 55 bits: the bits beyond 53 plus the remainder decide the rounding
    int extra = 0;
    while ((q >> (53 + extra)) != 0) ++extra;                 // extra = bits beyond 53 (1 or 2)
    const unsigned __int128 low = q & (((unsigned __int128)1 << extra) - 1);
    unsigned __int128 m = q >> extra;
    const unsigned __int128 half = (unsigned __int128)1 << (extra - 1);
    const bool above = low > half || (low == half && rem != 0);
    const bool tie = low == half && rem == 0;
    if (above || (tie && (m & 1))) ++m;                       // m may become 2^53: still exact as a double
    double r = (double)(uint64_t)m;
    int e = extra - sh;                                       // r * 2^e
    while (e > 0) { r *= 2.0; --e; }
    while (e < 0) { r *= 0.5; ++e; }
    return r;
}

// float(Decimal(value.replace(",", ""))) for value in [\d,.]+ (process_cached.py:141-143,
// 169-171): 0 = InvalidOperation (no digit or two dots), 1 = exact result in *out,
// 2 = more digits than this build carries (coefficient >= 2^64 or > 22 fraction digits):
// *out is then only an approximation and the caller flags it.
SG_HD int sg_to_float(const uint8_t* s, int lo, int hi, const LegacyTables& L, double* out) {
    uint64_t coef = 0;
    int frac = 0, dots = 0, ndig = 0;
    bool big = false;
    for (int p = lo; p < hi;) {
        uint32_t cp;
        const int n = sg_utf8_decode(s + p, hi - p, &cp);
        p += n;
        if (cp == ',') continue;
        if (cp == '.') { ++dots; continue; }
        const int d = sg_digit_value(cp, L.scan.t->digit_run_start, L.scan.t->digit_run_len);
        if (d < 0) return 0;
        ++ndig;
        if (coef > (0xFFFFFFFFFFFFFFFFull - (uint64_t)d) / 10) big = true;
        else coef = coef * 10 + (uint64_t)d;
        if (dots) ++frac;
    }
    if (dots > 1 || ndig == 0) return 0;
    if (big || frac > 22) { *out = (double)coef; return 2; }
    if (coef <= 0x1FFFFFFFFFFFFFull) {
        double p10 = 1.0;
        for (int k = 0; k < frac; ++k) p10 *= 10.0;     // exact up to 1e22
        *out = (double)coef / p10;            // both operands exact in binary64, one correctly rounded division
    } else {
        *out = sg_div_pow10_exact(coef, frac);
    }
    return 1;
}

struct LegacyResult {
    uint32_t status;        // LEG_*
    uint32_t lflags;        // LEG_F_*
    int strip_lo, strip_hi; // the stripped body inside the input
    int caps[2 * RX_MAX_GROUPS];   // group spans of the pattern that matched, relative to the stripped body, trimmed
                                   // where the reference strips; -1 = unset
    double amount, balance;
    uint32_t digest[8];     // SHA-256 of the stripped body (big-endian words)
};

SG_HD bool sg_legacy_numbers(const uint8_t* t, const int* caps, int g_amount, int g_balance, const LegacyTables& L,
                             LegacyResult* r) {
    const int ra = sg_to_float(t, caps[2 * g_amount], caps[2 * g_amount + 1], L, &r->amount);
    if (ra == 0) return false;
    const int rb = sg_to_float(t, caps[2 * g_balance], caps[2 * g_balance + 1], L, &r->balance);
    if (rb == 0) return false;
    if (ra == 2 || rb == 2) r->lflags |= LEG_F_INEXACT;
    return true;
}

// `mode` = 0 is process_sms_from_cache's per-entry work (process_cached.py:233-261).  The bits select what
// the reference's single-message functions do instead: parse_transaction_message / parse_credit_message
// (process_cached.py:150-204) apply ONE pattern to the message as given - no strip(), no OTP filter.
#define LEG_MODE_NO_STRIP 0x1
#define LEG_MODE_NO_OTP 0x2
#define LEG_MODE_NO_DEBIT 0x4
#define LEG_MODE_NO_CREDIT 0x8
SG_HD LegacyResult sg_legacy_process(const uint8_t* body, int len, const LegacyTables& L, uint32_t mode = 0) {
    LegacyResult r;
    r.status = LEG_FAILED; r.lflags = 0; r.amount = 0; r.balance = 0;
    for (int k = 0; k < 2 * RX_MAX_GROUPS; ++k) r.caps[k] = -1;
    int lo = 0, hi = len;
    if (!(mode & LEG_MODE_NO_STRIP)) sg_strip_span(body, &lo, &hi);      // process_cached.py:233
    r.strip_lo = lo; r.strip_hi = hi;
    const uint8_t* t = body + lo;
    const int tl = hi - lo;
    sg_sha256(t, tl, r.digest);                                          // :254 (computed for every body)
    // OTP / PASS= / CODE: in body_text.upper() (:236)
    uint32_t st = 0;
    for (int i = 0; i < tl && !(mode & LEG_MODE_NO_OTP); ++i) {
        const uint32_t cw = T_BYTE_CLASS(L.scan)[t[i]];
        if ((int32_t)cw < 0) {
            uint32_t cp, up[3];
            sg_utf8_decode(t + i, tl - i, &cp);
            const int nu = sg_upper_special(cp, up);
            if (nu == 0) st = sg_step(L.otp_next, st, CI_CLS_OTHER);
            for (int k = 0; k < nu; ++k) st = sg_step(L.otp_next, st, up[k]);
        } else st = sg_step(L.otp_next, st, sg_byte(cw, 0));
    }
    if (st == LEG_OTP_ACCEPT_OFF) { r.status = LEG_SKIPPED_OTP; return r; }
    int caps[2 * RX_MAX_GROUPS];
    int m = (mode & LEG_MODE_NO_DEBIT) ? 0 : sg_rx_exec(L.debit, t, tl, caps, L);   // :243 parse_transaction_message
    if (m < 0) r.lflags |= LEG_F_BUDGET;
    if (m == 1 && sg_legacy_numbers(t, caps, RX_DEBIT_G_AMOUNT, RX_DEBIT_G_BALANCE, L, &r)) {
        r.status = LEG_DEBIT;
        const int strip_groups[] = {RX_DEBIT_G_MERCHANT, RX_DEBIT_G_ADDRESS, RX_DEBIT_G_DATE, RX_DEBIT_G_TIME, RX_DEBIT_G_CURRENCY};
        for (int g : strip_groups) sg_strip_span(t, &caps[2 * g], &caps[2 * g + 1]);
        for (int k = 0; k < 2 * RX_MAX_GROUPS; ++k) r.caps[k] = caps[k];
        return r;
    }
    r.lflags &= ~LEG_F_INEXACT;
    m = (mode & LEG_MODE_NO_CREDIT) ? 0 : sg_rx_exec(L.credit, t, tl, caps, L);     // :248 parse_credit_message
    if (m < 0) r.lflags |= LEG_F_BUDGET;
    if (m == 1 && sg_legacy_numbers(t, caps, RX_CREDIT_G_AMOUNT, RX_CREDIT_G_BALANCE, L, &r)) {
        r.status = LEG_CREDIT;
        const int strip_groups[] = {RX_CREDIT_G_TYPE, RX_CREDIT_G_DATE, RX_CREDIT_G_TIME, RX_CREDIT_G_CURRENCY};
        for (int g : strip_groups) sg_strip_span(t, &caps[2 * g], &caps[2 * g + 1]);
        for (int k = 0; k < 2 * RX_MAX_GROUPS; ++k) r.caps[k] = caps[k];
        return r;
    }
    return r;
}

// ---------------------------------------------------------------------------
// save_to_pocketbase.py:34-43, `_dt_str(date, time_)`: datetime.strptime(f"{date} {time_}", f"{fmt} %H:%M")
// for fmt in %d.%m.%Y, %d/%m/%Y, %d-%m-%Y, %d.%m.%y, %d/%m/%y, %d-%m-%y; the first that parses wins.
// CPython turns such a format into the regular expression (Lib/_strptime.py, TimeRE)
//     (3[0-1]|[1-2]\d|0[1-9]|[1-9]| [1-9]) SEP (1[0-2]|0[1-9]|[1-9]) SEP (\d\d\d\d | \d\d) \s+
//     (2[0-3]|[0-1]\d|\d) : ([0-5]\d|\d)
// matches it at the start of the string (first alternative first, backtracking), demands that the match
// ends where the string ends, converts the groups with int() and builds a datetime (calendar validity,
// year >= 1; %y pivots at 69).  \d and \s are the Unicode classes, the bracketed digits are ASCII.
// Returns the packed datetime, 0 where every format raises.
// ---------------------------------------------------------------------------
struct DtCursor {
    const uint8_t* s;
    int len;
    const ScanTables* T;
    // code point at byte offset p: returns its length in bytes (0 at the end), *cp the code point
    SG_HD int at(int p, uint32_t* cp) const {
        if (p >= len) return 0;
        return sg_utf8_decode(s + p, len - p, cp);
    }
    // a \d at p: digit value or -1; *nb the bytes it occupies
    SG_HD int digit(int p, int* nb) const {
        uint32_t cp;
        *nb = at(p, &cp);
        if (*nb == 0) return -1;
        return sg_digit_value(cp, T->t->digit_run_start, T->t->digit_run_len);
    }
    SG_HD int ascii(int p) const { return p < len ? (int)s[p] : -1; }
};

// the alternatives of one numeric field, tried in the regular expression's order; alternative `k` of
// field `f` matched at `p` -> value and end offset, or false
SG_HD bool sg_dt_alt(const DtCursor& C, int field, int k, int p, int* val, int* end) {
    const int c0 = C.ascii(p), c1 = C.ascii(p + 1);
    int nb, d;
    switch (field) {
        case 0:     // %d
            if (k == 0) { if (c0 == '3' && (c1 == '0' || c1 == '1')) { *val = 30 + c1 - '0'; *end = p + 2; return true; } return false; }
            if (k == 1) { if (c0 == '1' || c0 == '2') { d = C.digit(p + 1, &nb); if (d >= 0) { *val = (c0 - '0') * 10 + d; *end = p + 1 + nb; return true; } } return false; }
            if (k == 2) { if (c0 == '0' && c1 >= '1' && c1 <= '9') { *val = c1 - '0'; *end = p + 2; return true; } return false; }
            if (k == 3) { if (c0 >= '1' && c0 <= '9') { *val = c0 - '0'; *end = p + 1; return true; } return false; }
            if (k == 4) { if (c0 == ' ' && c1 >= '1' && c1 <= '9') { *val = c1 - '0'; *end = p + 2; return true; } return false; }
            return false;
        case 1:     // %m
            if (k == 0) { if (c0 == '1' && c1 >= '0' && c1 <= '2') { *val = 10 + c1 - '0'; *end = p + 2; return true; } return false; }
            if (k == 1) { if (c0 == '0' && c1 >= '1' && c1 <= '9') { *val = c1 - '0'; *end = p + 2; return true; } return false; }
            if (k == 2) { if (c0 >= '1' && c0 <= '9') { *val = c0 - '0'; *end = p + 1; return true; } return false; }
            return false;
        case 2:     // %H
            if (k == 0) { if (c0 == '2' && c1 >= '0' && c1 <= '3') { *val = 20 + c1 - '0'; *end = p + 2; return true; } return false; }
            if (k == 1) { if (c0 == '0' || c0 == '1') { d = C.digit(p + 1, &nb); if (d >= 0) { *val = (c0 - '0') * 10 + d; *end = p + 1 + nb; return true; } } return false; }
            if (k == 2) { d = C.digit(p, &nb); if (d >= 0) { *val = d; *end = p + nb; return true; } return false; }
            return false;
        default:    // %M
            if (k == 0) { if (c0 >= '0' && c0 <= '5') { d = C.digit(p + 1, &nb); if (d >= 0) { *val = (c0 - '0') * 10 + d; *end = p + 1 + nb; return true; } } return false; }
            if (k == 1) { d = C.digit(p, &nb); if (d >= 0) { *val = d; *end = p + nb; return true; } return false; }
            return false;
    }
}

// one format: separator `sep`, `ydigits` 4 (%Y) or 2 (%y).  The first complete match of the regular
// expression in backtracking order decides; returns the packed datetime or 0.
SG_HD int64_t sg_dt_try(const DtCursor& C, int sep, int ydigits) {
    for (int kd = 0; kd < 5; ++kd) {
        int day, p1;
        if (!sg_dt_alt(C, 0, kd, 0, &day, &p1) || C.ascii(p1) != sep) continue;
        for (int km = 0; km < 3; ++km) {
            int mon, p2;
            if (!sg_dt_alt(C, 1, km, p1 + 1, &mon, &p2) || C.ascii(p2) != sep) continue;
            int year = 0, p3 = p2 + 1;
            bool ok = true;
            for (int k = 0; k < ydigits; ++k) {
                int nb;
                const int d = C.digit(p3, &nb);
                if (d < 0) { ok = false; break; }
                year = year * 10 + d;
                p3 += nb;
            }
            if (!ok) continue;
            int p4 = p3;                                   // \s+ (greedy; the hour starts with a digit, so
            for (;;) {                                     //  giving white space back never helps)
                uint32_t cp;
                const int nb = C.at(p4, &cp);
                if (nb == 0 || !sg_is_space(cp)) break;
                p4 += nb;
            }
            if (p4 == p3) continue;
            for (int kh = 0; kh < 3; ++kh) {
                int hour, p5;
                if (!sg_dt_alt(C, 2, kh, p4, &hour, &p5) || C.ascii(p5) != ':') continue;
                for (int kn = 0; kn < 2; ++kn) {
                    int minute, p6;
                    if (!sg_dt_alt(C, 3, kn, p5 + 1, &minute, &p6)) continue;
                    // re.match found its match: everything after this is strptime's own checking
                    if (p6 != C.len) return 0;                                  // "unconverted data remains"
                    if (ydigits == 2) year += (year < 69) ? 2000 : 1900;
                    if (year < 1) return 0;                                     // datetime.MINYEAR
                    if (day > sg_days_in_month(year, mon)) return 0;
                    return sg_pack_datetime(year, mon, day, hour, minute, 0, 0);
                }
            }
        }
    }
    return 0;
}

SG_HD int64_t sg_dt_str(const uint8_t* s, int len, const ScanTables& T) {
    DtCursor C;
    C.s = s; C.len = len; C.T = &T;
    const int seps[3] = {'.', '/', '-'};
    for (int yd = 4; yd >= 2; yd -= 2)
        for (int k = 0; k < 3; ++k) {
            const int64_t v = sg_dt_try(C, seps[k], yd);
            if (v) return v;
        }
    return 0;
}
