// Per-message scan logic of the keys kernel: skip-filter automata, body-date
// and card-form detection, the clean+mask rewrite, and SHA-1 / SHA-256.
// One message = one thread; `m` points at the message's bytes (shared memory
// on the device).  Reference semantics are cited per function.
#pragma once
#include "sg_common.cuh"

// All lookup tables of the scan in one block (copied to shared memory once per CTA).
struct DeviceTables {
    uint32_t byte_class[256];                    // byte0 = CI class*2, byte1 = CS class*2, byte2 = DM class*2, bit31 = special lead
    uint32_t digit_run_start[N_DIGIT_RUNS];
    uint16_t ci_next[CI_STATES * SCAN_STRIDE];   // byte offset of the next row
    uint16_t cs_next[CS_STATES * SCAN_STRIDE];
    uint16_t ci2_next[CI2_STATES * SCAN_STRIDE]; // first-pass automaton: worker needles + upper-cased case-sensitive ones
    uint16_t dm_next[DM_STATES * DM_STRIDE];
    uint8_t dm_has_card[DM_STATES];
    uint8_t digit_run_len[N_DIGIT_RUNS];
    uint8_t card_kmp[48];
};
static_assert(sizeof(DeviceTables) % 16 == 0, "tables are copied 16 bytes at a time");

inline void sg_fill_tables(DeviceTables* h) {
    for (int k = 0; k < 256; ++k) h->byte_class[k] = H_BYTE_CLASS[k];
    for (int k = 0; k < N_DIGIT_RUNS; ++k) { h->digit_run_start[k] = H_DIGIT_RUN_START[k]; h->digit_run_len[k] = H_DIGIT_RUN_LEN[k]; }
    for (int k = 0; k < CI_STATES * SCAN_STRIDE; ++k) h->ci_next[k] = H_CI_NEXT[k];
    for (int k = 0; k < CS_STATES * SCAN_STRIDE; ++k) h->cs_next[k] = H_CS_NEXT[k];
    for (int k = 0; k < CI2_STATES * SCAN_STRIDE; ++k) h->ci2_next[k] = H_CI2_NEXT[k];
    for (int k = 0; k < DM_STATES * DM_STRIDE; ++k) h->dm_next[k] = H_DM_NEXT[k];
    for (int k = 0; k < DM_STATES; ++k) h->dm_has_card[k] = H_DM_HAS_CARD[k];
    for (int k = 0; k < 48; ++k) h->card_kmp[k] = H_CARD_KMP[k];
}

// A view of the tables: one pointer, so it travels in registers.
struct ScanTables {
    const DeviceTables* t;
};
#define T_BYTE_CLASS(T) ((T).t->byte_class)
#define T_CI(T) (reinterpret_cast<const uint8_t*>((T).t->ci_next))
#define T_CS(T) (reinterpret_cast<const uint8_t*>((T).t->cs_next))
#define T_CI2(T) (reinterpret_cast<const uint8_t*>((T).t->ci2_next))
#define T_DM(T) (reinterpret_cast<const uint8_t*>((T).t->dm_next))

// one automaton step: tables hold the byte offset of the next row, classes come pre-multiplied by 2
SG_HD uint32_t sg_step(const uint8_t* tab, uint32_t state, uint32_t cls2) {
    return *reinterpret_cast<const uint16_t*>(tab + state + cls2);
}
SG_HD uint32_t sg_byte(uint32_t w, int k) {
#ifdef __CUDA_ARCH__
    return __byte_perm(w, 0, 0x4440 | k);
#else
    return (w >> (8 * k)) & 0xFF;
#endif
}

struct ScanResult {
    uint32_t flags;      // SG_F_*
    int32_t body_date;   // y<<9|m<<5|d or 0
};

// Symbols the CI automaton is fed for a code point whose str.upper() contains
// ASCII (17 code points, Unicode 15.0; see gen_tables.py).  Returns the count
// and up to three CI classes; 0 means "not special".
SG_HD int sg_upper_special(uint32_t cp, uint32_t cls[3]) {
    switch (cp) {
        case 0x00DF: cls[0] = CI_CLS_S; cls[1] = CI_CLS_S; return 2;               // ß -> SS
        case 0x0131: cls[0] = CI_CLS_I; return 1;                                  // ı -> I
        case 0x0149: cls[0] = CI_CLS_OTHER; cls[1] = CI_CLS_N; return 2;           // ŉ -> ʼN
        case 0x017F: cls[0] = CI_CLS_S; return 1;                                  // ſ -> S
        case 0x01F0: cls[0] = CI_CLS_OTHER; cls[1] = CI_CLS_OTHER; return 2;       // ǰ -> J + U+030C (J is in no needle)
        case 0x1E96: cls[0] = CI_CLS_H; cls[1] = CI_CLS_OTHER; return 2;           // ẖ
        case 0x1E97: cls[0] = CI_CLS_T; cls[1] = CI_CLS_OTHER; return 2;           // ẗ
        case 0x1E98: cls[0] = CI_CLS_OTHER; cls[1] = CI_CLS_OTHER; return 2;       // ẘ (W is in no needle)
        case 0x1E99: cls[0] = CI_CLS_Y; cls[1] = CI_CLS_OTHER; return 2;           // ẙ
        case 0x1E9A: cls[0] = CI_CLS_A; cls[1] = CI_CLS_OTHER; return 2;           // ẚ
        case 0xFB00: cls[0] = CI_CLS_F; cls[1] = CI_CLS_F; return 2;               // ﬀ
        case 0xFB01: cls[0] = CI_CLS_F; cls[1] = CI_CLS_I; return 2;               // ﬁ
        case 0xFB02: cls[0] = CI_CLS_F; cls[1] = CI_CLS_L; return 2;               // ﬂ
        case 0xFB03: cls[0] = CI_CLS_F; cls[1] = CI_CLS_F; cls[2] = CI_CLS_I; return 3;      // ﬃ
        case 0xFB04: cls[0] = CI_CLS_F; cls[1] = CI_CLS_F; cls[2] = CI_CLS_L; return 3;      // ﬄ
        case 0xFB05: cls[0] = CI_CLS_S; cls[1] = CI_CLS_T; return 2;               // ﬅ
        case 0xFB06: cls[0] = CI_CLS_S; cls[1] = CI_CLS_T; return 2;               // ﬆ
        default: return 0;
    }
}

// Step back `k` code points from byte offset `pos` (never below 0).
SG_HD int sg_back_cps(const uint8_t* m, int pos, int k) {
    for (; k > 0 && pos > 0; --k) {
        --pos;
        while (pos > 0 && (m[pos] & 0xC0) == 0x80) --pos;
    }
    return pos;
}

// The date `fix_broken_datetime` (libs/gemini_parser.py:67-104) would lift
// out of the body, given where the two regex forms first matched.  `lead` is
// the offset of the first byte of the match's last code point, `ncp` its
// length in code points (10 for dd.mm.yyyy, 8 for dd.mm.yy).  Implements the
// accept set of datetime.strptime's %d.%m.%Y / %d.%m.%y on a string that
// already matched \d{2}\.\d{2}\.\d{2,4} with Unicode digits: day
// 3[01]|[12]\d|0[1-9] and month 1[0-2]|0[1-9] need ASCII where the pattern
// names ASCII digits, the year accepts any Nd digits, then calendar validity.
#if defined(__CUDACC__)
__host__ __device__ __noinline__
#endif
static int32_t sg_body_date_at(const uint8_t* m, int len, int lead, int ncp, const ScanTables T) {
    uint64_t vals = 0;                                // 4 bits per code point, position k at bits 4k
    uint32_t ascm = 0;                                // bit k: code point k is ASCII
    int p = lead;
    if (lead >= ncp - 1) {
        // common case: the ncp bytes ending at `lead` are all ASCII, i.e. they ARE the match - digits at
        // 0,1,3,4,6.. and '.' at 2 and 5 are what the automaton matched: read the values straight off
        const uint8_t* q = m + lead - (ncp - 1);
        uint32_t any = q[0] | q[1] | q[2] | q[3] | q[4] | q[5] | q[6] | q[7];
        if (ncp == 10) any |= q[8] | q[9];
        if (!(any & 0x80u)) {
            const int d0 = q[0] - '0', d1 = q[1] - '0', m0 = q[3] - '0', m1 = q[4] - '0';
            const bool day_ok = (d0 == 3 && d1 <= 1) || d0 == 1 || d0 == 2 || (d0 == 0 && d1 >= 1);
            const bool mon_ok = (m0 == 1 && m1 <= 2) || (m0 == 0 && m1 >= 1);
            if (!day_ok || !mon_ok) return 0;
            const int day = d0 * 10 + d1, mon = m0 * 10 + m1;
            int year = (q[6] - '0') * 10 + (q[7] - '0');
            if (ncp == 10) {
                year = year * 100 + (q[8] - '0') * 10 + (q[9] - '0');
                if (year < 1) return 0;               // datetime.MINYEAR
            } else {
                year += (year < 69) ? 2000 : 1900;    // %y pivot (POSIX / _strptime)
            }
            if (day > sg_days_in_month(year, mon)) return 0;
            return (year << 9) | (mon << 5) | day;
        }
    }
    for (int k = ncp - 1; k >= 0; --k) {
        uint32_t cp;
        sg_utf8_decode(m + p, len - p, &cp);
        const int dv = sg_digit_value(cp, T.t->digit_run_start, T.t->digit_run_len);
        vals |= (uint64_t)(dv & 15) << (4 * k);
        ascm |= (cp < 0x80 ? 1u : 0u) << k;
        if (k) p = sg_back_cps(m, p, 1);
    }
    int v[10];
    bool asc[10];
#pragma unroll
    for (int k = 0; k < 10; ++k) { v[k] = (int)((vals >> (4 * k)) & 15); asc[k] = (ascm >> k) & 1; }
    const bool day_ok = (asc[0] && v[0] == 3 && asc[1] && v[1] <= 1) ||
                        (asc[0] && (v[0] == 1 || v[0] == 2)) ||
                        (asc[0] && v[0] == 0 && asc[1] && v[1] >= 1);
    const bool mon_ok = (asc[3] && v[3] == 1 && asc[4] && v[4] <= 2) ||
                        (asc[3] && v[3] == 0 && asc[4] && v[4] >= 1);
    if (!day_ok || !mon_ok) return 0;
    const int day = v[0] * 10 + v[1], mon = v[3] * 10 + v[4];
    int year;
    if (ncp == 10) {
        year = v[6] * 1000 + v[7] * 100 + v[8] * 10 + v[9];
        if (year < 1) return 0;                       // datetime.MINYEAR
    } else {
        year = v[6] * 10 + v[7];
        year += (year < 69) ? 2000 : 1900;            // %y pivot (POSIX / _strptime)
    }
    if (day > sg_days_in_month(year, mon)) return 0;
    return (year << 9) | (mon << 5) | day;
}

// Running state of the three automata over one body.
struct ScanState {
    uint32_t ci, cs, dm;     // row byte offsets in their tables
    uint32_t dm_event;       // next date-progress threshold the DM row offset has to reach
    int posB, posA;          // offset of the byte that completed the short / long date form
    uint32_t modified;       // U+00A0 or U+2022 seen
};

// A date-progress event (rows are sorted by progress): rare, at most twice per body.
SG_HD void sg_scan_event(ScanState& st, int i) {
    if (st.posB < 0) st.posB = i;
    if (st.dm >= DM_FIRST_A_OFF) { st.posA = i; st.dm_event = 0xFFFFFFFFu; }
    else st.dm_event = DM_FIRST_A_OFF;
}

// The scan comes in two forms.  FULL runs all three automata and yields every flag.  FIRST is the scan
// kernel's first pass: instead of the worker's upper-case automaton AND the case-sensitive one it steps a
// single upper-case automaton (worker needles + the upper-cased case-sensitive needles, CI2 in
// gen_tables.py) plus the date / card automaton.  If that automaton never accepts, no needle of either
// kind occurs (a case-sensitive occurrence upper-cases into an occurrence of its upper-cased form) and
// both filter flags are 0; if it accepts, the body is flagged SG_F_MAYBE and a FULL scan decides.  Only
// ~5 % of real bodies are flagged, so the common byte costs two table steps instead of three.

// One byte, any value.  A special lead byte (may upper-case into ASCII, be a
// Unicode digit, U+00A0 or U+2022) is decoded and its symbols fed explicitly;
// continuation bytes keep the CI / DM state, other lead bytes reset it.
template <bool FIRST>
#if defined(__CUDACC__)
__host__ __device__ __noinline__
#endif
static ScanState sg_scan_byte_careful(const uint8_t* m, int len, int i, const ScanTables T, ScanState st) {
    const uint8_t* ci_tab = FIRST ? T_CI2(T) : T_CI(T);
    const uint32_t cw = T_BYTE_CLASS(T)[m[i]];
    if (!FIRST) st.cs = sg_step(T_CS(T), st.cs, sg_byte(cw, 1));
    if ((int32_t)cw < 0) {
        uint32_t cp;
        sg_utf8_decode(m + i, len - i, &cp);
        uint32_t up[3];
        const int nu = sg_upper_special(cp, up);
        if (nu == 0) st.ci = sg_step(ci_tab, st.ci, CI_CLS_OTHER);
        for (int k = 0; k < nu; ++k) st.ci = sg_step(ci_tab, st.ci, up[k]);
        uint32_t dcls = DM_CLS_OTHER;
        if (cp == 0x2022) { dcls = DM_CLS_STAR; st.modified = 1; }
        else if (cp == 0xA0) st.modified = 1;
        else if (sg_digit_value(cp, T.t->digit_run_start, T.t->digit_run_len) >= 0) dcls = DM_CLS_DIGIT;
        st.dm = sg_step(T_DM(T), st.dm, dcls);
    } else {
        st.ci = sg_step(ci_tab, st.ci, sg_byte(cw, 0));
        st.dm = sg_step(T_DM(T), st.dm, sg_byte(cw, 2));
    }
    if (st.dm >= st.dm_event) sg_scan_event(st, i);
    return st;
}

// One pass over the body: both skip filters, NBSP / bullet / card-form
// detection (is fixed_body different?) and the first hits of the two body
// date forms.  worker.py:112-121, gemini_parser.py:198-203, :73-100, :131.
// Aligned 4-byte words that are pure ASCII take the branch-free fast path (three
// table steps per byte, two in the FIRST form); every other byte goes through sg_scan_byte_careful.
template <bool FIRST>
SG_HD ScanResult sg_scan_body_t(const uint8_t* m, int len, const ScanTables& T) {
    const uint8_t* ci_tab = FIRST ? T_CI2(T) : T_CI(T);
    ScanState st;
    st.ci = st.cs = st.dm = 0;
    st.dm_event = DM_FIRST_B_OFF;
    st.posB = st.posA = -1;
    st.modified = 0;
    int i = 0;
    const int head = (int)((4 - ((uintptr_t)m & 3)) & 3);
    for (; i < len && i < head; ++i) st = sg_scan_byte_careful<FIRST>(m, len, i, T, st);
    for (; i + 4 <= len; i += 4) {
        const uint32_t w4 = *reinterpret_cast<const uint32_t*>(m + i);
        if (w4 & 0x80808080u) {
            // A high byte that is not a special lead (bit 31 of its class word) is stepped from the class
            // table exactly as sg_scan_byte_careful's else-branch does: continuation bytes keep the CI / DM
            // state, ordinary leads feed OTHER.  Cyrillic, Armenian, Greek ... words therefore stay in the
            // branch-free loop; only a word holding a lead that may upper-case into ASCII, start a Unicode
            // digit, U+00A0 or U+2022 is taken byte by byte.
            const uint32_t any = T_BYTE_CLASS(T)[sg_byte(w4, 0)] | T_BYTE_CLASS(T)[sg_byte(w4, 1)] |
                                 T_BYTE_CLASS(T)[sg_byte(w4, 2)] | T_BYTE_CLASS(T)[sg_byte(w4, 3)];
            if ((int32_t)any < 0) {
#pragma unroll 1
                for (int k = 0; k < 4; ++k) st = sg_scan_byte_careful<FIRST>(m, len, i + k, T, st);
                continue;
            }
        }
        uint32_t ci = st.ci, cs = st.cs, dm = st.dm;
        const uint32_t dm_before = dm;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t cw = T_BYTE_CLASS(T)[sg_byte(w4, k)];
            ci = sg_step(ci_tab, ci, sg_byte(cw, 0));
            if (!FIRST) cs = sg_step(T_CS(T), cs, sg_byte(cw, 1));
            dm = sg_step(T_DM(T), dm, sg_byte(cw, 2));
        }
        st.ci = ci; st.cs = cs;
        if (dm >= st.dm_event) {
            // a date form completed inside this word: replay its four bytes one by one for the position
            st.dm = dm_before;
            for (int k = 0; k < 4; ++k) {
                st.dm = sg_step(T_DM(T), st.dm, sg_byte(T_BYTE_CLASS(T)[sg_byte(w4, k)], 2));
                if (st.dm >= st.dm_event) sg_scan_event(st, i + k);
            }
        } else st.dm = dm;
    }
    for (; i < len; ++i) st = sg_scan_byte_careful<FIRST>(m, len, i, T, st);
    ScanResult r;
    r.flags = 0;
    if (FIRST) {
        if (st.ci == CI2_ACCEPT_OFF) r.flags |= SG_F_MAYBE;
    } else {
        const bool parser_hit = st.cs == CS_ACCEPT_OFF;
        const bool daily = st.cs >= CS_DAILY_OFF;              // includes the parser-hit state
        if (st.ci == CI_ACCEPT_OFF || daily) r.flags |= SG_F_WORKER_SKIP;
        if (parser_hit) r.flags |= SG_F_PARSER_OTP;
    }
    if (st.modified || T.t->dm_has_card[st.dm / DM_ROW_BYTES]) r.flags |= SG_F_MODIFIED;
    int32_t bd = 0;
    if (st.posA >= 0) bd = sg_body_date_at(m, len, st.posA, 10, T);
    if (bd == 0 && st.posB >= 0) bd = sg_body_date_at(m, len, st.posB, 8, T);
    r.body_date = bd;
    if (bd) r.flags |= SG_F_BODY_DATE;
    return r;
}
SG_HD ScanResult sg_scan_body(const uint8_t* m, int len, const ScanTables& T) { return sg_scan_body_t<false>(m, len, T); }
SG_HD ScanResult sg_scan_first(const uint8_t* m, int len, const ScanTables& T) { return sg_scan_body_t<true>(m, len, T); }
// the two filter flags of a body the first pass flagged (kept out of line: it runs for few bodies)
#if defined(__CUDACC__)
__host__ __device__ __noinline__
#endif
static uint32_t sg_scan_filters(const uint8_t* m, int len, const ScanTables T) {
    return sg_scan_body_t<false>(m, len, T).flags & (SG_F_WORKER_SKIP | SG_F_PARSER_OTP);
}
// both passes one after the other, as the scan kernel composes them
SG_HD ScanResult sg_scan_two_pass(const uint8_t* m, int len, const ScanTables& T) {
    ScanResult r = sg_scan_first(m, len, T);
    if (r.flags & SG_F_MAYBE) r.flags = (r.flags & ~(uint32_t)SG_F_MAYBE) | sg_scan_filters(m, len, T);
    return r;
}

// clean + mask_card_number_with_prefix, in place (libs/gemini_parser.py:202-203,
// :121-137): U+00A0 -> ' ', U+2022 -> '*', then each leftmost non-overlapping
// \d{4}\*{3}(\d{4}) (Unicode \d) -> "CARD:" + group 1.  The output is never
// longer than the input, so the write cursor never passes the read cursor.
// Returns the new length.
SG_HD int sg_rewrite_fixed(uint8_t* m, int len, const ScanTables& T) {
    int r = 0, w = 0;
    uint32_t k = 0;                                   // card-form KMP progress
    while (r < len) {
        const uint32_t b = m[r];
        uint32_t cls;
        if (b < 0x80) {                               // ASCII: class straight from the byte-class word
            m[w++] = (uint8_t)b;
            ++r;
            cls = sg_byte(T_BYTE_CLASS(T)[b], 2) >> 1;   // DM class ids 0 other, 1 digit, 2 '.', 3 '*'
        } else {
            if ((int32_t)T_BYTE_CLASS(T)[b] >= 0 && b >= 0xC2 && b <= 0xF4) {
                // a lead that starts neither U+00A0, U+2022 nor a Unicode digit: one OTHER symbol, copied as is
                int n = (b >= 0xF0) ? 4 : (b >= 0xE0) ? 3 : 2;
                bool ok = n <= len - r;
                for (int j = 1; ok && j < n; ++j) ok = (m[r + j] & 0xC0) == 0x80;
                if (!ok) n = 1;                       // malformed: one byte, as sg_utf8_decode reports it
                for (int j = 0; j < n; ++j) m[w + j] = m[r + j];
                w += n;
                r += n;
                k = T.t->card_kmp[k * 4 + KMP_CLS_OTHER];
                continue;
            }
            uint32_t cp;
            const int n = sg_utf8_decode(m + r, len - r, &cp);
            if (cp == 0xA0) { m[w++] = ' '; cls = KMP_CLS_OTHER; }
            else if (cp == 0x2022) { m[w++] = '*'; cls = KMP_CLS_STAR; }
            else {
                for (int j = 0; j < n; ++j) m[w + j] = m[r + j];
                w += n;
                cls = (sg_digit_value(cp, T.t->digit_run_start, T.t->digit_run_len) >= 0) ? KMP_CLS_DIGIT : KMP_CLS_OTHER;
            }
            r += n;
        }
        k = T.t->card_kmp[k * 4 + cls];
        if (k == 11) {
            const int g = sg_back_cps(m, w, 4);       // start of the captured group
            const int s = sg_back_cps(m, g - 3, 4);   // start of the match
            const int glen = w - g;
            m[s] = 'C'; m[s + 1] = 'A'; m[s + 2] = 'R'; m[s + 3] = 'D'; m[s + 4] = ':';
            for (int j = 0; j < glen; ++j) m[s + 5 + j] = m[g + j];   // moves left by >= 2
            w = s + 5 + glen;
            k = 0;
        }
    }
    return w;
}

// ---------------------------------------------------------------------------
// SHA-1 / SHA-256 (FIPS 180-4) over a byte string in place, big-endian words
// assembled from unaligned bytes.  hashlib.sha1(body.encode()).hexdigest() is
// the XML idempotency key (services/xml_watcher/watcher.py:45); hashlib.sha256
// of the rewritten body is the Gemini cache key (libs/gemini_parser.py:207).
// ---------------------------------------------------------------------------
SG_HD uint32_t sg_rotl(uint32_t x, int n) {
#ifdef __CUDA_ARCH__
    return __funnelshift_l(x, x, n);
#else
    return (x << n) | (x >> (32 - n));
#endif
}
SG_HD uint32_t sg_rotr(uint32_t x, int n) {
#ifdef __CUDA_ARCH__
    return __funnelshift_r(x, x, n);
#else
    return (x >> n) | (x << (32 - n));
#endif
}

// Word reader over an arbitrarily aligned byte string: big-endian 32-bit word
// number `j` of the string.  The caller guarantees 4 readable bytes past any
// in-range word start rounded down to 4 (staging buffers are padded).
struct WordReader {
    const uint32_t* base;   // 4-byte aligned address at or below the first byte
    uint32_t shift;         // (first byte address & 3) * 8
    SG_HD void init(const uint8_t* m) {
        // pointer arithmetic on `m` itself (no integer round trip) keeps the
        // compiler's address-space inference: LDS for staged tiles, LDG otherwise
        const uint32_t mis = (uint32_t)((uintptr_t)m & 3);
        base = (const uint32_t*)(m - mis);
        shift = mis * 8;
    }
    SG_HD uint32_t le(int j) const {
        const uint32_t lo = base[j], hi = base[j + 1];
#ifdef __CUDA_ARCH__
        return __funnelshift_r(lo, hi, shift);
#else
        return shift ? ((lo >> shift) | (hi << (32 - shift))) : lo;
#endif
    }
    SG_HD uint32_t be(int j) const {
        const uint32_t lo = base[j], hi = base[j + 1];
#ifdef __CUDA_ARCH__
        const uint32_t le = __funnelshift_r(lo, hi, shift);
        return __byte_perm(le, 0, 0x0123);
#else
        const uint32_t le = shift ? ((lo >> shift) | (hi << (32 - shift))) : lo;
        return ((le & 0xFF) << 24) | ((le & 0xFF00) << 8) | ((le >> 8) & 0xFF00) | (le >> 24);
#endif
    }
};

// Word `j` of the padded message (0x80, zeros, 64-bit bit length), branch-free:
// k = message bytes left at this word; k >= 4 keeps the word, 0 <= k < 4 keeps k
// bytes and plants the 0x80 terminator, k < 0 is zero fill; the last two words
// carry the bit length.
#ifndef SG_OLD_PAD
// `endw` = index of the word that holds the 0x80 terminator: words before it are message words, it keeps the
// message's last len & 3 bytes, words after it are zero; the last two words carry the bit length.  A compare and
// two selects per word (the form below it needs a dozen instructions per word).
SG_HD uint32_t sg_padded_word(const WordReader& R, int len, int j, int nwords) {
    const int endw = len >> 2;
    const uint32_t sh = 8u * (uint32_t)(len & 3);
    uint32_t w = 0;
    if (j <= endw) {
        const uint32_t raw = R.be(j);
        w = j < endw ? raw : ((raw & ~(0xFFFFFFFFu >> sh)) | (0x80000000u >> sh));
    }
    if (j == nwords - 1) w |= (uint32_t)len << 3;
    if (j == nwords - 2) w |= (uint32_t)len >> 29;
    return w;
}
#else
SG_HD uint32_t sg_padded_word(const WordReader& R, int len, int j, int nwords) {
    const int k = len - j * 4;
    const uint32_t raw = R.be(j);
    const uint32_t kk = (uint32_t)(k < 0 ? 0 : (k > 4 ? 4 : k));
    const uint32_t keep = kk == 0 ? 0u : (0xFFFFFFFFu << (32 - 8 * kk));      // kk in 1..4
    uint32_t w = raw & keep;
    if (k >= 0 && k < 4) w |= 0x80000000u >> (8 * kk);
    if (j == nwords - 1) w |= (uint32_t)len << 3;
    if (j == nwords - 2) w |= (uint32_t)len >> 29;
    return w;
}
#endif

SG_HD int sg_sha_blocks(int len) { return (len + 9 + 63) >> 6; }

SG_HD void sg_sha1_compress(uint32_t h[5], uint32_t w[16]) {
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4];
#pragma unroll
    for (int i = 0; i < 80; ++i) {
        uint32_t wi;
        if (i < 16) wi = w[i];
        else {
            wi = sg_rotl(w[(i - 3) & 15] ^ w[(i - 8) & 15] ^ w[(i - 14) & 15] ^ w[i & 15], 1);
            w[i & 15] = wi;
        }
        uint32_t f, k;
        if (i < 20) { f = (b & c) | (~b & d); k = 0x5A827999u; }
        else if (i < 40) { f = b ^ c ^ d; k = 0x6ED9EBA1u; }
        else if (i < 60) { f = (b & c) | (b & d) | (c & d); k = 0x8F1BBCDCu; }
        else { f = b ^ c ^ d; k = 0xCA62C1D6u; }
        const uint32_t t = sg_rotl(a, 5) + f + e + k + wi;
        e = d; d = c; c = sg_rotl(b, 30); b = a; a = t;
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e;
}

SG_HD void sg_sha1(const uint8_t* m, int len, uint32_t h[5]) {
    h[0] = 0x67452301u; h[1] = 0xEFCDAB89u; h[2] = 0x98BADCFEu; h[3] = 0x10325476u; h[4] = 0xC3D2E1F0u;
    WordReader R;
    R.init(m);
    const int nblk = sg_sha_blocks(len), nwords = nblk * 16;
    for (int blk = 0; blk < nblk; ++blk) {
        uint32_t w[16];
        if ((blk + 1) * 64 <= len) {
#pragma unroll
            for (int j = 0; j < 16; ++j) w[j] = R.be(blk * 16 + j);
        } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) w[j] = sg_padded_word(R, len, blk * 16 + j, nwords);
        }
        sg_sha1_compress(h, w);
    }
}

// ---------------------------------------------------------------------------
// MD5 (RFC 1321): the idempotency key of a device-sourced message,
// hashlib.md5(message.encode('utf-8')).hexdigest() (libs/models.py:97-109, used at
// services/api_gateway/main.py:113).  Little-endian words straight from the bytes.
// ---------------------------------------------------------------------------
#define SG_MD5_K_INIT { \
    0xd76aa478u, 0xe8c7b756u, 0x242070dbu, 0xc1bdceeeu, 0xf57c0fafu, 0x4787c62au, 0xa8304613u, 0xfd469501u, \
    0x698098d8u, 0x8b44f7afu, 0xffff5bb1u, 0x895cd7beu, 0x6b901122u, 0xfd987193u, 0xa679438eu, 0x49b40821u, \
    0xf61e2562u, 0xc040b340u, 0x265e5a51u, 0xe9b6c7aau, 0xd62f105du, 0x02441453u, 0xd8a1e681u, 0xe7d3fbc8u, \
    0x21e1cde6u, 0xc33707d6u, 0xf4d50d87u, 0x455a14edu, 0xa9e3e905u, 0xfcefa3f8u, 0x676f02d9u, 0x8d2a4c8au, \
    0xfffa3942u, 0x8771f681u, 0x6d9d6122u, 0xfde5380cu, 0xa4beea44u, 0x4bdecfa9u, 0xf6bb4b60u, 0xbebfbc70u, \
    0x289b7ec6u, 0xeaa127fau, 0xd4ef3085u, 0x04881d05u, 0xd9d4d039u, 0xe6db99e5u, 0x1fa27cf8u, 0xc4ac5665u, \
    0xf4292244u, 0x432aff97u, 0xab9423a7u, 0xfc93a039u, 0x655b59c3u, 0x8f0ccc92u, 0xffeff47du, 0x85845dd1u, \
    0x6fa87e4fu, 0xfe2ce6e0u, 0xa3014314u, 0x4e0811a1u, 0xf7537e82u, 0xbd3af235u, 0x2ad7d2bbu, 0xeb86d391u}

// word j of the MD5-padded message (0x80, zeros, 64-bit little-endian bit length), branch-free like
// sg_padded_word but little-endian: the first message byte is the LOW byte of word 0
SG_HD uint32_t sg_md5_word(const WordReader& R, int len, int j, int nwords) {
    const int k = len - j * 4;
    const uint32_t raw = k > 0 ? R.le(j) : 0u;       // padding words are not read: MD5 runs straight off global memory,
                                                     // and the caller owes only 8 readable bytes past the last body
    const uint32_t kk = (uint32_t)(k < 0 ? 0 : (k > 4 ? 4 : k));
    const uint32_t keep = kk == 0 ? 0u : (0xFFFFFFFFu >> (32 - 8 * kk));
    uint32_t w = raw & keep;
    if (k >= 0 && k < 4) w |= 0x80u << (8 * kk);
    if (j == nwords - 2) w |= (uint32_t)len << 3;
    if (j == nwords - 1) w |= (uint32_t)len >> 29;
    return w;
}

SG_HD void sg_md5(const uint8_t* m, int len, uint32_t h[4]) {
    const uint32_t K[64] = SG_MD5_K_INIT;
    const int S[16] = {7, 12, 17, 22, 5, 9, 14, 20, 4, 11, 16, 23, 6, 10, 15, 21};
    h[0] = 0x67452301u; h[1] = 0xefcdab89u; h[2] = 0x98badcfeu; h[3] = 0x10325476u;
    WordReader R;
    R.init(m);
    const int nblk = sg_sha_blocks(len), nwords = nblk * 16;   // same padding rule as SHA: 1 + 8 bytes
    for (int blk = 0; blk < nblk; ++blk) {
        uint32_t w[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) w[j] = sg_md5_word(R, len, blk * 16 + j, nwords);
        uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
#pragma unroll
        for (int i = 0; i < 64; ++i) {
            uint32_t f;
            int g;
            if (i < 16) { f = (b & c) | (~b & d); g = i; }
            else if (i < 32) { f = (d & b) | (~d & c); g = (5 * i + 1) & 15; }
            else if (i < 48) { f = b ^ c ^ d; g = (3 * i + 5) & 15; }
            else { f = c ^ (b | ~d); g = (7 * i) & 15; }
            const uint32_t t = a + f + K[i] + w[g];
            a = d; d = c; c = b;
            b = b + sg_rotl(t, S[(i >> 4) * 4 + (i & 3)]);
        }
        h[0] += a; h[1] += b; h[2] += c; h[3] += d;
    }
}

#define SG_K256_INIT { \
    0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u, 0xab1c5ed5u, \
    0xd807aa98u, 0x12835b01u, 0x243185beu, 0x550c7dc3u, 0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u, \
    0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau, \
    0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u, 0x06ca6351u, 0x14292967u, \
    0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u, \
    0xa2bfe8a1u, 0xa81a664bu, 0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u, \
    0x19a4c116u, 0x1e376c08u, 0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu, 0x682e6ff3u, \
    0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u, 0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u}
static const uint32_t h_k256[64] = SG_K256_INIT;
#if defined(__CUDACC__)
static __constant__ uint32_t c_k256[64] = SG_K256_INIT;   // constant-bank operands once the rounds are unrolled
#endif
#ifdef __CUDA_ARCH__
#define SG_K256 c_k256
#else
#define SG_K256 h_k256
#endif

// Rotations through the FMA pipe (variant SG_SHA256_IMAD): x * 2^(32-n) as a 64-bit product holds x >> n in
// its high word and x << (32-n) in its low word, and the two are bit-disjoint, so rotr(x, n) = hi ^ lo.  One
// IMAD.WIDE (fma pipe, otherwise ~8 % busy in this kernel) replaces one SHF (alu pipe, the kernel's limiter);
// the XORs of the halves fold into 3-input LOP3s.  The multipliers sit in constant memory so that ptxas
// cannot turn the multiplication back into a shift.
#if defined(__CUDACC__) && defined(SG_SHA256_IMAD)
static __constant__ uint32_t c_pow2[33] = {
    1u << 0, 1u << 1, 1u << 2, 1u << 3, 1u << 4, 1u << 5, 1u << 6, 1u << 7, 1u << 8, 1u << 9, 1u << 10, 1u << 11,
    1u << 12, 1u << 13, 1u << 14, 1u << 15, 1u << 16, 1u << 17, 1u << 18, 1u << 19, 1u << 20, 1u << 21, 1u << 22,
    1u << 23, 1u << 24, 1u << 25, 1u << 26, 1u << 27, 1u << 28, 1u << 29, 1u << 30, 1u << 31, 0};
#endif
#if defined(__CUDA_ARCH__) && defined(SG_SHA256_IMAD)
__device__ __forceinline__ uint2 sg_mw(uint32_t x, int n) {          // {x << (32-n), x >> n}
    unsigned long long p;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(p) : "r"(x), "r"(c_pow2[32 - n]));
    return make_uint2((uint32_t)p, (uint32_t)(p >> 32));
}
__device__ __forceinline__ uint32_t sg_big_sigma(uint32_t x, int n0, int n1, int n2) {
    const uint2 a = sg_mw(x, n0), b = sg_mw(x, n1), c = sg_mw(x, n2);
    return (a.x ^ a.y ^ b.x) ^ (b.y ^ c.x ^ c.y);
}
__device__ __forceinline__ uint32_t sg_small_sigma(uint32_t x, int n0, int n1, int sh) {
    const uint2 a = sg_mw(x, n0), b = sg_mw(x, n1), c = sg_mw(x, sh);
    return (a.x ^ a.y ^ b.x) ^ (b.y ^ c.y);
}
#define SG_BIG_SIGMA(x, n0, n1, n2) sg_big_sigma(x, n0, n1, n2)
#define SG_SMALL_SIGMA(x, n0, n1, sh) sg_small_sigma(x, n0, n1, sh)
#else
#define SG_BIG_SIGMA(x, n0, n1, n2) (sg_rotr(x, n0) ^ sg_rotr(x, n1) ^ sg_rotr(x, n2))
#define SG_SMALL_SIGMA(x, n0, n1, sh) (sg_rotr(x, n0) ^ sg_rotr(x, n1) ^ ((x) >> (sh)))
#endif

SG_HD void sg_sha256_compress(uint32_t h[8], uint32_t w[16]) {
    // 4 x 16 rounds: the 16-round body is unrolled (static schedule indices), the outer loop is not,
    // which keeps the kernel's instruction footprint inside the instruction cache
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
#pragma unroll 1
    for (int r = 0; r < 64; r += 16) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            uint32_t wi;
            if (r == 0) wi = w[j];
            else {
                const uint32_t w15 = w[(j + 1) & 15], w2 = w[(j + 14) & 15];
                const uint32_t s0 = SG_SMALL_SIGMA(w15, 7, 18, 3);
                const uint32_t s1 = SG_SMALL_SIGMA(w2, 17, 19, 10);
                wi = w[j] + s0 + w[(j + 9) & 15] + s1;
                w[j] = wi;
            }
            const uint32_t S1 = SG_BIG_SIGMA(e, 6, 11, 25);
            const uint32_t ch = (e & f) ^ (~e & g);
            const uint32_t S0 = SG_BIG_SIGMA(a, 2, 13, 22);
            const uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
            const uint32_t t1 = hh + S1 + ch + SG_K256[r + j] + wi;
            const uint32_t t2 = S0 + mj;
            hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
        }
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
}

SG_HD void sg_sha256(const uint8_t* m, int len, uint32_t h[8]) {
    h[0] = 0x6a09e667u; h[1] = 0xbb67ae85u; h[2] = 0x3c6ef372u; h[3] = 0xa54ff53au;
    h[4] = 0x510e527fu; h[5] = 0x9b05688cu; h[6] = 0x1f83d9abu; h[7] = 0x5be0cd19u;
    WordReader R;
    R.init(m);
    const int nblk = sg_sha_blocks(len), nwords = nblk * 16;
    for (int blk = 0; blk < nblk; ++blk) {
        uint32_t w[16];
        if ((blk + 1) * 64 <= len) {
#pragma unroll
            for (int j = 0; j < 16; ++j) w[j] = R.be(blk * 16 + j);
        } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) w[j] = sg_padded_word(R, len, blk * 16 + j, nwords);
        }
        sg_sha256_compress(h, w);
    }
}

// SHA-1 and SHA-256 of the same byte string in one walk: each 64-byte block is
// loaded and byte-swapped once and fed to both compressions.
SG_HD void sg_sha1_sha256(const uint8_t* m, int len, uint32_t h1[5], uint32_t h2[8]) {
    h1[0] = 0x67452301u; h1[1] = 0xEFCDAB89u; h1[2] = 0x98BADCFEu; h1[3] = 0x10325476u; h1[4] = 0xC3D2E1F0u;
    h2[0] = 0x6a09e667u; h2[1] = 0xbb67ae85u; h2[2] = 0x3c6ef372u; h2[3] = 0xa54ff53au;
    h2[4] = 0x510e527fu; h2[5] = 0x9b05688cu; h2[6] = 0x1f83d9abu; h2[7] = 0x5be0cd19u;
    WordReader R;
    R.init(m);
    const int nblk = sg_sha_blocks(len), nwords = nblk * 16;
    for (int blk = 0; blk < nblk; ++blk) {
        uint32_t w[16], v[16];
        if ((blk + 1) * 64 <= len) {
#pragma unroll
            for (int j = 0; j < 16; ++j) w[j] = R.be(blk * 16 + j);
        } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) w[j] = sg_padded_word(R, len, blk * 16 + j, nwords);
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = w[j];
        sg_sha1_compress(h1, w);
        sg_sha256_compress(h2, v);
    }
}

// One walk over a byte string feeding SHA-1 (mode bit 0) and / or SHA-256 (mode bit 1); a digest whose bit
// is clear is left untouched.  The scan kernel has exactly one inlined copy of this loop and runs it once
// for an untouched body (mode 3) and twice for a rewritten one (mode 1 before the rewrite, mode 2 after):
// the hashing code is most of the kernel's instruction footprint, and one copy stays in the instruction cache.
SG_HD void sg_hash_pass(const uint8_t* m, int len, uint32_t mode, uint32_t h1[5], uint32_t h2[8]) {
    if (mode & 1u) {
        h1[0] = 0x67452301u; h1[1] = 0xEFCDAB89u; h1[2] = 0x98BADCFEu; h1[3] = 0x10325476u; h1[4] = 0xC3D2E1F0u;
    }
    if (mode & 2u) {
        h2[0] = 0x6a09e667u; h2[1] = 0xbb67ae85u; h2[2] = 0x3c6ef372u; h2[3] = 0xa54ff53au;
        h2[4] = 0x510e527fu; h2[5] = 0x9b05688cu; h2[6] = 0x1f83d9abu; h2[7] = 0x5be0cd19u;
    }
    WordReader R;
    R.init(m);
    const int nblk = sg_sha_blocks(len), nwords = nblk * 16;
#if defined(__CUDACC__)
#pragma unroll 1
#endif
    for (int blk = 0; blk < nblk; ++blk) {
#ifndef SG_RELOAD_BLOCK
        uint32_t w[16], v[16];
        if ((blk + 1) * 64 <= len) {
#pragma unroll
            for (int j = 0; j < 16; ++j) w[j] = R.be(blk * 16 + j);
        } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) w[j] = sg_padded_word(R, len, blk * 16 + j, nwords);
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = w[j];
        if (mode & 1u) sg_sha1_compress(h1, w);
        if (mode & 2u) sg_sha256_compress(h2, v);
#else
        // variant: the block is read from the staged bytes once per digest instead of being kept in 16 more
        // registers across the first compression (fewer registers -> more warps per SM, 48 more instructions per block)
#pragma unroll 1
        for (uint32_t which = 1; which <= 2; ++which) {
            if (!(mode & which)) continue;
            uint32_t w[16];
            if ((blk + 1) * 64 <= len) {
#pragma unroll
                for (int j = 0; j < 16; ++j) w[j] = R.be(blk * 16 + j);
            } else {
#pragma unroll
                for (int j = 0; j < 16; ++j) w[j] = sg_padded_word(R, len, blk * 16 + j, nwords);
            }
            if (which == 1) sg_sha1_compress(h1, w); else sg_sha256_compress(h2, w);
        }
#endif
    }
}

#if defined(__CUDACC__)
__host__ __device__ __noinline__
#endif
static void sg_sha256_outofline(const uint8_t* m, int len, uint32_t h[8]) { sg_sha256(m, len, h); }
#if defined(__CUDACC__)
__host__ __device__ __noinline__
#endif
static void sg_sha1_outofline(const uint8_t* m, int len, uint32_t h[5]) { sg_sha1(m, len, h); }
