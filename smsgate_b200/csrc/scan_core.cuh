// Per-message scan logic of the keys kernel: skip-filter automata, body-date
// and card-form detection, the clean+mask rewrite, and SHA-1 / SHA-256.
// One message = one thread; `m` points at the message's bytes (shared memory
// on the device).  Reference semantics are cited per function.
#pragma once
#include "sg_common.cuh"

struct ScanTables {
    const uint16_t* byte_class;   // [256]
    const uint8_t* ci_next;       // [CI_STATES * SCAN_STRIDE]
    const uint8_t* cs_next;       // [CS_STATES * SCAN_STRIDE]
    const uint16_t* dm_next;      // [DM_STATES * DM_STRIDE]
    const uint32_t* digit_run_start;
    const uint8_t* digit_run_len;
    const uint8_t* card_kmp;      // [12 * 4]
};

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
        case 0xFB02: cls[0] = CI_CLS_F; cls[1] = CI_CLS_OTHER; return 2;           // ﬂ (L is in no needle)
        case 0xFB03: cls[0] = CI_CLS_F; cls[1] = CI_CLS_F; cls[2] = CI_CLS_I; return 3;      // ﬃ
        case 0xFB04: cls[0] = CI_CLS_F; cls[1] = CI_CLS_F; cls[2] = CI_CLS_OTHER; return 3;  // ﬄ
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
SG_HD int32_t sg_body_date_at(const uint8_t* m, int len, int lead, int ncp, const ScanTables& T) {
    uint64_t vals = 0;                                // 4 bits per code point, position k at bits 4k
    uint32_t ascm = 0;                                // bit k: code point k is ASCII
    int p = lead;
    for (int k = ncp - 1; k >= 0; --k) {
        uint32_t cp;
        sg_utf8_decode(m + p, len - p, &cp);
        const int dv = sg_digit_value(cp, T.digit_run_start, T.digit_run_len);
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

// One pass over the body: both skip filters, NBSP / bullet / card-form
// detection (is fixed_body different?) and the first hits of the two body
// date forms.  worker.py:112-121, gemini_parser.py:198-203, :73-100, :131.
SG_HD ScanResult sg_scan_body(const uint8_t* m, int len, const ScanTables& T) {
    uint32_t ci = 0, cs = 0, dm = 0;
    int nB = 0, nA = 0;              // bytes consumed before the short / long date form completed
    uint32_t modified = 0, nonascii = 0;
    for (int i = 0; i < len; ++i) {
        const uint32_t b = m[i];
        const uint32_t cw = T.byte_class[b];
        cs = T.cs_next[cs * SCAN_STRIDE + ((cw >> 5) & 31)];
        if (cw & 0x8000) {
            // lead byte of a sequence that may upper-case into ASCII, be a
            // Unicode digit, or be U+2022: decode and feed symbols explicitly.
            uint32_t cp;
            sg_utf8_decode(m + i, len - i, &cp);
            uint32_t up[3];
            const int nu = sg_upper_special(cp, up);
            if (nu == 0) ci = T.ci_next[ci * SCAN_STRIDE + CI_CLS_OTHER];
            for (int k = 0; k < nu; ++k) ci = T.ci_next[ci * SCAN_STRIDE + up[k]];
            uint32_t dcls = DM_CLS_OTHER;
            if (cp == 0x2022) { dcls = DM_CLS_STAR; modified = 1; }
            else if (sg_digit_value(cp, T.digit_run_start, T.digit_run_len) >= 0) dcls = DM_CLS_DIGIT;
            dm = T.dm_next[(dm & DM_ROW_MASK) | dcls];
        } else {
            ci = T.ci_next[ci * SCAN_STRIDE + (cw & 31)];
            dm = T.dm_next[(dm & DM_ROW_MASK) | ((cw >> 10) & 7)];
        }
        nonascii |= b;
        nB += (dm & DM_FLAG_B) ? 0 : 1;
        nA += (dm & DM_FLAG_A) ? 0 : 1;
    }
    ScanResult r;
    r.flags = 0;
    r.body_date = 0;
    // CS state layout: block = W | N<<1 (CS_BLOCK states each), then the parser-hit states
    const bool parser_hit = cs >= CS_ACC0;
    const uint32_t blk = cs / CS_BLOCK;
    const bool daily = !parser_hit && (blk & 1);
    const bool nbsp = parser_hit ? (cs == CS_ACC1) : ((blk >> 1) & 1);
    if (ci == CI_ACCEPT || daily || parser_hit) r.flags |= SG_F_WORKER_SKIP;
    if (parser_hit) r.flags |= SG_F_PARSER_OTP;
    if (modified || nbsp || (dm & DM_FLAG_C)) r.flags |= SG_F_MODIFIED;
    if (nonascii & 0x80) r.flags |= SG_F_NON_ASCII;
    int32_t bd = 0;
    if (dm & DM_FLAG_A) bd = sg_body_date_at(m, len, nA, 10, T);
    if (bd == 0 && (dm & DM_FLAG_B)) bd = sg_body_date_at(m, len, nB, 8, T);
    r.body_date = bd;
    if (bd) r.flags |= SG_F_BODY_DATE;
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
        uint32_t cp;
        const int n = sg_utf8_decode(m + r, len - r, &cp);
        uint32_t cls;
        if (cp == 0xA0) { m[w++] = ' '; cls = DM_CLS_OTHER; }
        else if (cp == 0x2022) { m[w++] = '*'; cls = DM_CLS_STAR; }
        else {
            for (int j = 0; j < n; ++j) m[w + j] = m[r + j];
            w += n;
            cls = (cp == '*') ? DM_CLS_STAR
                : (sg_digit_value(cp, T.digit_run_start, T.digit_run_len) >= 0) ? DM_CLS_DIGIT : DM_CLS_OTHER;
        }
        r += n;
        k = T.card_kmp[k * 4 + cls];
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

// Word `j` of the padded message (0x80, zeros, 64-bit bit length).
SG_HD uint32_t sg_padded_word(const WordReader& R, int len, int j, int nwords) {
    const int off = j * 4;
    if (off + 4 <= len) return R.be(j);
    uint32_t w = 0;
    if (off <= len) {
        const int keep = len - off;                   // 0..3 message bytes in this word
        if (keep) w = R.be(j) & (0xFFFFFFFFu << (32 - 8 * keep));
        w |= 0x80u << (24 - 8 * keep);
    }
    if (j == nwords - 1) w |= (uint32_t)len << 3;
    if (j == nwords - 2) w |= (uint32_t)len >> 29;
    return w;
}

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

SG_HD void sg_sha256_compress(uint32_t h[8], uint32_t w[16]) {
    // round constants as immediates: the loop is fully unrolled
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
#pragma unroll
    for (int i = 0; i < 64; ++i) {
        uint32_t wi;
        if (i < 16) wi = w[i];
        else {
            const uint32_t w15 = w[(i - 15) & 15], w2 = w[(i - 2) & 15];
            const uint32_t s0 = sg_rotr(w15, 7) ^ sg_rotr(w15, 18) ^ (w15 >> 3);
            const uint32_t s1 = sg_rotr(w2, 17) ^ sg_rotr(w2, 19) ^ (w2 >> 10);
            wi = w[i & 15] + s0 + w[(i - 7) & 15] + s1;
            w[i & 15] = wi;
        }
        const uint32_t S1 = sg_rotr(e, 6) ^ sg_rotr(e, 11) ^ sg_rotr(e, 25);
        const uint32_t ch = (e & f) ^ (~e & g);
        const uint32_t t1 = hh + S1 + ch + SG_K256[i] + wi;
        const uint32_t S0 = sg_rotr(a, 2) ^ sg_rotr(a, 13) ^ sg_rotr(a, 22);
        const uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
        const uint32_t t2 = S0 + mj;
        hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
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
