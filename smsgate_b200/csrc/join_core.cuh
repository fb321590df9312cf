// Per-message logic of the join/route kernel: decimal canonicalisation, the
// strict cached-date form, card normalisation and the routing decision.
#pragma once
#include "sg_common.cuh"

struct DecResult {
    int64_t coef;     // |coefficient|
    int32_t exp;      // Decimal exponent (<= 0)
    uint32_t err;     // SG_DEC_*
    uint32_t sign;    // 1 = negative (also for -0)
};

// parse_ambiguous_decimal (libs/decimal_utils.py:4-63) on a UTF-8 string,
// result as Decimal.as_tuple() with the digit tuple folded into an int64:
//   strip() (Unicode whitespace) and drop ASCII spaces; empty -> Decimal("0.0");
//   both '.' and ',' present -> the later one is the decimal point, the other is dropped;
//   only ',' -> more than one: all dropped, exactly one: it is the point;
//   only '.' -> all but the last dropped;
//   then every char outside [0-9.-] is dropped and Decimal() must accept the rest,
//   i.e. -?digits[.digits] with at least one digit (a '-' anywhere else, or a
//   lone '.', is InvalidOperation -> ValueError).
// Two passes over at most a few dozen bytes: the general form (any UTF-8).
#if defined(__CUDACC__)
__host__ __device__ __noinline__
#endif
static DecResult sg_parse_decimal_general(const uint8_t* s, int len) {
    DecResult r;
    r.coef = 0; r.exp = 0; r.err = SG_DEC_OK; r.sign = 0;
    // pass 1: is there anything but whitespace; where are the last '.' and ','
    int last_dot = -1, last_comma = -1, n_dot = 0, n_comma = 0;
    bool any = false;
    for (int i = 0; i < len;) {
        const uint32_t b = s[i];
        if (b < 0x80) {                                    // ASCII: no decoding, white space is 9-13 and 28-32
            if (b == '.') { last_dot = i; ++n_dot; }
            else if (b == ',') { last_comma = i; ++n_comma; }
            if (!((b >= 9 && b <= 13) || (b >= 0x1C && b <= 0x20))) any = true;
            ++i;
            continue;
        }
        uint32_t cp;
        const int n = sg_utf8_decode(s + i, len - i, &cp);
        if (!sg_is_space(cp)) any = true;
        i += n;
    }
    if (!any) { r.exp = -1; return r; }                    // Decimal("0.0")
    // which byte offset is THE decimal point (-1: none)
    int point;
    if (last_dot >= 0 && last_comma >= 0) point = last_comma > last_dot ? last_comma : last_dot;
    else if (last_comma >= 0) point = (n_comma > 1) ? -1 : last_comma;
    else point = last_dot;                                 // -1 when absent
    // With both separators present the reference turns EVERY occurrence of the
    // later kind into '.', not just the last one ("1.234,56,78" -> "1234.56.78").
    const uint8_t point_char = (point >= 0) ? s[point] : 0;
    const bool all_of_kind = (last_dot >= 0 && last_comma >= 0);
    // pass 2: Decimal syntax over the kept characters
    bool seen_digit = false, seen_point = false, seen_any = false;
    int frac = 0;
    bool overflow = false;
    uint64_t coef = 0;
    for (int i = 0; i < len; ++i) {
        const uint8_t c = s[i];
        bool is_point = false;
        if (c == point_char && point >= 0) is_point = all_of_kind ? true : (i == point);
        if (c >= '0' && c <= '9') {
            const uint32_t d = c - '0';
            // coef * 10 + d would pass 2^64 - 1 = 1844674407370955161 * 10 + 5
            if (coef > 1844674407370955161ull - (d > 5 ? 1u : 0u)) overflow = true;
            coef = coef * 10 + d;
            if (seen_point) ++frac;
            seen_digit = true; seen_any = true;
        } else if (is_point) {
            if (seen_point) { r.err = SG_DEC_SYNTAX; return r; }
            seen_point = true; seen_any = true;
        } else if (c == '-') {
            if (seen_any) { r.err = SG_DEC_SYNTAX; return r; }
            r.sign = 1; seen_any = true;
        }
        // every other byte (letters, the dropped separator kind, spaces,
        // non-ASCII bytes) is removed by the final re.sub
    }
    if (!seen_digit) { r.err = SG_DEC_SYNTAX; r.sign = 0; return r; }
    if (overflow || coef > 0x7FFFFFFFFFFFFFFFull) { r.err = SG_DEC_OVERFLOW; return r; }
    r.coef = (int64_t)coef;
    r.exp = -frac;
    return r;
}

// The same in ONE pass for the strings that occur in practice - pure ASCII - and the general form for anything
// else.  Which separator is the decimal point is only known at the end of the string, but nothing that is
// accumulated on the way depends on it: the coefficient is all digits in order whichever separators are dropped,
// and the exponent is the number of digits after the last '.' or after the last ',' (both are counted).  The
// syntax rules of Decimal() reduce to: at least one digit; at most one '-', before every digit and before the
// decimal point; at most one decimal point among the kept characters - which, by the reference's replacement
// rules, fails exactly when both kinds occur and the later kind occurs more than once.
SG_HD DecResult sg_parse_decimal(const uint8_t* s, int len) {
    DecResult r;
    r.coef = 0; r.exp = 0; r.err = SG_DEC_OK; r.sign = 0;
    int n_dot = 0, n_comma = 0, last_dot = -1, last_comma = -1, after_dot = 0, after_comma = 0;
    int first_digit = -1, minus_pos = -1, n_minus = 0;
    bool any = false, overflow = false;
    uint64_t coef = 0;
    for (int i = 0; i < len; ++i) {
        const uint32_t c = s[i];
        if (c >= 0x80) return sg_parse_decimal_general(s, len);
        const uint32_t d = c - '0';
        if (d <= 9) {
            if (coef > 1844674407370955161ull - (d > 5 ? 1u : 0u)) overflow = true;
            coef = coef * 10 + d;
            ++after_dot; ++after_comma;
            if (first_digit < 0) first_digit = i;
            any = true;
        } else if (c == '.') { ++n_dot; last_dot = i; after_dot = 0; any = true; }
        else if (c == ',') { ++n_comma; last_comma = i; after_comma = 0; any = true; }
        else if (c == '-') { ++n_minus; if (minus_pos < 0) minus_pos = i; any = true; }
        else if (!((c >= 9 && c <= 13) || (c >= 0x1C && c <= 0x20))) any = true;
    }
    if (!any) { r.exp = -1; return r; }                    // Decimal("0.0")
    int point = -1, frac = 0;
    bool two_points = false;
    if (n_dot && n_comma) {
        if (last_comma > last_dot) { point = last_comma; frac = after_comma; two_points = n_comma > 1; }
        else { point = last_dot; frac = after_dot; two_points = n_dot > 1; }
    } else if (n_comma) {
        if (n_comma == 1) { point = last_comma; frac = after_comma; }
    } else if (n_dot) { point = last_dot; frac = after_dot; }
    // with both kinds present every occurrence of the later kind is a point: the '-' has to precede the FIRST of
    // them, but then there is only one (two_points is an error anyway), so `point` is that one
    const bool minus_bad = n_minus > 1 || (n_minus == 1 && ((first_digit >= 0 && first_digit < minus_pos) ||
                                                            (point >= 0 && point < minus_pos)));
    if (two_points || minus_bad || first_digit < 0) { r.err = SG_DEC_SYNTAX; return r; }
    r.sign = n_minus ? 1u : 0u;
    if (overflow || coef > 0x7FFFFFFFFFFFFFFFull) { r.err = SG_DEC_OVERFLOW; return r; }
    r.coef = (int64_t)coef;
    r.exp = -frac;
    return r;
}

// datetime.strptime(s, '%d.%m.%y %H:%M') restricted to the ASCII subset this
// build evaluates on the GPU (libs/gemini_parser.py:116): d{1,2}.d{1,2}.dd,
// one or more of " \t\n\r\v\f", d{1,2}:d{1,2}, with the directive regexes'
// value ranges (_strptime: %d 3[01]|[12]\d|0[1-9]|[1-9], %m 1[0-2]|0[1-9]|[1-9],
// %H 2[0-3]|[01]\d|\d, %M [0-5]\d|\d) and calendar validity.  Anything else
// returns -1 and is resolved at cache-load time by the host loader with the
// reference's own strptime -> dateutil chain.
SG_HD int64_t sg_parse_strict_datetime(const uint8_t* s, int len) {
    int i = 0;
    auto dig = [&](int k) -> int { return (k < len && s[k] >= '0' && s[k] <= '9') ? s[k] - '0' : -1; };
    // day
    int d0 = dig(i), d1 = dig(i + 1), day;
    if (d0 < 0) return -1;
    if (d1 >= 0) { day = d0 * 10 + d1; i += 2; if (day < 1 || day > 31) return -1; }
    else { day = d0; i += 1; if (day < 1) return -1; }
    if (i >= len || s[i] != '.') return -1;
    ++i;
    int m0 = dig(i), m1 = dig(i + 1), mon;
    if (m0 < 0) return -1;
    if (m1 >= 0) { mon = m0 * 10 + m1; i += 2; if (mon < 1 || mon > 12) return -1; }
    else { mon = m0; i += 1; if (mon < 1) return -1; }
    if (i >= len || s[i] != '.') return -1;
    ++i;
    const int y0 = dig(i), y1 = dig(i + 1);
    if (y0 < 0 || y1 < 0) return -1;
    int year = y0 * 10 + y1;
    year += (year < 69) ? 2000 : 1900;
    i += 2;
    int ws = 0;
    while (i < len && (s[i] == ' ' || (s[i] >= 9 && s[i] <= 13))) { ++i; ++ws; }
    if (!ws) return -1;
    int h0 = dig(i), h1 = dig(i + 1), hour;
    if (h0 < 0) return -1;
    if (h1 >= 0) { hour = h0 * 10 + h1; i += 2; if (hour > 23) return -1; }
    else { hour = h0; i += 1; }
    if (i >= len || s[i] != ':') return -1;
    ++i;
    int n0 = dig(i), n1 = dig(i + 1), minute;
    if (n0 < 0) return -1;
    if (n1 >= 0) { minute = n0 * 10 + n1; i += 2; if (minute > 59) return -1; }
    else { minute = n0; i += 1; }
    if (i != len) return -1;                               // "unconverted data remains"
    if (day > sg_days_in_month(year, mon)) return -1;
    return sg_pack_datetime(year, mon, day, hour, minute, 0, 0);
}

// card.replace('*','').replace(' ',''), first four characters
// (libs/gemini_parser.py:232-234).  Writes at most 16 bytes to out16 (zero
// padded) and returns (code points kept, bytes kept).
SG_HD void sg_normalise_card(const uint8_t* s, int len, uint8_t out16[16], int* ncp, int* nbytes) {
    int cps = 0, nb = 0;
    for (int k = 0; k < 16; ++k) out16[k] = 0;
    for (int i = 0; i < len && cps < 4;) {
        uint32_t cp;
        const int n = sg_utf8_decode(s + i, len - i, &cp);
        if (cp != '*' && cp != ' ') {
            for (int j = 0; j < n; ++j) out16[nb + j] = s[i + j];
            nb += n;
            ++cps;
        }
        i += n;
    }
    *ncp = cps;
    *nbytes = nb;
}

// What Postgres stores in a Numeric(14,2) column (db/models.py:21,27): the value rounded to two
// decimals, halves away from zero; more than 12 integer digits is a numeric overflow (the
// reference swallows that error, services/pb_writer/upsert.py:32-33).  Returns |cents|, sets
// *overflow.  Derived from PostgreSQL's documented numeric rounding, not from reference code.
SG_HD int64_t sg_numeric_14_2(const DecResult& d, bool* overflow) {
    *overflow = false;
    if (d.err != SG_DEC_OK) { *overflow = d.err == SG_DEC_OVERFLOW; return 0; }
    uint64_t c = (uint64_t)d.coef;
    const int e = d.exp;                       // <= 0
    if (e >= -2) {
        for (int k = 0; k < 2 + e; ++k) {
            if (c > 99999999999999ull) { *overflow = true; return 0; }
            c *= 10;
        }
    } else {
        const int k = -e - 2;                  // digits to drop
        if (k > 19) c = 0;                     // coef < 10^19: rounds to zero
        else {
            uint64_t div = 1;
            for (int j = 0; j < k; ++j) div *= 10;          // k <= 19 fits uint64
            const uint64_t q = c / div, r = c % div;
            c = q + ((r >= div - r) ? 1 : 0);               // 2r >= div without overflow
        }
    }
    if (c > 99999999999999ull) { *overflow = true; return 0; }
    return (int64_t)c;
}

struct EntryView {              // one cache entry as the loader laid it out
    const uint8_t* amount; int amount_len;
    const uint8_t* balance; int balance_len;
    const uint8_t* date; int date_len;
    const uint8_t* card; int card_len;
    uint32_t eflags;
    int64_t date_packed;
};

struct RouteResult {
    uint32_t status;
    DecResult amount, balance;
    int64_t date;               // packed datetime, 0 when none
    uint8_t card[16];
    int card_bytes;
    int64_t amount_cents, balance_cents;    // |value| as stored by Numeric(14,2)
    uint32_t cents_overflow;                // bit 0 amount, bit 1 balance
};

// worker.py:112-189 + gemini_parser.py:224-247 + writer.py:65-84 for one
// message whose cache probe has already been done (`hit`, entry `e`).
// `phone_dt`: for a message whose raw.date is a unix timestamp (device-sourced / DLQ messages), the naive
// Asia/Yerevan datetime parse_unix_timestamp(int(raw.date)) gives (gemini_parser.py:139-188), packed; 0 when
// int(raw.date) or the conversion raises (always the case for XML messages: raw.date is ISO text there).
SG_HD RouteResult sg_route(uint32_t mflags, int32_t body_date, bool hit, const EntryView& e, int64_t now_packed,
                           int64_t phone_dt = 0) {
    RouteResult r;
    r.status = SG_ST_SKIPPED_FILTER;
    r.amount.coef = 0; r.amount.exp = 0; r.amount.err = 0; r.amount.sign = 0;
    r.balance = r.amount;
    r.date = 0;
    r.card_bytes = 0;
    r.amount_cents = r.balance_cents = 0;
    r.cents_overflow = 0;
    for (int k = 0; k < 16; ++k) r.card[k] = 0;
    int card_cps = 0;
    bool fail = false;
    bool tz_aware = false;
    if (hit) {
        // canonical amounts are reported for every hit, whatever the routing says
        r.amount = sg_parse_decimal(e.amount, e.amount_len);
        r.balance = sg_parse_decimal(e.balance, e.balance_len);
        bool ov = false;
        r.amount_cents = sg_numeric_14_2(r.amount, &ov);
        r.cents_overflow |= ov ? 1u : 0u;
        r.balance_cents = sg_numeric_14_2(r.balance, &ov);
        r.cents_overflow |= ov ? 2u : 0u;
        if (e.eflags & SG_E_CARD_STR) sg_normalise_card(e.card, e.card_len, r.card, &card_cps, &r.card_bytes);
        // --- date (gemini_parser.py:226-230)
        int64_t dt = 0;
        bool has_dt = false;
        if (e.eflags & SG_E_DATE_STRICT) { dt = sg_parse_strict_datetime(e.date, e.date_len); has_dt = dt > 0; }
        else if (e.eflags & SG_E_DATE_HOST) { dt = e.date_packed; has_dt = true; }
        else if ((e.eflags & SG_E_DATE_PHONE) && phone_dt > 0) { dt = phone_dt; has_dt = true; }   // :227-229
        if (has_dt) {
            if (body_date) r.date = sg_replace_date(dt, body_date);
            else { r.date = dt; tz_aware = (e.eflags & SG_E_DATE_TZ) != 0; }
        } else if (!(e.eflags & SG_E_DATE_KEY) || body_date || !(e.eflags & SG_E_DATE_NOBODY)) {
            fail = true;        // KeyError / AttributeError on .time() / pydantic rejects the raw value
        } else {
            r.date = e.date_packed;
            tz_aware = (e.eflags & SG_E_DATE_TZ) != 0;
        }
        if (!(e.eflags & SG_E_CARD_STR)) fail = true;                          // :232 AttributeError
        if (r.amount.err == SG_DEC_SYNTAX || r.balance.err == SG_DEC_SYNTAX) fail = true;   // :235-236
        if (!(e.eflags & SG_E_SCHEMA_OK)) fail = true;                         // :237
        if (r.amount.sign && (r.amount.coef != 0 || r.amount.err == SG_DEC_OVERFLOW)) fail = true;   // ge=0
    }
    if (mflags & SG_F_WORKER_SKIP) { r.status = SG_ST_SKIPPED_FILTER; return r; }
    if (mflags & SG_F_PARSER_OTP) { r.status = SG_ST_UNMATCHED; return r; }
    if (!hit) { r.status = SG_ST_CACHE_MISS; return r; }
    if (!(e.eflags & SG_E_IS_DICT) || !(e.eflags & SG_E_HAS_TXN)) { r.status = SG_ST_FAILED_EXC; return r; }
    if (fail) { r.status = SG_ST_UNMATCHED; return r; }
    if (card_cps < 4) { r.status = SG_ST_BROKEN; return r; }                   // :246-247
    if (tz_aware) { r.status = SG_ST_ESCAPED; return r; }                      // naive/aware compare raises (worker.py:174), outside every try
    if (r.date > now_packed) { r.status = SG_ST_FUTURE_DATE; return r; }       // worker.py:174
    r.status = (e.eflags & SG_E_MERCHANT) ? SG_ST_UPSERT : SG_ST_PARSED_NO_MERCHANT;   // writer.py:70
    return r;
}
