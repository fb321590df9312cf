// Shared helpers for the replay kernels: UTF-8 decoding, Unicode digit lookup,
// packed date arithmetic.  Everything here is __host__ __device__ so that the
// same source is unit-tested on the CPU by tests/hostcheck (test-only library;
// the shipped library launches these only from kernels).
#pragma once
#include <stdint.h>
#include "scan_tables.h"

#if defined(__CUDACC__)
#define SG_HD __host__ __device__ __forceinline__
#define SG_D __device__ __forceinline__
#else
#define SG_HD inline
#define SG_D inline
#endif

// message flag bits (flags[] output of the scan kernel; include/smsgate_b200.h)
#define SG_F_WORKER_SKIP 0x01   // services/parser_worker/worker.py:112-121
#define SG_F_PARSER_OTP  0x02   // libs/gemini_parser.py:198-200
#define SG_F_MODIFIED    0x04   // fixed_body != body (libs/gemini_parser.py:202-203)
#define SG_F_BODY_DATE   0x08   // fix_broken_datetime finds a usable date (libs/gemini_parser.py:73-100)
#define SG_F_MAYBE       0x80   // internal to the scan kernel, never stored: the first pass saw a needle's upper-cased form

// status codes (status[] output of the join/route kernel)
#define SG_ST_UPSERT 0
#define SG_ST_PARSED_NO_MERCHANT 1
#define SG_ST_SKIPPED_FILTER 2
#define SG_ST_CACHE_MISS 3
#define SG_ST_UNMATCHED 4
#define SG_ST_BROKEN 5
#define SG_ST_FUTURE_DATE 6
#define SG_ST_FAILED_EXC 7
#define SG_ST_ESCAPED 8

// per cache entry flags, computed by the loader (smsgate_b200/gemini_cache.py)
#define SG_E_IS_DICT      0x0001  // cached value is a dict
#define SG_E_HAS_TXN      0x0002  // 'txn_type' key present (else the except-handler itself raises, gemini_parser.py:239)
#define SG_E_SCHEMA_OK    0x0004  // ParsedSmsCore would accept the non-computed fields (libs/llm_core.py:9-19)
#define SG_E_CARD_STR     0x0008  // 'card' is a str (gemini_parser.py:232)
#define SG_E_MERCHANT     0x0010  // merchant truthy (services/pb_writer/writer.py:70)
#define SG_E_DATE_STRICT  0x0020  // date string is in the strict dd.mm.yy HH:MM form: parsed on the GPU
#define SG_E_DATE_HOST    0x0040  // date_packed holds what strptime/dateutil produced at load time
#define SG_E_DATE_NOBODY  0x0080  // no datetime object, but pydantic accepts the raw value when the body has no date
#define SG_E_DATE_TZ      0x0100  // that datetime is timezone-aware
#define SG_E_DATE_KEY     0x0200  // 'date' key present
#define SG_E_DATE_PHONE   0x0400  // dateutil: "String does not contain a date" -> the phone's timestamp, int(raw.date) (gemini_parser.py:227-229)

// decimal result flags: bits 0-1 amount error, 2-3 balance error, 4 amount sign, 5 balance sign
#define SG_DEC_OK 0
#define SG_DEC_SYNTAX 1
#define SG_DEC_OVERFLOW 2

// Decode one UTF-8 sequence at p (at most `avail` bytes).  Returns its length
// and writes the code point; a malformed sequence decodes as length 1 with
// cp = 0xFFFFFFFF (Python's str.encode() never produces one).
SG_HD int sg_utf8_decode(const uint8_t* p, int avail, uint32_t* cp) {
    uint32_t b0 = p[0];
    if (b0 < 0x80) { *cp = b0; return 1; }
    int n = (b0 >= 0xF0) ? 4 : (b0 >= 0xE0) ? 3 : 2;
    if (b0 < 0xC2 || b0 > 0xF4 || n > avail) { *cp = 0xFFFFFFFFu; return 1; }
    uint32_t c = b0 & (0x7Fu >> n);
    for (int k = 1; k < n; ++k) {
        uint32_t b = p[k];
        if ((b & 0xC0) != 0x80) { *cp = 0xFFFFFFFFu; return 1; }
        c = (c << 6) | (b & 0x3F);
    }
    *cp = c;
    return n;
}

// Value 0-9 of a Unicode Nd code point (Python `\d`, int()), -1 otherwise.
// Every one of the 64 runs starts at a digit zero (gen_tables.py asserts it).
SG_HD int sg_digit_value(uint32_t cp, const uint32_t* run_start, const uint8_t* run_len) {
    if (cp < 0x80) return (cp >= '0' && cp <= '9') ? (int)(cp - '0') : -1;
    int lo = 0, hi = N_DIGIT_RUNS - 1;
    while (lo < hi) {                       // last run whose start <= cp
        int mid = (lo + hi + 1) >> 1;
        if (run_start[mid] <= cp) lo = mid; else hi = mid - 1;
    }
    uint32_t s = run_start[lo];
    if (cp < s || cp >= s + run_len[lo]) return -1;
    return (int)((cp - s) % 10);
}

// str.isspace() / re `\s` / str.strip() set (29 code points, Unicode 15.0)
SG_HD bool sg_is_space(uint32_t cp) {
    if (cp < 0x80) return (cp >= 9 && cp <= 13) || (cp >= 0x1C && cp <= 0x20);
    return cp == 0x85 || cp == 0xA0 || cp == 0x1680 || (cp >= 0x2000 && cp <= 0x200A) ||
           cp == 0x2028 || cp == 0x2029 || cp == 0x202F || cp == 0x205F || cp == 0x3000;
}

SG_HD int sg_days_in_month(int y, int m) {
    const int d = 31 - ((m == 4 || m == 6 || m == 9 || m == 11) ? 1 : 0);
    if (m != 2) return d;
    return ((y % 4 == 0 && y % 100 != 0) || y % 400 == 0) ? 29 : 28;
}

// int64 packing of a naive datetime, order preserving:
// y<<46 | m<<42 | d<<37 | H<<32 | M<<26 | S<<20 | microsecond
SG_HD int64_t sg_pack_datetime(int y, int mo, int d, int h, int mi, int s, int us) {
    return ((int64_t)y << 46) | ((int64_t)mo << 42) | ((int64_t)d << 37) | ((int64_t)h << 32) |
           ((int64_t)mi << 26) | ((int64_t)s << 20) | (int64_t)us;
}
// body date: y<<9 | m<<5 | d ; replace the date part of a packed datetime
SG_HD int64_t sg_replace_date(int64_t packed, int32_t ymd) {
    const int64_t time_mask = ((int64_t)1 << 37) - 1;
    return ((int64_t)ymd << 37) | (packed & time_mask);
}
