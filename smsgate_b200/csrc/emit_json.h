// Bulk emit of what the reference sends downstream for a replayed message (second half):
//   mode 0  the `sms.parsed` payload, `ParsedSMS.model_dump_json()` (libs/nats_utils.py:127, worker.py:183)
//   mode 1  the record `upsert_parsed_sms` hands to PocketBase (libs/pocketbase.py:296-309), as compact JSON
// assembled as BYTES straight from the replay's columns and the natively read cache (cache_reader.h): no
// pydantic object, no Python per record.  Host code; two passes (count, then write) like the XML emitter.
//
// Text rules restated from what pydantic-core 2.x / Python produce (pinned by tests against model_dump_json):
//   str        JSON string, UTF-8 kept as is; only '"', '\\', and C0 controls are escaped (\b \f \n \r \t, \u00XX)
//   Decimal    str(Decimal): digits with the point at exponent, scientific ("1E-7") when exponent + digits <= -6
//   datetime   isoformat(): YYYY-MM-DDTHH:MM:SS[.ffffff] (naive; aware dates never reach an upsert)
//   currency   upper-cased (libs/models.py:93-95) - ASCII here, a non-ASCII currency is left to the caller
//   address    "null" -> "" (libs/gemini_parser.py:243-244)
// A row this code does not render (non-ASCII currency, a cache field of an unexpected type, a coefficient
// overflow) is flagged in `fallback` and the caller renders it through pydantic.
#pragma once
#include <stdint.h>
#include <string.h>

#include <string>
#include <thread>
#include <vector>

namespace sgemit {

struct In {
    // per message (batch order)
    const uint8_t* msg_id;            // [n*20]
    const uint8_t* status;            // [n]
    const uint8_t* flags;             // [n] SMSGATE_F_*
    const uint8_t* dec_flags;         // [n]
    const int32_t* slot;              // [n]
    const int64_t* amount_coef;       // [n]
    const int64_t* balance_coef;
    const int32_t* amount_exp;
    const int32_t* balance_exp;
    const int64_t* date;              // [n] packed
    const uint8_t* card;              // [n*16]
    const uint8_t* card_len;          // [n]
    const int64_t* body_offsets;      // [n+1]
    const uint8_t* body_bytes;
    const int32_t* fixed_len;         // [n]
    const uint8_t* fixed_bytes;       // fixed_body at body_offsets[i] where F_MODIFIED
    const int64_t* sender_offsets;    // [n+1]
    const uint8_t* sender_bytes;
    const char* device_id;            // one value for the batch, NULL -> null
    // per cache entry (cache_reader.h: txn_type, currency, merchant, city, address)
    const int64_t* text_offsets;      // [5m+1]
    const uint8_t* text_bytes;
    const uint8_t* text_kind;         // [5m]
    int64_t record_stride;            // 0: the arrays above are separate columns; 64: msg_id..date are fields of smsgate_record
};

struct Sink {
    uint8_t* out;                     // NULL = count only
    int64_t n = 0;
    void put(const void* p, size_t k) {
        if (out) memcpy(out + n, p, k);
        n += (int64_t)k;
    }
    void lit(const char* s) { put(s, strlen(s)); }
    void ch(char c) { put(&c, 1); }
};

inline void json_string(Sink& o, const uint8_t* p, int64_t len) {
    static const char* hex = "0123456789abcdef";
    o.ch('"');
    int64_t run = 0;
    for (int64_t i = 0; i < len; ++i) {
        const uint8_t c = p[i];
        const char* esc = nullptr;
        char u[7];
        if (c == '"') esc = "\\\"";
        else if (c == '\\') esc = "\\\\";
        else if (c < 0x20) {
            if (c == '\b') esc = "\\b";
            else if (c == '\f') esc = "\\f";
            else if (c == '\n') esc = "\\n";
            else if (c == '\r') esc = "\\r";
            else if (c == '\t') esc = "\\t";
            else { u[0] = '\\'; u[1] = 'u'; u[2] = '0'; u[3] = '0'; u[4] = hex[c >> 4]; u[5] = hex[c & 15]; u[6] = 0; esc = u; }
        }
        if (esc) {
            if (i > run) o.put(p + run, (size_t)(i - run));
            o.lit(esc);
            run = i + 1;
        }
    }
    if (len > run) o.put(p + run, (size_t)(len - run));
    o.ch('"');
}

// str(Decimal((sign, digits(coef), exp))) - CPython's Decimal.__str__ (Lib/_pydecimal.py), exp <= 0 here
inline void decimal_text(Sink& o, int sign, int64_t coef, int32_t exp) {
    char d[24];
    int nd = 0;
    uint64_t c = (uint64_t)coef;
    do { d[nd++] = (char)('0' + c % 10); c /= 10; } while (c);
    for (int a = 0, b = nd - 1; a < b; ++a, --b) { const char t = d[a]; d[a] = d[b]; d[b] = t; }
    if (sign) o.ch('-');
    const int64_t left = (int64_t)exp + nd;
    const int64_t dot = (exp <= 0 && left > -6) ? left : 1;
    if (dot <= 0) {
        o.lit("0.");
        for (int64_t k = 0; k < -dot; ++k) o.ch('0');
        o.put(d, (size_t)nd);
    } else if (dot >= nd) {
        o.put(d, (size_t)nd);
        for (int64_t k = 0; k < dot - nd; ++k) o.ch('0');
    } else {
        o.put(d, (size_t)dot);
        o.ch('.');
        o.put(d + dot, (size_t)(nd - dot));
    }
    if (left != dot) {
        const std::string e = std::to_string(left - dot);
        o.ch('E');
        if (left - dot >= 0) o.ch('+');
        o.put(e.data(), e.size());
    }
}

inline void two(Sink& o, int v) { o.ch((char)('0' + v / 10)); o.ch((char)('0' + v % 10)); }
// datetime.isoformat(sep): packed y<<46 | m<<42 | d<<37 | H<<32 | M<<26 | S<<20 | us
inline void iso_datetime(Sink& o, int64_t p, char sep) {
    const int y = (int)(p >> 46), mo = (int)((p >> 42) & 15), d = (int)((p >> 37) & 31), h = (int)((p >> 32) & 31),
              mi = (int)((p >> 26) & 63), s = (int)((p >> 20) & 63), us = (int)(p & 0xFFFFF);
    o.ch((char)('0' + y / 1000 % 10)); o.ch((char)('0' + y / 100 % 10)); o.ch((char)('0' + y / 10 % 10)); o.ch((char)('0' + y % 10));
    o.ch('-'); two(o, mo); o.ch('-'); two(o, d); o.ch(sep); two(o, h); o.ch(':'); two(o, mi); o.ch(':'); two(o, s);
    if (us) {
        char f[7];
        int v = us;
        for (int k = 5; k >= 0; --k) { f[k] = (char)('0' + v % 10); v /= 10; }
        o.ch('.');
        o.put(f, 6);
    }
}

// a cache text field as JSON: null / string; returns false for an unexpected type (caller falls back)
inline bool text_field(Sink& o, const In& in, int64_t e, int j, bool upper, bool null_word_to_empty) {
    const uint8_t kind = in.text_kind[5 * e + j];
    if (kind == 1) { o.lit("null"); return true; }
    if (kind != 2) return false;
    const uint8_t* p = in.text_bytes + in.text_offsets[5 * e + j];
    const int64_t len = in.text_offsets[5 * e + j + 1] - in.text_offsets[5 * e + j];
    if (null_word_to_empty && len == 4 && !memcmp(p, "null", 4)) { o.lit("\"\""); return true; }
    if (!upper) { json_string(o, p, len); return true; }
    uint8_t up[64];
    if (len > (int64_t)sizeof(up)) return false;
    for (int64_t k = 0; k < len; ++k) {
        if (p[k] >= 0x80) return false;                    // str.upper() of non-ASCII text: Python's tables
        up[k] = (p[k] >= 'a' && p[k] <= 'z') ? (uint8_t)(p[k] - 32) : p[k];
    }
    if (len == 0) { o.lit("\"\""); return true; }          // `v.upper() if v else v`
    json_string(o, up, len);
    return true;
}

template <typename T>
inline T field(const In& in, const T* col, int64_t i, int64_t rec_off) {
    if (!in.record_stride) return col[i];
    T v;
    memcpy(&v, reinterpret_cast<const uint8_t*>(in.msg_id) + i * in.record_stride + rec_off, sizeof(T));
    return v;
}

// one message; returns false (and writes nothing that counts) when the row is left to the caller
inline bool emit_one(Sink& o, const In& in, int64_t i, int mode) {
    const int64_t start = o.n;
    const bool rec = in.record_stride != 0;
    const uint8_t* id = rec ? in.msg_id + i * in.record_stride : in.msg_id + i * 20;
    const uint8_t dec = rec ? field<uint8_t>(in, nullptr, i, 22) : in.dec_flags[i];
    const uint8_t mfl = rec ? field<uint8_t>(in, nullptr, i, 21) : in.flags[i];
    const int32_t e = rec ? field<int32_t>(in, nullptr, i, 24) : in.slot[i];
    const int32_t aexp = rec ? field<int32_t>(in, nullptr, i, 28) : in.amount_exp[i];
    const int32_t bexp = rec ? field<int32_t>(in, nullptr, i, 32) : in.balance_exp[i];
    const int64_t acoef = rec ? field<int64_t>(in, nullptr, i, 40) : in.amount_coef[i];
    const int64_t bcoef = rec ? field<int64_t>(in, nullptr, i, 48) : in.balance_coef[i];
    const int64_t date = rec ? field<int64_t>(in, nullptr, i, 56) : in.date[i];
    auto bail = [&]() { o.n = start; return false; };
    if (e < 0 || (dec & 0x0F)) return bail();              // no cache entry / a decimal error or overflow
    static const char* hexd = "0123456789abcdef";
    char hex[40];
    for (int k = 0; k < 20; ++k) { hex[2 * k] = hexd[id[k] >> 4]; hex[2 * k + 1] = hexd[id[k] & 15]; }
    const uint8_t* body = (mfl & 0x04) ? in.fixed_bytes + in.body_offsets[i] : in.body_bytes + in.body_offsets[i];
    const int64_t blen = (mfl & 0x04) ? in.fixed_len[i] : in.body_offsets[i + 1] - in.body_offsets[i];
    const uint8_t* snd = in.sender_bytes + in.sender_offsets[i];
    const int64_t slen = in.sender_offsets[i + 1] - in.sender_offsets[i];
    if (mode == 0) {
        o.lit("{\"msg_id\":\""); o.put(hex, 40); o.lit("\",\"device_id\":");
        if (in.device_id) json_string(o, reinterpret_cast<const uint8_t*>(in.device_id), (int64_t)strlen(in.device_id)); else o.lit("null");
        o.lit(",\"sender\":"); json_string(o, snd, slen);
        o.lit(",\"date\":\""); iso_datetime(o, date, 'T'); o.lit("\",\"raw_body\":"); json_string(o, body, blen);
        o.lit(",\"txn_type\":"); if (!text_field(o, in, e, 0, false, false)) return bail();
        o.lit(",\"amount\":\""); decimal_text(o, (dec >> 4) & 1, acoef, aexp);
        o.lit("\",\"currency\":"); if (!text_field(o, in, e, 1, true, false)) return bail();
        o.lit(",\"card\":"); json_string(o, in.card + i * 16, in.card_len[i]);
        o.lit(",\"merchant\":"); if (!text_field(o, in, e, 2, false, false)) return bail();
        o.lit(",\"city\":"); if (!text_field(o, in, e, 3, false, false)) return bail();
        o.lit(",\"address\":"); if (!text_field(o, in, e, 4, false, true)) return bail();
        o.lit(",\"balance\":\""); decimal_text(o, (dec >> 5) & 1, bcoef, bexp);
        o.lit("\",\"parser_version\":\"llm-0.2.0\"}");
    } else {
        o.lit("{\"msg_id\":\""); o.put(hex, 40);
        o.lit("\",\"original_body\":"); json_string(o, body, blen);
        o.lit(",\"sender\":"); json_string(o, snd, slen);
        o.lit(",\"datetime\":\""); iso_datetime(o, date, 'T');
        o.lit("\",\"card\":"); json_string(o, in.card + i * 16, in.card_len[i]);
        o.lit(",\"amount\":\""); decimal_text(o, (dec >> 4) & 1, acoef, aexp);
        o.lit("\",\"currency\":"); if (!text_field(o, in, e, 1, true, false)) return bail();
        o.lit(",\"balance\":\""); decimal_text(o, (dec >> 5) & 1, bcoef, bexp);
        o.lit("\",\"merchant\":"); if (!text_field(o, in, e, 2, false, false)) return bail();
        o.lit(",\"address\":"); if (!text_field(o, in, e, 4, false, true)) return bail();
        o.lit(",\"city\":"); if (!text_field(o, in, e, 3, false, false)) return bail();
        o.lit(",\"txn_type\":"); if (!text_field(o, in, e, 0, false, false)) return bail();
        o.ch('}');
    }
    return true;
}

// The selected rows of a batch: lengths, offsets and (with out_bytes) the bytes.  Rows are independent: a sizing pass
// leaves every row's length in out_offsets[k + 1], a prefix sum turns the lengths into offsets, the writing pass puts
// each row at its offset.  Both passes run on `threads` host threads (0: one per 32 K rows, at most 8 and at most the
// host's cores).  Returns false when `capacity` is too small for the bytes (`needed` says how many).
inline bool emit_many(const In& I, int mode, int64_t n_sel, const int64_t* sel, int64_t* out_offsets, uint8_t* out_bytes,
                      int64_t capacity, uint8_t* fallback, int64_t* needed, int threads) {
    if (threads <= 0) {
        const unsigned hw = std::thread::hardware_concurrency();
        int64_t t = n_sel / 32768;
        if (t > 8) t = 8;
        if (hw && t > (int64_t)hw) t = hw;
        threads = (int)t;
    }
    if (threads < 1) threads = 1;
    if ((int64_t)threads > n_sel) threads = n_sel > 0 ? (int)n_sel : 1;
    auto in_chunks = [&](auto&& body) {                      // body(first row, one past the last row) on each thread
        if (threads == 1) { body((int64_t)0, n_sel); return; }
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t)
            pool.emplace_back([&, t] { body(n_sel * t / threads, n_sel * (t + 1) / threads); });
        for (auto& th : pool) th.join();
    };
    out_offsets[0] = 0;
    in_chunks([&](int64_t k0, int64_t k1) {
        for (int64_t k = k0; k < k1; ++k) {
            Sink count{nullptr};
            emit_one(count, I, sel[k], mode);
            out_offsets[k + 1] = count.n;
        }
    });
    for (int64_t k = 0; k < n_sel; ++k) out_offsets[k + 1] += out_offsets[k];
    const int64_t total = out_offsets[n_sel];
    if (needed) *needed = total;
    if (!out_bytes) return true;                             // sizing call
    if (capacity < total) return false;
    in_chunks([&](int64_t k0, int64_t k1) {
        for (int64_t k = k0; k < k1; ++k) {
            // a row the sizing pass declined has an empty span and must not be started again: what it writes before
            // it gives up would land in its neighbour's span (or past the end of the buffer)
            const bool declined = out_offsets[k + 1] == out_offsets[k];
            if (!declined) {
                Sink w{out_bytes + out_offsets[k]};
                emit_one(w, I, sel[k], mode);
            }
            if (fallback) fallback[k] = declined ? 1 : 0;
        }
    });
    return true;
}

}  // namespace sgemit
