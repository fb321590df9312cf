// Native reader of the reference's on-disk Gemini response cache.
//
// The reference probes `diskcache.Cache(".gemini_cache")` once per message: `cache_key in cache` and
// `cache[cache_key]` are two SQL queries and one unpickle each (libs/gemini_parser.py:33, :208-209).  For a
// bulk replay the whole cache is needed once, as flat arrays (smsgate_cache_load).  This file reads
// diskcache 5.6.3's SQLite file directly - table Cache(rowid, key, raw, store_time, expire_time, access_time,
// access_count, tag, size, mode, filename, value) - and decodes the pickled response dicts itself, so that no
// Python object is created per entry: a 10 M-entry cache is a few seconds of C++ instead of minutes of
// `pickle.loads` + dict walking.
//
// What is decoded: the subset of pickle protocols 2-5 that a dict of str keys with str / None / bool / int /
// float values produces (PROTO, FRAME, EMPTY_DICT, MARK, SHORT_BINUNICODE, BINUNICODE, BINUNICODE8, NONE,
// NEWTRUE, NEWFALSE, BININT, BININT1, BININT2, LONG1, BINFLOAT, MEMOIZE, BINPUT, LONG_BINPUT, BINGET,
// LONG_BINGET, SETITEM, SETITEMS, STOP).  Anything else - another opcode, a nested container, a value kept in
// a side file because it exceeds 32 KiB, a float amount (its str() is Python's shortest-repr algorithm) - marks
// the entry `needs_host`, and the Python layer runs exactly its old per-entry code for those (counted).
//
// SQLite is used through dlopen("libsqlite3.so.0") with the handful of prototypes below: the image has the
// library (Python's _sqlite3 links it) but no development header.
#pragma once
#include <dlfcn.h>
#include <stdint.h>
#include <string.h>

#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

namespace sgcache {

struct Sqlite {
    void* so = nullptr;
    int (*open_v2)(const char*, void**, int, const char*) = nullptr;
    int (*prepare_v2)(void*, const char*, int, void**, const char**) = nullptr;
    int (*bind_double)(void*, int, double) = nullptr;
    int (*bind_int64)(void*, int, long long) = nullptr;
    int (*step)(void*) = nullptr;
    const void* (*column_blob)(void*, int) = nullptr;
    int (*column_bytes)(void*, int) = nullptr;
    const unsigned char* (*column_text)(void*, int) = nullptr;
    long long (*column_int64)(void*, int) = nullptr;
    int (*column_type)(void*, int) = nullptr;
    int (*finalize)(void*) = nullptr;
    int (*close)(void*) = nullptr;
    const char* (*errmsg)(void*) = nullptr;
    int (*busy_timeout)(void*, int) = nullptr;
    bool load(std::string* err) {
        if (so) return true;
        for (const char* name : {"libsqlite3.so.0", "libsqlite3.so"}) {
            so = dlopen(name, RTLD_NOW | RTLD_LOCAL);
            if (so) break;
        }
        if (!so) { *err = "libsqlite3 not found (dlopen)"; return false; }
#define SG_SYM(field, sym)                                             \
    field = reinterpret_cast<decltype(field)>(dlsym(so, sym));         \
    if (!field) { *err = std::string("libsqlite3 lacks ") + sym; return false; }
        SG_SYM(open_v2, "sqlite3_open_v2") SG_SYM(prepare_v2, "sqlite3_prepare_v2") SG_SYM(bind_double, "sqlite3_bind_double") SG_SYM(bind_int64, "sqlite3_bind_int64")
        SG_SYM(step, "sqlite3_step") SG_SYM(column_blob, "sqlite3_column_blob") SG_SYM(column_bytes, "sqlite3_column_bytes")
        SG_SYM(column_text, "sqlite3_column_text") SG_SYM(column_int64, "sqlite3_column_int64")
        SG_SYM(column_type, "sqlite3_column_type") SG_SYM(finalize, "sqlite3_finalize") SG_SYM(close, "sqlite3_close")
        SG_SYM(errmsg, "sqlite3_errmsg") SG_SYM(busy_timeout, "sqlite3_busy_timeout")
#undef SG_SYM
        return true;
    }
};

// ---- one decoded pickle item
enum Kind : uint8_t { K_MISSING = 0, K_NONE = 1, K_STR = 2, K_OTHER = 3, K_TRUE = 4, K_FALSE = 5, K_INT = 6, K_FLOAT = 7, K_DICT = 8, K_MARK = 9 };
struct Item {
    uint8_t kind = K_MISSING;
    const char* p = nullptr;   // K_STR: UTF-8 bytes inside the blob
    uint32_t len = 0;
    long long ival = 0;        // K_INT
};

// the nine fields of the response schema (libs/gemini_parser.py:46-61), in the order used below
enum Field { F_TXN = 0, F_DATE, F_AMOUNT, F_CURRENCY, F_CARD, F_MERCHANT, F_CITY, F_ADDRESS, F_BALANCE, F_COUNT };
static const char* const FIELD_NAMES[F_COUNT] = {"txn_type", "date", "amount", "currency", "card", "merchant", "city", "address", "balance"};

struct Decoded {
    bool is_dict = false;
    bool unsupported = false;  // the blob holds something this decoder does not read
    Item f[F_COUNT];
};

inline int field_of(const Item& k) {
    if (k.kind != K_STR) return -1;
    for (int i = 0; i < F_COUNT; ++i)
        if (strlen(FIELD_NAMES[i]) == k.len && memcmp(FIELD_NAMES[i], k.p, k.len) == 0) return i;
    return -1;
}

// Decode one pickled value.  Only a top-level dict with scalar values is read; the first thing outside the
// subset sets `unsupported`.
inline Decoded decode_pickle(const uint8_t* b, size_t n) {
    Decoded d;
    std::vector<Item> stack, memo;
    stack.reserve(32);
    memo.reserve(32);
    auto fail = [&]() { d.unsupported = true; return d; };
    auto memo_put = [&](size_t idx) {
        if (stack.empty()) return false;
        if (memo.size() <= idx) memo.resize(idx + 1);
        memo[idx] = stack.back();
        return true;
    };
    auto set_pair = [&](const Item& k, const Item& v) {
        const int fi = field_of(k);
        if (fi >= 0) d.f[fi] = v;
        if (k.kind != K_STR) d.unsupported = true;          // a non-str key: leave the entry to Python
    };
    size_t i = 0;
    while (i < n) {
        const uint8_t op = b[i++];
        switch (op) {
            case 0x80: if (i + 1 > n) return fail(); ++i; break;                                   // PROTO
            case 0x95: if (i + 8 > n) return fail(); i += 8; break;                                // FRAME
            case '}': { Item it; it.kind = K_DICT; stack.push_back(it); break; }                   // EMPTY_DICT
            case '(': { Item it; it.kind = K_MARK; stack.push_back(it); break; }                   // MARK
            case 0x8c: case 'X': case 0x8d: {                                                      // *BINUNICODE*
                uint64_t len = 0;
                const int w = op == 0x8c ? 1 : op == 'X' ? 4 : 8;
                if (i + w > n) return fail();
                for (int k = 0; k < w; ++k) len |= (uint64_t)b[i + k] << (8 * k);
                i += w;
                if (len > n - i || len > 0x7fffffffu) return fail();
                Item it; it.kind = K_STR; it.p = reinterpret_cast<const char*>(b + i); it.len = (uint32_t)len;
                stack.push_back(it);
                i += len;
                break;
            }
            case 'N': { Item it; it.kind = K_NONE; stack.push_back(it); break; }
            case 0x88: { Item it; it.kind = K_TRUE; stack.push_back(it); break; }
            case 0x89: { Item it; it.kind = K_FALSE; stack.push_back(it); break; }
            case 'K': case 'M': case 'J': {                                                        // BININT1 / 2 / BININT
                const int w = op == 'K' ? 1 : op == 'M' ? 2 : 4;
                if (i + w > n) return fail();
                uint32_t v = 0;
                for (int k = 0; k < w; ++k) v |= (uint32_t)b[i + k] << (8 * k);
                i += w;
                Item it; it.kind = K_INT; it.ival = op == 'J' ? (long long)(int32_t)v : (long long)v;
                stack.push_back(it);
                break;
            }
            case 0x8a: {                                                                           // LONG1
                if (i + 1 > n) return fail();
                const int w = b[i++];
                if (i + w > n) return fail();
                Item it; it.kind = K_INT;
                if (w > 8) it.kind = K_OTHER;                                                       // beyond int64: Python
                else {
                    unsigned long long v = 0;
                    for (int k = 0; k < w; ++k) v |= (unsigned long long)b[i + k] << (8 * k);
                    if (w > 0 && w < 8 && (b[i + w - 1] & 0x80)) v |= ~0ull << (8 * w);              // sign extension
                    it.ival = (long long)v;
                }
                i += w;
                stack.push_back(it);
                break;
            }
            case 'G': { if (i + 8 > n) return fail(); i += 8; Item it; it.kind = K_FLOAT; stack.push_back(it); break; }
            case 0x94: if (!memo_put(memo.size())) return fail(); break;                            // MEMOIZE
            case 'q': if (i + 1 > n || !memo_put(b[i])) return fail(); i += 1; break;               // BINPUT
            case 'r': {                                                                            // LONG_BINPUT
                if (i + 4 > n) return fail();
                const uint32_t idx = b[i] | (b[i + 1] << 8) | (b[i + 2] << 16) | ((uint32_t)b[i + 3] << 24);
                i += 4;
                if (idx > (1u << 20) || !memo_put(idx)) return fail();
                break;
            }
            case 'h': case 'j': {                                                                  // BINGET / LONG_BINGET
                const int w = op == 'h' ? 1 : 4;
                if (i + w > n) return fail();
                uint32_t idx = 0;
                for (int k = 0; k < w; ++k) idx |= (uint32_t)b[i + k] << (8 * k);
                i += w;
                if (idx >= memo.size() || memo[idx].kind == K_MISSING) return fail();
                stack.push_back(memo[idx]);
                break;
            }
            case 's': {                                                                            // SETITEM
                if (stack.size() < 3 || stack[stack.size() - 3].kind != K_DICT || stack.size() != 3) return fail();
                set_pair(stack[stack.size() - 2], stack.back());
                stack.resize(stack.size() - 2);
                break;
            }
            case 'u': {                                                                            // SETITEMS
                size_t m = stack.size();
                while (m > 0 && stack[m - 1].kind != K_MARK) --m;
                if (m < 2 || stack[m - 2].kind != K_DICT || m != 2 || ((stack.size() - m) & 1)) return fail();
                for (size_t k = m; k + 1 < stack.size(); k += 2) set_pair(stack[k], stack[k + 1]);
                stack.resize(m - 1);
                break;
            }
            case '.':                                                                              // STOP
                if (stack.size() != 1) return fail();
                if (stack[0].kind == K_DICT) d.is_dict = true;
                else if (stack[0].kind == K_MARK) return fail();
                return d;                                  // a top-level scalar: "not a dict", decided natively
            default:
                return fail();
        }
        if (d.unsupported) return d;
        // a container other than the top-level dict never gets here (its opcode is not in the switch);
        // a dict nested as a VALUE would be EMPTY_DICT on a non-empty stack:
        if (op == '}' && stack.size() != 1) return fail();
    }
    return fail();                                         // no STOP
}

// ---- the whole cache, decoded
struct Table {
    int64_t m = 0, skipped = 0, fallbacks = 0;
    std::vector<uint8_t> keys;            // [m * 32]
    std::vector<int64_t> rowids;          // [m]
    std::vector<int64_t> val_offsets;     // [4m + 1] str(amount), str(balance), date, card
    std::vector<uint8_t> val_bytes;
    std::vector<uint16_t> eflags;         // [m]  schema facts; the date flags are added by the caller
    std::vector<uint8_t> needs_host;      // [m]  1 = decode this entry in Python, 2 = only its date
    std::vector<int64_t> text_offsets;    // [5m + 1] txn_type, currency, merchant, city, address
    std::vector<uint8_t> text_bytes;
    std::vector<uint8_t> text_kind;       // [5m]  Kind of each of those five
    std::string err;
};

inline int hexval(unsigned char c) {
    if (c >= '0' && c <= '9') return c - '0';
    if (c >= 'a' && c <= 'f') return c - 'a' + 10;
    return -1;
}

// eflags bits as in include/smsgate_b200.h
constexpr uint16_t EF_IS_DICT = 0x1, EF_HAS_TXN = 0x2, EF_SCHEMA_OK = 0x4, EF_CARD_STR = 0x8, EF_MERCHANT = 0x10, EF_DATE_KEY = 0x200;

// str(v) of a scalar into `out`; false when only Python knows (float repr, big ints, other objects)
inline bool py_str(const Item& v, std::vector<uint8_t>* out) {
    switch (v.kind) {
        case K_STR: out->insert(out->end(), v.p, v.p + v.len); return true;
        case K_NONE: { const char* s = "None"; out->insert(out->end(), s, s + 4); return true; }
        case K_TRUE: { const char* s = "True"; out->insert(out->end(), s, s + 4); return true; }
        case K_FALSE: { const char* s = "False"; out->insert(out->end(), s, s + 5); return true; }
        case K_INT: { const std::string s = std::to_string(v.ival); out->insert(out->end(), s.begin(), s.end()); return true; }
        default: return false;
    }
}

inline void append_entry(Table* T, const uint8_t key32[32], int64_t rowid, const Decoded& d, bool force_host) {
    uint16_t f = 0;
    uint8_t host = force_host ? 1 : 0;
    const size_t v0 = T->val_bytes.size(), t0 = T->text_bytes.size();
    int64_t vo[4] = {0, 0, 0, 0};
    if (!host && d.unsupported) host = 1;
    if (!host && d.is_dict) {
        f |= EF_IS_DICT;
        if (d.f[F_TXN].kind != K_MISSING) f |= EF_HAS_TXN;
        bool ok = true;
        for (int k = 0; k < F_COUNT; ++k) ok = ok && d.f[k].kind != K_MISSING;
        if (ok) {
            const Item& t = d.f[F_TXN];
            ok = t.kind == K_STR && ((t.len == 5 && !memcmp(t.p, "debit", 5)) || (t.len == 6 && !memcmp(t.p, "credit", 6)) ||
                                     (t.len == 3 && !memcmp(t.p, "otp", 3)) || (t.len == 7 && !memcmp(t.p, "unknown", 7)));
            for (int k : {F_CURRENCY, F_MERCHANT, F_CITY, F_ADDRESS}) ok = ok && (d.f[k].kind == K_STR || d.f[k].kind == K_NONE);
        }
        if (ok) f |= EF_SCHEMA_OK;
        if (d.f[F_AMOUNT].kind != K_MISSING && !py_str(d.f[F_AMOUNT], &T->val_bytes)) host = 1;
        vo[0] = (int64_t)(T->val_bytes.size() - v0);
        if (d.f[F_BALANCE].kind != K_MISSING && !py_str(d.f[F_BALANCE], &T->val_bytes)) host = 1;
        vo[1] = (int64_t)(T->val_bytes.size() - v0);
        if (d.f[F_DATE].kind != K_MISSING) {
            f |= EF_DATE_KEY;
            const Item& dt = d.f[F_DATE];
            bool gpu = dt.kind == K_STR && dt.len <= 32;
            for (uint32_t k = 0; gpu && k < dt.len; ++k) gpu = (unsigned char)dt.p[k] < 0x80;
            if (gpu) T->val_bytes.insert(T->val_bytes.end(), dt.p, dt.p + dt.len);
            else if (!host) host = 2;                                      // the reference's strptime / dateutil chain
        }
        vo[2] = (int64_t)(T->val_bytes.size() - v0);
        if (d.f[F_CARD].kind == K_STR) {
            f |= EF_CARD_STR;
            T->val_bytes.insert(T->val_bytes.end(), d.f[F_CARD].p, d.f[F_CARD].p + d.f[F_CARD].len);
        }
        vo[3] = (int64_t)(T->val_bytes.size() - v0);
        if (d.f[F_MERCHANT].kind == K_STR && d.f[F_MERCHANT].len > 0) f |= EF_MERCHANT;
    }
    if (host == 1) {                                                        // Python redoes the whole entry
        T->val_bytes.resize(v0);
        vo[0] = vo[1] = vo[2] = vo[3] = 0;
        f = 0;
        ++T->fallbacks;
    }
    T->keys.insert(T->keys.end(), key32, key32 + 32);
    T->rowids.push_back(rowid);
    for (int k = 0; k < 4; ++k) T->val_offsets.push_back((int64_t)v0 + vo[k]);
    T->eflags.push_back(f);
    T->needs_host.push_back(host);
    const int text_fields[5] = {F_TXN, F_CURRENCY, F_MERCHANT, F_CITY, F_ADDRESS};
    for (int k = 0; k < 5; ++k) {
        const Item& it = d.f[text_fields[k]];
        uint8_t kind = host == 1 ? (uint8_t)K_OTHER : it.kind;
        if (kind == K_STR) T->text_bytes.insert(T->text_bytes.end(), it.p, it.p + it.len);
        T->text_offsets.push_back((int64_t)T->text_bytes.size());
        T->text_kind.push_back(kind);
    }
    (void)t0;
    ++T->m;
}

// Read the rows of <dir>/cache.db whose rowid lies in [lo, hi] (lo > hi: all of them) into *T, on a connection of
// its own.  Entries: string keys (TEXT, raw = 1) of 64 lower-case hex digits whose expire_time is NULL or later
// than `now`, in rowid order - what `key in cache` answers True for and the order diskcache iterates in.
// T->val_offsets holds four ENDS per entry, T->text_offsets starts with 0.
inline bool read_rows(Sqlite& S, const char* dir, double now, long long lo, long long hi, Table* T) {
    const std::string uri = std::string("file:") + dir + "/cache.db?mode=ro";
    void* db = nullptr;
    if (S.open_v2(uri.c_str(), &db, 0x1 /*READONLY*/ | 0x40 /*URI*/, nullptr) != 0) {
        T->err = std::string("cannot open ") + dir + "/cache.db: " + (db ? S.errmsg(db) : "out of memory");
        if (db) S.close(db);
        return false;
    }
    S.busy_timeout(db, 60000);
    void* st = nullptr;
    const bool ranged = lo <= hi;
    const std::string sql = std::string("SELECT rowid, key, mode, filename, value FROM Cache WHERE raw = 1 AND typeof(key) = 'text' "
                                        "AND (expire_time IS NULL OR expire_time > ?1)") +
                            (ranged ? " AND rowid BETWEEN ?2 AND ?3" : "") + " ORDER BY rowid";
    if (S.prepare_v2(db, sql.c_str(), -1, &st, nullptr) != 0) {
        T->err = std::string("not a diskcache database: ") + S.errmsg(db);
        S.close(db);
        return false;
    }
    S.bind_double(st, 1, now);
    if (ranged) { S.bind_int64(st, 2, lo); S.bind_int64(st, 3, hi); }
    T->val_offsets.clear();
    T->text_offsets.assign(1, 0);
    int rc;
    while ((rc = S.step(st)) == 100) {
        const long long rowid = S.column_int64(st, 0);
        const unsigned char* key = S.column_text(st, 1);
        const int klen = S.column_bytes(st, 1);
        uint8_t k32[32];
        bool hex = klen == 64;
        for (int k = 0; hex && k < 32; ++k) {
            const int hi4 = hexval(key[2 * k]), lo4 = hexval(key[2 * k + 1]);
            hex = hi4 >= 0 && lo4 >= 0;
            k32[k] = (uint8_t)((hi4 << 4) | lo4);
        }
        if (!hex) { ++T->skipped; continue; }
        const long long mode = S.column_int64(st, 2);
        const bool in_file = S.column_type(st, 3) != 5;                     // filename NOT NULL: value lives in a side file
        Decoded d;
        if (mode == 4 && !in_file && S.column_type(st, 4) == 4) {           // MODE_PICKLE, inline blob
            const uint8_t* blob = static_cast<const uint8_t*>(S.column_blob(st, 4));
            d = decode_pickle(blob, (size_t)S.column_bytes(st, 4));
            append_entry(T, k32, rowid, d, false);                          // spans point into the blob: consumed before the next step
            continue;
        }
        if (mode == 1 && !in_file) {                                         // MODE_RAW: a str / bytes / number stored as is - not a dict
            append_entry(T, k32, rowid, d, false);
            continue;
        }
        append_entry(T, k32, rowid, d, true);                                // side files, other modes: Python's Disk.fetch
    }
    S.finalize(st);
    const bool ok = rc == 101;
    if (!ok) T->err = std::string("reading cache.db failed: ") + S.errmsg(db);
    S.close(db);
    return ok;
}

// lowest and highest rowid of the table; false if it is empty or unreadable (the caller then reads it in one piece)
inline bool rowid_span(Sqlite& S, const char* dir, long long* lo, long long* hi) {
    const std::string uri = std::string("file:") + dir + "/cache.db?mode=ro";
    void* db = nullptr;
    if (S.open_v2(uri.c_str(), &db, 0x1 | 0x40, nullptr) != 0) {
        if (db) S.close(db);
        return false;
    }
    S.busy_timeout(db, 60000);
    void* st = nullptr;
    bool ok = false;
    if (S.prepare_v2(db, "SELECT min(rowid), max(rowid) FROM Cache", -1, &st, nullptr) == 0) {
        if (S.step(st) == 100 && S.column_type(st, 0) == 1) {
            *lo = S.column_int64(st, 0);
            *hi = S.column_int64(st, 1);
            ok = true;
        }
        S.finalize(st);
    }
    S.close(db);
    return ok;
}

template <class V>
inline void append_vec(V* dst, const V& src) { dst->insert(dst->end(), src.begin(), src.end()); }

// Read <dir>/cache.db into *T.  A large table is read as `threads` contiguous rowid ranges at once, each on its own
// read-only connection, and the parts are joined in rowid order: SQLite's row decoding and the pickle decoding are
// the cost (2.2 us per entry on one core), both per row.  threads <= 0: one per 64 K rows, at most 8 and at most the
// host's cores.  The ranges are separate read transactions: the cache is expected not to be written meanwhile (a
// bulk replay is an offline job; a writer committing between two ranges' snapshots shows in one and not the other).
inline bool read_diskcache(const char* dir, double now, Table* T, int threads = 0) {
    static Sqlite S;
    static std::mutex load_mutex;
    {
        std::lock_guard<std::mutex> g(load_mutex);
        if (!S.load(&T->err)) return false;
    }
    long long lo = 0, hi = -1;
    if (threads != 1 && rowid_span(S, dir, &lo, &hi)) {
        const long long span = hi - lo + 1;
        if (threads <= 0) {
            const unsigned hw = std::thread::hardware_concurrency();
            long long t = span / 65536;
            if (t > 8) t = 8;
            if (hw && t > (long long)hw) t = hw;
            threads = t < 1 ? 1 : (int)t;
        }
        if ((long long)threads > span) threads = (int)span;
    } else threads = 1;
    bool ok = true;
    if (threads <= 1) {
        ok = read_rows(S, dir, now, 0, -1, T);
    } else {
        std::vector<Table> parts((size_t)threads);
        std::vector<char> oks((size_t)threads, 0);
        std::vector<std::thread> pool;
        const long long span = hi - lo + 1;
        for (int t = 0; t < threads; ++t) {
            const long long a = lo + span * t / threads, b = lo + span * (t + 1) / threads - 1;
            pool.emplace_back([&, t, a, b] {
                try {
                    oks[(size_t)t] = read_rows(S, dir, now, a, b, &parts[(size_t)t]) ? 1 : 0;
                } catch (const std::bad_alloc&) {
                    parts[(size_t)t].err = "out of host memory";
                }
            });
        }
        for (auto& th : pool) th.join();
        T->val_offsets.clear();
        T->text_offsets.assign(1, 0);
        for (int t = 0; t < threads; ++t) {
            Table& P = parts[(size_t)t];
            if (!oks[(size_t)t]) { ok = false; if (T->err.empty()) T->err = P.err; continue; }
            const int64_t vbase = (int64_t)T->val_bytes.size(), tbase = (int64_t)T->text_bytes.size();
            T->m += P.m; T->skipped += P.skipped; T->fallbacks += P.fallbacks;
            append_vec(&T->keys, P.keys); append_vec(&T->rowids, P.rowids); append_vec(&T->eflags, P.eflags);
            append_vec(&T->needs_host, P.needs_host); append_vec(&T->text_kind, P.text_kind);
            append_vec(&T->val_bytes, P.val_bytes); append_vec(&T->text_bytes, P.text_bytes);
            for (int64_t v : P.val_offsets) T->val_offsets.push_back(v + vbase);
            for (size_t k = 1; k < P.text_offsets.size(); ++k) T->text_offsets.push_back(P.text_offsets[k] + tbase);
            P = Table();                                    // free the part's buffers before the next one is copied
        }
    }
    // val_offsets were pushed as 4 per entry (ends of the four strings); turn them into 4m+1 starts/ends
    std::vector<int64_t> vo(4 * T->m + 1, 0);
    for (int64_t e = 0; e < 4 * T->m; ++e) vo[e + 1] = T->val_offsets[e];
    T->val_offsets.swap(vo);
    return ok;
}

}  // namespace sgcache
