/*
 * smsgate_b200 — C ABI of the B200 bulk-replay library.
 *
 * The reference (vpuhoff/smsgate) is pure Python and has no FFI seam; its seam
 * is a handful of Python functions.  Each entry point below names the reference
 * code it stands in for (paths relative to the reference root).  The Python
 * drop-in (the modules of smsgate_b200/) binds these with ctypes; INTEGRATION.md shows the
 * stub a maintainer of the reference would add.
 *
 * Conventions
 *   - Every function returns 0 on success or a negative SMSGATE_E_* code and
 *     never throws; smsgate_last_error() gives the text.  Per-message problems
 *     are data (flags / status / decimal error bits), not errors.
 *   - A context is bound to one GPU and one CUDA stream and is not re-entrant;
 *     different contexts are independent.
 *   - "d_" pointers are device memory owned by the caller; "h_" pointers are
 *     host memory.
 *   - Every d_bytes buffer (bodies, XML text) needs 8 readable bytes past its last byte: the kernels that read
 *     a body straight from global memory (MD5 keys, the legacy path, bodies larger than the keys kernel's staging
 *     buffer) fetch aligned 4-byte words and may touch the word after the one holding the last byte.  Buffers the
 *     library allocates itself (the host-buffer calls) have that slack; any CUDA allocation that is not filled to
 *     its last byte has it too.
 *   - Which calls allocate.  The replay path proper allocates NOTHING per call: smsgate_scan_keys,
 *     smsgate_join_route, smsgate_dedup, smsgate_join_route_dedup, smsgate_replay_device, smsgate_replay_host*
 *     (all forms), smsgate_fetch_columns, smsgate_device_keys, smsgate_cross_source, the smsgate_dedup_export /
 *     _merge / _apply and smsgate_p2p_* exchange steps, smsgate_dedup_static - their scratch lives in the context
 *     and is sized once by smsgate_reserve / smsgate_reserve_merge / smsgate_p2p_reserve / _reserve_filter.
 *     Set-up and batch-helper calls DO allocate (and free) device memory inside the call: smsgate_create,
 *     smsgate_reserve*, smsgate_cache_load, smsgate_decimal_batch, smsgate_strict_date_batch / _entries,
 *     smsgate_dt_str_batch, smsgate_legacy_parse*, smsgate_xml_scan / _attrs_sizes.  smsgate_dcache_* and
 *     smsgate_emit_json are host code and use host memory only.
 *   - A batch is Arrow-style: int64 offsets[n+1] + uint8 bytes[offsets[n]],
 *     UTF-8, offsets[0] may be non-zero, `bytes` must be 16-byte aligned.
 */
#ifndef SMSGATE_B200_H
#define SMSGATE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SMSGATE_ABI_VERSION 2

/* error codes */
#define SMSGATE_OK 0
#define SMSGATE_E_CUDA (-1)
#define SMSGATE_E_ARG (-2)
#define SMSGATE_E_NOMEM (-3)
#define SMSGATE_E_STATE (-4)
#define SMSGATE_E_LIMIT (-5)
#define SMSGATE_E_XML (-6)         /* the document is not well-formed: what ElementTree reports as ParseError */
#define SMSGATE_E_UNSUPPORTED (-7) /* well-formedness undecided: a construct outside the supported XML dialect */

/* flags[] bits written by the scan */
#define SMSGATE_F_WORKER_SKIP 0x01 /* services/parser_worker/worker.py:112-121 */
#define SMSGATE_F_PARSER_OTP 0x02  /* libs/gemini_parser.py:198-200 */
#define SMSGATE_F_MODIFIED 0x04    /* clean/mask changed the body (libs/gemini_parser.py:202-203) */
#define SMSGATE_F_BODY_DATE 0x08   /* the body carries a usable dd.mm.yy[yy] (libs/gemini_parser.py:73-100) */

/* status[] values: where the reference would have sent the message */
#define SMSGATE_ST_UPSERT 0             /* sms.parsed, truthy merchant: upserted (services/pb_writer/writer.py:70-74) */
#define SMSGATE_ST_PARSED_NO_MERCHANT 1 /* sms.parsed, writer acks without upsert (writer.py:70,75) */
#define SMSGATE_ST_SKIPPED_FILTER 2     /* worker skip filter: acked, nothing published (worker.py:112-126) */
#define SMSGATE_ST_CACHE_MISS 3         /* would call Gemini (libs/gemini_parser.py:210-222); offline: DLQ set */
#define SMSGATE_ST_UNMATCHED 4          /* parse_sms_llm returned None -> sms.failed (worker.py:151-158) */
#define SMSGATE_ST_BROKEN 5             /* BrokenMessage: acked, dropped (worker.py:136-140) */
#define SMSGATE_ST_FUTURE_DATE 6        /* parsed.date > now (worker.py:174): the branch means to publish to sms.failed, but its
                                           `msg.data.decode(...)` (:177) runs on a str and raises - run as written, the handler
                                           dies there and nothing is published or acked (tests/golden/worker_flow.json) */
#define SMSGATE_ST_FAILED_EXC 7         /* parse_sms_llm raised something else -> sms.failed (worker.py:141-149) */
#define SMSGATE_ST_ESCAPED 8            /* parsed, but the date is timezone-aware: `parsed.date > datetime.now()` (worker.py:174)
                                           raises TypeError outside every try - nothing is published or acked */

/* per cache entry flags supplied by the loader (see smsgate_cache_load) */
#define SMSGATE_E_IS_DICT 0x0001
#define SMSGATE_E_HAS_TXN 0x0002
#define SMSGATE_E_SCHEMA_OK 0x0004
#define SMSGATE_E_CARD_STR 0x0008
#define SMSGATE_E_MERCHANT 0x0010
#define SMSGATE_E_DATE_STRICT 0x0020
#define SMSGATE_E_DATE_HOST 0x0040
#define SMSGATE_E_DATE_NOBODY 0x0080
#define SMSGATE_E_DATE_TZ 0x0100
#define SMSGATE_E_DATE_KEY 0x0200
#define SMSGATE_E_DATE_PHONE 0x0400 /* the cached date holds no date text: the phone's timestamp stands in (below) */

/* dec_flags[] layout: bits 0-1 amount error, 2-3 balance error, bit 4 amount sign, bit 5 balance sign,
 * bit 6 / bit 7: amount / balance does not fit Numeric(14,2) (more than 12 integer digits) */
#define SMSGATE_DEC_OK 0
#define SMSGATE_DEC_SYNTAX 1   /* parse_ambiguous_decimal raises ValueError */
#define SMSGATE_DEC_OVERFLOW 2 /* |coefficient| > INT64_MAX: representation limit of this library */

typedef struct smsgate_ctx smsgate_ctx;

/* Per-message outputs.  Any pointer may be NULL to skip that column, except
 * that later stages need the columns of earlier ones when run separately.
 * Device pointers for the *_device calls, host pointers for smsgate_replay_host. */
typedef struct smsgate_outputs {
    uint8_t* msg_id;      /* [n*20] SHA-1(body) digest bytes (services/xml_watcher/watcher.py:45) */
    uint8_t* cache_key;   /* [n*32] SHA-256(fixed_body) digest bytes (libs/gemini_parser.py:207) */
    uint8_t* flags;       /* [n] SMSGATE_F_* */
    int32_t* body_date;   /* [n] y<<9|m<<5|d lifted from the body, 0 if none */
    int32_t* fixed_len;   /* [n] byte length of fixed_body */
    uint8_t* fixed_bytes; /* [same size as the input bytes] fixed_body of message i at offsets[i],
                             valid only where flags has SMSGATE_F_MODIFIED; raw_body otherwise is the body */
    int32_t* slot;        /* [n] index of the matching cache entry, -1 on a miss */
    uint8_t* status;      /* [n] SMSGATE_ST_* */
    int64_t* amount_coef; /* [n] |coefficient| of parse_ambiguous_decimal(str(amount)) (libs/decimal_utils.py:4-63) */
    int64_t* balance_coef;
    int32_t* amount_exp;  /* [n] Decimal exponent */
    int32_t* balance_exp;
    uint8_t* dec_flags;   /* [n] */
    int64_t* date;        /* [n] resolved naive datetime, packed y<<46|m<<42|d<<37|H<<32|M<<26|S<<20|us; 0 if none */
    uint8_t* card;        /* [n*16] first four characters of the normalised card, UTF-8, zero padded */
    uint8_t* card_len;    /* [n] bytes used in card */
    uint8_t* survivor;    /* [n] 1 = last occurrence of its msg_id among upserts: the record that survives */
    uint8_t* first_seen;  /* [n] 1 = first occurrence: an insert rather than a patch */
    int64_t* last_idx;    /* [n] global index of the surviving occurrence, -1 if the message is not upserted */
    int64_t* amount_cents;  /* [n] |amount| as a Numeric(14,2) column stores it (db/models.py:21): two decimals, halves
                               away from zero; sign is dec_flags bit 4 */
    int64_t* balance_cents; /* [n] same for the balance (db/models.py:27), sign in dec_flags bit 5 */
} smsgate_outputs;

/* The result of one message as one 64-byte record (smsgate_replay_host_records): the same values as
 * the columns of the same names in smsgate_outputs.  The fields a consumer of the reference's
 * parse_sms_llm / upsert path needs per message travel together, which is also how they are read. */
typedef struct smsgate_record {
    uint8_t msg_id[20];
    uint8_t status, flags, dec_flags, card_len;
    int32_t slot;
    int32_t amount_exp, balance_exp;
    int32_t body_date;
    int64_t amount_coef, balance_coef;
    int64_t date;
} smsgate_record;

/* ---- lifecycle ---------------------------------------------------------- */
int smsgate_abi_version(void);
/* Create a context on CUDA device `device`.  cuda_stream is the cudaStream_t
 * every copy and kernel of this context is issued on (NULL = CUDA's default
 * stream).  No reference counterpart. */
int smsgate_create(int device, void* cuda_stream, smsgate_ctx** out);
void smsgate_destroy(smsgate_ctx* ctx);
/* Text of the last error on this context (or of the last failed create when ctx is NULL). */
const char* smsgate_last_error(const smsgate_ctx* ctx);
/* Size the context's scratch (dedup table, staging for the host call) for
 * batches of up to max_msgs messages / max_bytes body bytes.  Allocates. */
int smsgate_reserve(smsgate_ctx* ctx, int64_t max_msgs, int64_t max_bytes);
/* Queue everything this context does from now on onto another cudaStream_t of its device.  The caller
 * orders the streams (events): this is how a multi-GPU replay runs the dedup exchange of batch k on a
 * side stream while the keys kernel of batch k+1 runs on the main one (smsgate_b200/dist.py).
 * No reference counterpart (the reference is single-threaded asyncio, services/xml_watcher/watcher.py:78). */
int smsgate_set_stream(smsgate_ctx* ctx, void* cuda_stream);
/* How many CTAs smsgate_scan_keys may use: 0 = one per SM (default).  Fewer leaves SMs for kernels of
 * another stream (the exchange above) while the keys kernel, which otherwise fills every SM, runs. */
int smsgate_set_scan_ctas(smsgate_ctx* ctx, int ctas);
/* Kernels launched by this context since creation (for bench accounting). */
int64_t smsgate_launch_count(const smsgate_ctx* ctx);

/* ---- the Gemini response cache ------------------------------------------
 * Replaces `cache = diskcache.Cache(".gemini_cache")` + `cache_key in cache` +
 * `cache[cache_key]` (libs/gemini_parser.py:33, :208-209) by a device hash
 * table.  All pointers are HOST memory; the call copies and builds.
 *   keys        [m*32]   SHA-256 digests (the hex keys of the cache, decoded)
 *   val_offsets [4*m+1]  Arrow offsets into val_bytes; entry e owns strings
 *                        4e..4e+3 = str(amount), str(balance), date, card
 *   eflags      [m]      SMSGATE_E_* (schema facts the loader established)
 *   date_packed [m]      packed datetime for SMSGATE_E_DATE_HOST / _NOBODY entries
 */
int smsgate_cache_load(smsgate_ctx* ctx, int64_t m, const uint8_t* h_keys, const int32_t* h_val_offsets,
                       const uint8_t* h_val_bytes, const uint16_t* h_eflags, const int64_t* h_date_packed);
int64_t smsgate_cache_size(const smsgate_ctx* ctx);
/* Let `ctx` probe the table `owner` loaded instead of loading its own copy (two contexts on one GPU that
 * keep two batches in flight, see smsgate_set_stream).  `owner` must outlive the sharing and must not
 * reload its cache while `ctx` uses it; a later smsgate_cache_load / _cache_share on `ctx` ends the sharing.
 * Replaces the same reference code as smsgate_cache_load: one `.gemini_cache` per process
 * (libs/gemini_parser.py:33). */
int smsgate_cache_share(smsgate_ctx* ctx, const smsgate_ctx* owner);

/* The reference's `.gemini_cache` directory read natively: diskcache 5.6.3's SQLite file and its pickled
 * response dicts decoded in C++ (smsgate_b200/csrc/cache_reader.h), no Python object per entry.  Replaces
 * `cache = diskcache.Cache(".gemini_cache")` + the per-message `cache_key in cache` / `cache[cache_key]`
 * (libs/gemini_parser.py:33, :208-209) by one pass that yields the arrays smsgate_cache_load takes.
 *   _open   reads <cache_dir>/cache.db (read-only); `now` = time.time() for diskcache's expiry rule.  A large table
 *           is read as up to eight rowid ranges at once, each on a read-only connection of its own (environment
 *           variable SMSGATE_DCACHE_THREADS overrides the count; 1 = one connection); the cache is expected not
 *           to be written while it is read
 *   _sizes  m entries, bytes of the four device strings, bytes of the five host-side strings, keys skipped
 *           because they are not 64 lower-case hex digits, entries the decoder leaves to the caller
 *   _fill   keys[m*32], val_offsets[4m+1], val_bytes, eflags[m] (schema facts; date flags are the caller's:
 *           smsgate_strict_date_entries), rowids[m], needs_host[m] (1 = decode this entry yourself, by rowid;
 *           2 = only its date needs the reference's strptime/dateutil chain), and for building payloads on
 *           the host text_offsets[5m+1] + text_bytes + text_kind[5m] = txn_type, currency, merchant, city,
 *           address (kind: 0 missing, 1 None, 2 str, 3 other).  Any pointer may be NULL.
 * Host memory only; no GPU involved. */
typedef struct smsgate_dcache smsgate_dcache;
int smsgate_dcache_open(const char* cache_dir, double now, smsgate_dcache** out, char* errbuf, int errbuf_len);
int smsgate_dcache_sizes(const smsgate_dcache* h, int64_t* m, int64_t* value_bytes, int64_t* text_bytes,
                         int64_t* skipped_keys, int64_t* needs_host);
int smsgate_dcache_fill(const smsgate_dcache* h, uint8_t* keys, int32_t* val_offsets, uint8_t* val_bytes, uint16_t* eflags,
                        int64_t* rowids, uint8_t* needs_host, int64_t* text_offsets, uint8_t* text_bytes, uint8_t* text_kind);
void smsgate_dcache_close(smsgate_dcache* h);
/* smsgate_strict_date_batch over the date string of every entry in smsgate_cache_load's layout (string 4e+2):
 * h_packed[e] = packed datetime, or -1 (empty / not in the strict form). */
int smsgate_strict_date_entries(smsgate_ctx* ctx, int64_t m, const int32_t* h_val_offsets, const uint8_t* h_val_bytes,
                                int64_t* h_packed);

/* ---- batch helpers (host buffers in and out) ----------------------------- */
/* libs.decimal_utils.parse_ambiguous_decimal over n strings (libs/decimal_utils.py:4-63).
 * flags[i]: bits 0-1 SMSGATE_DEC_*, bit 4 sign. */
int smsgate_decimal_batch(smsgate_ctx* ctx, int64_t n, const int32_t* h_offsets, const uint8_t* h_bytes,
                          int64_t* h_coef, int32_t* h_exp, uint8_t* h_flags);
/* datetime.strptime(s, '%d.%m.%y %H:%M') for the ASCII subset evaluated on the
 * GPU (libs/gemini_parser.py:116); out[i] = packed datetime or -1 when the
 * string needs the reference's strptime/dateutil chain on the host. */
int smsgate_strict_date_batch(smsgate_ctx* ctx, int64_t n, const int32_t* h_offsets, const uint8_t* h_bytes,
                              int64_t* h_packed);
/* `_dt_str(date, time_)` of save_to_pocketbase.py:34-43 over a ragged batch of the strings
 * f"{date} {time_}": packed datetime (same packing as smsgate_outputs.date) of the first of the six
 * formats %d.%m.%Y, %d/%m/%Y, %d-%m-%Y, %d.%m.%y, %d/%m/%y, %d-%m-%y (+ " %H:%M") that
 * datetime.strptime accepts, 0 where it returns None.  Host arrays; allocates temporaries. */
int smsgate_dt_str_batch(smsgate_ctx* ctx, int64_t n, const int32_t* h_offsets, const uint8_t* h_bytes,
                         int64_t* h_packed);

/* ---- XML backup -> ragged batch on the GPU -----------------------------------
 * Replaces, for a whole backup file at once, `ET.parse(path)` + `root.findall("sms")` +
 * `elem.get("body" / "date" / "address")` + `int(date)` of `_iter_sms`
 * (services/xml_watcher/watcher.py:37-44; the same loop in read_xml.py:25-37).
 * smsgate_xml_scan takes the file bytes (host memory), tokenizes and validates them on the
 * GPU and reports how many top-level <sms> elements there are and how many bytes their decoded
 * body / address attributes take; smsgate_xml_emit then writes the batch into caller-owned
 * device buffers: body_offsets[n+1] + body_bytes (exactly the input of smsgate_scan_keys),
 * sender_offsets[n+1] + sender_bytes, date_ms[n] and mflags[n].  The scan keeps its
 * intermediate state in the context until emit or smsgate_xml_release.
 * Dialect: UTF-8, declaration / processing instructions, comments without '<', elements
 * nested at most 64 deep, at most 64 attributes per element, the five predefined entities
 * and numeric character references.  SMSGATE_E_XML = not well-formed (err_pos = byte
 * offset); SMSGATE_E_UNSUPPORTED = DOCTYPE / CDATA / another encoding / '<' in a comment:
 * hand the file to a general parser. */
#define SMSGATE_XML_HAS_BODY 0x01    /* the element has a body attribute (else elem.get gives "") */
#define SMSGATE_XML_HAS_ADDRESS 0x02
#define SMSGATE_XML_DATE_BAD 0x04    /* int(date) raises ValueError */
#define SMSGATE_XML_DATE_BIG 0x08    /* int(date) is valid but beyond int64 */
int smsgate_xml_scan(smsgate_ctx* ctx, const uint8_t* h_xml, int64_t nbytes, int64_t* n_out, int64_t* body_bytes,
                     int64_t* sender_bytes, int64_t* err_pos);
int smsgate_xml_emit(smsgate_ctx* ctx, uint8_t* d_body_bytes, int64_t* d_body_offsets, uint8_t* d_sender_bytes,
                     int64_t* d_sender_offsets, int64_t* d_date_ms, uint8_t* d_mflags);
int smsgate_xml_release(smsgate_ctx* ctx);
/* Between smsgate_xml_scan and smsgate_xml_emit / _release: EVERY attribute of every top-level <sms>, names as
 * written and values decoded like the ones above - what `elem.attrib` holds in read_xml.parse_and_cache_sms
 * (read_xml.py:32-41).  _attrs_sizes reports the totals; _attrs_emit fills caller-owned device buffers:
 * attr_offsets[n+1] (message i owns attributes attr_offsets[i] .. attr_offsets[i+1], in document order),
 * name_offsets[A+1] + name_bytes, value_offsets[A+1] + value_bytes.  Allocates temporaries in the context. */
int smsgate_xml_attrs_sizes(smsgate_ctx* ctx, int64_t* n_attrs, int64_t* name_bytes, int64_t* value_bytes);
int smsgate_xml_attrs_emit(smsgate_ctx* ctx, int64_t* d_attr_offsets, int64_t* d_name_offsets, uint8_t* d_name_bytes,
                           int64_t* d_value_offsets, uint8_t* d_value_bytes);

/* ---- the legacy offline replay (host buffers in and out) --------------------
 * process_cached.process_sms_from_cache (process_cached.py:210-268) for n source-cache
 * bodies at once: strip (:233), OTP filter on body.upper() (:236), the debit pattern
 * TRANSACTION_RE via parse_transaction_message (:98-120, :150-177), else the credit
 * pattern CREDIT_PAYMENT_RE via parse_credit_message (:122-135, :180-204), `_to_decimal`
 * -> float (:141-143) and the SHA-256 of the stripped body that keys the result caches
 * (:254).  Both patterns run on the GPU in a backtracking matcher that tries
 * alternatives in the order Python's `re` does. */
#define SMSGATE_LEG_SKIPPED_OTP 0 /* status "skipped_otp" */
#define SMSGATE_LEG_DEBIT 1       /* "processed", goes to the purchase cache */
#define SMSGATE_LEG_CREDIT 2      /* "processed", goes to the credit cache */
#define SMSGATE_LEG_FAILED 3      /* "failed_to_parse" */
#define SMSGATE_LEG_F_BUDGET 0x01  /* matcher budget (4 M steps / 128 pending choices) ran out: treat as failed */
#define SMSGATE_LEG_F_INEXACT 0x02 /* amount or balance beyond 2^53 / 22 fraction digits: float only approximate */
#define SMSGATE_LEG_GROUPS 10
/* group numbers inside caps[]: debit 1 merchant 2 city 3 address 4 date 5 time 6 card 7 amount 8 currency 9 balance;
 * credit 1 type 2 date 3 time 4 card 5 amount 6 currency 7 balance */
typedef struct smsgate_legacy_outputs {
    uint8_t* status;    /* [n] SMSGATE_LEG_* */
    uint8_t* lflags;    /* [n] SMSGATE_LEG_F_* */
    int32_t* strip;     /* [n*2] byte span of body.strip() inside the body */
    int32_t* caps;      /* [n*2*SMSGATE_LEG_GROUPS] byte spans of the groups inside the stripped body, already
                           trimmed where the reference calls .strip(); -1 = group did not take part */
    double* amount;     /* [n] float(_to_decimal(amount)) */
    double* balance;    /* [n] */
    uint8_t* body_hash; /* [n*32] SHA-256 of the stripped body */
} smsgate_legacy_outputs;
int smsgate_legacy_parse(smsgate_ctx* ctx, int64_t n, const int64_t* h_offsets, const uint8_t* h_bytes,
                         const smsgate_legacy_outputs* h_out);
/* The same kernel doing what the reference's single-message functions do: parse_transaction_message
 * (process_cached.py:150-177) = NO_STRIP | NO_OTP | NO_CREDIT, parse_credit_message (:180-204) =
 * NO_STRIP | NO_OTP | NO_DEBIT - one pattern, applied to the message as given.  mode 0 = smsgate_legacy_parse. */
#define SMSGATE_LEG_MODE_NO_STRIP 0x1
#define SMSGATE_LEG_MODE_NO_OTP 0x2
#define SMSGATE_LEG_MODE_NO_DEBIT 0x4
#define SMSGATE_LEG_MODE_NO_CREDIT 0x8
int smsgate_legacy_parse_mode(smsgate_ctx* ctx, int64_t n, const int64_t* h_offsets, const uint8_t* h_bytes,
                              uint32_t mode, const smsgate_legacy_outputs* h_out);

/* ---- the replay stages, device pointers ----------------------------------- */
/* Stage 1 — keys and scans: msg_id (watcher.py:45), worker skip filter
 * (worker.py:112-121), parser OTP filter (gemini_parser.py:198-200), clean +
 * card masking (:202-203, :121-137), cache key (:207), body date (:73-100).
 * Writes msg_id, cache_key, flags, body_date, fixed_len, fixed_bytes. */
int smsgate_scan_keys(smsgate_ctx* ctx, const uint8_t* d_bytes, const int64_t* d_offsets, int64_t n,
                      const smsgate_outputs* out);
/* Stage 2 — cache join, amounts, date, card, routing (gemini_parser.py:208-247,
 * decimal_utils.py:4-63, worker.py:136-189, writer.py:65-84).  Reads
 * cache_key, flags, body_date; writes slot, status, amount_*, balance_*,
 * dec_flags, date, card, card_len.  now_packed is the frozen `datetime.now()`. */
int smsgate_join_route(smsgate_ctx* ctx, int64_t n, int64_t now_packed, const smsgate_outputs* out);
/* Messages that carry a unix timestamp in raw.date (device-sourced messages, re-parsed DLQ entries): when the
 * cached date holds no date text, parse_sms_llm falls back to parse_unix_timestamp(int(raw.date),
 * tz="Asia/Yerevan", aware=False) (libs/gemini_parser.py:227-229, :139-188).  d_phone_dt[n] = that naive
 * datetime, packed like smsgate_outputs.date, or 0 where int(raw.date) / the conversion raises; the pointer is
 * used by every later smsgate_join_route / smsgate_join_route_dedup on this context until it is reset with
 * NULL.  Without it (XML replay: raw.date is ISO text, int() raises) such messages end as UNMATCHED. */
int smsgate_set_phone_dates(smsgate_ctx* ctx, const int64_t* d_phone_dt);
/* Stage 3 — upsert-by-msg_id (libs/pocketbase.py:168-200,
 * services/pb_writer/upsert.py:19-31): among messages with status UPSERT the
 * last occurrence of a msg_id survives, the first one is the insert.  Reads
 * msg_id, status; writes survivor, first_seen, last_idx (= base_index + local). */
int smsgate_dedup(smsgate_ctx* ctx, int64_t n, int64_t base_index, const smsgate_outputs* out);
/* Stages 2 and 3 in one call (what smsgate_replay_device runs after stage 1): the join kernel inserts each
 * upsert into the dedup table as it routes it.  Same columns, same values as smsgate_join_route followed by
 * smsgate_dedup. */
int smsgate_join_route_dedup(smsgate_ctx* ctx, int64_t n, int64_t now_packed, int64_t base_index,
                             const smsgate_outputs* out);
/* Device-source keys.  d_md5[n*16] = MD5(UTF-8 body) of every message: the msg_id the
 * API gateway assigns, get_md5_hash (libs/models.py:97-109) as called at services/api_gateway/main.py:113.
 * Device pointers; d_bytes needs 8 readable bytes past the last body, d_md5 16-byte alignment. */
int smsgate_device_keys(smsgate_ctx* ctx, const uint8_t* d_bytes, const int64_t* d_offsets, int64_t n, uint8_t* d_md5);
/* Bodies the reference stores twice because the two sources key them differently (SHA-1 of the body for
 * XML, services/xml_watcher/watcher.py:45; MD5 for device messages, main.py:113): d_match[i] = index of
 * the first of the n_xml XML messages whose body equals device message i's, or -1.  Bodies are compared
 * through their SHA-1 (the msg_id column of smsgate_scan_keys on either batch).  Uses the dedup scratch:
 * n_xml must fit smsgate_reserve().  No counterpart in the reference (it never notices). */
int smsgate_cross_source(smsgate_ctx* ctx, int64_t n_xml, const uint8_t* d_xml_sha1, int64_t n_dev,
                         const uint8_t* d_dev_sha1, int32_t* d_match);

/* All three stages on one batch. */
int smsgate_replay_device(smsgate_ctx* ctx, const uint8_t* d_bytes, const int64_t* d_offsets, int64_t n,
                          int64_t now_packed, int64_t base_index, const smsgate_outputs* out);

/* ---- global upsert-by-msg_id across GPUs -----------------------------------
 * The reference gets global idempotency from one PocketBase/Postgres unique key
 * on msg_id (db/models.py:16, services/pb_writer/upsert.py:19-31).  With the
 * backup sharded by index over `world` GPUs the same guarantee needs ONE exchange:
 * after smsgate_dedup on its shard, every rank exports one 32-byte record per
 * locally distinct msg_id {key[20], first48, last48 (global indices)}, grouped by
 * the rank that owns the key's 32-bit prefix ((prefix * world) >> 32); the records
 * travel in one all-to-all (NCCL over NVLink, done by the caller, e.g.
 * torch.distributed.all_to_all_single); the owner merges them and answers each
 * record with {global first, global last}; the answers return on the mirrored
 * all-to-all and smsgate_dedup_apply rewrites survivor / first_seen / last_idx. */
#define SMSGATE_MAX_WORLD 64
#define SMSGATE_RECORD_BYTES 32
/* Writes the records (destination-contiguous) to d_records [<= n*32], the local slot
 * of each record to d_rec_slot [<= n] and the per-destination counts to HOST
 * h_counts[world].  Synchronises the stream (the counts are the split sizes). */
int smsgate_dedup_export(smsgate_ctx* ctx, int64_t n, int64_t base_index, int world, const smsgate_outputs* out,
                         uint8_t* d_records, int32_t* d_rec_slot, int64_t* h_counts);
/* Size the owner-side merge table for up to max_records received records.  Allocates. */
int smsgate_reserve_merge(smsgate_ctx* ctx, int64_t max_records);
/* Owner side: merge n_recv received records, write {first, last} per record to d_verdicts [n_recv*2]. */
int smsgate_dedup_merge(smsgate_ctx* ctx, int64_t n_recv, const uint8_t* d_records, int64_t* d_verdicts);
/* Home side: d_verdicts [n_rec*2] in the order of the exported records; rewrites
 * out->survivor, out->first_seen, out->last_idx of the n local messages. */
int smsgate_dedup_apply(smsgate_ctx* ctx, int64_t n, int64_t base_index, int64_t n_rec, const int32_t* d_rec_slot,
                        const int64_t* d_verdicts, const smsgate_outputs* out);

/* ---- the same exchange over peer memory (NVLink P2P) ------------------------
 * One process per GPU; every rank maps every other rank's inbox, arrival counters and
 * verdict arrays (e.g. through torch.distributed._symmetric_memory) and passes the
 * mapped pointers in `smsgate_p2p`.  smsgate_p2p_scatter stores the export records
 * straight into the owners' inboxes (one remote atomic per block and destination reserves
 * the range), smsgate_p2p_merge stores a verdict straight into the home rank's gfl array
 * whenever another rank's copy of the key changes it (and sets the slot's bit in the home rank's override
 * bitmap), smsgate_p2p_resolve then rewrites survivor / first_seen / last_idx of exactly those messages: the
 * columns must already hold the local verdicts (smsgate_dedup / smsgate_join_route_dedup).
 * The caller puts a cross-GPU barrier between scatter and merge and between merge and
 * resolve; nothing is copied through host memory and no count is exchanged.
 * `parity` alternates 0/1 from one batch to the next (two arrival counters). */
#define SMSGATE_P2P_RECORD_BYTES 48
typedef struct smsgate_p2p {
    int32_t world, rank;
    uint8_t* inbox[SMSGATE_MAX_WORLD];    /* [capacity * SMSGATE_P2P_RECORD_BYTES] per rank */
    int32_t* in_count[SMSGATE_MAX_WORLD]; /* [2] per rank, zero before the first batch */
    int64_t* gfl[SMSGATE_MAX_WORLD];      /* [2 * (pow2 >= 2 * max batch)] per rank: {first, last} per local slot, 16-byte aligned */
    int64_t capacity;                     /* records per inbox; world * max batch never overflows */
    /* prefilter (optional, bitmap_words = 0 switches it off): only keys that two or more ranks hold are exchanged */
    uint32_t* stage[SMSGATE_MAX_WORLD];   /* [world * (bitmap_words / world + 12)] words per rank: staged bitmap slices */
    uint32_t* collide[SMSGATE_MAX_WORLD]; /* [bitmap_words] words per rank: bits set on two or more ranks */
    int64_t bitmap_words;                 /* power of two in [64, 2^26]; about 16 bits per key of the whole job is plenty */
    int32_t* flags[SMSGATE_MAX_WORLD];    /* [SMSGATE_MAX_WORLD] int32 per rank, zero before the first use: arrival flags of
                                             smsgate_p2p_barrier (NULL when the caller brings its own barrier) */
    uint32_t* ovr[SMSGATE_MAX_WORLD];     /* [2 * ovr_words] per rank, zero before the first batch: one bit per local slot (one half
                                             per parity) set by the owner that answered the slot with a different global first /
                                             last; smsgate_p2p_resolve rewrites only those */
    int64_t ovr_words;                    /* >= gfl_slots / 32 */
    int64_t gfl_slots;                    /* slots every gfl array holds (a pair of int64 each); a batch whose table would
                                             need more is refused with SMSGATE_E_LIMIT instead of storing out of bounds */
} smsgate_p2p;
/* Size the owner-side table for inboxes of `capacity` records.  Allocates. */
int smsgate_p2p_reserve(smsgate_ctx* ctx, int64_t capacity);
int smsgate_p2p_scatter(smsgate_ctx* ctx, int64_t n, int64_t base_index, int parity, const smsgate_outputs* out,
                        const smsgate_p2p* p);
int smsgate_p2p_merge(smsgate_ctx* ctx, int parity, const smsgate_p2p* p);
/* The prefilter, run before smsgate_p2p_scatter when the descriptor carries bitmaps: _filter_push marks this
 * rank's keys in a bitmap and stores slice w of it into rank w's staging area; [barrier]; _filter_reduce turns
 * the staged slices into "held by two or more ranks" and stores that slice into every rank's `collide`;
 * [barrier]; smsgate_p2p_scatter then sends only keys whose bit is set.  False positives cost a record, there
 * are no false negatives.  _reserve_filter allocates this rank's private bitmap. */
int smsgate_p2p_reserve_filter(smsgate_ctx* ctx, int64_t bitmap_words);
int smsgate_p2p_filter_push(smsgate_ctx* ctx, int64_t n, const smsgate_outputs* out, const smsgate_p2p* p);
int smsgate_p2p_filter_reduce(smsgate_ctx* ctx, const smsgate_p2p* p);
/* The cross-GPU barrier the steps above need between them, queued on the context's stream like any kernel (no
 * host synchronisation): every rank calls it with the same `epoch`, which must grow by one from call to call.
 * It returns (on the GPU) once every rank's earlier stores into peer memory are visible here.  One tiny block,
 * so it is scheduled even while another stream's kernels fill the SMs.  A rank that waits ~10 s gives up and
 * raises the overflow flag (reported as 2 by smsgate_p2p_resolve): the ranks have to reach a barrier within that
 * time of each other. */
int smsgate_p2p_barrier(smsgate_ctx* ctx, const smsgate_p2p* p, int epoch);
/* Let the join kernel's fused upsert insert (smsgate_join_route_dedup, smsgate_replay_host*) mark this rank's
 * prefilter bitmap as it goes, so that smsgate_p2p_filter_push has no pass of its own over the batch left
 * (bitmap_words must equal the descriptor's; 0 switches it off). */
int smsgate_p2p_fuse_filter(smsgate_ctx* ctx, int64_t bitmap_words);
/* The whole exchange of one batch in one call: filter_push, barrier, filter_reduce, barrier (when the descriptor
 * has bitmaps), scatter, barrier, merge, barrier, resolve - barriers numbered epoch_base+1 ....  Returns the number
 * of barriers used (>= 0) or a negative SMSGATE_E_* code. */
int smsgate_p2p_exchange(smsgate_ctx* ctx, int64_t n, int64_t base_index, int parity, int epoch_base,
                         const smsgate_outputs* out, const smsgate_p2p* p);
/* h_overflow (optional, host): set to 1 if an inbox ever overflowed, 2 if a barrier timed out; passing it synchronises. */
int smsgate_p2p_resolve(smsgate_ctx* ctx, int64_t n, int64_t base_index, int parity, const smsgate_outputs* out,
                        const smsgate_p2p* p, int* h_overflow);

/* ---- the replay, host buffers --------------------------------------------
 * What `_iter_sms` -> worker -> parse_sms_llm -> pb_writer do for a whole
 * backup, called with the batch in host memory: copies in, runs the three
 * stages, copies every non-NULL output column back and synchronises.
 * total_bytes = h_offsets[n].  Needs smsgate_reserve() to have covered n. */
int smsgate_replay_host(smsgate_ctx* ctx, const uint8_t* h_bytes, const int64_t* h_offsets, int64_t n,
                        int64_t now_packed, int64_t base_index, const smsgate_outputs* h_out);

/* Tuning of the host call's pipeline, no counterpart in the reference: every chunk takes 1/divisor of what is left
 * of the batch (at least min_chunk messages), and the offsets travel either with their chunk or all at once ahead
 * of the bodies.  Defaults 3, 131072, 1 (the best of a sweep at 1 M messages).  At most 16 chunks: the last one takes
 * the rest. */
int smsgate_set_host_pipeline(smsgate_ctx* ctx, int divisor, int64_t min_chunk, int offsets_first);
/* Diagnostics, no counterpart in the reference: where the time of the last smsgate_replay_host()
 * went.  Fills ms[] with GPU times relative to the start of that call: ms[0..c) = host->device copy
 * of chunk k complete, ms[c..2c) = kernels of chunk k complete, ms[2c] = dedup complete,
 * ms[2c+1] = last byte stored to the host buffers.  Returns c (the number of chunks) or a negative
 * SMSGATE_E_* code; `capacity` must be at least 2c+2 (34 is always enough). */
int smsgate_replay_host_timeline(smsgate_ctx* ctx, float* ms, int capacity);
/* The same call in two halves, for a multi-GPU replay that has to run the global
 * dedup exchange on the device-resident columns in between: _begin queues the
 * chunked copy-in, stages 1-3 and the copy-out of every column except survivor /
 * first_seen / last_idx and returns without waiting; smsgate_device_view gives the
 * device columns of the pending batch; _finish copies the three dedup columns out
 * and synchronises.  Host buffers should be pinned for the copies to overlap. */
int smsgate_replay_host_begin(smsgate_ctx* ctx, const uint8_t* h_bytes, const int64_t* h_offsets, int64_t n,
                              int64_t now_packed, int64_t base_index, const smsgate_outputs* h_out);
int smsgate_device_view(smsgate_ctx* ctx, int64_t n, smsgate_outputs* d_out);
int smsgate_replay_host_finish(smsgate_ctx* ctx, const smsgate_outputs* h_out);

/* smsgate_replay_host with packed results: h_records[n] receives one smsgate_record per message and
 * h_verdict[n] the dedup verdict of message i as one integer: -1 when the message is not upserted, else
 * (last_idx << 2) | (first_seen << 1) | survivor.  Same work, same values as smsgate_replay_host; the
 * difference is the transport: one device->host transfer per chunk instead of one per column and chunk,
 * which is what the end-to-end time is sensitive to.  The _begin / _finish halves are for the multi-GPU
 * exchange, as above.  No counterpart in the reference (it has no batch API). */
int smsgate_replay_host_records(smsgate_ctx* ctx, const uint8_t* h_bytes, const int64_t* h_offsets, int64_t n,
                                int64_t now_packed, int64_t base_index, smsgate_record* h_records, int64_t* h_verdict);
int smsgate_replay_host_records_begin(smsgate_ctx* ctx, const uint8_t* h_bytes, const int64_t* h_offsets, int64_t n,
                                      int64_t now_packed, int64_t base_index, smsgate_record* h_records);
int smsgate_replay_host_records_finish(smsgate_ctx* ctx, int64_t* h_verdict);
/* The same with ONE byte of verdict per message instead of eight: 0 = the message is not upserted, else
 * 0x80 | first_seen << 1 | survivor.  The index of the surviving occurrence (last_idx) stays on the device and
 * can be fetched on request (smsgate_fetch_columns).  8 MB less to bring back per 1 M messages, all of it after
 * the last kernel: this is the call the end-to-end figure uses. */
int smsgate_replay_host_records8(smsgate_ctx* ctx, const uint8_t* h_bytes, const int64_t* h_offsets, int64_t n,
                                 int64_t now_packed, int64_t base_index, smsgate_record* h_records, uint8_t* h_verdict8);
int smsgate_replay_host_records8_finish(smsgate_ctx* ctx, uint8_t* h_verdict8);
/* Columns of the LAST smsgate_replay_host* batch that the packed record does not carry (card, cache_key,
 * fixed_len, fixed_bytes, amount_cents, balance_cents) or any other column, copied to the non-NULL host
 * pointers of h_out on request: the device keeps every column of a batch until the next replay or reserve
 * on this context.  n must be that batch's size; SMSGATE_E_STATE when the batch is gone.  Synchronises.
 * This is what lets `parse_sms_llm`-style materialisation (libs/gemini_parser.py:249-266) work on a result
 * that came back as records. */
int smsgate_fetch_columns(smsgate_ctx* ctx, int64_t n, const smsgate_outputs* h_out);

/* ---- what goes downstream, in bulk -------------------------------------------
 * For the selected messages of a replayed batch (normally the survivors), the bytes the reference would send:
 *   mode 0  the sms.parsed payload, ParsedSMS.model_dump_json() (libs/nats_utils.py:127, services/parser_worker/worker.py:183)
 *   mode 1  the record upsert_parsed_sms hands to PocketBase (libs/pocketbase.py:296-309), as compact JSON
 * assembled on the host from the replay's columns (or packed records) and the natively read cache texts
 * (smsgate_dcache_fill: text_offsets / text_bytes / text_kind) - no pydantic object per record.  Two calls:
 * with out_bytes = NULL it reports `needed` (and already fills out_offsets[n_sel+1]); then out_bytes[capacity] is filled and
 * fallback[k] = 1 marks a row this code leaves to the caller (non-ASCII currency, a cache field of another type,
 * an overflowed coefficient): its span is empty.  Host memory only; rows are independent, so large selections are
 * sized and written by up to eight host threads (environment variable SMSGATE_EMIT_THREADS overrides the count). */
typedef struct smsgate_emit_input {
    const smsgate_record* records;   /* packed records, or NULL to use the columns below */
    const uint8_t* msg_id; const uint8_t* status; const uint8_t* flags; const uint8_t* dec_flags; const int32_t* slot;
    const int64_t* amount_coef; const int64_t* balance_coef; const int32_t* amount_exp; const int32_t* balance_exp;
    const int64_t* date;
    const uint8_t* card; const uint8_t* card_len;                 /* always columns (the record does not carry the card) */
    const int64_t* body_offsets; const uint8_t* body_bytes;       /* the batch */
    const int32_t* fixed_len; const uint8_t* fixed_bytes;         /* raw_body where SMSGATE_F_MODIFIED */
    const int64_t* sender_offsets; const uint8_t* sender_bytes;   /* RawSMS.sender per message, ragged */
    const char* device_id;                                        /* one value for the batch, NULL -> null */
    const int64_t* text_offsets; const uint8_t* text_bytes; const uint8_t* text_kind;   /* per cache entry */
} smsgate_emit_input;
int smsgate_emit_json(const smsgate_emit_input* in, int mode, int64_t n_sel, const int64_t* sel, int64_t* out_offsets,
                      uint8_t* out_bytes, int64_t capacity, uint8_t* fallback, int64_t* needed);

/* ---- upsert-by-msg_id for callers that must not allocate -------------------
 * No counterpart in the reference (it has no device memory to manage).
 *
 * Same verdicts as smsgate_dedup (survivor / first_seen / last_idx from msg_id
 * and status), but without a context and without a single allocation: no
 * cudaMalloc, no cudaFree, no library call that allocates.  The hash table it
 * needs is a fixed array compiled into the library, so this is the call for
 * code that runs where the allocator is off limits - a replay loop whose memory
 * was budgeted and handed out at start-up, a host process in which another
 * component owns the CUDA allocator, a section that must not stall or fail in
 * an allocation.  The caller passes device pointers for inputs and outputs
 * (any output may be NULL) and the stream to run on.
 *
 * Limits and rules
 *   - at most SMSGATE_DEDUP_STATIC_MAX messages per call; a larger n is refused
 *     with SMSGATE_E_LIMIT and a message in errbuf (split the batch, or use
 *     smsgate_dedup with smsgate_reserve);
 *   - the built-in table takes 20 MiB on every GPU that loads the library,
 *     for the life of the process;
 *   - the first call on a given device does the set-up for that device: it
 *     looks up the table's device address (cudaGetSymbolAddress), once.  Make
 *     that first call before any stream capture begins on the device;
 *   - the call makes `device` the calling thread's current CUDA device and
 *     leaves it so; cuda_stream must be a stream of that device (NULL = its
 *     default stream);
 *   - it only queues work: the table reset and the two kernels run on
 *     cuda_stream and the outputs are complete once that stream has reached
 *     them (synchronise the stream, or order later work on it, before reading);
 *   - all calls on one device share the table: each call resets the part it
 *     uses, and two calls whose GPU work overlaps (different streams, different
 *     threads) corrupt each other - serialise them;
 *   - errbuf/errbuf_len (optional) receive the error text.
 */
#define SMSGATE_DEDUP_STATIC_MAX (1 << 20)
int smsgate_dedup_static(int device, void* cuda_stream, int64_t n, int64_t base_index, const uint8_t* d_msg_id,
                         const uint8_t* d_status, uint8_t* d_survivor, uint8_t* d_first_seen, int64_t* d_last_idx,
                         char* errbuf, int errbuf_len);

#ifdef __cplusplus
}
#endif
#endif /* SMSGATE_B200_H */
