// TEST ONLY — never loaded by the product.  Compiles the __host__ __device__
// cores of the CUDA kernels (smsgate_b200/csrc/*_core.cuh) for the CPU with g++
// so that `pytest -m "not gpu"` can check the device logic, bit for bit, against
// the oracle and the golden vectors in a container without a GPU.  The shipped
// library (libsmsgate_b200.so) has no CPU path; this file is not part of it.
#include <stdint.h>
#include <string.h>
#include <string>
#include <unordered_map>
#include <vector>

#include "join_core.cuh"
#include "legacy_core.cuh"
#include "scan_core.cuh"

static ScanTables host_tables() {
    static DeviceTables tab;
    static bool ready = false;
    if (!ready) { sg_fill_tables(&tab); ready = true; }
    ScanTables T;
    T.t = &tab;
    return T;
}

static void put_be(uint8_t* dst, const uint32_t* h, int nwords) {
    for (int k = 0; k < nwords; ++k) {
        dst[4 * k] = h[k] >> 24; dst[4 * k + 1] = h[k] >> 16; dst[4 * k + 2] = h[k] >> 8; dst[4 * k + 3] = h[k];
    }
}

extern "C" int hc_scan_keys(const uint8_t* bytes, const int64_t* offsets, int64_t n, uint8_t* msg_id,
                            uint8_t* cache_key, uint8_t* flags, int32_t* body_date, int32_t* fixed_len,
                            uint8_t* fixed_bytes) {
    const ScanTables T = host_tables();
    std::vector<uint8_t> buf;
    for (int64_t i = 0; i < n; ++i) {
        const int len = (int)(offsets[i + 1] - offsets[i]);
        // an unaligned copy with slack, like a message inside a staged tile
        buf.assign((size_t)len + 3 + 16, 0xEE);
        uint8_t* m = buf.data() + (offsets[i] & 3) + ((4 - ((uintptr_t)buf.data() & 3)) & 3);
        memcpy(m, bytes + offsets[i], len);
        ScanResult sr = sg_scan_body(m, len, T);
        uint32_t h1[5], h2[8];
        sg_sha1(m, len, h1);
        int flen = len;
        if (sr.flags & SG_F_MODIFIED) flen = sg_rewrite_fixed(m, len, T);
        sg_sha256(m, flen, h2);
        put_be(msg_id + i * 20, h1, 5);
        put_be(cache_key + i * 32, h2, 8);
        flags[i] = (uint8_t)sr.flags;
        body_date[i] = sr.body_date;
        fixed_len[i] = flen;
        memcpy(fixed_bytes + offsets[i], m, flen);
    }
    return 0;
}

extern "C" int hc_decimal(int64_t n, const int32_t* offsets, const uint8_t* bytes, int64_t* coef, int32_t* exp,
                          uint8_t* flags) {
    for (int64_t i = 0; i < n; ++i) {
        DecResult r = sg_parse_decimal(bytes + offsets[i], offsets[i + 1] - offsets[i]);
        coef[i] = r.coef; exp[i] = r.exp; flags[i] = (uint8_t)(r.err | (r.sign << 4));
    }
    return 0;
}

extern "C" int hc_strict_date(int64_t n, const int32_t* offsets, const uint8_t* bytes, int64_t* packed) {
    for (int64_t i = 0; i < n; ++i) packed[i] = sg_parse_strict_datetime(bytes + offsets[i], offsets[i + 1] - offsets[i]);
    return 0;
}

extern "C" int hc_join_route(int64_t n, const uint8_t* cache_key, const uint8_t* flags, const int32_t* body_date,
                             int64_t now_packed, int64_t m, const uint8_t* keys, const int32_t* val_offsets,
                             const uint8_t* val_bytes, const uint16_t* eflags, const int64_t* date_packed,
                             int32_t* slot, uint8_t* status, int64_t* amount_coef, int64_t* balance_coef,
                             int32_t* amount_exp, int32_t* balance_exp, uint8_t* dec_flags, int64_t* date,
                             uint8_t* card, uint8_t* card_len, int64_t* amount_cents, int64_t* balance_cents) {
    std::unordered_map<std::string, int> idx;
    for (int64_t e = 0; e < m; ++e) idx.emplace(std::string((const char*)keys + e * 32, 32), (int)e);
    for (int64_t i = 0; i < n; ++i) {
        auto it = idx.find(std::string((const char*)cache_key + i * 32, 32));
        const int e = it == idx.end() ? -1 : it->second;
        EntryView ev;
        memset(&ev, 0, sizeof(ev));
        if (e >= 0) {
            const int32_t* o = val_offsets + (int64_t)e * 4;
            ev.amount = val_bytes + o[0]; ev.amount_len = o[1] - o[0];
            ev.balance = val_bytes + o[1]; ev.balance_len = o[2] - o[1];
            ev.date = val_bytes + o[2]; ev.date_len = o[3] - o[2];
            ev.card = val_bytes + o[3]; ev.card_len = o[4] - o[3];
            ev.eflags = eflags[e];
            ev.date_packed = date_packed[e];
        }
        RouteResult r = sg_route(flags[i], body_date[i], e >= 0, ev, now_packed);
        slot[i] = e; status[i] = (uint8_t)r.status;
        amount_coef[i] = r.amount.coef; balance_coef[i] = r.balance.coef;
        amount_exp[i] = r.amount.exp; balance_exp[i] = r.balance.exp;
        dec_flags[i] = (uint8_t)(r.amount.err | (r.balance.err << 2) | (r.amount.sign << 4) | (r.balance.sign << 5) |
                                 (r.cents_overflow << 6));
        amount_cents[i] = r.amount_cents; balance_cents[i] = r.balance_cents;
        date[i] = r.date;
        memcpy(card + i * 16, r.card, 16);
        card_len[i] = (uint8_t)r.card_bytes;
    }
    return 0;
}

extern "C" int hc_legacy(int64_t n, const int64_t* offsets, const uint8_t* bytes, uint8_t* status, uint8_t* lflags,
                         int32_t* strip, int32_t* caps, double* amount, double* balance, uint8_t* body_hash) {
    LegacyTables L;
    L.debit = H_RX_DEBIT; L.credit = H_RX_CREDIT; L.word_lo = H_RX_WORD_LO; L.word_hi = H_RX_WORD_HI;
    L.ci_exc = H_RX_CI_EXC; L.otp_next = (const uint8_t*)H_LEG_OTP_NEXT;
    L.scan = host_tables();
    for (int64_t i = 0; i < n; ++i) {
        LegacyResult r = sg_legacy_process(bytes + offsets[i], (int)(offsets[i + 1] - offsets[i]), L);
        status[i] = (uint8_t)r.status; lflags[i] = (uint8_t)r.lflags;
        strip[2 * i] = r.strip_lo; strip[2 * i + 1] = r.strip_hi;
        for (int k = 0; k < 2 * RX_MAX_GROUPS; ++k) caps[i * 2 * RX_MAX_GROUPS + k] = r.caps[k];
        amount[i] = r.amount; balance[i] = r.balance;
        put_be(body_hash + i * 32, r.digest, 8);
    }
    return 0;
}
