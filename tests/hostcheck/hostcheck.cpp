// TEST ONLY — never loaded by the product.  Compiles the __host__ __device__
// cores of the CUDA kernels (smsgate_b200/csrc/*_core.cuh) for the CPU with g++
// so that `pytest -m "not gpu"` can check the device logic, bit for bit, against
// the oracle and the golden vectors in a container without a GPU.  The shipped
// library (libsmsgate_b200.so) has no CPU path; this file is not part of it.
#include <cstring>
#include <stdint.h>
#include <string.h>
#include <string>
#include <unordered_map>
#include <vector>

#include "join_core.cuh"
#include "legacy_core.cuh"
#include "scan_core.cuh"
#include "xml_core.cuh"

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
        // the scan kernel's composition (first pass, full pass for flagged bodies) must equal the full pass
        ScanResult sr = sg_scan_two_pass(m, len, T);
        const ScanResult full = sg_scan_body(m, len, T);
        if (sr.flags != full.flags || sr.body_date != full.body_date) return -100 - (int)i;
        uint32_t h1[5], h2[8];
        sg_sha1(m, len, h1);
        // the kernel's hashing loop (sg_hash_pass, one copy, three modes) must agree with the plain digests
        uint32_t p1[5], p2[8], q1[5] = {1, 2, 3, 4, 5}, q2[8];
        sg_hash_pass(m, len, 3u, p1, p2);                       // untouched body: both digests in one walk
        sg_hash_pass(m, len, 1u, q1, q2);                       // rewritten body, first pass: SHA-1 only
        uint32_t o2[8];
        sg_sha256(m, len, o2);
        if (memcmp(p1, h1, 20) || memcmp(q1, h1, 20) || memcmp(p2, o2, 32)) return -200000 - (int)i;
        int flen = len;
        if (sr.flags & SG_F_MODIFIED) flen = sg_rewrite_fixed(m, len, T);
        sg_sha256(m, flen, h2);
        sg_hash_pass(m, flen, 2u, q1, q2);                      // second pass: SHA-256 only, SHA-1 left alone
        if (memcmp(q2, h2, 32) || memcmp(q1, h1, 20)) return -300000 - (int)i;
        put_be(msg_id + i * 20, h1, 5);
        put_be(cache_key + i * 32, h2, 8);
        flags[i] = (uint8_t)sr.flags;
        body_date[i] = sr.body_date;
        fixed_len[i] = flen;
        memcpy(fixed_bytes + offsets[i], m, flen);
    }
    return 0;
}

extern "C" int hc_md5(const uint8_t* bytes, const int64_t* offsets, int64_t n, uint8_t* md5) {
    for (int64_t i = 0; i < n; ++i) {
        uint32_t h[4];
        sg_md5(bytes + offsets[i], (int)(offsets[i + 1] - offsets[i]), h);
        memcpy(md5 + i * 16, h, 16);
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

extern "C" int hc_dt_str(int64_t n, const int32_t* offsets, const uint8_t* bytes, int64_t* packed) {
    const ScanTables T = host_tables();
    for (int64_t i = 0; i < n; ++i) packed[i] = sg_dt_str(bytes + offsets[i], offsets[i + 1] - offsets[i], T);
    return 0;
}

extern "C" int hc_join_route(int64_t n, const uint8_t* cache_key, const uint8_t* flags, const int32_t* body_date,
                             int64_t now_packed, int64_t m, const uint8_t* keys, const int32_t* val_offsets,
                             const uint8_t* val_bytes, const uint16_t* eflags, const int64_t* date_packed,
                             int32_t* slot, uint8_t* status, int64_t* amount_coef, int64_t* balance_coef,
                             int32_t* amount_exp, int32_t* balance_exp, uint8_t* dec_flags, int64_t* date,
                             uint8_t* card, uint8_t* card_len, int64_t* amount_cents, int64_t* balance_cents,
                             const int64_t* phone_dt) {
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
        RouteResult r = sg_route(flags[i], body_date[i], e >= 0, ev, now_packed, phone_dt ? phone_dt[i] : 0);
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
                         int32_t* strip, int32_t* caps, double* amount, double* balance, uint8_t* body_hash,
                         uint32_t mode) {
    LegacyTables L;
    L.debit = H_RX_DEBIT; L.credit = H_RX_CREDIT; L.word_lo = H_RX_WORD_LO; L.word_hi = H_RX_WORD_HI;
    L.ci_exc = H_RX_CI_EXC; L.otp_next = (const uint8_t*)H_LEG_OTP_NEXT;
    L.scan = host_tables();
    for (int64_t i = 0; i < n; ++i) {
        LegacyResult r = sg_legacy_process(bytes + offsets[i], (int)(offsets[i + 1] - offsets[i]), L, mode);
        status[i] = (uint8_t)r.status; lflags[i] = (uint8_t)r.lflags;
        strip[2 * i] = r.strip_lo; strip[2 * i + 1] = r.strip_hi;
        for (int k = 0; k < 2 * RX_MAX_GROUPS; ++k) caps[i * 2 * RX_MAX_GROUPS + k] = r.caps[k];
        amount[i] = r.amount; balance[i] = r.balance;
        put_be(body_hash + i * 32, r.digest, 8);
    }
    return 0;
}

// Sequential stand-in for the XML extraction kernels (xml_extract.cuh): same per-tag functions,
// same checks, in document order.  Returns 0 ok, 1 malformed, 2 unsupported; *n = messages.
extern "C" int hc_xml_extract(const uint8_t* x, int64_t nx, int64_t* n_out, uint8_t* body_bytes, int64_t* body_off,
                              uint8_t* sender_bytes, int64_t* sender_off, int64_t* date_ms, uint8_t* mflags) {
    const ScanTables ST = host_tables();
    std::vector<int64_t> pos;
    for (int64_t i = 0; i < nx; ++i) if (x[i] == '<') pos.push_back(i);
    const int64_t T = (int64_t)pos.size();
    *n_out = 0;
    if (T == 0) return 1;
    std::vector<XmlTag> tags(T);
    int worst = 0;
    int64_t worst_pos = nx + 1;
    auto report = [&](int64_t p, int code) { if (p < worst_pos) { worst_pos = p; worst = code; } };
    for (int64_t t = 0; t < T; ++t) {
        tags[t] = xml_parse_tag(x, nx, pos[t], t + 1 < T ? pos[t + 1] : nx);
        if (tags[t].err) report(pos[t], tags[t].err);
    }
    std::vector<int64_t> depth(T + 1, 0);
    for (int64_t t = 0; t < T; ++t) depth[t + 1] = depth[t] + (tags[t].kind == XT_OPEN ? 1 : tags[t].kind == XT_CLOSE ? -1 : 0);
    int roots = 0;
    std::vector<int64_t> sms, stack;
    for (int64_t t = 0; t < T; ++t) {
        const int kind = tags[t].kind;
        if (kind == XT_CLOSE && depth[t] <= 0) report(pos[t], XML_MALFORMED);
        if ((kind == XT_OPEN || kind == XT_SELFCLOSE) && depth[t] == 0) ++roots;
        if ((kind == XT_OPEN || kind == XT_SELFCLOSE) && depth[t] == 1 && tags[t].name_len == 3 &&
            memcmp(x + pos[t] + tags[t].name_off, "sms", 3) == 0) sms.push_back(t);
        const int64_t a = tags[t].end + 1, b = t + 1 < T ? pos[t + 1] : nx;
        if (a < b && xml_check_text(x, a, b, depth[t + 1] == 0) != XML_OK) report(a, XML_MALFORMED);
        if (t == 0) {
            int64_t p = 0;
            if (pos[0] >= 3 && x[0] == 0xEF && x[1] == 0xBB && x[2] == 0xBF) p = 3;
            if (xml_check_text(x, p, pos[0], true) != XML_OK) report(p, XML_MALFORMED);
        }
        if (xml_misplaced_decl(x, pos[t], tags[t].end, kind)) report(pos[t], XML_MALFORMED);
        if (t == T - 1 && depth[t + 1] != 0) report(nx, XML_MALFORMED);
        if (!tags[t].err && kind == XT_OPEN) stack.push_back(t);
        if (!tags[t].err && kind == XT_CLOSE) {
            if (stack.empty()) report(pos[t], XML_MALFORMED);
            else {
                const int64_t o = stack.back(); stack.pop_back();
                const int64_t a0 = pos[o] + tags[o].name_off, b0 = pos[t] + tags[t].name_off;
                if (!xml_span_eq(x, a0, a0 + tags[o].name_len, b0, b0 + tags[t].name_len)) report(pos[t], XML_MALFORMED);
            }
        }
    }
    if (worst) return worst;
    if (roots != 1) return 1;
    int64_t bo = 0, so = 0;
    for (size_t i = 0; i < sms.size(); ++i) {
        const int64_t t = sms[i], lim = tags[t].end + 1;
        int64_t p = pos[t] + tags[t].name_off + tags[t].name_len;
        bool first = true;
        uint32_t fl = 0;
        int64_t bl = 0, sl = 0, dms = 0;
        for (;;) {
            XmlAttr a;
            const int r = xml_next_attr(x, &p, lim, &a, first);
            if (r <= 0) break;
            first = false;
            if (xml_span_is(x, a.name, a.name_end, "body", 4)) { xml_attr_value(x, a.val, lim, a.quote, body_bytes + bo, &bl); fl |= 1; }
            else if (xml_span_is(x, a.name, a.name_end, "address", 7)) { xml_attr_value(x, a.val, lim, a.quote, sender_bytes + so, &sl); fl |= 2; }
            else if (xml_span_is(x, a.name, a.name_end, "date", 4)) {
                int64_t dl = 0;
                xml_attr_value(x, a.val, lim, a.quote, nullptr, &dl);
                uint8_t buf[96];
                if (dl > 96) fl |= 4;
                else {
                    xml_attr_value(x, a.val, lim, a.quote, buf, &dl);
                    const int ok = xml_python_int(buf, (int)dl, ST.t->digit_run_start, ST.t->digit_run_len, &dms);
                    if (ok == 0) fl |= 4;
                    if (ok == 2) fl |= 8;
                }
            }
        }
        body_off[i] = bo; sender_off[i] = so; date_ms[i] = dms; mflags[i] = (uint8_t)fl;
        bo += bl; so += sl;
    }
    body_off[sms.size()] = bo; sender_off[sms.size()] = so;
    *n_out = (int64_t)sms.size();
    return 0;
}
