// TEST ONLY - never part of the product.  The two host functions of the library that use threads - the cache reader
// reading rowid ranges on several connections (csrc/cache_reader.h) and the payload emitter (csrc/emit_json.h) - built
// with ThreadSanitizer and run with more threads than there is work for comfort.  argv[1]: a diskcache directory,
// argv[2]: a directory of column dumps (tests/test_host_cpp_sanitized.py writes both).
#include <stdio.h>
#include <stdlib.h>
#include <string>
#include <vector>
#include "cache_reader.h"
#include "emit_json.h"

static std::vector<uint8_t> rd(const std::string& dir, const char* name) {
    std::string p = dir + "/" + name;
    FILE* f = fopen(p.c_str(), "rb");
    if (!f) { fprintf(stderr, "missing %s\n", p.c_str()); exit(2); }
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    std::vector<uint8_t> v((size_t)sz);
    if (sz && fread(v.data(), 1, (size_t)sz, f) != (size_t)sz) exit(3);
    fclose(f);
    return v;
}

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    sgcache::Table one, many;
    if (!sgcache::read_diskcache(argv[1], 1e18 * 0 + 1.0, &one, 1)) { fprintf(stderr, "%s\n", one.err.c_str()); return 3; }
    if (!sgcache::read_diskcache(argv[1], 1.0, &many, 7)) { fprintf(stderr, "%s\n", many.err.c_str()); return 3; }
    const bool same = one.m == many.m && one.keys == many.keys && one.rowids == many.rowids && one.val_offsets == many.val_offsets &&
                      one.val_bytes == many.val_bytes && one.eflags == many.eflags && one.needs_host == many.needs_host &&
                      one.text_offsets == many.text_offsets && one.text_bytes == many.text_bytes && one.text_kind == many.text_kind;
    printf("reader: %lld entries, one connection == seven: %s\n", (long long)one.m, same ? "yes" : "NO");

    std::string d = argv[2];
    auto nn = rd(d, "n"); const int64_t n = ((int64_t*)nn.data())[0];
    auto msg_id = rd(d, "msg_id"), status = rd(d, "status"), flags = rd(d, "flags"), dec = rd(d, "dec_flags"), slot = rd(d, "slot"),
         ac = rd(d, "amount_coef"), bc = rd(d, "balance_coef"), ae = rd(d, "amount_exp"), be = rd(d, "balance_exp"), date = rd(d, "date"),
         card = rd(d, "card"), cl = rd(d, "card_len"), fl = rd(d, "fixed_len"), fb = rd(d, "fixed_bytes"), bo = rd(d, "body_offsets"),
         bb = rd(d, "body_bytes"), so = rd(d, "sender_offsets"), sb = rd(d, "sender_bytes"), to = rd(d, "text_offsets"), tb = rd(d, "text_bytes"),
         tk = rd(d, "text_kind");
    sgemit::In I{};
    I.msg_id = msg_id.data(); I.status = status.data(); I.flags = flags.data(); I.dec_flags = dec.data(); I.slot = (const int32_t*)slot.data();
    I.amount_coef = (const int64_t*)ac.data(); I.balance_coef = (const int64_t*)bc.data(); I.amount_exp = (const int32_t*)ae.data();
    I.balance_exp = (const int32_t*)be.data(); I.date = (const int64_t*)date.data(); I.card = card.data(); I.card_len = cl.data();
    I.body_offsets = (const int64_t*)bo.data(); I.body_bytes = bb.data(); I.fixed_len = (const int32_t*)fl.data(); I.fixed_bytes = fb.data();
    I.sender_offsets = (const int64_t*)so.data(); I.sender_bytes = sb.data(); I.device_id = "xml_backup";
    I.text_offsets = (const int64_t*)to.data(); I.text_bytes = tb.data(); I.text_kind = tk.data(); I.record_stride = 0;
    std::vector<int64_t> sel;
    for (int64_t i = 0; i < n; ++i) if (status[i] == 0 || status[i] == 1) sel.push_back(i);
    bool emit_same = true;
    for (int mode = 0; mode < 2; ++mode) {
        std::vector<std::vector<uint8_t>> bytes(2);
        std::vector<std::vector<int64_t>> offs(2);
        std::vector<std::vector<uint8_t>> fbk(2);
        for (int v = 0; v < 2; ++v) {
            const int threads = v ? 6 : 1;
            offs[v].assign(sel.size() + 1, 0);
            fbk[v].assign(sel.size() + 1, 0);
            int64_t needed = 0;
            sgemit::emit_many(I, mode, (int64_t)sel.size(), sel.data(), offs[v].data(), nullptr, 0, nullptr, &needed, threads);
            bytes[v].assign((size_t)needed, 0);
            if (!sgemit::emit_many(I, mode, (int64_t)sel.size(), sel.data(), offs[v].data(), bytes[v].data(), needed, fbk[v].data(), &needed, threads))
                return 4;
        }
        emit_same = emit_same && bytes[0] == bytes[1] && offs[0] == offs[1] && fbk[0] == fbk[1];
    }
    printf("emitter: %zu rows, one thread == six: %s\n", sel.size(), emit_same ? "yes" : "NO");
    return same && emit_same ? 0 : 1;
}
