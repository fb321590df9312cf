// TEST ONLY - never part of the product.  The payload emitter (csrc/emit_json.h) run row by row into heap buffers of
// exactly the size its sizing pass announced, under AddressSanitizer and UBSan, on columns dumped by
// tests/test_host_cpp_sanitized.py.
#include <stdio.h>
#include <stdlib.h>
#include <string>
#include <vector>
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
    std::string d = argv[1];
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
    long rows = 0, declined = 0;
    for (int mode = 0; mode < 2; ++mode)
        for (int64_t i = 0; i < n; ++i) {
            if (status[i] != 0 && status[i] != 1) continue;                 // UPSERT / PARSED_NO_MERCHANT
            sgemit::Sink c{nullptr};
            sgemit::emit_one(c, I, i, mode);
            if (c.n == 0) { ++declined; continue; }
            uint8_t* buf = (uint8_t*)malloc((size_t)c.n);                   // exact size: ASan sees any byte past it
            sgemit::Sink w{buf};
            sgemit::emit_one(w, I, i, mode);
            if (w.n != c.n) { printf("SIZE MISMATCH row %lld\n", (long long)i); return 1; }
            free(buf);
            ++rows;
        }
    printf("rows %ld declined %ld\n", rows, declined);
    return 0;
}
