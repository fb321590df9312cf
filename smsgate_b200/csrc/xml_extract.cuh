// GPU extraction of the <sms> elements of an XML backup into the ragged batch the replay
// kernels take (replaces ET.parse + root.findall("sms") + elem.get of
// services/xml_watcher/watcher.py:37-44 and read_xml.py:25-37).  Included by smsgate_b200.cu.
//
// Stages (all data parallel, ordered by prefix sums):
//   '<' positions -> one thread per tag parses and validates it -> depth by prefix sum ->
//   text / structure checks and selection of root's direct <sms> children -> per message
//   sizes of the decoded body / address, Python int() of date -> offsets by prefix sum ->
//   decode into the output buffers.
#pragma once
#include "xml_core.cuh"

// ---------------------------------------------------------------- exclusive prefix sum, int32 -> int64
constexpr int SCAN_ITEMS = 1024;     // per block: 256 threads x 4
__global__ void __launch_bounds__(256) k_scan_sums(const int32_t* __restrict__ in, int64_t n, int64_t* __restrict__ sums) {
    __shared__ long long red[256];
    const int64_t base = (int64_t)blockIdx.x * SCAN_ITEMS;
    long long s = 0;
    for (int k = 0; k < 4; ++k) {
        const int64_t i = base + threadIdx.x * 4 + k;
        if (i < n) s += in[i];
    }
    red[threadIdx.x] = s;
    __syncthreads();
    for (int d = 128; d > 0; d >>= 1) {
        if (threadIdx.x < d) red[threadIdx.x] += red[threadIdx.x + d];
        __syncthreads();
    }
    if (threadIdx.x == 0) sums[blockIdx.x] = red[0];
}
__global__ void __launch_bounds__(1024) k_scan_top(int64_t* __restrict__ sums, int64_t nb, int64_t* __restrict__ total) {
    // one block, exclusive scan in place, chunks of 1024 carried sequentially
    __shared__ long long buf[1024];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int64_t c = 0; c < nb; c += 1024) {
        const int64_t i = c + threadIdx.x;
        const long long v = i < nb ? sums[i] : 0;
        buf[threadIdx.x] = v;
        __syncthreads();
        for (int d = 1; d < 1024; d <<= 1) {
            const long long t = threadIdx.x >= d ? buf[threadIdx.x - d] : 0;
            __syncthreads();
            buf[threadIdx.x] += t;
            __syncthreads();
        }
        if (i < nb) sums[i] = carry + buf[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry += buf[1023];
        __syncthreads();
    }
    if (threadIdx.x == 0 && total) *total = carry;
}
__global__ void __launch_bounds__(256) k_scan_apply(const int32_t* __restrict__ in, int64_t n, const int64_t* __restrict__ sums,
                                                    int64_t* __restrict__ out) {
    __shared__ long long part[256];
    const int64_t base = (int64_t)blockIdx.x * SCAN_ITEMS;
    long long v[4], s = 0;
    for (int k = 0; k < 4; ++k) {
        const int64_t i = base + threadIdx.x * 4 + k;
        v[k] = i < n ? in[i] : 0;
        s += v[k];
    }
    part[threadIdx.x] = s;
    __syncthreads();
    for (int d = 1; d < 256; d <<= 1) {
        const long long t = threadIdx.x >= d ? part[threadIdx.x - d] : 0;
        __syncthreads();
        part[threadIdx.x] += t;
        __syncthreads();
    }
    long long run = sums[blockIdx.x] + part[threadIdx.x] - s;
    for (int k = 0; k < 4; ++k) {
        const int64_t i = base + threadIdx.x * 4 + k;
        if (i < n) out[i] = run;
        run += v[k];
    }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 255) out[n] = run;   // out has n + 1 entries: the total
}

// ---------------------------------------------------------------- stage kernels
constexpr int XML_CHUNK = 4096;      // bytes per block when looking for '<'
__global__ void __launch_bounds__(256) k_xml_count_lt(const uint8_t* __restrict__ x, int64_t nx, int32_t* __restrict__ counts) {
    __shared__ int red[256];
    const int64_t base = (int64_t)blockIdx.x * XML_CHUNK + threadIdx.x * 16;
    int c = 0;
    for (int k = 0; k < 16; ++k) c += (base + k < nx && x[base + k] == '<') ? 1 : 0;
    red[threadIdx.x] = c;
    __syncthreads();
    for (int d = 128; d > 0; d >>= 1) {
        if (threadIdx.x < d) red[threadIdx.x] += red[threadIdx.x + d];
        __syncthreads();
    }
    if (threadIdx.x == 0) counts[blockIdx.x] = red[0];
}
__global__ void __launch_bounds__(256) k_xml_write_lt(const uint8_t* __restrict__ x, int64_t nx, const int64_t* __restrict__ offs,
                                                      int64_t* __restrict__ tag_pos) {
    __shared__ int part[256];
    const int64_t base = (int64_t)blockIdx.x * XML_CHUNK + threadIdx.x * 16;
    int c = 0;
    for (int k = 0; k < 16; ++k) c += (base + k < nx && x[base + k] == '<') ? 1 : 0;
    part[threadIdx.x] = c;
    __syncthreads();
    for (int d = 1; d < 256; d <<= 1) {
        const int t = threadIdx.x >= d ? part[threadIdx.x - d] : 0;
        __syncthreads();
        part[threadIdx.x] += t;
        __syncthreads();
    }
    int64_t w = offs[blockIdx.x] + part[threadIdx.x] - c;
    for (int k = 0; k < 16; ++k)
        if (base + k < nx && x[base + k] == '<') tag_pos[w++] = base + k;
}

struct XmlTags {                 // SoA over the T tags of the file
    const int64_t* pos;
    int64_t* end;
    int32_t* name_off;
    int32_t* name_len;
    uint8_t* kind;
    int32_t* delta;              // +1 open, -1 close, 0 otherwise
    int32_t* is_sms;             // direct <sms> child of the root (filled after the depth scan)
    int32_t* is_block;           // open or close tag (for the name-matching pass)
};

__device__ __forceinline__ void xml_report(unsigned long long* err, int64_t pos, int code) {
    atomicMin(err, ((unsigned long long)pos << 2) | (unsigned long long)code);
}

__global__ void __launch_bounds__(128) k_xml_parse_tags(const uint8_t* __restrict__ x, int64_t nx, int64_t T, XmlTags G,
                                                        unsigned long long* err) {
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= T) return;
    const int64_t pos = G.pos[t];
    const int64_t next_lt = t + 1 < T ? G.pos[t + 1] : nx;
    const XmlTag tag = xml_parse_tag(x, nx, pos, next_lt);
    G.end[t] = tag.end;
    G.name_off[t] = tag.name_off;
    G.name_len[t] = tag.name_len;
    G.kind[t] = tag.kind;
    G.delta[t] = tag.kind == XT_OPEN ? 1 : tag.kind == XT_CLOSE ? -1 : 0;
    G.is_block[t] = (tag.kind == XT_OPEN || tag.kind == XT_CLOSE) ? 1 : 0;
    if (tag.err) xml_report(err, pos, tag.err);
}

__global__ void __launch_bounds__(128) k_xml_structure(const uint8_t* __restrict__ x, int64_t nx, int64_t T, XmlTags G,
                                                       const int64_t* __restrict__ depth, unsigned long long* err,
                                                       int* root_count) {
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= T) return;
    const int64_t pos = G.pos[t];
    const int kind = G.kind[t];
    const int64_t d0 = depth[t], d1 = depth[t + 1];
    int sel = 0;
    if (kind == XT_CLOSE && d0 <= 0) xml_report(err, pos, XML_MALFORMED);          // closes nothing
    if ((kind == XT_OPEN || kind == XT_SELFCLOSE) && d0 == 0) atomicAdd(root_count, 1);
    if ((kind == XT_OPEN || kind == XT_SELFCLOSE) && d0 == 1 && G.name_len[t] == 3) {
        const uint8_t* nm = x + pos + G.name_off[t];
        sel = nm[0] == 's' && nm[1] == 'm' && nm[2] == 's';
    }
    G.is_sms[t] = sel;
    // the character data that follows this tag
    const int64_t a = G.end[t] + 1, b = t + 1 < T ? G.pos[t + 1] : nx;
    if (a < b && xml_check_text(x, a, b, d1 == 0) != XML_OK) xml_report(err, a, XML_MALFORMED);
    if (t == 0) {                                                                  // before the first tag: BOM, white space
        int64_t p = 0;
        if (pos >= 3 && x[0] == 0xEF && x[1] == 0xBB && x[2] == 0xBF) p = 3;
        if (xml_check_text(x, p, pos, true) != XML_OK) xml_report(err, p, XML_MALFORMED);
        // an XML declaration may only sit at the very start
    }
    if (xml_misplaced_decl(x, pos, G.end[t], kind)) xml_report(err, pos, XML_MALFORMED);
    if (t == T - 1 && d1 != 0) xml_report(err, nx, XML_MALFORMED);                 // something is still open
}

// ordered compaction of the indices whose flag is set (offsets come from the prefix sum of the flags)
__global__ void __launch_bounds__(256) k_xml_compact(int64_t T, const int32_t* __restrict__ flag, const int64_t* __restrict__ offs,
                                                     int64_t* __restrict__ out) {
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t < T && flag[t]) out[offs[t]] = t;
}

// open / close names must nest: a tiny sequential pass over the non-self-closing element tags only
// (two of them in a plain SMS backup; grows with <mms> subtrees)
__global__ void k_xml_match_names(const uint8_t* __restrict__ x, XmlTags G, const int64_t* __restrict__ blocks, int64_t nb,
                                  unsigned long long* err) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int64_t stack[64];
    int sp = 0;
    for (int64_t k = 0; k < nb; ++k) {
        const int64_t t = blocks[k];
        if (G.kind[t] == XT_OPEN) {
            if (sp >= 64) { xml_report(err, G.pos[t], XML_UNSUPPORTED); return; }
            stack[sp++] = t;
        } else {
            if (sp == 0) { xml_report(err, G.pos[t], XML_MALFORMED); return; }
            const int64_t o = stack[--sp];
            const int64_t a = G.pos[o] + G.name_off[o], b = G.pos[t] + G.name_off[t];
            if (!xml_span_eq(x, a, a + G.name_len[o], b, b + G.name_len[t])) { xml_report(err, G.pos[t], XML_MALFORMED); return; }
        }
    }
}

struct XmlSms {                  // SoA over the n selected <sms> elements
    int32_t* body_len;
    int32_t* sender_len;
    int64_t* date_ms;
    uint8_t* mflags;             // SMSGATE_XML_* bits
};
#define XMLF_HAS_BODY 0x01
#define XMLF_HAS_ADDRESS 0x02
#define XMLF_DATE_BAD 0x04       // int(date) raises ValueError
#define XMLF_DATE_BIG 0x08       // int(date) does not fit int64

__device__ __forceinline__ int xml_find_attrs(const uint8_t* x, int64_t pos, int name_off, int name_len, int64_t lim,
                                              XmlAttr* body, XmlAttr* date, XmlAttr* addr) {
    int found = 0;
    int64_t p = pos + name_off + name_len;
    bool first = true;
    for (;;) {
        XmlAttr a;
        const int r = xml_next_attr(x, &p, lim, &a, first);
        if (r <= 0) break;
        first = false;
        if (xml_span_is(x, a.name, a.name_end, "body", 4)) { *body = a; found |= 1; }
        else if (xml_span_is(x, a.name, a.name_end, "date", 4)) { *date = a; found |= 2; }
        else if (xml_span_is(x, a.name, a.name_end, "address", 7)) { *addr = a; found |= 4; }
    }
    return found;
}

__global__ void __launch_bounds__(128) k_xml_sms_sizes(const uint8_t* __restrict__ x, int64_t n, const int64_t* __restrict__ sms_tag,
                                                       XmlTags G, const DeviceTables* __restrict__ tab, XmlSms S) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t t = sms_tag[i], pos = G.pos[t], lim = G.end[t] + 1;
    XmlAttr body, date, addr;
    const int found = xml_find_attrs(x, pos, G.name_off[t], G.name_len[t], lim, &body, &date, &addr);
    int64_t bl = 0, sl = 0, dms = 0;
    uint32_t fl = 0;
    if (found & 1) { xml_attr_value(x, body.val, lim, body.quote, nullptr, &bl); fl |= XMLF_HAS_BODY; }
    if (found & 4) { xml_attr_value(x, addr.val, lim, addr.quote, nullptr, &sl); fl |= XMLF_HAS_ADDRESS; }
    if (found & 2) {
        int64_t dl = 0;
        xml_attr_value(x, date.val, lim, date.quote, nullptr, &dl);
        uint8_t buf[96];
        if (dl > (int64_t)sizeof(buf)) fl |= XMLF_DATE_BAD;       // longer than any int64 literal with padding: rejected
        else {
            xml_attr_value(x, date.val, lim, date.quote, buf, &dl);
            const int ok = xml_python_int(buf, (int)dl, tab->digit_run_start, tab->digit_run_len, &dms);
            if (ok == 0) fl |= XMLF_DATE_BAD;
            if (ok == 2) fl |= XMLF_DATE_BIG;
        }
    }
    S.body_len[i] = (int32_t)bl;
    S.sender_len[i] = (int32_t)sl;
    S.date_ms[i] = dms;
    S.mflags[i] = (uint8_t)fl;
}

__global__ void __launch_bounds__(128) k_xml_sms_emit(const uint8_t* __restrict__ x, int64_t n, const int64_t* __restrict__ sms_tag,
                                                      XmlTags G, const int64_t* __restrict__ body_off,
                                                      const int64_t* __restrict__ sender_off, uint8_t* __restrict__ body_bytes,
                                                      uint8_t* __restrict__ sender_bytes) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t t = sms_tag[i], pos = G.pos[t], lim = G.end[t] + 1;
    XmlAttr body, date, addr;
    const int found = xml_find_attrs(x, pos, G.name_off[t], G.name_len[t], lim, &body, &date, &addr);
    int64_t l;
    if (found & 1) xml_attr_value(x, body.val, lim, body.quote, body_bytes + body_off[i], &l);
    if (found & 4) xml_attr_value(x, addr.val, lim, addr.quote, sender_bytes + sender_off[i], &l);
}

// ---------------------------------------------------------------- every attribute of every <sms>
// read_xml.parse_and_cache_sms stores `elem.attrib` - the whole attribute dict of each <sms>
// (read_xml.py:32-41) - so that path needs all names and decoded values, not only body / address / date.
// Same two-pass shape as above: per message counts and byte totals, prefix sums, then the emit pass.
__global__ void __launch_bounds__(128) k_xml_attr_sizes(const uint8_t* __restrict__ x, int64_t n, const int64_t* __restrict__ sms_tag,
                                                        XmlTags G, int32_t* __restrict__ cnt, int32_t* __restrict__ nbytes,
                                                        int32_t* __restrict__ vbytes) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t t = sms_tag[i], pos = G.pos[t], lim = G.end[t] + 1;
    int64_t p = pos + G.name_off[t] + G.name_len[t];
    bool first = true;
    int c = 0;
    int64_t nb = 0, vb = 0;
    for (;;) {
        XmlAttr a;
        if (xml_next_attr(x, &p, lim, &a, first) <= 0) break;      // the tag was validated by k_xml_parse_tags
        first = false;
        int64_t l = 0;
        xml_attr_value(x, a.val, lim, a.quote, nullptr, &l);
        ++c;
        nb += a.name_end - a.name;
        vb += l;
    }
    cnt[i] = c;
    nbytes[i] = (int32_t)nb;
    vbytes[i] = (int32_t)vb;
}

__global__ void __launch_bounds__(128) k_xml_attr_emit(const uint8_t* __restrict__ x, int64_t n, const int64_t* __restrict__ sms_tag,
                                                       XmlTags G, const int64_t* __restrict__ attr_off,
                                                       const int64_t* __restrict__ name_base, const int64_t* __restrict__ val_base,
                                                       int64_t* __restrict__ name_offsets, uint8_t* __restrict__ name_bytes,
                                                       int64_t* __restrict__ val_offsets, uint8_t* __restrict__ val_bytes) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t t = sms_tag[i], pos = G.pos[t], lim = G.end[t] + 1;
    int64_t p = pos + G.name_off[t] + G.name_len[t];
    bool first = true;
    int64_t a_idx = attr_off[i], nw = name_base[i], vw = val_base[i];
    for (;;) {
        XmlAttr a;
        if (xml_next_attr(x, &p, lim, &a, first) <= 0) break;
        first = false;
        name_offsets[a_idx] = nw;
        val_offsets[a_idx] = vw;
        for (int64_t k = a.name; k < a.name_end; ++k) name_bytes[nw++] = x[k];
        int64_t l = 0;
        xml_attr_value(x, a.val, lim, a.quote, val_bytes + vw, &l);
        vw += l;
        ++a_idx;
    }
    if (i == n - 1) { name_offsets[a_idx] = nw; val_offsets[a_idx] = vw; }
}
