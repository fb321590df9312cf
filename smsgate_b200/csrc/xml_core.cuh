// XML backup tokenizer: per-tag logic shared by the GPU kernels (xml_extract.cuh)
// and the CPU check build.  It covers the dialect SMS Backup & Restore writes and
// `_iter_sms` reads (services/xml_watcher/watcher.py:35-54, read_xml.py:25-37):
// UTF-8, an XML declaration / processing instructions / comments without '<' in
// them, elements with attributes, the five predefined entities and numeric
// character references.  Constructs outside the dialect (DOCTYPE, CDATA, other
// encodings, '<' inside a comment) are reported as XML_UNSUPPORTED so the caller
// can hand the file to the host parser; malformed input inside the dialect is
// XML_MALFORMED, which is what expat / ElementTree raise ParseError for.
//
// Every '<' outside comments starts markup (a literal '<' is illegal in attribute
// values and text), which is what makes the tokenization data parallel: one thread
// per '<'.
#pragma once
#include "sg_common.cuh"

#define XML_OK 0
#define XML_MALFORMED 1
#define XML_UNSUPPORTED 2

#define XT_OPEN 0
#define XT_CLOSE 1
#define XT_SELFCLOSE 2
#define XT_PI 3
#define XT_COMMENT 4

struct XmlTag {
    int64_t end;          // offset of the closing '>'
    int32_t name_off;     // name start relative to the tag's '<'
    int32_t name_len;
    uint8_t kind;         // XT_*
    uint8_t err;          // XML_*
};

SG_HD bool xml_is_ws(uint32_t c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
// Names: the ASCII subset is decided here.  A colon (namespace processing: ElementTree resolves
// prefixes and renames the elements) or a non-ASCII name character (the Unicode NameChar tables)
// sends the document to the general parser instead.
SG_HD bool xml_name_start(uint32_t c) { return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || c == '_'; }
SG_HD bool xml_name_char(uint32_t c) { return xml_name_start(c) || (c >= '0' && c <= '9') || c == '-' || c == '.'; }
SG_HD bool xml_name_foreign(uint32_t c) { return c == ':' || c >= 0x80; }

// XML 1.0 Char production
SG_HD bool xml_legal_char(uint32_t cp) {
    return cp == 0x9 || cp == 0xA || cp == 0xD || (cp >= 0x20 && cp <= 0xD7FF) || (cp >= 0xE000 && cp <= 0xFFFD) ||
           (cp >= 0x10000 && cp <= 0x10FFFF);
}

// A reference starting at x[p] == '&'.  On success returns its length in the input and the code
// point it stands for; 0 when it is not a well-formed reference to a legal character.
SG_HD int xml_reference(const uint8_t* x, int64_t p, int64_t lim, uint32_t* cp) {
    if (p + 2 >= lim) return 0;
    if (x[p + 1] == '#') {
        int64_t q = p + 2;
        uint32_t v = 0;
        int nd = 0;
        if (x[q] == 'x') {
            ++q;
            for (; q < lim && x[q] != ';'; ++q, ++nd) {
                const uint32_t c = x[q];
                uint32_t d;
                if (c >= '0' && c <= '9') d = c - '0';
                else if (c >= 'a' && c <= 'f') d = c - 'a' + 10;
                else if (c >= 'A' && c <= 'F') d = c - 'A' + 10;
                else return 0;
                v = v > 0x10FFFF ? v : v * 16 + d;
            }
        } else {
            for (; q < lim && x[q] != ';'; ++q, ++nd) {
                const uint32_t c = x[q];
                if (c < '0' || c > '9') return 0;
                v = v > 0x10FFFF ? v : v * 10 + (c - '0');
            }
        }
        if (q >= lim || nd == 0 || !xml_legal_char(v)) return 0;
        *cp = v;
        return (int)(q - p + 1);
    }
    // amp lt gt quot apos
    const int64_t left = lim - p;
    if (left >= 4 && x[p + 1] == 'l' && x[p + 2] == 't' && x[p + 3] == ';') { *cp = '<'; return 4; }
    if (left >= 4 && x[p + 1] == 'g' && x[p + 2] == 't' && x[p + 3] == ';') { *cp = '>'; return 4; }
    if (left >= 5 && x[p + 1] == 'a' && x[p + 2] == 'm' && x[p + 3] == 'p' && x[p + 4] == ';') { *cp = '&'; return 5; }
    if (left >= 6 && x[p + 1] == 'q' && x[p + 2] == 'u' && x[p + 3] == 'o' && x[p + 4] == 't' && x[p + 5] == ';') { *cp = '"'; return 6; }
    if (left >= 6 && x[p + 1] == 'a' && x[p + 2] == 'p' && x[p + 3] == 'o' && x[p + 4] == 's' && x[p + 5] == ';') { *cp = '\''; return 6; }
    return 0;
}

SG_HD int xml_utf8_len(uint32_t cp) { return cp < 0x80 ? 1 : cp < 0x800 ? 2 : cp < 0x10000 ? 3 : 4; }
SG_HD int xml_utf8_put(uint8_t* out, uint32_t cp) {
    if (cp < 0x80) { out[0] = (uint8_t)cp; return 1; }
    if (cp < 0x800) { out[0] = (uint8_t)(0xC0 | cp >> 6); out[1] = (uint8_t)(0x80 | (cp & 63)); return 2; }
    if (cp < 0x10000) {
        out[0] = (uint8_t)(0xE0 | cp >> 12); out[1] = (uint8_t)(0x80 | ((cp >> 6) & 63)); out[2] = (uint8_t)(0x80 | (cp & 63));
        return 3;
    }
    out[0] = (uint8_t)(0xF0 | cp >> 18); out[1] = (uint8_t)(0x80 | ((cp >> 12) & 63));
    out[2] = (uint8_t)(0x80 | ((cp >> 6) & 63)); out[3] = (uint8_t)(0x80 | (cp & 63));
    return 4;
}

// One raw character of character data or of an attribute value at x[p] (not '&', not '<'):
// returns its byte length, 0 if it is not a legal, well-formed UTF-8 XML character.
SG_HD int xml_raw_char(const uint8_t* x, int64_t p, int64_t lim) {
    const uint32_t b = x[p];
    if (b < 0x80) return (b >= 0x20 || b == 9 || b == 10 || b == 13) ? 1 : 0;
    uint32_t cp;
    const int64_t avail = lim - p;
    const int n = sg_utf8_decode(x + p, (int)(avail > 4 ? 4 : avail), &cp);
    if (cp == 0xFFFFFFFFu) return 0;
    // overlong forms, surrogates, noncharacters
    if ((n == 2 && cp < 0x80) || (n == 3 && cp < 0x800) || (n == 4 && cp < 0x10000)) return 0;
    return xml_legal_char(cp) ? n : 0;
}

// Walk an attribute value from x[p] (just after the opening quote `q`) to the closing quote.
// Validates it and, when `out` is given, writes the value the parser reports: references
// resolved, literal tab / newline / carriage return turned into spaces with CR LF counting
// once (XML 1.0 2.11 + 3.3.3).  Returns the offset of the closing quote, -1 if malformed;
// *outlen = bytes of the reported value.
SG_HD int64_t xml_attr_value(const uint8_t* x, int64_t p, int64_t lim, uint8_t q, uint8_t* out, int64_t* outlen) {
    int64_t w = 0;
    while (p < lim) {
        const uint8_t b = x[p];
        if (b == q) { *outlen = w; return p; }
        if (b == '<') return -1;
        if (b == '&') {
            uint32_t cp;
            const int n = xml_reference(x, p, lim, &cp);
            if (!n) return -1;
            if (out) w += xml_utf8_put(out + w, cp); else w += xml_utf8_len(cp);
            p += n;
            continue;
        }
        if (b == '\r') {                                   // CR LF and lone CR are one line end -> one space
            if (out) out[w] = ' ';
            ++w;
            p += (p + 1 < lim && x[p + 1] == '\n') ? 2 : 1;
            continue;
        }
        if (b == '\n' || b == '\t') { if (out) out[w] = ' '; ++w; ++p; continue; }
        const int n = xml_raw_char(x, p, lim);
        if (!n) return -1;
        if (out) for (int k = 0; k < n; ++k) out[w + k] = x[p + k];
        w += n;
        p += n;
    }
    return -1;
}

// Iterator over the attributes of a start tag.  `p` starts just after the element name.
struct XmlAttr {
    int64_t name, name_end;   // [name, name_end)
    int64_t val, val_end;     // value between the quotes
    uint8_t quote;
};
// returns 1 and fills `a` for the next attribute, 0 at the end of the tag (p then sits on '/' or '>'), -1 if malformed,
// -2 if the name needs the general parser (namespaces, non-ASCII names)
SG_HD int xml_next_attr(const uint8_t* x, int64_t* pp, int64_t lim, XmlAttr* a, bool first) {
    int64_t p = *pp;
    int64_t ws = 0;
    while (p < lim && xml_is_ws(x[p])) { ++p; ++ws; }
    if (p >= lim) return -1;
    if (x[p] == '/' || x[p] == '>') { *pp = p; return 0; }
    if (!ws) return -1;                                    // attributes are separated by white space
    if (xml_name_foreign(x[p])) return -2;
    if (!xml_name_start(x[p])) return -1;
    a->name = p;
    while (p < lim && xml_name_char(x[p])) ++p;
    if (p < lim && xml_name_foreign(x[p])) return -2;
    a->name_end = p;
    if (a->name_end - a->name == 5 && x[a->name] == 'x' && x[a->name + 1] == 'm' && x[a->name + 2] == 'l' &&
        x[a->name + 3] == 'n' && x[a->name + 4] == 's') return -2;       // a default namespace renames every element
    while (p < lim && xml_is_ws(x[p])) ++p;
    if (p >= lim || x[p] != '=') return -1;
    ++p;
    while (p < lim && xml_is_ws(x[p])) ++p;
    if (p >= lim || (x[p] != '"' && x[p] != '\'')) return -1;
    a->quote = x[p];
    a->val = p + 1;
    int64_t outlen;
    const int64_t e = xml_attr_value(x, p + 1, lim, a->quote, nullptr, &outlen);
    if (e < 0) return -1;
    a->val_end = e;
    *pp = e + 1;
    return 1;
}

SG_HD bool xml_span_eq(const uint8_t* x, int64_t a0, int64_t a1, int64_t b0, int64_t b1) {
    if (a1 - a0 != b1 - b0) return false;
    for (int64_t k = 0; k < a1 - a0; ++k) if (x[a0 + k] != x[b0 + k]) return false;
    return true;
}
SG_HD bool xml_span_is(const uint8_t* x, int64_t a0, int64_t a1, const char* lit, int n) {
    if (a1 - a0 != n) return false;
    for (int k = 0; k < n; ++k) if (x[a0 + k] != (uint8_t)lit[k]) return false;
    return true;
}

// Parse the markup starting at the '<' at x[pos].  `next_lt` is the offset of the next '<' (or nx):
// nothing of this tag may extend past it, except a comment, which must not contain one.
SG_HD XmlTag xml_parse_tag(const uint8_t* x, int64_t nx, int64_t pos, int64_t next_lt) {
    XmlTag t;
    t.end = pos; t.name_off = 0; t.name_len = 0; t.kind = XT_OPEN; t.err = XML_MALFORMED;
    const int64_t lim = next_lt;
    int64_t p = pos + 1;
    if (p >= lim) return t;
    if (x[p] == '?') {                                     // processing instruction / XML declaration
        t.kind = XT_PI;
        int64_t q = p + 1;
        if (q < lim && xml_name_foreign(x[q])) { t.err = XML_UNSUPPORTED; return t; }
        if (q >= lim || !xml_name_start(x[q])) return t;   // the target is a Name
        const int64_t n0 = q;
        while (q < lim && xml_name_char(x[q])) ++q;
        if (q < lim && xml_name_foreign(x[q])) { t.err = XML_UNSUPPORTED; return t; }
        const int64_t n1 = q;
        if (q >= lim) return t;
        if (!xml_is_ws(x[q]) && !(x[q] == '?' && q + 1 < lim && x[q + 1] == '>')) return t;   // Name, then S or "?>"
        int64_t close = -1;
        for (int64_t k = q; k + 1 < lim; ++k) if (x[k] == '?' && x[k + 1] == '>') { close = k; break; }
        if (close < 0) { t.err = XML_UNSUPPORTED; return t; }   // maybe a '<' inside the data: not decidable per tag
        for (int64_t k = q; k < close;) {                  // the data: legal characters
            const int n = xml_raw_char(x, k, close);
            if (!n) return t;
            k += n;
        }
        t.end = close + 1;
        const bool xml3 = n1 - n0 == 3 && (x[n0] | 0x20) == 'x' && (x[n0 + 1] | 0x20) == 'm' && (x[n0 + 2] | 0x20) == 'l';
        if (!xml3) { t.err = XML_OK; return t; }
        if (!(x[n0] == 'x' && x[n0 + 1] == 'm' && x[n0 + 2] == 'l')) return t;      // "XML", "Xml" are reserved
        // XMLDecl ::= '<?xml' VersionInfo EncodingDecl? SDDecl? S? '?>'
        int stage = 0;                                     // 0 expect version, 1 encoding or standalone, 2 standalone, 3 nothing
        int64_t k = n1;
        for (;;) {
            int64_t ws = 0;
            while (k < close && xml_is_ws(x[k])) { ++k; ++ws; }
            if (k >= close) break;
            if (!ws) return t;
            const int64_t a0 = k;
            while (k < close && x[k] >= 'a' && x[k] <= 'z') ++k;
            const int64_t a1 = k;
            while (k < close && xml_is_ws(x[k])) ++k;
            if (k >= close || x[k] != '=') return t;
            ++k;
            while (k < close && xml_is_ws(x[k])) ++k;
            if (k >= close || (x[k] != '"' && x[k] != '\'')) return t;
            const uint8_t qu = x[k];
            const int64_t v0 = ++k;
            while (k < close && x[k] != qu) ++k;
            if (k >= close) return t;
            const int64_t v1 = k++;
            if (stage == 0 && xml_span_is(x, a0, a1, "version", 7)) {
                // expat accepts any run of [A-Za-z0-9_.-] as the version number, the empty one included
                for (int64_t c = v0; c < v1; ++c) {
                    const uint8_t ch = x[c];
                    if (!((ch >= 'a' && ch <= 'z') || (ch >= 'A' && ch <= 'Z') || (ch >= '0' && ch <= '9') || ch == '_' ||
                          ch == '.' || ch == '-')) return t;
                }
                stage = 1;
            } else if (stage == 1 && xml_span_is(x, a0, a1, "encoding", 8)) {
                const bool utf8 = v1 - v0 == 5 && (x[v0] | 0x20) == 'u' && (x[v0 + 1] | 0x20) == 't' && (x[v0 + 2] | 0x20) == 'f' &&
                                  x[v0 + 3] == '-' && x[v0 + 4] == '8';
                if (!utf8) { t.err = XML_UNSUPPORTED; return t; }
                stage = 2;
            } else if ((stage == 1 || stage == 2) && xml_span_is(x, a0, a1, "standalone", 10)) {
                if (!(xml_span_is(x, v0, v1, "yes", 3) || xml_span_is(x, v0, v1, "no", 2))) return t;
                stage = 3;
            } else return t;
        }
        if (stage == 0) return t;
        t.err = XML_OK;
        return t;
    }
    if (x[p] == '!') {
        if (p + 2 < nx && x[p + 1] == '-' && x[p + 2] == '-') {
            t.kind = XT_COMMENT;
            for (int64_t q = p + 3; q + 2 < lim; ++q)
                if (x[q] == '-' && x[q + 1] == '-') {
                    if (x[q + 2] != '>') return t;          // "--" inside a comment is not well-formed
                    for (int64_t k = p + 3; k < q;) {
                        const int n = xml_raw_char(x, k, q);
                        if (!n) return t;
                        k += n;
                    }
                    t.end = q + 2; t.err = XML_OK; return t;
                }
            t.err = XML_UNSUPPORTED;                       // no "-->" before the next '<': a '<' inside the comment
            return t;
        }
        t.err = XML_UNSUPPORTED;                           // DOCTYPE, CDATA
        return t;
    }
    if (x[p] == '/') {
        t.kind = XT_CLOSE;
        ++p;
        if (p < lim && xml_name_foreign(x[p])) { t.err = XML_UNSUPPORTED; return t; }
        if (p >= lim || !xml_name_start(x[p])) return t;
        t.name_off = (int32_t)(p - pos);
        while (p < lim && xml_name_char(x[p])) ++p;
        if (p < lim && xml_name_foreign(x[p])) { t.err = XML_UNSUPPORTED; return t; }
        t.name_len = (int32_t)(p - pos) - t.name_off;
        while (p < lim && xml_is_ws(x[p])) ++p;
        if (p >= lim || x[p] != '>') return t;
        t.end = p; t.err = XML_OK;
        return t;
    }
    if (xml_name_foreign(x[p])) { t.err = XML_UNSUPPORTED; return t; }
    if (!xml_name_start(x[p])) return t;
    t.name_off = (int32_t)(p - pos);
    while (p < lim && xml_name_char(x[p])) ++p;
    if (p < lim && xml_name_foreign(x[p])) { t.err = XML_UNSUPPORTED; return t; }
    t.name_len = (int32_t)(p - pos) - t.name_off;
    // attributes: well-formed, no duplicates
    int64_t names[2 * 64];
    int na = 0;
    bool first = true;
    for (;;) {
        XmlAttr a;
        const int r = xml_next_attr(x, &p, lim, &a, first);
        if (r == -2) { t.err = XML_UNSUPPORTED; return t; }
        if (r < 0) return t;
        if (r == 0) break;
        first = false;
        for (int k = 0; k < na; ++k) if (xml_span_eq(x, names[2 * k], names[2 * k + 1], a.name, a.name_end)) return t;
        if (na < 64) { names[2 * na] = a.name; names[2 * na + 1] = a.name_end; ++na; }
        else { t.err = XML_UNSUPPORTED; return t; }
    }
    if (x[p] == '/') {
        if (p + 1 >= lim || x[p + 1] != '>') return t;
        t.kind = XT_SELFCLOSE; t.end = p + 1;
    } else {
        t.kind = XT_OPEN; t.end = p;
    }
    t.err = XML_OK;
    return t;
}

// <?xml ...?> is only allowed as the very first thing in the document (after an optional BOM)
SG_HD bool xml_misplaced_decl(const uint8_t* x, int64_t pos, int64_t end, int kind) {
    if (kind != XT_PI || end - pos < 5) return false;
    const bool is_decl = x[pos + 2] == 'x' && x[pos + 3] == 'm' && x[pos + 4] == 'l' && (xml_is_ws(x[pos + 5]) || x[pos + 5] == '?');
    if (!is_decl) return false;
    const bool bom = pos == 3 && x[0] == 0xEF && x[1] == 0xBB && x[2] == 0xBF;
    return !(pos == 0 || bom);
}

// Character data between two tags, x[a, b): legal characters, well-formed references, no "]]>".
// outside_root: only white space may appear.
SG_HD int xml_check_text(const uint8_t* x, int64_t a, int64_t b, bool outside_root) {
    for (int64_t p = a; p < b;) {
        const uint8_t c = x[p];
        if (outside_root) { if (!xml_is_ws(c)) return XML_MALFORMED; ++p; continue; }
        if (c == '&') {
            uint32_t cp;
            const int n = xml_reference(x, p, b, &cp);
            if (!n) return XML_MALFORMED;
            p += n;
            continue;
        }
        if (c == '>' && p >= a + 2 && x[p - 1] == ']' && x[p - 2] == ']') return XML_MALFORMED;
        const int n = xml_raw_char(x, p, b);
        if (!n) return XML_MALFORMED;
        p += n;
    }
    return XML_OK;
}

// int(text) as Python evaluates it for the `date` attribute (watcher.py:42): optional white space,
// optional sign, decimal digits of any Unicode Nd block with single underscores between digits.
// ok: 1 = value in *out, 0 = ValueError, 2 = valid but does not fit int64.
SG_HD int xml_python_int(const uint8_t* s, int len, const uint32_t* run_start, const uint8_t* run_len, int64_t* out) {
    int a = 0, b = len;
    while (a < b) {
        uint32_t cp; const int n = sg_utf8_decode(s + a, b - a, &cp);
        if (!sg_is_space(cp)) break;
        a += n;
    }
    while (b > a) {
        int p = b - 1;
        while (p > a && (s[p] & 0xC0) == 0x80) --p;
        uint32_t cp; sg_utf8_decode(s + p, b - p, &cp);
        if (!sg_is_space(cp)) break;
        b = p;
    }
    bool neg = false;
    if (a < b && (s[a] == '+' || s[a] == '-')) { neg = s[a] == '-'; ++a; }
    uint64_t v = 0;
    bool big = false, prev_digit = false;
    int nd = 0;
    while (a < b) {
        uint32_t cp; const int n = sg_utf8_decode(s + a, b - a, &cp);
        a += n;
        if (cp == '_') { if (!prev_digit) return 0; prev_digit = false; continue; }
        const int d = sg_digit_value(cp, run_start, run_len);
        if (d < 0) return 0;
        if (v > (0x7FFFFFFFFFFFFFFFull - (uint64_t)d) / 10) big = true; else v = v * 10 + (uint64_t)d;
        prev_digit = true;
        ++nd;
    }
    if (nd == 0 || !prev_digit) return 0;
    if (big) return 2;
    *out = neg ? -(int64_t)v : (int64_t)v;
    return 1;
}
