// json.hpp — minimal JSON value + parser + a serializer that is byte-compatible with Go's
// encoding/json.Marshal as the reference uses it (nats_llm_studio.go:207-217): object keys of
// Go maps sorted, HTML-escaping on (<, >, & -> <, >, &), U+2028/U+2029 escaped,
// invalid UTF-8 replaced by U+FFFD, embedded raw JSON compacted.  Parser error strings follow
// Go's decoder for the cases the handlers surface ("invalid JSON in ChatModel: <err>", :340).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <map>
#include <memory>
#include <string>
#include <utility>
#include <vector>

namespace b200 {

struct Json {
    enum Kind { Null, Bool, Num, Str, Arr, Obj, Raw } kind = Null;
    bool b = false; double num = 0; bool is_int = false; long long i = 0;
    std::string s;                                        // Str, or Raw (pre-serialised JSON)
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;        // insertion order kept
    bool sorted = false;                                  // serialise keys sorted (Go map semantics)

    static Json null() { return Json(); }
    static Json boolean(bool v) { Json j; j.kind = Bool; j.b = v; return j; }
    static Json integer(long long v) { Json j; j.kind = Num; j.is_int = true; j.i = v; j.num = (double)v; return j; }
    static Json number(double v) { Json j; j.kind = Num; j.num = v; return j; }
    static Json string(std::string v) { Json j; j.kind = Str; j.s = std::move(v); return j; }
    static Json raw(std::string v) { Json j; j.kind = Raw; j.s = std::move(v); return j; }
    static Json array() { Json j; j.kind = Arr; return j; }
    static Json object(bool go_map = false) { Json j; j.kind = Obj; j.sorted = go_map; return j; }
    Json& set(const std::string& k, Json v) { for (auto& kv : obj) if (kv.first == k) { kv.second = std::move(v); return *this; } obj.emplace_back(k, std::move(v)); return *this; }
    Json& push(Json v) { arr.push_back(std::move(v)); return *this; }
    const Json* get(const std::string& k) const { for (auto& kv : obj) if (kv.first == k) return &kv.second; return nullptr; }
    // Go's decoder matches struct fields case-insensitively (exact match preferred)
    const Json* get_fold(const std::string& k) const {
        if (auto p = get(k)) return p;
        for (auto& kv : obj) { if (kv.first.size() != k.size()) continue; bool eq = true; for (size_t i = 0; i < k.size(); i++) if (tolower((unsigned char)kv.first[i]) != tolower((unsigned char)k[i])) { eq = false; break; } if (eq) return &kv.second; }
        return nullptr;
    }
    const char* go_kind() const { switch (kind) { case Null: return "null"; case Bool: return "bool"; case Num: return "number"; case Str: return "string"; case Arr: return "array"; default: return "object"; } }
};

// ------------------------------------------------------------------------------------ serialise
inline void json_escape(const std::string& s, std::string& out) {
    static const char* hex = "0123456789abcdef";
    out.push_back('"');
    size_t i = 0, n = s.size();
    while (i < n) {
        unsigned char c = (unsigned char)s[i];
        if (c < 0x80) {
            if (c == '"' || c == '\\') { out.push_back('\\'); out.push_back((char)c); }
            else if (c == '\n') out += "\\n"; else if (c == '\r') out += "\\r"; else if (c == '\t') out += "\\t";
            else if (c < 0x20 || c == '<' || c == '>' || c == '&') { out += "\\u00"; out.push_back(hex[c >> 4]); out.push_back(hex[c & 15]); }
            else out.push_back((char)c);
            i++; continue;
        }
        // decode one UTF-8 sequence; invalid -> U+FFFD (one byte consumed), as Go does
        int len = 0; uint32_t cp = 0;
        if (c >= 0xC2 && c <= 0xDF) { len = 2; cp = c & 0x1F; } else if (c >= 0xE0 && c <= 0xEF) { len = 3; cp = c & 0x0F; } else if (c >= 0xF0 && c <= 0xF4) { len = 4; cp = c & 0x07; }
        bool ok = len != 0 && i + len <= n;
        for (int k = 1; ok && k < len; k++) { unsigned char cc = (unsigned char)s[i + k]; if ((cc & 0xC0) != 0x80) ok = false; else cp = (cp << 6) | (cc & 0x3F); }
        if (ok && ((len == 3 && (cp < 0x800 || (cp >= 0xD800 && cp <= 0xDFFF))) || (len == 4 && (cp < 0x10000 || cp > 0x10FFFF)))) ok = false;
        if (!ok) { out += "\\ufffd"; i++; continue; }
        if (cp == 0x2028 || cp == 0x2029) { out += cp == 0x2028 ? "\\u2028" : "\\u2029"; }
        else out.append(s, i, len);
        i += len;
    }
    out.push_back('"');
}

// compact raw JSON (drop insignificant whitespace) with Go's HTML escaping inside string literals
inline void json_compact_raw(const std::string& raw, std::string& out) {
    bool in_str = false;
    for (size_t i = 0; i < raw.size(); i++) {
        unsigned char c = (unsigned char)raw[i];
        if (in_str) {
            if (c == '\\' && i + 1 < raw.size()) { out.push_back('\\'); out.push_back(raw[++i]); continue; }
            if (c == '"') in_str = false;
            if (c == '<' || c == '>' || c == '&') { static const char* hex = "0123456789abcdef"; out += "\\u00"; out.push_back(hex[c >> 4]); out.push_back(hex[c & 15]); continue; }
            if (c == 0xE2 && i + 2 < raw.size() && (unsigned char)raw[i + 1] == 0x80 && ((unsigned char)raw[i + 2] & 0xFE) == 0xA8) { out += (raw[i + 2] & 1) ? "\\u2029" : "\\u2028"; i += 2; continue; }
            out.push_back((char)c);
        } else {
            if (c == ' ' || c == '\t' || c == '\n' || c == '\r') continue;
            if (c == '"') in_str = true;
            out.push_back((char)c);
        }
    }
}

inline void json_write(const Json& j, std::string& out) {
    switch (j.kind) {
        case Json::Null: out += "null"; break;
        case Json::Bool: out += j.b ? "true" : "false"; break;
        case Json::Num:
            if (j.is_int) out += std::to_string(j.i);
            else if (!std::isfinite(j.num)) out += "0";
            else { char buf[40]; snprintf(buf, sizeof buf, "%.6g", j.num); out += buf; }
            break;
        case Json::Str: json_escape(j.s, out); break;
        case Json::Raw: json_compact_raw(j.s, out); break;
        case Json::Arr:
            out.push_back('[');
            for (size_t i = 0; i < j.arr.size(); i++) { if (i) out.push_back(','); json_write(j.arr[i], out); }
            out.push_back(']'); break;
        case Json::Obj: {
            out.push_back('{');
            std::vector<const std::pair<std::string, Json>*> items; for (auto& kv : j.obj) items.push_back(&kv);
            if (j.sorted) std::sort(items.begin(), items.end(), [](auto a, auto b) { return a->first < b->first; });
            bool first = true;
            for (auto kv : items) { if (!first) out.push_back(','); first = false; json_escape(kv->first, out); out.push_back(':'); json_write(kv->second, out); }
            out.push_back('}');
        } break;
    }
}
inline std::string json_dump(const Json& j) { std::string s; json_write(j, s); return s; }

// ------------------------------------------------------------------------------------ parse
struct JsonParser {
    const std::string& t; size_t p = 0; std::string err;
    explicit JsonParser(const std::string& text) : t(text) {}
    static std::string quote_char(unsigned char c) {
        if (c == '\'') return "'\\''"; if (c == '"') return "'\"'";
        if (c >= 0x20 && c < 0x7f) return std::string("'") + (char)c + "'";
        char b[16]; snprintf(b, sizeof b, c < 0x80 ? "'\\x%02x'" : "'\\u00%02x'", c); return b;
    }
    void ws() { while (p < t.size() && (t[p] == ' ' || t[p] == '\t' || t[p] == '\n' || t[p] == '\r')) p++; }
    bool fail_at(const std::string& ctx) {
        if (p >= t.size()) err = "unexpected end of JSON input";
        else err = "invalid character " + quote_char((unsigned char)t[p]) + " " + ctx;
        return false;
    }
    bool parse(Json& out) {
        ws(); if (!value(out, "looking for beginning of value")) return false;
        ws(); if (p < t.size()) { err = "invalid character " + quote_char((unsigned char)t[p]) + " after top-level value"; return false; }
        return true;
    }
    bool value(Json& out, const char* ctx) {
        ws(); if (p >= t.size()) return fail_at(ctx);
        char c = t[p];
        if (c == '{') return object(out);
        if (c == '[') return array(out);
        if (c == '"') { out.kind = Json::Str; return str(out.s); }
        if (c == 't') return lit("true", out, Json::Bool, true);
        if (c == 'f') return lit("false", out, Json::Bool, false);
        if (c == 'n') return lit("null", out, Json::Null, false);
        if (c == '-' || (c >= '0' && c <= '9')) return number(out);
        return fail_at(ctx);
    }
    bool lit(const char* w, Json& out, Json::Kind k, bool b) {
        for (size_t i = 0; w[i]; i++) {
            if (p >= t.size()) { err = "unexpected end of JSON input"; return false; }
            if (t[p] != w[i]) { err = "invalid character " + quote_char((unsigned char)t[p]) + " in literal " + w + " (expecting " + quote_char((unsigned char)w[i]) + ")"; return false; }
            p++;
        }
        out.kind = k; out.b = b; return true;
    }
    bool number(Json& out) {
        size_t s = p; if (t[p] == '-') p++;
        if (p >= t.size()) return fail_at("in numeric literal");
        if (t[p] == '0') p++; else if (t[p] >= '1' && t[p] <= '9') { while (p < t.size() && isdigit((unsigned char)t[p])) p++; } else return fail_at("in numeric literal");
        bool isint = true;
        if (p < t.size() && t[p] == '.') { isint = false; p++; if (p >= t.size() || !isdigit((unsigned char)t[p])) return fail_at("after decimal point in numeric literal"); while (p < t.size() && isdigit((unsigned char)t[p])) p++; }
        if (p < t.size() && (t[p] == 'e' || t[p] == 'E')) { isint = false; p++; if (p < t.size() && (t[p] == '+' || t[p] == '-')) p++; if (p >= t.size() || !isdigit((unsigned char)t[p])) return fail_at("in exponent of numeric literal"); while (p < t.size() && isdigit((unsigned char)t[p])) p++; }
        out.kind = Json::Num; out.num = strtod(t.substr(s, p - s).c_str(), nullptr);
        if (isint && p - s < 18) { out.is_int = true; out.i = strtoll(t.substr(s, p - s).c_str(), nullptr, 10); }
        return true;
    }
    static void put_utf8(uint32_t cp, std::string& o) {
        if (cp < 0x80) o.push_back((char)cp);
        else if (cp < 0x800) { o.push_back((char)(0xC0 | (cp >> 6))); o.push_back((char)(0x80 | (cp & 0x3F))); }
        else if (cp < 0x10000) { o.push_back((char)(0xE0 | (cp >> 12))); o.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); o.push_back((char)(0x80 | (cp & 0x3F))); }
        else { o.push_back((char)(0xF0 | (cp >> 18))); o.push_back((char)(0x80 | ((cp >> 12) & 0x3F))); o.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); o.push_back((char)(0x80 | (cp & 0x3F))); }
    }
    bool hex4(uint32_t& v) {
        v = 0;
        for (int i = 0; i < 4; i++) { if (p >= t.size()) { err = "unexpected end of JSON input"; return false; } char c = t[p]; int d = c >= '0' && c <= '9' ? c - '0' : c >= 'a' && c <= 'f' ? c - 'a' + 10 : c >= 'A' && c <= 'F' ? c - 'A' + 10 : -1; if (d < 0) return fail_at("in \\u hexadecimal character escape"); v = v * 16 + d; p++; }
        return true;
    }
    bool str(std::string& out) {
        out.clear(); p++;
        for (;;) {
            if (p >= t.size()) { err = "unexpected end of JSON input"; return false; }
            unsigned char c = (unsigned char)t[p];
            if (c == '"') { p++; return true; }
            if (c < 0x20) return fail_at("in string literal");
            if (c == '\\') {
                p++; if (p >= t.size()) { err = "unexpected end of JSON input"; return false; }
                char e = t[p];
                switch (e) {
                    case '"': out.push_back('"'); p++; break; case '\\': out.push_back('\\'); p++; break; case '/': out.push_back('/'); p++; break;
                    case 'b': out.push_back('\b'); p++; break; case 'f': out.push_back('\f'); p++; break; case 'n': out.push_back('\n'); p++; break;
                    case 'r': out.push_back('\r'); p++; break; case 't': out.push_back('\t'); p++; break;
                    case 'u': {
                        p++; uint32_t cp; if (!hex4(cp)) return false;
                        if (cp >= 0xD800 && cp <= 0xDBFF && p + 1 < t.size() && t[p] == '\\' && t[p + 1] == 'u') {
                            size_t save = p; p += 2; uint32_t lo; if (!hex4(lo)) return false;
                            if (lo >= 0xDC00 && lo <= 0xDFFF) cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00); else { p = save; cp = 0xFFFD; }
                        } else if (cp >= 0xD800 && cp <= 0xDFFF) cp = 0xFFFD;
                        put_utf8(cp, out);
                    } break;
                    default: return fail_at("in string escape code");
                }
                continue;
            }
            out.push_back((char)c); p++;
        }
    }
    bool array(Json& out) {
        out.kind = Json::Arr; p++; ws();
        if (p < t.size() && t[p] == ']') { p++; return true; }
        for (;;) {
            Json v; if (!value(v, "looking for beginning of value")) return false;
            out.arr.push_back(std::move(v)); ws();
            if (p >= t.size()) { err = "unexpected end of JSON input"; return false; }
            if (t[p] == ',') { p++; continue; }
            if (t[p] == ']') { p++; return true; }
            return fail_at("after array element");
        }
    }
    bool object(Json& out) {
        out.kind = Json::Obj; p++; ws();
        if (p < t.size() && t[p] == '}') { p++; return true; }
        for (;;) {
            ws(); if (p >= t.size()) { err = "unexpected end of JSON input"; return false; }
            if (t[p] != '"') return fail_at("looking for beginning of object key string");
            std::string k; if (!str(k)) return false;
            ws(); if (p >= t.size()) { err = "unexpected end of JSON input"; return false; }
            if (t[p] != ':') return fail_at("after object key");
            p++;
            Json v; if (!value(v, "looking for beginning of value")) return false;
            out.set(k, std::move(v)); ws();
            if (p >= t.size()) { err = "unexpected end of JSON input"; return false; }
            if (t[p] == ',') { p++; continue; }
            if (t[p] == '}') { p++; return true; }
            return fail_at("after object key:value pair");
        }
    }
};

}  // namespace b200
