// json.h -- a small JSON value, parser and serialiser for the OpenAI-style request/response bodies the
// API gateway forwards (ramalama-models/helm-chart/templates/api-gateway.yaml:63-104).  Header only.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <string>
#include <utility>
#include <vector>

namespace b2json {

struct Value;
using Array = std::vector<Value>;
using Object = std::vector<std::pair<std::string, Value>>;  // insertion ordered

struct Value {
    enum Kind { Null, Bool, Number, String, Arr, Obj } kind = Null;
    bool b = false;
    double num = 0;
    std::string str;
    std::shared_ptr<Array> arr;
    std::shared_ptr<Object> obj;

    Value() = default;
    Value(std::nullptr_t) {}
    Value(bool v) : kind(Bool), b(v) {}
    Value(int v) : kind(Number), num(v) {}
    Value(int64_t v) : kind(Number), num((double)v) {}
    Value(uint64_t v) : kind(Number), num((double)v) {}
    Value(double v) : kind(Number), num(v) {}
    Value(const char* s) : kind(String), str(s) {}
    Value(std::string s) : kind(String), str(std::move(s)) {}
    static Value array() { Value v; v.kind = Arr; v.arr = std::make_shared<Array>(); return v; }
    static Value object() { Value v; v.kind = Obj; v.obj = std::make_shared<Object>(); return v; }

    bool is_null() const { return kind == Null; }
    bool is_string() const { return kind == String; }
    bool is_number() const { return kind == Number; }
    bool is_bool() const { return kind == Bool; }
    bool is_array() const { return kind == Arr; }
    bool is_object() const { return kind == Obj; }

    // object access; returns a shared null for missing keys
    const Value& operator[](const std::string& key) const {
        static const Value null_value;
        if (kind != Obj) return null_value;
        for (auto& kv : *obj)
            if (kv.first == key) return kv.second;
        return null_value;
    }
    Value& set(const std::string& key, Value v) {
        if (kind != Obj) { kind = Obj; obj = std::make_shared<Object>(); }
        for (auto& kv : *obj)
            if (kv.first == key) { kv.second = std::move(v); return kv.second; }
        obj->emplace_back(key, std::move(v));
        return obj->back().second;
    }
    void push(Value v) {
        if (kind != Arr) { kind = Arr; arr = std::make_shared<Array>(); }
        arr->push_back(std::move(v));
    }
    size_t size() const { return kind == Arr ? arr->size() : kind == Obj ? obj->size() : 0; }
    const Value& at(size_t i) const { return (*arr)[i]; }

    double number_or(double d) const { return kind == Number ? num : d; }
    int64_t int_or(int64_t d) const { return kind == Number ? (int64_t)num : d; }
    bool bool_or(bool d) const { return kind == Bool ? b : d; }
    std::string string_or(const std::string& d) const { return kind == String ? str : d; }
};

// length of the well-formed UTF-8 sequence starting at s[i], or 0 if the bytes there are not valid UTF-8
inline size_t utf8_valid_len(const std::string& s, size_t i) {
    const unsigned char c = (unsigned char)s[i];
    size_t n = c < 0x80 ? 1 : (c >= 0xC2 && c < 0xE0) ? 2 : (c >= 0xE0 && c < 0xF0) ? 3 : (c >= 0xF0 && c < 0xF5) ? 4 : 0;
    if (n == 0 || i + n > s.size()) return 0;
    for (size_t k = 1; k < n; k++)
        if (((unsigned char)s[i + k] & 0xC0) != 0x80) return 0;
    if (n == 3) {
        const unsigned char d = (unsigned char)s[i + 1];
        if ((c == 0xE0 && d < 0xA0) || (c == 0xED && d >= 0xA0)) return 0;   // overlong / surrogate
    }
    if (n == 4) {
        const unsigned char d = (unsigned char)s[i + 1];
        if ((c == 0xF0 && d < 0x90) || (c == 0xF4 && d >= 0x90)) return 0;
    }
    return n;
}

// A model can emit byte tokens that do not form valid UTF-8; JSON must be valid UTF-8, so such bytes become U+FFFD.
inline void escape_to(const std::string& s, std::string& out) {
    out.push_back('"');
    for (size_t i = 0; i < s.size(); i++) {
        const unsigned char c = (unsigned char)s[i];
        if (c >= 0x80) {
            const size_t n = utf8_valid_len(s, i);
            if (n == 0) { out += "\\ufffd"; continue; }
            out.append(s, i, n);
            i += n - 1;
            continue;
        }
        switch (c) {
            case '"': out += "\\\""; break;
            case '\\': out += "\\\\"; break;
            case '\n': out += "\\n"; break;
            case '\r': out += "\\r"; break;
            case '\t': out += "\\t"; break;
            case '\b': out += "\\b"; break;
            case '\f': out += "\\f"; break;
            default:
                if (c < 0x20) {
                    char buf[8];
                    snprintf(buf, sizeof buf, "\\u%04x", c);
                    out += buf;
                } else {
                    out.push_back((char)c);
                }
        }
    }
    out.push_back('"');
}

inline void dump_to(const Value& v, std::string& out) {
    switch (v.kind) {
        case Value::Null: out += "null"; break;
        case Value::Bool: out += v.b ? "true" : "false"; break;
        case Value::Number: {
            char buf[40];
            if (std::isfinite(v.num) && v.num == std::floor(v.num) && std::fabs(v.num) < 9.0e15) snprintf(buf, sizeof buf, "%lld", (long long)v.num);
            else if (std::isfinite(v.num)) snprintf(buf, sizeof buf, "%.9g", v.num);
            else snprintf(buf, sizeof buf, "null");
            out += buf;
            break;
        }
        case Value::String: escape_to(v.str, out); break;
        case Value::Arr:
            out.push_back('[');
            for (size_t i = 0; i < v.arr->size(); i++) {
                if (i) out.push_back(',');
                dump_to((*v.arr)[i], out);
            }
            out.push_back(']');
            break;
        case Value::Obj:
            out.push_back('{');
            for (size_t i = 0; i < v.obj->size(); i++) {
                if (i) out.push_back(',');
                escape_to((*v.obj)[i].first, out);
                out.push_back(':');
                dump_to((*v.obj)[i].second, out);
            }
            out.push_back('}');
            break;
    }
}
inline std::string dump(const Value& v) {
    std::string s;
    dump_to(v, s);
    return s;
}

// ---- parser: returns false and fills err on malformed input ----
struct Parser {
    const char* p;
    const char* end;
    std::string err;
    int depth = 0;

    void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) p++; }
    bool fail(const char* m) { if (err.empty()) err = m; return false; }

    static void utf8_append(uint32_t cp, std::string& out) {
        if (cp < 0x80) out.push_back((char)cp);
        else if (cp < 0x800) { out.push_back((char)(0xC0 | (cp >> 6))); out.push_back((char)(0x80 | (cp & 0x3F))); }
        else if (cp < 0x10000) { out.push_back((char)(0xE0 | (cp >> 12))); out.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out.push_back((char)(0x80 | (cp & 0x3F))); }
        else { out.push_back((char)(0xF0 | (cp >> 18))); out.push_back((char)(0x80 | ((cp >> 12) & 0x3F))); out.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out.push_back((char)(0x80 | (cp & 0x3F))); }
    }
    bool hex4(uint32_t* out) {
        if (end - p < 4) return false;
        uint32_t v = 0;
        for (int i = 0; i < 4; i++) {
            char c = *p++;
            v <<= 4;
            if (c >= '0' && c <= '9') v |= c - '0';
            else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
            else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
            else return false;
        }
        *out = v;
        return true;
    }
    bool string(std::string& out) {
        if (p >= end || *p != '"') return fail("expected string");
        p++;
        while (p < end && *p != '"') {
            if (*p == '\\') {
                if (++p >= end) return fail("bad escape");
                char c = *p++;
                switch (c) {
                    case '"': out.push_back('"'); break;
                    case '\\': out.push_back('\\'); break;
                    case '/': out.push_back('/'); break;
                    case 'b': out.push_back('\b'); break;
                    case 'f': out.push_back('\f'); break;
                    case 'n': out.push_back('\n'); break;
                    case 'r': out.push_back('\r'); break;
                    case 't': out.push_back('\t'); break;
                    case 'u': {
                        uint32_t cp;
                        if (!hex4(&cp)) return fail("bad \\u escape");
                        if (cp >= 0xD800 && cp < 0xDC00 && end - p >= 6 && p[0] == '\\' && p[1] == 'u') {
                            p += 2;
                            uint32_t lo;
                            if (!hex4(&lo)) return fail("bad surrogate");
                            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                        }
                        utf8_append(cp, out);
                        break;
                    }
                    default: return fail("bad escape");
                }
            } else {
                out.push_back(*p++);
            }
        }
        if (p >= end) return fail("unterminated string");
        p++;
        return true;
    }
    bool value(Value& v) {
        if (++depth > 64) return fail("nesting too deep");
        ws();
        if (p >= end) return fail("unexpected end");
        bool ok = true;
        if (*p == '{') {
            p++;
            v = Value::object();
            ws();
            if (p < end && *p == '}') { p++; }
            else {
                for (;;) {
                    ws();
                    std::string k;
                    if (!string(k)) { ok = false; break; }
                    ws();
                    if (p >= end || *p != ':') { ok = fail("expected ':'"); break; }
                    p++;
                    Value child;
                    if (!value(child)) { ok = false; break; }
                    v.obj->emplace_back(std::move(k), std::move(child));
                    ws();
                    if (p < end && *p == ',') { p++; continue; }
                    if (p < end && *p == '}') { p++; break; }
                    ok = fail("expected ',' or '}'");
                    break;
                }
            }
        } else if (*p == '[') {
            p++;
            v = Value::array();
            ws();
            if (p < end && *p == ']') { p++; }
            else {
                for (;;) {
                    Value child;
                    if (!value(child)) { ok = false; break; }
                    v.arr->push_back(std::move(child));
                    ws();
                    if (p < end && *p == ',') { p++; continue; }
                    if (p < end && *p == ']') { p++; break; }
                    ok = fail("expected ',' or ']'");
                    break;
                }
            }
        } else if (*p == '"') {
            v.kind = Value::String;
            ok = string(v.str);
        } else if (end - p >= 4 && !std::char_traits<char>::compare(p, "true", 4)) { v = Value(true); p += 4; }
        else if (end - p >= 5 && !std::char_traits<char>::compare(p, "false", 5)) { v = Value(false); p += 5; }
        else if (end - p >= 4 && !std::char_traits<char>::compare(p, "null", 4)) { v = Value(); p += 4; }
        else {
            char* e = nullptr;
            std::string tmp(p, (size_t)std::min<ptrdiff_t>(end - p, 64));
            double d = strtod(tmp.c_str(), &e);
            if (e == tmp.c_str()) ok = fail("unexpected character");
            else { v = Value(d); p += e - tmp.c_str(); }
        }
        depth--;
        return ok;
    }
};

inline bool parse(const std::string& text, Value& out, std::string* err = nullptr) {
    Parser ps{text.data(), text.data() + text.size(), {}, 0};
    bool ok = ps.value(out);
    if (ok) {
        ps.ws();
        if (ps.p != ps.end) ok = ps.fail("trailing characters");
    }
    if (!ok && err) *err = ps.err;
    return ok;
}

}  // namespace b2json
