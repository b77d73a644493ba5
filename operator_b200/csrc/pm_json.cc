// Minimal JSON DOM.  See pm_json.h.
#include "pm_json.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace pm {

const JsonValue *JsonValue::get(const char *key) const {
    if (type != Object) return nullptr;
    for (auto &kv : obj)
        if (kv.first == key) return &kv.second;
    return nullptr;
}

namespace {

struct P {
    const char *s;
    size_t n, i = 0;
    std::string err;
    int depth = 0;

    bool fail(const char *m) {
        if (err.empty()) err = m;
        return false;
    }
    void ws() {
        while (i < n && (s[i] == ' ' || s[i] == '\t' || s[i] == '\n' || s[i] == '\r')) i++;
    }
    static void put_utf8(uint32_t cp, std::string &o) {
        if (cp < 0x80) {
            o.push_back((char)cp);
        } else if (cp < 0x800) {
            o.push_back((char)(0xC0 | (cp >> 6)));
            o.push_back((char)(0x80 | (cp & 0x3F)));
        } else if (cp < 0x10000) {
            o.push_back((char)(0xE0 | (cp >> 12)));
            o.push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
            o.push_back((char)(0x80 | (cp & 0x3F)));
        } else {
            o.push_back((char)(0xF0 | (cp >> 18)));
            o.push_back((char)(0x80 | ((cp >> 12) & 0x3F)));
            o.push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
            o.push_back((char)(0x80 | (cp & 0x3F)));
        }
    }
    bool hex4(uint32_t &v) {
        if (i + 4 > n) return fail("bad \\u escape");
        v = 0;
        for (int k = 0; k < 4; k++) {
            char c = s[i++];
            int d = (c >= '0' && c <= '9') ? c - '0' : (c >= 'a' && c <= 'f') ? c - 'a' + 10 : (c >= 'A' && c <= 'F') ? c - 'A' + 10 : -1;
            if (d < 0) return fail("bad \\u escape");
            v = v * 16 + d;
        }
        return true;
    }
    bool string(std::string &o) {
        if (i >= n || s[i] != '"') return fail("expected string");
        i++;
        for (;;) {
            // bulk-copy the run up to the next quote or backslash: logs are megabytes
            const char *q = s + i;
            size_t left = n - i, k = 0;
            while (k < left && q[k] != '"' && q[k] != '\\') k++;
            o.append(q, k);
            i += k;
            if (i >= n) return fail("unterminated string");
            if (s[i] == '"') {
                i++;
                return true;
            }
            i++;
            if (i >= n) return fail("unterminated string");
            char c = s[i++];
            switch (c) {
                case '"': o.push_back('"'); break;
                case '\\': o.push_back('\\'); break;
                case '/': o.push_back('/'); break;
                case 'b': o.push_back('\b'); break;
                case 'f': o.push_back('\f'); break;
                case 'n': o.push_back('\n'); break;
                case 'r': o.push_back('\r'); break;
                case 't': o.push_back('\t'); break;
                case 'u': {
                    uint32_t v;
                    if (!hex4(v)) return false;
                    if (v >= 0xD800 && v <= 0xDBFF && i + 1 < n && s[i] == '\\' && s[i + 1] == 'u') {
                        size_t save = i;
                        i += 2;
                        uint32_t lo;
                        if (!hex4(lo)) return false;
                        if (lo >= 0xDC00 && lo <= 0xDFFF)
                            v = 0x10000 + ((v - 0xD800) << 10) + (lo - 0xDC00);
                        else
                            i = save;
                    }
                    if (v >= 0xD800 && v <= 0xDFFF) v = 0xFFFD;   // String.getBytes(UTF_8) would give '?'
                    put_utf8(v, o);
                    break;
                }
                default: return fail("bad escape");
            }
        }
    }
    bool value(JsonValue &v) {
        if (++depth > 200) return fail("nesting too deep");
        ws();
        if (i >= n) return fail("unexpected end");
        v.raw_begin = i;
        char c = s[i];
        bool ok = true;
        if (c == '{') {
            v.type = JsonValue::Object;
            i++;
            ws();
            if (i < n && s[i] == '}') {
                i++;
            } else {
                for (;;) {
                    ws();
                    std::string k;
                    if (!string(k)) { ok = false; break; }
                    ws();
                    if (i >= n || s[i] != ':') { ok = fail("expected ':'"); break; }
                    i++;
                    v.obj.emplace_back(std::move(k), JsonValue());
                    if (!value(v.obj.back().second)) { ok = false; break; }
                    ws();
                    if (i < n && s[i] == ',') { i++; continue; }
                    if (i < n && s[i] == '}') { i++; break; }
                    ok = fail("expected ',' or '}'");
                    break;
                }
            }
        } else if (c == '[') {
            v.type = JsonValue::Array;
            i++;
            ws();
            if (i < n && s[i] == ']') {
                i++;
            } else {
                for (;;) {
                    v.arr.emplace_back();
                    if (!value(v.arr.back())) { ok = false; break; }
                    ws();
                    if (i < n && s[i] == ',') { i++; continue; }
                    if (i < n && s[i] == ']') { i++; break; }
                    ok = fail("expected ',' or ']'");
                    break;
                }
            }
        } else if (c == '"') {
            v.type = JsonValue::String;
            ok = string(v.str);
        } else if (c == 't' && n - i >= 4 && !memcmp(s + i, "true", 4)) {
            v.type = JsonValue::Bool, v.b = true, i += 4;
        } else if (c == 'f' && n - i >= 5 && !memcmp(s + i, "false", 5)) {
            v.type = JsonValue::Bool, v.b = false, i += 5;
        } else if (c == 'n' && n - i >= 4 && !memcmp(s + i, "null", 4)) {
            v.type = JsonValue::Null, i += 4;
        } else if (c == '-' || (c >= '0' && c <= '9')) {
            size_t j = i;
            bool integral = true;
            if (s[j] == '-') j++;
            while (j < n && ((s[j] >= '0' && s[j] <= '9') || s[j] == '.' || s[j] == 'e' || s[j] == 'E' || s[j] == '+' || s[j] == '-')) {
                if (s[j] == '.' || s[j] == 'e' || s[j] == 'E') integral = false;
                j++;
            }
            std::string t(s + i, j - i);
            char *end = nullptr;
            v.type = JsonValue::Number;
            v.num = strtod(t.c_str(), &end);
            if (end == t.c_str() || *end) ok = fail("bad number");
            v.is_int = integral && t.size() < 19;
            if (v.is_int) v.inum = strtoll(t.c_str(), nullptr, 10);
            i = j;
        } else {
            ok = fail("unexpected character");
        }
        v.raw_end = i;
        depth--;
        return ok;
    }
};

}  // namespace

bool json_parse(const char *data, size_t n, JsonValue &out, std::string &err, size_t *err_off) {
    P p{data, n};
    bool ok = p.value(out);
    if (ok) {
        p.ws();
        if (p.i != n) ok = p.fail("trailing characters");
    }
    if (!ok) {
        err = p.err;
        if (err_off) *err_off = p.i;
    }
    return ok;
}

void json_escape(const std::string &s, std::string &out) {
    out.push_back('"');
    for (unsigned char c : s) {
        switch (c) {
            case '"': out += "\\\""; break;
            case '\\': out += "\\\\"; break;
            case '\n': out += "\\n"; break;
            case '\r': out += "\\r"; break;
            case '\t': out += "\\t"; break;
            default:
                if (c < 0x20) {
                    char b[8];
                    snprintf(b, sizeof b, "\\u%04x", c);
                    out += b;
                } else {
                    out.push_back((char)c);
                }
        }
    }
    out.push_back('"');
}

std::string json_number(double v) {
    if (!std::isfinite(v)) return "null";
    char b[40];
    for (int prec = 15; prec <= 17; prec++) {
        snprintf(b, sizeof b, "%.*g", prec, v);
        if (strtod(b, nullptr) == v) break;
    }
    return b;
}

}  // namespace pm
