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

// Index of the quote that closes the string whose first content byte is s[j] (n if there is none).
// memchr does the walking: a quote is real when an even number of backslashes stands before it.
static size_t string_end(const char *s, size_t n, size_t j) {
    const size_t first = j;
    while (j < n) {
        const char *q = static_cast<const char *>(memchr(s + j, '"', n - j));
        if (!q) return n;
        const size_t e = (size_t)(q - s);
        size_t k = e;
        while (k > first && s[k - 1] == '\\') k--;
        if (((e - k) & 1) == 0) return e;
        j = e + 1;
    }
    return n;
}

bool json_find_top_level_string(const char *s, size_t n, const char *key, size_t *begin, size_t *end) {
    const size_t klen = strlen(key);
    int depth = 0;
    size_t i = 0;
    bool expect_key = false;
    while (i < n) {
        char c = s[i];
        if (c == '"') {
            const size_t j = string_end(s, n, i + 1);
            if (j >= n) return false;
            if (depth == 1 && expect_key) {
                bool is_key = (j - i - 1 == klen) && !memcmp(s + i + 1, key, klen);
                size_t k = j + 1;
                while (k < n && (s[k] == ' ' || s[k] == '\t' || s[k] == '\n' || s[k] == '\r')) k++;
                if (k < n && s[k] == ':') {
                    k++;
                    while (k < n && (s[k] == ' ' || s[k] == '\t' || s[k] == '\n' || s[k] == '\r')) k++;
                    if (is_key) {
                        if (k >= n || s[k] != '"') return false;
                        const size_t e = string_end(s, n, k + 1);
                        if (e >= n) return false;
                        *begin = k + 1;
                        *end = e;
                        return true;
                    }
                    expect_key = false;
                    i = k;
                    continue;
                }
            }
            i = j + 1;
            continue;
        }
        if (c == '{' || c == '[') {
            depth++;
            expect_key = (c == '{' && depth == 1);
        } else if (c == '}' || c == ']') {
            depth--;
        } else if (c == ',' && depth == 1) {
            expect_key = true;
        }
        i++;
    }
    return false;
}

size_t json_decode_string(const char *data, size_t n, char *out) {
    std::string tmp;
    std::string quoted;
    // reuse the parser's string routine on a quoted copy only when escapes are present; the common
    // log has few escapes per kilobyte, so copy runs between backslashes directly
    size_t o = 0, i = 0;
    while (i < n) {
        const char *q = (const char *)memchr(data + i, '\\', n - i);
        size_t run = q ? (size_t)(q - (data + i)) : n - i;
        memcpy(out + o, data + i, run);
        o += run;
        i += run;
        if (i >= n) break;
        if (i + 1 >= n) return (size_t)-1;
        char c = data[i + 1];
        i += 2;
        switch (c) {
            case '"': out[o++] = '"'; break;
            case '\\': out[o++] = '\\'; break;
            case '/': out[o++] = '/'; break;
            case 'b': out[o++] = '\b'; break;
            case 'f': out[o++] = '\f'; break;
            case 'n': out[o++] = '\n'; break;
            case 'r': out[o++] = '\r'; break;
            case 't': out[o++] = '\t'; break;
            case 'u': {
                auto hex4 = [&](size_t at, uint32_t &v) {
                    if (at + 4 > n) return false;
                    v = 0;
                    for (int k = 0; k < 4; k++) {
                        char h = data[at + k];
                        int d = (h >= '0' && h <= '9') ? h - '0' : (h >= 'a' && h <= 'f') ? h - 'a' + 10 : (h >= 'A' && h <= 'F') ? h - 'A' + 10 : -1;
                        if (d < 0) return false;
                        v = v * 16 + d;
                    }
                    return true;
                };
                uint32_t v;
                if (!hex4(i, v)) return (size_t)-1;
                i += 4;
                if (v >= 0xD800 && v <= 0xDBFF && i + 6 <= n && data[i] == '\\' && data[i + 1] == 'u') {
                    uint32_t lo;
                    if (hex4(i + 2, lo) && lo >= 0xDC00 && lo <= 0xDFFF) {
                        v = 0x10000 + ((v - 0xD800) << 10) + (lo - 0xDC00);
                        i += 6;
                    }
                }
                if (v >= 0xD800 && v <= 0xDFFF) v = 0xFFFD;
                tmp.clear();
                if (v < 0x80) {
                    out[o++] = (char)v;
                } else if (v < 0x800) {
                    out[o++] = (char)(0xC0 | (v >> 6));
                    out[o++] = (char)(0x80 | (v & 0x3F));
                } else if (v < 0x10000) {
                    out[o++] = (char)(0xE0 | (v >> 12));
                    out[o++] = (char)(0x80 | ((v >> 6) & 0x3F));
                    out[o++] = (char)(0x80 | (v & 0x3F));
                } else {
                    out[o++] = (char)(0xF0 | (v >> 18));
                    out[o++] = (char)(0x80 | ((v >> 12) & 0x3F));
                    out[o++] = (char)(0x80 | ((v >> 6) & 0x3F));
                    out[o++] = (char)(0x80 | (v & 0x3F));
                }
                break;
            }
            default: return (size_t)-1;
        }
    }
    return o;
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
