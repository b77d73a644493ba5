// Java-regex subset parser -> AST.  See pm_regex.h and SPEC.md section 3.
#include "pm_regex.h"

#include <algorithm>
#include <cstring>

#include "podmortem_scan.h"

namespace pm {

std::vector<CpRange> normalize_ranges(std::vector<CpRange> r) {
    std::sort(r.begin(), r.end());
    std::vector<CpRange> out;
    for (auto &x : r) {
        if (!out.empty() && x.first <= out.back().second + 1)
            out.back().second = std::max(out.back().second, x.second);
        else
            out.push_back(x);
    }
    return out;
}

std::vector<CpRange> negate_ranges(const std::vector<CpRange> &in) {
    std::vector<CpRange> out;
    uint32_t prev = 0;
    for (auto &x : normalize_ranges(in)) {
        if (x.first > prev) out.push_back({prev, x.first - 1});
        prev = x.second + 1;
    }
    if (prev <= kMaxCp) out.push_back({prev, kMaxCp});
    return out;
}

static std::vector<CpRange> casefold_ascii(const std::vector<CpRange> &in) {
    std::vector<CpRange> r = in;
    for (auto &x : in) {
        uint32_t a = std::max<uint32_t>(x.first, 0x41), b = std::min<uint32_t>(x.second, 0x5A);
        if (a <= b) r.push_back({a + 32, b + 32});
        a = std::max<uint32_t>(x.first, 0x61), b = std::min<uint32_t>(x.second, 0x7A);
        if (a <= b) r.push_back({a - 32, b - 32});
    }
    return normalize_ranges(r);
}

namespace {

const std::vector<CpRange> kDigit = {{0x30, 0x39}};
const std::vector<CpRange> kWord = {{0x30, 0x39}, {0x41, 0x5A}, {0x5F, 0x5F}, {0x61, 0x7A}};
const std::vector<CpRange> kSpace = {{0x09, 0x0D}, {0x20, 0x20}};

struct Posix {
    const char *name;
    std::vector<CpRange> r;
};
const Posix kPosix[] = {
    {"Alpha", {{0x41, 0x5A}, {0x61, 0x7A}}},
    {"Digit", {{0x30, 0x39}}},
    {"Alnum", {{0x30, 0x39}, {0x41, 0x5A}, {0x61, 0x7A}}},
    {"Upper", {{0x41, 0x5A}}},
    {"Lower", {{0x61, 0x7A}}},
    {"Punct", {{0x21, 0x2F}, {0x3A, 0x40}, {0x5B, 0x60}, {0x7B, 0x7E}}},
    {"Space", {{0x09, 0x0D}, {0x20, 0x20}}},
    {"XDigit", {{0x30, 0x39}, {0x41, 0x46}, {0x61, 0x66}}},
};

struct Flags {
    bool ci = false, dotall = false;
};

struct Fail {
    int status, offset;
    std::string msg;
};

class Parser {
  public:
    explicit Parser(const std::string &src) : raw_(src) {}

    NodePtr run() {
        decode();
        unquote();
        if (cps_.empty()) fail(PM_E_UNSUPPORTED, 0, "empty regex");
        Flags fl;
        NodePtr n = alt(fl);
        if (i_ < cps_.size()) fail(PM_E_SYNTAX, off(), "unmatched ')'");
        return n;
    }

  private:
    const std::string &raw_;
    std::vector<uint32_t> cps_;
    std::vector<int> offs_;
    size_t i_ = 0;

    [[noreturn]] void fail(int status, int offset, const std::string &msg) { throw Fail{status, offset, msg}; }

    int off() const { return i_ < offs_.size() ? offs_[i_] : (int)raw_.size(); }
    int off_at(size_t k) const { return k < offs_.size() ? offs_[k] : (int)raw_.size(); }
    // 0 means end of input: a NUL can not occur in a JSON-borne regex that reached us as a C++ string
    // except as a real character, which is handled because we compare positions, not values.
    bool eof(size_t k = 0) const { return i_ + k >= cps_.size(); }
    uint32_t peek(size_t k = 0) const { return eof(k) ? 0xFFFFFFFFu : cps_[i_ + k]; }
    bool is(char c, size_t k = 0) const { return !eof(k) && cps_[i_ + k] == (uint32_t)(unsigned char)c; }

    void decode() {
        const unsigned char *s = (const unsigned char *)raw_.data();
        size_t n = raw_.size(), i = 0;
        while (i < n) {
            uint32_t c = s[i];
            int len = c < 0x80 ? 1 : (c >> 5) == 6 ? 2 : (c >> 4) == 14 ? 3 : (c >> 3) == 30 ? 4 : 0;
            if (!len || i + len > n) fail(PM_E_SYNTAX, (int)i, "regex is not valid UTF-8");
            uint32_t cp = len == 1 ? c : c & (0xFF >> (len + 1));
            for (int k = 1; k < len; k++) {
                if ((s[i + k] & 0xC0) != 0x80) fail(PM_E_SYNTAX, (int)i, "regex is not valid UTF-8");
                cp = (cp << 6) | (s[i + k] & 0x3F);
            }
            static const uint32_t mins[5] = {0, 0, 0x80, 0x800, 0x10000};
            if (cp < mins[len] || cp > kMaxCp || (cp >= 0xD800 && cp <= 0xDFFF))
                fail(PM_E_SYNTAX, (int)i, "regex is not valid UTF-8");
            cps_.push_back(cp);
            offs_.push_back((int)i);
            i += len;
        }
    }

    // java.util.regex.Pattern rewrites the pattern text before parsing it (Pattern.RemoveQEQuoting;
    // "Quotation" in the Pattern Javadoc): from the first \Q on, quoted characters are replaced IN
    // PLACE by individually escaped ones.  So a quantifier after \E binds to the last quoted
    // character only (`\Qab\E?c` is `ab?c`, `\Qab\E{2}` is `abb`) and quoting also works inside a
    // character class.  ASCII letters are copied as they are, a digit that opens a quotation becomes
    // \x3N (it must not extend an escape in front of it), any other ASCII character gets a backslash.
    // offs_ keeps pointing at the source bytes, so error offsets are unchanged.
    void unquote() {
        const size_t n = cps_.size();
        size_t i = 0;
        while (i + 1 < n) {
            if (cps_[i] != '\\')
                i += 1;
            else if (cps_[i + 1] != 'Q')
                i += 2;
            else
                break;
        }
        if (i + 1 >= n) return;
        std::vector<uint32_t> out(cps_.begin(), cps_.begin() + i);
        std::vector<int> origin(offs_.begin(), offs_.begin() + i);
        auto put = [&](uint32_t c, size_t src) {
            out.push_back(c);
            origin.push_back(offs_[src]);
        };
        i += 2;
        bool in_quote = true, begin = true;
        while (i < n) {
            const size_t src = i;
            const uint32_t c = cps_[i++];
            if (c >= 128 || isalpha_ascii(c)) {
                put(c, src);
            } else if (isdig(c)) {
                if (begin) put('\\', src), put('x', src), put('3', src);
                put(c, src);
            } else if (c != '\\') {
                if (in_quote) put('\\', src);
                put(c, src);
            } else if (in_quote) {
                if (i < n && cps_[i] == 'E') {
                    i++;
                    in_quote = false;
                } else {
                    put('\\', src), put('\\', src);
                }
            } else {
                if (i < n && cps_[i] == 'Q') {
                    i++;
                    in_quote = begin = true;
                    continue;
                }
                put(c, src);
                if (i < n) put(cps_[i], i), i++;
            }
            begin = false;
        }
        cps_.swap(out);
        offs_.swap(origin);
    }

    static NodePtr mk(NodeKind k, int pos) {
        NodePtr n(new Node());
        n->kind = k;
        n->pos = pos;
        return n;
    }

    NodePtr alt(Flags &fl) {
        int start = off();
        std::vector<NodePtr> alts;
        alts.push_back(cat(fl));
        while (is('|')) {
            i_++;
            alts.push_back(cat(fl));
        }
        if (alts.size() == 1) return std::move(alts[0]);
        NodePtr n = mk(NodeKind::Alt, start);
        n->kids = std::move(alts);
        return n;
    }

    NodePtr cat(Flags &fl) {
        int start = off();
        std::vector<NodePtr> items;
        while (!eof() && !is('|') && !is(')')) {
            NodePtr n = repeat(fl);
            if (n) items.push_back(std::move(n));
        }
        if (items.empty()) return mk(NodeKind::Empty, start);
        if (items.size() == 1) return std::move(items[0]);
        NodePtr n = mk(NodeKind::Cat, start);
        n->kids = std::move(items);
        return n;
    }

    static bool is_assertion(const Node &n) {
        return n.kind == NodeKind::Bol || n.kind == NodeKind::EolJava || n.kind == NodeKind::End ||
               n.kind == NodeKind::WordB || n.kind == NodeKind::NotWordB;
    }

    NodePtr repeat(Flags &fl) {
        int at = off();
        NodePtr atom_n = atom(fl);
        if (!atom_n) return nullptr;
        int lo, hi;
        if (is('*')) {
            lo = 0, hi = -1, i_++;
        } else if (is('+')) {
            lo = 1, hi = -1, i_++;
        } else if (is('?')) {
            lo = 0, hi = 1, i_++;
        } else if (is('{')) {
            braces(lo, hi);
        } else {
            return atom_n;
        }
        if (is_assertion(*atom_n)) fail(PM_E_UNSUPPORTED, at, "quantified assertion");
        bool lazy = false;
        if (is('?')) {
            lazy = true, i_++;
        } else if (is('+')) {
            fail(PM_E_UNSUPPORTED, off(), "possessive quantifier");
        }
        if (is('*') || is('+') || is('?') || is('{')) fail(PM_E_SYNTAX, off(), "dangling quantifier");
        NodePtr r = mk(NodeKind::Rep, at);
        r->lo = lo, r->hi = hi, r->lazy = lazy;
        r->kids.push_back(std::move(atom_n));
        return r;
    }

    static bool isdig(uint32_t c) { return c >= '0' && c <= '9'; }

    void braces(int &lo, int &hi) {
        int at = off();
        i_++;
        size_t j = i_;
        long v = 0;
        while (!eof() && isdig(peek())) v = std::min<long>(v * 10 + (peek() - '0'), 1 << 20), i_++;
        if (j == i_) fail(PM_E_UNSUPPORTED, at, "'{' is not a valid quantifier");
        lo = (int)v;
        hi = lo;
        if (is(',')) {
            i_++;
            j = i_;
            v = 0;
            while (!eof() && isdig(peek())) v = std::min<long>(v * 10 + (peek() - '0'), 1 << 20), i_++;
            hi = i_ > j ? (int)v : -1;
        }
        if (!is('}')) fail(PM_E_UNSUPPORTED, at, "'{' is not a valid quantifier");
        i_++;
        if (hi >= 0 && hi < lo) fail(PM_E_SYNTAX, at, "repetition bounds out of order");
        if (lo > kMaxRepeat || hi > kMaxRepeat) fail(PM_E_TOO_COMPLEX, at, "repetition bound > 256");
    }

    NodePtr lit(uint32_t cp, const Flags &fl, int at) {
        if (cp >= 0xD800 && cp <= 0xDFFF) fail(PM_E_UNSUPPORTED, at, "surrogate code point");
        NodePtr n = mk(NodeKind::Lit, at);
        n->cp = cp;
        n->ci = fl.ci && ((cp >= 0x41 && cp <= 0x5A) || (cp >= 0x61 && cp <= 0x7A));
        return n;
    }

    NodePtr set(std::vector<CpRange> r, const Flags &fl, int at, bool negate) {
        r = normalize_ranges(std::move(r));
        if (fl.ci) r = casefold_ascii(r);
        if (negate) r = negate_ranges(r);
        NodePtr n = mk(NodeKind::Set, at);
        n->ranges = std::move(r);
        return n;
    }

    NodePtr atom(Flags &fl) {
        int at = off();
        if (is('(')) return group(fl);
        if (is('[')) return klass(fl);
        if (is('.')) {
            i_++;
            NodePtr n = mk(NodeKind::Dot, at);
            n->dotall = fl.dotall;
            return n;
        }
        if (is('^')) {
            i_++;
            return mk(NodeKind::Bol, at);
        }
        if (is('$')) {
            i_++;
            return mk(NodeKind::EolJava, at);
        }
        if (is('*') || is('+') || is('?')) fail(PM_E_SYNTAX, at, "dangling quantifier");
        if (is('{')) fail(PM_E_UNSUPPORTED, at, "'{' is not a valid quantifier");
        if (is('\\')) return escape(fl);
        uint32_t c = peek();
        i_++;
        return lit(c, fl, at);
    }

    static bool isalnum_ascii(uint32_t c) {
        return (c >= '0' && c <= '9') || (c >= 'A' && c <= 'Z') || (c >= 'a' && c <= 'z');
    }
    static bool isalpha_ascii(uint32_t c) { return (c >= 'A' && c <= 'Z') || (c >= 'a' && c <= 'z'); }

    NodePtr group(Flags &fl) {
        int at = off();
        i_++;
        Flags inner = fl;
        if (is('?')) {
            i_++;
            if (is(':')) {
                i_++;
            } else if (is('<')) {
                if (is('=', 1) || is('!', 1)) fail(PM_E_UNSUPPORTED, at, "look-behind");
                i_++;
                size_t j = i_;
                while (!eof() && isalnum_ascii(peek())) i_++;
                if (j == i_ || !isalpha_ascii(cps_[j]) || !is('>')) fail(PM_E_SYNTAX, at, "bad group name");
                i_++;
            } else if (is('=') || is('!')) {
                fail(PM_E_UNSUPPORTED, at, "look-ahead");
            } else if (is('>')) {
                fail(PM_E_UNSUPPORTED, at, "atomic group");
            } else {
                bool on = true;
                Flags tmp = fl;
                for (;;) {
                    if (eof()) fail(PM_E_SYNTAX, at, "unterminated group");
                    if (is('i'))
                        tmp.ci = on;
                    else if (is('s'))
                        tmp.dotall = on;
                    else if (is('-') && on)
                        on = false;
                    else if (is(')') || is(':'))
                        break;
                    else
                        fail(PM_E_UNSUPPORTED, at, "inline flag outside {i,s}");
                    i_++;
                }
                if (is(')')) {
                    i_++;
                    fl = tmp;   // applies to the rest of the enclosing group
                    return nullptr;
                }
                i_++;
                inner = tmp;
            }
        }
        NodePtr n = alt(inner);
        if (!is(')')) fail(PM_E_SYNTAX, at, "unterminated group");
        i_++;
        return n;
    }

    uint32_t hex(int n, int at) {
        uint32_t v = 0;
        for (int k = 0; k < n; k++) {
            if (eof()) fail(PM_E_SYNTAX, at, "bad hex escape");
            uint32_t c = peek();
            int d = (c >= '0' && c <= '9') ? c - '0' : (c >= 'a' && c <= 'f') ? c - 'a' + 10 : (c >= 'A' && c <= 'F') ? c - 'A' + 10 : -1;
            if (d < 0) fail(PM_E_SYNTAX, at, "bad hex escape");
            v = v * 16 + d;
            i_++;
        }
        return v;
    }

    // After the backslash: escapes that denote one code point.  Returns false if not one.
    bool escape_cp(int at, uint32_t &out) {
        if (eof()) return false;
        switch (peek()) {
            case 't': out = 9; i_++; return true;
            case 'n': out = 10; i_++; return true;
            case 'r': out = 13; i_++; return true;
            case 'f': out = 12; i_++; return true;
            case 'a': out = 7; i_++; return true;
            case 'e': out = 27; i_++; return true;
            case 'x': {
                i_++;
                if (is('{')) {
                    i_++;
                    size_t j = i_;
                    uint32_t v = 0;
                    while (!eof()) {
                        uint32_t c = peek();
                        int d = (c >= '0' && c <= '9') ? c - '0' : (c >= 'a' && c <= 'f') ? c - 'a' + 10 : (c >= 'A' && c <= 'F') ? c - 'A' + 10 : -1;
                        if (d < 0) break;
                        v = v * 16 + d;
                        i_++;
                        if (i_ - j > 6) fail(PM_E_SYNTAX, at, "bad hex escape");
                    }
                    if (j == i_ || !is('}')) fail(PM_E_SYNTAX, at, "bad hex escape");
                    i_++;
                    if (v > kMaxCp) fail(PM_E_SYNTAX, at, "code point out of range");
                    out = v;
                    return true;
                }
                out = hex(2, at);
                return true;
            }
            case 'u': i_++; out = hex(4, at); return true;
            default: return false;
        }
    }

    std::vector<CpRange> posix(int at, bool &neg) {
        neg = is('P');
        i_++;
        if (!is('{')) fail(PM_E_UNSUPPORTED, at, "\\p without braces");
        i_++;
        std::string name;
        while (!eof() && !is('}')) {
            name.push_back(peek() < 128 ? (char)peek() : '?');
            i_++;
        }
        if (eof()) fail(PM_E_SYNTAX, at, "unterminated \\p{");
        i_++;
        for (auto &p : kPosix)
            if (name == p.name) return p.r;
        fail(PM_E_UNSUPPORTED, at, "property \\p{" + name + "}");
    }

    NodePtr escape(Flags &fl) {
        int at = off();
        i_++;
        if (eof()) fail(PM_E_SYNTAX, at, "trailing backslash");
        uint32_t cp;
        if (escape_cp(at, cp)) return lit(cp, fl, at);
        uint32_t c = peek();
        switch (c) {
            case 'd': case 'w': case 's': case 'D': case 'W': case 'S': {
                i_++;
                const std::vector<CpRange> &base = (c == 'd' || c == 'D') ? kDigit : (c == 'w' || c == 'W') ? kWord : kSpace;
                NodePtr n = mk(NodeKind::Set, at);
                n->ranges = (c < 'a') ? negate_ranges(base) : normalize_ranges(base);
                return n;
            }
            case 'p': case 'P': {
                bool neg;
                auto r = posix(at, neg);
                return set(r, fl, at, neg);
            }
            case 'b':
                i_++;
                if (is('{')) fail(PM_E_UNSUPPORTED, at, "\\b{..}");
                return mk(NodeKind::WordB, at);
            case 'B': i_++; return mk(NodeKind::NotWordB, at);
            case 'A': i_++; return mk(NodeKind::Bol, at);
            case 'Z': i_++; return mk(NodeKind::EolJava, at);
            case 'z': i_++; return mk(NodeKind::End, at);
            default: break;
        }
        if (c < 128 && std::strchr("GRXhHvVNck0123456789", (int)c)) fail(PM_E_UNSUPPORTED, at, std::string("escape \\") + (char)c);
        if (isalpha_ascii(c)) fail(PM_E_SYNTAX, at, std::string("illegal escape \\") + (char)c);
        i_++;
        return lit(c, fl, at);
    }

    // One class item: a code point (returns true, sets cp) or a set (returns false, fills r).
    bool class_item(int at_class, uint32_t &cp, std::vector<CpRange> &r) {
        int p = off();
        if (eof()) fail(PM_E_SYNTAX, at_class, "unterminated class");
        if (is('[')) fail(PM_E_UNSUPPORTED, p, "nested class");
        if (is('&') && is('&', 1)) fail(PM_E_UNSUPPORTED, p, "class intersection");
        if (is('\\')) {
            i_++;
            if (eof()) fail(PM_E_SYNTAX, p, "trailing backslash");
            if (escape_cp(p, cp)) return true;
            uint32_t e = peek();
            if (e == 'd' || e == 'w' || e == 's' || e == 'D' || e == 'W' || e == 'S') {
                i_++;
                const std::vector<CpRange> &base = (e == 'd' || e == 'D') ? kDigit : (e == 'w' || e == 'W') ? kWord : kSpace;
                r = (e < 'a') ? negate_ranges(base) : base;
                return false;
            }
            if (e == 'p' || e == 'P') {
                bool neg;
                r = posix(p, neg);
                if (neg) r = negate_ranges(r);
                return false;
            }
            if (isalnum_ascii(e)) fail(PM_E_UNSUPPORTED, p, std::string("escape \\") + (char)e + " in class");
            i_++;
            cp = e;
            return true;
        }
        cp = peek();
        i_++;
        return true;
    }

    NodePtr klass(Flags &fl) {
        int at = off();
        i_++;
        bool neg = false;
        if (is('^')) neg = true, i_++;
        if (is(']')) fail(PM_E_UNSUPPORTED, at, "']' first in class");
        std::vector<CpRange> ranges;
        for (;;) {
            if (is(']')) {
                i_++;
                break;
            }
            uint32_t cp = 0;
            std::vector<CpRange> r;
            bool single = class_item(at, cp, r);
            if (single && is('-') && !eof(1) && !is(']', 1)) {
                i_++;
                int p2 = off();
                uint32_t cp2 = 0;
                std::vector<CpRange> r2;
                if (!class_item(at, cp2, r2)) fail(PM_E_SYNTAX, p2, "bad range end");
                if (cp2 < cp) fail(PM_E_SYNTAX, p2, "range out of order");
                ranges.push_back({cp, cp2});
            } else if (single) {
                ranges.push_back({cp, cp});
            } else {
                ranges.insert(ranges.end(), r.begin(), r.end());
            }
        }
        for (auto &x : ranges)
            if (x.first <= 0xDFFF && x.second >= 0xD800 && (x.first >= 0xD800 || x.second <= 0xDFFF))
                fail(PM_E_UNSUPPORTED, at, "surrogate code point");
        return set(std::move(ranges), fl, at, neg);
    }
};

void check_tail(const Node &n, bool tail) {
    switch (n.kind) {
        case NodeKind::EolJava:
            if (!tail) throw Fail{PM_E_UNSUPPORTED, n.pos, "'$' followed by a consuming item"};
            break;
        case NodeKind::Cat:
            for (size_t i = 0; i < n.kids.size(); i++) {
                long rest = 0;
                for (size_t j = i + 1; j < n.kids.size(); j++) rest += node_width(*n.kids[j]).second;
                check_tail(*n.kids[i], tail && rest == 0);
            }
            break;
        case NodeKind::Alt:
            for (auto &k : n.kids) check_tail(*k, tail);
            break;
        case NodeKind::Rep: {
            bool single = n.hi >= 0 && n.hi <= 1;
            check_tail(*n.kids[0], tail && (single || node_width(*n.kids[0]).second == 0));
            break;
        }
        default: break;
    }
}

}  // namespace

std::pair<int, int> node_width(const Node &n) {
    switch (n.kind) {
        case NodeKind::Lit: case NodeKind::Set: case NodeKind::Dot: return {1, 1};
        case NodeKind::Cat: {
            long lo = 0, hi = 0;
            for (auto &k : n.kids) {
                auto w = node_width(*k);
                lo = std::min<long>(kInf, lo + w.first);
                hi = std::min<long>(kInf, hi + w.second);
            }
            return {(int)lo, (int)hi};
        }
        case NodeKind::Alt: {
            int lo = kInf, hi = 0;
            for (auto &k : n.kids) {
                auto w = node_width(*k);
                lo = std::min(lo, w.first);
                hi = std::max(hi, w.second);
            }
            return {lo, hi};
        }
        case NodeKind::Rep: {
            auto w = node_width(*n.kids[0]);
            long hi = (n.hi < 0) ? (w.second > 0 ? kInf : 0) : std::min<long>(kInf, (long)w.second * n.hi);
            long lo = std::min<long>(kInf, (long)w.first * n.lo);
            return {(int)lo, (int)hi};
        }
        default: return {0, 0};
    }
}

NodePtr parse_regex(const std::string &src, RegexError &err) {
    try {
        Parser p(src);
        NodePtr n = p.run();
        check_tail(*n, true);
        if (node_width(*n).first == 0) throw Fail{PM_E_UNSUPPORTED, 0, "regex can match the empty string"};
        return n;
    } catch (const Fail &f) {
        err.status = f.status;
        err.offset = f.offset;
        err.message = f.msg;
        return nullptr;
    }
}

}  // namespace pm
