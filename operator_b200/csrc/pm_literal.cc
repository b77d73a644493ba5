// Required-literal analysis.  See pm_literal.h.
#include "pm_literal.h"

#include <algorithm>

namespace pm {

namespace {

constexpr size_t kMaxExact = 16;   // members of an exact set / cross product
constexpr size_t kMaxReq = 32;     // members of a required set
constexpr size_t kMaxLen = 64;     // bytes per literal
constexpr int kQualityCap = 24;

struct Info {
    bool has_exact = false;
    std::vector<Literal> exact;   // every string the node can match, when finite and small
    std::vector<Literal> req;     // best required set (empty: none)
};

int quality(const std::vector<Literal> &s) {
    if (s.empty()) return 0;
    size_t q = kMaxLen + 1;
    for (auto &l : s) q = std::min(q, l.bytes.size());
    return (int)q;
}

bool better(const std::vector<Literal> &a, const std::vector<Literal> &b) {
    int qa = std::min(quality(a), kQualityCap), qb = std::min(quality(b), kQualityCap);
    if (qa != qb) return qa > qb;
    return a.size() < b.size();
}

void dedup(std::vector<Literal> &s) {
    std::sort(s.begin(), s.end());
    s.erase(std::unique(s.begin(), s.end()), s.end());
}

void consider(std::vector<Literal> &best, std::vector<Literal> cand) {
    dedup(cand);
    if (cand.empty() || cand.size() > kMaxReq || quality(cand) == 0) return;
    if (best.empty() || better(cand, best)) best = std::move(cand);
}

bool cross(const std::vector<Literal> &a, const std::vector<Literal> &b, std::vector<Literal> &out) {
    if (a.size() * b.size() > kMaxExact) return false;
    out.clear();
    for (auto &x : a)
        for (auto &y : b) {
            if (x.bytes.size() + y.bytes.size() > kMaxLen) return false;
            Literal l = x;
            l.bytes.insert(l.bytes.end(), y.bytes.begin(), y.bytes.end());
            l.mask.insert(l.mask.end(), y.mask.begin(), y.mask.end());
            out.push_back(std::move(l));
        }
    dedup(out);
    return true;
}

Literal cp_literal(uint32_t cp, bool ci) {
    Literal l;
    if (cp < 0x80) {
        l.bytes.push_back(ci ? (uint8_t)(cp | 0x20) : (uint8_t)cp);
        l.mask.push_back(ci ? 0x20 : 0);
    } else {
        uint8_t b[4];
        int n;
        if (cp < 0x800) {
            b[0] = 0xC0 | (cp >> 6), b[1] = 0x80 | (cp & 0x3F), n = 2;
        } else if (cp < 0x10000) {
            b[0] = 0xE0 | (cp >> 12), b[1] = 0x80 | ((cp >> 6) & 0x3F), b[2] = 0x80 | (cp & 0x3F), n = 3;
        } else {
            b[0] = 0xF0 | (cp >> 18), b[1] = 0x80 | ((cp >> 12) & 0x3F), b[2] = 0x80 | ((cp >> 6) & 0x3F), b[3] = 0x80 | (cp & 0x3F), n = 4;
        }
        for (int i = 0; i < n; i++) l.bytes.push_back(b[i]), l.mask.push_back(0);
    }
    return l;
}

Info analyse(const Node &n) {
    Info r;
    switch (n.kind) {
        case NodeKind::Lit:
            r.has_exact = true;
            r.exact.push_back(cp_literal(n.cp, n.ci));
            r.req = r.exact;
            break;
        case NodeKind::Set: {
            size_t count = 0;
            for (auto &x : n.ranges) count += (size_t)(x.second - x.first) + 1;
            if (count == 0 || count > 4) break;
            std::vector<uint32_t> cps;
            for (auto &x : n.ranges)
                for (uint32_t c = x.first; c <= x.second; c++) cps.push_back(c);
            std::vector<bool> used(cps.size(), false);
            for (size_t i = 0; i < cps.size(); i++) {
                if (used[i]) continue;
                uint32_t c = cps[i];
                bool letter = (c >= 0x41 && c <= 0x5A) || (c >= 0x61 && c <= 0x7A);
                bool paired = false;
                if (letter)
                    for (size_t j = i + 1; j < cps.size(); j++)
                        if (!used[j] && cps[j] == (c ^ 0x20)) used[j] = paired = true;
                r.exact.push_back(cp_literal(c, paired));
            }
            r.has_exact = true;
            r.req = r.exact;
            break;
        }
        case NodeKind::Dot: break;
        case NodeKind::Empty: case NodeKind::Bol: case NodeKind::EolJava: case NodeKind::End:
        case NodeKind::WordB: case NodeKind::NotWordB:
            // zero-width: matches exactly the empty string ('$' may swallow one terminator, so it
            // is NOT exact: treat it as opaque)
            if (n.kind != NodeKind::EolJava) {
                r.has_exact = true;
                r.exact.push_back(Literal());
            }
            break;
        case NodeKind::Cat: {
            std::vector<Literal> cur{Literal()}, tmp;
            bool all_exact = true;
            for (auto &k : n.kids) {
                Info ki = analyse(*k);
                consider(r.req, ki.req);
                if (ki.has_exact && cross(cur, ki.exact, tmp)) {
                    cur.swap(tmp);
                } else {
                    all_exact = false;
                    consider(r.req, cur);
                    if (ki.has_exact && ki.exact.size() <= kMaxExact)
                        cur = ki.exact;
                    else
                        cur.assign(1, Literal());
                }
            }
            consider(r.req, cur);
            if (all_exact) {
                r.has_exact = true;
                r.exact = cur;
            }
            break;
        }
        case NodeKind::Alt: {
            bool all_exact = true, all_req = true;
            std::vector<Literal> ex, rq;
            for (auto &k : n.kids) {
                Info ki = analyse(*k);
                if (ki.has_exact)
                    ex.insert(ex.end(), ki.exact.begin(), ki.exact.end());
                else
                    all_exact = false;
                if (!ki.req.empty())
                    rq.insert(rq.end(), ki.req.begin(), ki.req.end());
                else
                    all_req = false;
            }
            dedup(ex);
            if (all_exact && ex.size() <= kMaxExact) {
                r.has_exact = true;
                r.exact = ex;
            }
            if (all_req) consider(r.req, rq);
            if (r.has_exact) consider(r.req, r.exact);
            break;
        }
        case NodeKind::Rep: {
            Info ci = analyse(*n.kids[0]);
            if (n.lo == 0) {
                if (n.hi == 1 && ci.has_exact && ci.exact.size() < kMaxExact) {
                    r.has_exact = true;
                    r.exact = ci.exact;
                    r.exact.push_back(Literal());
                    dedup(r.exact);
                }
                break;
            }
            r.req = ci.req;
            if (ci.has_exact) {
                std::vector<Literal> pw{Literal()}, tmp;
                int k = 0;
                while (k < n.lo && cross(pw, ci.exact, tmp)) pw.swap(tmp), k++;
                if (k > 0) consider(r.req, pw);
                if (k == n.lo && n.hi == n.lo) {
                    r.has_exact = true;
                    r.exact = pw;
                }
            }
            break;
        }
    }
    return r;
}

}  // namespace

namespace {

// Only nodes whose match is a fixed byte string chosen from a finite set, with nothing asserted about
// the text around it (no anchors, no word boundaries: analyse() counts those as "the empty string").
bool plain_strings_only(const Node &n) {
    switch (n.kind) {
        case NodeKind::Lit: return n.cp < 0x80 && n.cp != '\n' && n.cp != '\r';
        case NodeKind::Set:
            for (auto &x : n.ranges)
                if (x.second >= 0x80 || (x.first <= '\n' && '\n' <= x.second) || (x.first <= '\r' && '\r' <= x.second)) return false;
            return true;
        case NodeKind::Empty: return true;
        case NodeKind::Cat: case NodeKind::Alt: case NodeKind::Rep:
            for (auto &k : n.kids)
                if (!plain_strings_only(*k)) return false;
            return true;
        default: return false;
    }
}

}  // namespace

bool literal_hit_is_match(const Node &ast, const std::vector<Literal> &chosen) {
    if (chosen.empty() || !plain_strings_only(ast)) return false;
    Info i = analyse(ast);
    if (!i.has_exact) return false;
    std::vector<Literal> ex = i.exact, ch = chosen;
    dedup(ex);
    dedup(ch);
    return ex == ch;
}

std::vector<Literal> required_literals(const Node &ast, int *min_len) {
    Info i = analyse(ast);
    std::vector<Literal> best = i.req;
    if (i.has_exact) consider(best, i.exact);
    if (min_len) *min_len = quality(best);
    return best;
}

int byte_rarity(uint8_t b) {
    // Rough -log2 P(byte) for container logs (timestamps, levels, Java identifiers, stack frames).
    if (b == ' ') return 3;
    if (b >= 'a' && b <= 'z') {
        switch (b) {
            case 'e': case 't': case 'a': case 'o': case 'i': case 'n': case 's': case 'r': return 4;
            case 'l': case 'c': case 'd': case 'u': case 'm': case 'p': return 5;
            case 'j': case 'q': case 'x': case 'z': case 'k': return 8;
            default: return 6;
        }
    }
    if (b >= '0' && b <= '9') return 4;
    if (b >= 'A' && b <= 'Z') return (b == 'I' || b == 'E' || b == 'T' || b == 'S' || b == 'C' || b == 'A') ? 6 : 7;
    switch (b) {
        case '.': case ':': case '-': return 4;
        case '/': case '=': case '[': case ']': case '(': case ')': case ',': case '_': return 5;
        case '\t': return 6;
        default: break;
    }
    if (b >= 0x80) return 10;
    return 8;
}

}  // namespace pm
