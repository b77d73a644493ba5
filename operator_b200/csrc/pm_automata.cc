// NFA / DFA construction.  See pm_automata.h.
#include "pm_automata.h"

#include <algorithm>
#include <map>
#include <unordered_map>

namespace pm {

// ------------------------------------------------------------------ UTF-8 range compiler
namespace {

struct ByteRange {
    uint8_t lo, hi;
};
using Utf8Seq = std::vector<ByteRange>;

int utf8_encode(uint32_t cp, uint8_t *b) {
    if (cp < 0x80) {
        b[0] = (uint8_t)cp;
        return 1;
    }
    if (cp < 0x800) {
        b[0] = 0xC0 | (cp >> 6), b[1] = 0x80 | (cp & 0x3F);
        return 2;
    }
    if (cp < 0x10000) {
        b[0] = 0xE0 | (cp >> 12), b[1] = 0x80 | ((cp >> 6) & 0x3F), b[2] = 0x80 | (cp & 0x3F);
        return 3;
    }
    b[0] = 0xF0 | (cp >> 18), b[1] = 0x80 | ((cp >> 12) & 0x3F), b[2] = 0x80 | ((cp >> 6) & 0x3F), b[3] = 0x80 | (cp & 0x3F);
    return 4;
}

// Splits [lo, hi] into ranges whose UTF-8 encodings differ only by per-byte intervals.
void utf8_split(uint32_t lo, uint32_t hi, std::vector<Utf8Seq> &out) {
    if (lo > hi) return;
    if (lo <= 0xDFFF && hi >= 0xD800) {   // surrogates are not encodable
        if (lo < 0xD800) utf8_split(lo, 0xD7FF, out);
        if (hi > 0xDFFF) utf8_split(0xE000, hi, out);
        return;
    }
    static const uint32_t bounds[3] = {0x7F, 0x7FF, 0xFFFF};
    for (uint32_t b : bounds)
        if (lo <= b && hi > b) {
            utf8_split(lo, b, out);
            utf8_split(b + 1, hi, out);
            return;
        }
    if (hi < 0x80) {
        out.push_back({{(uint8_t)lo, (uint8_t)hi}});
        return;
    }
    for (int i = 1; i < 4; i++) {
        uint32_t m = (1u << (6 * i)) - 1;
        if ((lo & ~m) != (hi & ~m)) {
            if ((lo & m) != 0) {
                utf8_split(lo, lo | m, out);
                utf8_split((lo | m) + 1, hi, out);
                return;
            }
            if ((hi & m) != m) {
                utf8_split(lo, (hi & ~m) - 1, out);
                utf8_split(hi & ~m, hi, out);
                return;
            }
        }
    }
    uint8_t a[4], b[4];
    int n = utf8_encode(lo, a);
    utf8_encode(hi, b);
    Utf8Seq s;
    for (int i = 0; i < n; i++) s.push_back({a[i], b[i]});
    out.push_back(s);
}

const std::vector<CpRange> kLineTerm = {{0x0A, 0x0A}, {0x0D, 0x0D}, {0x85, 0x85}, {0x2028, 0x2029}};

struct Builder {
    Nfa &nfa;
    int budget;
    bool over = false;

    int add_state() {
        if ((int)nfa.st.size() >= budget) {
            over = true;
            return 0;
        }
        nfa.st.emplace_back();
        return (int)nfa.st.size() - 1;
    }
    void eps(int a, int b, uint8_t cond = A_NONE) { nfa.st[a].eps.push_back({b, cond}); }
    void tr(int a, uint8_t lo, uint8_t hi, int b) { nfa.st[a].tr.push_back({lo, hi, b}); }

    void add_ranges(const std::vector<CpRange> &ranges, int s, int e) {
        std::vector<Utf8Seq> seqs;
        for (auto &r : ranges) utf8_split(r.first, r.second, seqs);
        // share the common trailing "any continuation byte" states to keep the NFA small
        std::map<std::vector<std::pair<int, int>>, int> suffix;   // remaining byte ranges -> state
        for (auto &q : seqs) {
            int cur = s;
            for (size_t i = 0; i < q.size(); i++) {
                if (i + 1 == q.size()) {
                    tr(cur, q[i].lo, q[i].hi, e);
                    break;
                }
                std::vector<std::pair<int, int>> key;
                for (size_t k = i + 1; k < q.size(); k++) key.push_back({q[k].lo, q[k].hi});
                auto it = suffix.find(key);
                int nxt;
                if (it != suffix.end()) {
                    nxt = it->second;
                    tr(cur, q[i].lo, q[i].hi, nxt);
                    cur = -1;
                    break;
                }
                nxt = add_state();
                if (over) return;
                suffix[key] = nxt;
                tr(cur, q[i].lo, q[i].hi, nxt);
                cur = nxt;
            }
        }
    }

    // Builds the fragment for `n` between fresh states; returns {start, end}.
    std::pair<int, int> build(const Node &n) {
        int s = add_state(), e = add_state();
        if (over) return {0, 0};
        switch (n.kind) {
            case NodeKind::Empty: eps(s, e); break;
            case NodeKind::Lit: {
                if (n.ci) {
                    tr(s, (uint8_t)(n.cp | 0x20), (uint8_t)(n.cp | 0x20), e);
                    tr(s, (uint8_t)(n.cp & ~0x20u), (uint8_t)(n.cp & ~0x20u), e);
                } else {
                    add_ranges({{n.cp, n.cp}}, s, e);
                }
                break;
            }
            case NodeKind::Set: add_ranges(n.ranges, s, e); break;
            case NodeKind::Dot:
                if (n.dotall)
                    add_ranges({{0, kMaxCp}}, s, e);
                else
                    add_ranges(negate_ranges(kLineTerm), s, e);
                break;
            case NodeKind::Bol: eps(s, e, A_BOL); break;
            case NodeKind::End: eps(s, e, A_END); break;
            case NodeKind::WordB: eps(s, e, A_WORDB); nfa.uses_word_boundary = true; break;
            case NodeKind::NotWordB: eps(s, e, A_NWORDB); nfa.uses_word_boundary = true; break;
            case NodeKind::EolJava: {
                // '$': optionally one final \r, U+0085, U+2028 or U+2029, then end of line.  Only
                // allowed where nothing consuming follows (checked by the parser), so consuming
                // the terminator does not change whether the line matches.
                int m = add_state();
                if (over) break;
                eps(s, m);
                add_ranges({{0x0D, 0x0D}, {0x85, 0x85}, {0x2028, 0x2029}}, s, m);
                eps(m, e, A_END);
                break;
            }
            case NodeKind::Cat: {
                int cur = s;
                for (auto &k : n.kids) {
                    auto f = build(*k);
                    if (over) return {0, 0};
                    eps(cur, f.first);
                    cur = f.second;
                }
                eps(cur, e);
                break;
            }
            case NodeKind::Alt:
                for (auto &k : n.kids) {
                    auto f = build(*k);
                    if (over) return {0, 0};
                    eps(s, f.first);
                    eps(f.second, e);
                }
                break;
            case NodeKind::Rep: {
                const Node &c = *n.kids[0];
                int cur = s;
                for (int i = 0; i < n.lo; i++) {
                    auto f = build(c);
                    if (over) return {0, 0};
                    eps(cur, f.first);
                    cur = f.second;
                }
                if (n.hi < 0) {
                    auto f = build(c);
                    if (over) return {0, 0};
                    eps(cur, f.first);
                    eps(f.second, f.first);
                    eps(f.second, e);
                    eps(cur, e);
                } else {
                    for (int i = n.lo; i < n.hi; i++) {
                        auto f = build(c);
                        if (over) return {0, 0};
                        eps(cur, f.first);
                        eps(cur, e);
                        cur = f.second;
                    }
                    eps(cur, e);
                }
                break;
            }
        }
        return {s, e};
    }
};

inline bool is_word_byte(int b) {
    return (b >= '0' && b <= '9') || (b >= 'A' && b <= 'Z') || (b >= 'a' && b <= 'z') || b == '_';
}

}  // namespace

bool Nfa::add(const Node &ast, int unit, int max_states) {
    Builder b{*this, max_states};
    auto f = b.build(ast);
    if (b.over) return false;
    st[f.second].match = unit;
    starts.push_back(f.first);
    return true;
}

// ------------------------------------------------------------------ subset construction
namespace {

struct Key {
    std::vector<int> kernel;   // sorted NFA states before closure
    std::vector<int> acc;      // units accepted on entering this state
    uint8_t prev_word = 0, bol = 0, terminal = 0;
    bool operator==(const Key &o) const {
        return prev_word == o.prev_word && bol == o.bol && terminal == o.terminal && kernel == o.kernel && acc == o.acc;
    }
};
struct KeyHash {
    size_t operator()(const Key &k) const {
        uint64_t h = 1469598103934665603ull ^ (k.prev_word | (k.bol << 1) | (k.terminal << 2));
        for (int v : k.kernel) h = (h ^ (uint64_t)v) * 1099511628211ull;
        h = (h ^ 0x9E37u) * 1099511628211ull;
        for (int v : k.acc) h = (h ^ (uint64_t)v) * 1099511628211ull;
        return (size_t)h;
    }
};

}  // namespace

bool build_search_dfa(const Nfa &nfa, bool early_accept, int max_states, Dfa &out) {
    // ---- byte classes: cut points at every transition boundary (and word/non-word edges)
    bool cut[257] = {false};
    cut[0] = cut[256] = true;
    for (auto &s : nfa.st)
        for (auto &t : s.tr) cut[t.lo] = true, cut[(int)t.hi + 1] = true;
    if (nfa.uses_word_boundary) {
        for (int b = 0; b < 256; b++)
            if (is_word_byte(b) != is_word_byte(b == 0 ? 1 : b - 1) || b == 0) cut[b] = true;
    }
    // merge intervals with identical behaviour: signature = set of (state, tr index) hit
    std::vector<int> rep;   // representative byte per class
    {
        std::map<std::vector<int>, int> sig2cls;
        int b = 0;
        while (b < 256) {
            int e = b + 1;
            while (!cut[e]) e++;
            std::vector<int> sig;
            sig.push_back(nfa.uses_word_boundary ? (int)is_word_byte(b) : 0);
            int idx = 0;
            for (auto &s : nfa.st)
                for (auto &t : s.tr) {
                    if (t.lo <= b && b <= t.hi) sig.push_back(idx);
                    idx++;
                }
            auto it = sig2cls.find(sig);
            int c;
            if (it == sig2cls.end()) {
                c = (int)rep.size();
                sig2cls[sig] = c;
                rep.push_back(b);
            } else {
                c = it->second;
            }
            for (int x = b; x < e; x++) out.classmap[x] = (uint8_t)c;
            b = e;
        }
    }
    const int ncls = (int)rep.size();
    const int stride = ncls + 1;
    out.ncls = ncls;

    std::unordered_map<Key, int, KeyHash> ids;
    std::vector<Key> keys;
    std::vector<uint32_t> next;
    auto intern = [&](Key &&k) -> int {
        auto it = ids.find(k);
        if (it != ids.end()) return it->second;
        int id = (int)keys.size();
        ids.emplace(k, id);
        keys.push_back(std::move(k));
        return id;
    };

    std::vector<int> start_kernel = nfa.starts;
    std::sort(start_kernel.begin(), start_kernel.end());
    {
        Key k0;
        k0.kernel = start_kernel;
        k0.bol = 1;
        intern(std::move(k0));
    }
    int sticky = -1;   // early-accept sink

    std::vector<int> stack;
    std::vector<int> closures[3];
    std::vector<uint32_t> mark(nfa.st.size(), 0);
    uint32_t epoch = 0;

    for (size_t cur = 0; cur < keys.size(); cur++) {
        if ((int)keys.size() > max_states) return false;
        next.resize((cur + 1) * stride);
        // note: keys may reallocate inside the loop body, so copy what we need
        const std::vector<int> kernel = keys[cur].kernel;
        const bool pw = keys[cur].prev_word, bol = keys[cur].bol, terminal = keys[cur].terminal;
        const bool is_sticky = early_accept && !keys[cur].acc.empty();
        bool have[3] = {false, false, false};
        for (int c = 0; c <= ncls; c++) {
            if (terminal || is_sticky) {   // after END-OF-LINE / after acceptance: absorbing
                next[cur * stride + c] = (uint32_t)cur;
                continue;
            }
            const bool eol = (c == ncls);
            const int b = eol ? -1 : rep[c];
            const bool nw = !eol && is_word_byte(b);
            // closure under the assertions that hold between the previous and the next symbol;
            // it depends on the next symbol only through (end-of-line, word-ness): 3 variants
            const int variant = eol ? 2 : ((nw && nfa.uses_word_boundary) ? 1 : 0);
            if (!have[variant]) {
                have[variant] = true;
                std::vector<int> &cl = closures[variant];
                epoch++;
                cl.clear();
                stack.assign(kernel.begin(), kernel.end());
                for (int s : stack) mark[s] = epoch;
                while (!stack.empty()) {
                    int s = stack.back();
                    stack.pop_back();
                    cl.push_back(s);
                    for (auto &e : nfa.st[s].eps) {
                        bool ok = true;
                        switch (e.cond) {
                            case A_BOL: ok = bol; break;
                            case A_END: ok = eol; break;
                            case A_WORDB: ok = (pw != nw); break;
                            case A_NWORDB: ok = (pw == nw); break;
                            default: break;
                        }
                        if (ok && mark[e.to] != epoch) {
                            mark[e.to] = epoch;
                            stack.push_back(e.to);
                        }
                    }
                }
            }
            const std::vector<int> &closure = closures[variant];
            Key nk;
            for (int s : closure)
                if (nfa.st[s].match >= 0) nk.acc.push_back(nfa.st[s].match);
            std::sort(nk.acc.begin(), nk.acc.end());
            nk.acc.erase(std::unique(nk.acc.begin(), nk.acc.end()), nk.acc.end());
            if (early_accept && !nk.acc.empty()) {
                if (sticky < 0) {
                    Key sk;
                    sk.acc = nk.acc;
                    sk.terminal = 0;
                    sticky = intern(std::move(sk));
                }
                next[cur * stride + c] = (uint32_t)sticky;
                continue;
            }
            if (eol) {
                nk.terminal = 1;
            } else {
                for (int s : closure)
                    for (auto &t : nfa.st[s].tr)
                        if (t.lo <= b && b <= t.hi) nk.kernel.push_back(t.to);
                nk.kernel.insert(nk.kernel.end(), start_kernel.begin(), start_kernel.end());
                std::sort(nk.kernel.begin(), nk.kernel.end());
                nk.kernel.erase(std::unique(nk.kernel.begin(), nk.kernel.end()), nk.kernel.end());
                nk.prev_word = nw;
            }
            int id = intern(std::move(nk));
            next[cur * stride + c] = (uint32_t)id;
        }
    }
    const int n = (int)keys.size();
    if (n > max_states) return false;

    // ---- Moore minimisation: initial partition by accept list, refine by successor blocks
    std::vector<int> block(n);
    {
        std::map<std::vector<int>, int> m;
        for (int s = 0; s < n; s++) {
            auto it = m.find(keys[s].acc);
            if (it == m.end()) it = m.emplace(keys[s].acc, (int)m.size()).first;
            block[s] = it->second;
        }
    }
    for (;;) {
        std::map<std::vector<int>, int> m;
        std::vector<int> nb(n);
        std::vector<int> sig(stride + 1);
        for (int s = 0; s < n; s++) {
            sig[0] = block[s];
            for (int c = 0; c < stride; c++) sig[c + 1] = block[next[(size_t)s * stride + c]];
            auto it = m.find(sig);
            if (it == m.end()) it = m.emplace(sig, (int)m.size()).first;
            nb[s] = it->second;
        }
        int before = *std::max_element(block.begin(), block.end()) + 1;
        block.swap(nb);
        if ((int)m.size() == before) break;
    }
    int nblocks = *std::max_element(block.begin(), block.end()) + 1;
    // representative + renumber: start block -> 0, non-accepting first, accepting last
    std::vector<int> repst(nblocks, -1);
    for (int s = 0; s < n; s++)
        if (repst[block[s]] < 0) repst[block[s]] = s;
    std::vector<int> order;
    order.push_back(block[0]);
    for (int bk = 0; bk < nblocks; bk++)
        if (bk != block[0] && keys[repst[bk]].acc.empty()) order.push_back(bk);
    int first_accept = (int)order.size();
    for (int bk = 0; bk < nblocks; bk++)
        if (bk != block[0] && !keys[repst[bk]].acc.empty()) order.push_back(bk);
    // (the start state can not be accepting: nullable regexes are rejected by the parser)
    std::vector<int> newid(nblocks);
    for (int i = 0; i < nblocks; i++) newid[order[i]] = i;
    out.nstates = nblocks;
    out.first_accept = first_accept;
    out.next.assign((size_t)nblocks * stride, 0);
    out.accept.clear();
    for (int i = 0; i < nblocks; i++) {
        int s = repst[order[i]];
        for (int c = 0; c < stride; c++) out.next[(size_t)i * stride + c] = (uint16_t)newid[block[next[(size_t)s * stride + c]]];
        if (i >= first_accept) out.accept.push_back(keys[s].acc);
    }
    return nblocks <= 65535;
}

}  // namespace pm
