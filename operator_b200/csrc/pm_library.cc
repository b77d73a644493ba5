// Pattern-library compiler.  See pm_library.h.
#include "pm_library.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>

#include "pm_json.h"

namespace pm {

namespace {

constexpr int kUnitMaxStates = 4096;       // verify DFA of one regex
constexpr int kUnitNfaBudget = 20000;
constexpr int kGroupTableBytes = 64 * 1024;   // combined DFA table: shared memory next to two 64 KiB text tiles (k_dfa_scan)
constexpr int kGroupMaxStates = 8192;
constexpr int kGroupNfaBudget = 400000;

const char *kStackRegex = "^(?:\\s+at\\s|Caused by:|\\s+\\.\\.\\.\\s\\d+\\s(?:more|common frames omitted)|\\s+Suppressed:)";
const char *kSeverities[5] = {"INFO", "LOW", "MEDIUM", "HIGH", "CRITICAL"};
const double kSevMult[5] = {0.2, 0.4, 0.6, 0.8, 1.0};

struct CompileFail {
    int status, pattern, secondary, offset;
    std::string id, msg;
};

struct Section {
    std::vector<uint8_t> bytes;
    template <class T>
    size_t push(const T &v) {
        size_t at = bytes.size();
        bytes.resize(at + sizeof(T));
        memcpy(bytes.data() + at, &v, sizeof(T));
        return at;
    }
    size_t push_raw(const void *p, size_t n) {
        size_t at = bytes.size();
        bytes.resize(at + n);
        if (n) memcpy(bytes.data() + at, p, n);
        return at;
    }
};

std::string str_of(const JsonValue *v, const char *dflt = "") {
    if (!v) return dflt;
    if (v->type == JsonValue::String) return v->str;
    if (v->type == JsonValue::Number) return json_number(v->num);
    return dflt;
}

bool num_of(const JsonValue *v, double &out) {
    if (!v || v->type != JsonValue::Number) return false;
    out = v->num;
    return true;
}

// Cost of sampling the 7-byte window at `o`: expected filter hits per text word (lower is
// better); windows that touch a case-insensitive position are pushed back because they force
// the whole library into fold mode.
double window_cost(const Literal &l, size_t o, int granule) {
    double cost = 0;
    for (int k = 0; k < granule; k++) {
        int bits = 0;
        for (int j = 0; j < kGramLen; j++) bits += byte_rarity(l.bytes[o + k + j]);
        cost += std::ldexp(1.0, -bits);
    }
    for (size_t j = o; j < o + (size_t)window_bytes(granule); j++)
        if (l.mask[j]) cost += 1.0;
    return cost;
}

uint32_t gram_at(const Literal &l, size_t o, bool fold) {
    uint32_t g = 0;
    for (int j = 0; j < kGramLen; j++) g |= (uint32_t)l.bytes[o + j] << (8 * j);
    return fold ? (g | 0x20202020u) : g;
}

struct Unit {
    Dfa dfa;
};

void compile_unit(const std::string &rx, int pattern, int secondary, const std::string &id, Dfa &dfa, NodePtr *ast_out) {
    RegexError err;
    NodePtr ast = parse_regex(rx, err);
    if (!ast) throw CompileFail{err.status, pattern, secondary, err.offset, id, err.message};
    Nfa nfa;
    if (!nfa.add(*ast, 0, kUnitNfaBudget))
        throw CompileFail{PM_E_TOO_COMPLEX, pattern, secondary, -1, id, "regex expands to too many NFA states"};
    if (!build_search_dfa(nfa, /*early_accept=*/true, kUnitMaxStates, dfa))
        throw CompileFail{PM_E_TOO_COMPLEX, pattern, secondary, -1, id, "DFA state limit exceeded"};
    if (ast_out) *ast_out = std::move(ast);
}

struct GroupBuild {
    Dfa dfa;
    ScanTable scan;
    std::vector<int> patterns;
};

// Packs tier-2 primaries into as few combined DFAs as fit the shared-memory budget.
void build_groups(const std::vector<int> &pats, const std::vector<NodePtr> &asts, const std::vector<PatternInfo> &info,
                  std::vector<GroupBuild> &out) {
    if (pats.empty()) return;
    Nfa nfa;
    bool ok = true;
    for (size_t u = 0; u < pats.size() && ok; u++) ok = nfa.add(*asts[pats[u]], (int)u, kGroupNfaBudget);
    GroupBuild g;
    if (ok) ok = build_search_dfa(nfa, /*early_accept=*/false, kGroupMaxStates, g.dfa);
    if (ok) ok = build_scan_table(g.dfa, (size_t)kGroupTableBytes, g.scan);
    if (ok) {
        g.patterns = pats;
        out.push_back(std::move(g));
        return;
    }
    if (pats.size() == 1)
        throw CompileFail{PM_E_TOO_COMPLEX, pats[0], -1, -1, info[pats[0]].id,
                          "regex has no literal of 7+ bytes and its scan DFA exceeds the shared-memory budget"};
    std::vector<int> a(pats.begin(), pats.begin() + pats.size() / 2), b(pats.begin() + pats.size() / 2, pats.end());
    build_groups(a, asts, info, out);
    build_groups(b, asts, info, out);
}

size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

}  // namespace

bool build_scan_table(const Dfa &d, size_t max_bytes, ScanTable &out) {
    out = ScanTable();
    const int n = d.nstates, st = d.stride(), eolc = d.ncls;
    // terminal twins: one per distinct END-OF-LINE target; their rows are the start-of-line row
    std::map<int, int> twin;   // original EOL target -> index among twins
    for (int s = 0; s < n; s++) twin.emplace((int)d.next[(size_t)s * st + eolc], 0);
    int k = 0;
    for (auto &kv : twin) kv.second = k++;
    const int total = n + (int)twin.size();
    auto accepting = [&](int s) { return s < n ? s >= d.first_accept : false; };
    std::vector<int> twin_of(twin.size());
    for (auto &kv : twin) twin_of[kv.second] = kv.first;
    auto is_acc = [&](int s) { return s < n ? accepting(s) : accepting(twin_of[s - n]); };
    // renumber: state 0 first, non-accepting, then accepting
    std::vector<int> order, newid(total);
    order.push_back(0);
    for (int s = 1; s < total; s++)
        if (!is_acc(s)) order.push_back(s);
    out.first_accept = (int)order.size();
    for (int s = 1; s < total; s++)
        if (is_acc(s)) order.push_back(s);
    for (int i = 0; i < total; i++) newid[order[i]] = i;
    out.nstates = total;
    if (total > 65535) return false;
    // transition of (old numbering) state s on raw byte b
    auto step = [&](int s, int b) -> int {
        int row = s < n ? s : 0;   // twins behave like the start-of-line state
        if (b == '\n') return n + twin[(int)d.next[(size_t)row * st + eolc]];
        return (int)d.next[(size_t)row * st + d.classmap[b]];
    };
    out.wide = (size_t)total * 256 * 2 <= max_bytes && (size_t)total * 256 < 65536;
    if (out.wide) {
        out.stride = 256;
        out.tab.resize((size_t)total * 256);
        for (int i = 0; i < total; i++)
            for (int b = 0; b < 256; b++) out.tab[(size_t)i * 256 + b] = (uint16_t)(newid[step(order[i], b)] * 256);
    } else {
        out.stride = d.ncls + 1;
        if ((size_t)total * out.stride * 2 > max_bytes || (size_t)total * out.stride >= 65536) return false;
        memcpy(out.cmap, d.classmap, 256);
        out.cmap[(int)'\n'] = (uint8_t)eolc;
        // representative byte per class
        std::vector<int> repb(d.ncls, -1);
        for (int b = 255; b >= 0; b--)
            if (b != '\n') repb[d.classmap[b]] = b;
        out.tab.assign((size_t)total * out.stride, 0);
        for (int i = 0; i < total; i++) {
            for (int c = 0; c < d.ncls; c++)
                if (repb[c] >= 0) out.tab[(size_t)i * out.stride + c] = (uint16_t)(newid[step(order[i], repb[c])] * out.stride);
            out.tab[(size_t)i * out.stride + eolc] = (uint16_t)(newid[step(order[i], '\n')] * out.stride);
        }
    }
    out.mid_state = newid[step(0, 0x01)];
    for (int i = out.first_accept; i < total; i++) {
        int s = order[i];
        int orig = s < n ? s : twin_of[s - n];
        out.accept.push_back(d.accept[orig - d.first_accept]);
    }
    return true;
}

void build_filter(const std::vector<std::vector<Literal>> &literals, int fold_mode, const GramCostFn *cost, int granule, FilterBuild &out) {
    out = FilterBuild();
    out.granule = granule;
    const size_t win = (size_t)window_bytes(granule);
    struct Chosen {
        int pattern;
        const Literal *lit;
        size_t o;
    };
    std::vector<Chosen> chosen;
    bool fold = fold_mode == 1;
    for (size_t i = 0; i < literals.size(); i++)
        for (auto &l : literals[i]) {
            size_t best_o = 0;
            double best = 1e300;
            for (size_t o = 0; o + win <= l.bytes.size(); o++) {
                bool masked = false;
                for (size_t j = o; j < o + win; j++) masked |= l.mask[j] != 0;
                double c;
                if (cost) {
                    c = 1e-12 * window_cost(l, o, granule);   // static model only breaks ties
                    for (int k = 0; k < granule; k++) c += (*cost)(gram_at(l, o + k, fold_mode == 1));
                    if (masked && fold_mode == 0) c = 1e299;   // would need fold mode, which is fixed
                } else {
                    c = window_cost(l, o, granule);   // penalises masked windows (they force fold mode)
                    if (masked && fold_mode == 0) c = 1e299;
                }
                if (c < best) best = c, best_o = o;
            }
            for (size_t j = best_o; j < best_o + win; j++)
                if (l.mask[j]) fold = true;
            chosen.push_back({(int)i, &l, best_o});
        }
    out.fold = fold;
    std::map<uint32_t, std::vector<Probe>> by_gram;
    for (auto &c : chosen) {
        // 4-byte aligned, both halves padded to whole words with 0xFF: k_confirm compares word-wise and
        // (text | 0xFF) == 0xFF holds for whatever follows the literal in the text
        while (out.litpool.size() % 4) out.litpool.push_back(0xFF);
        uint32_t lit_off = (uint32_t)out.litpool.size();
        out.litpool.insert(out.litpool.end(), c.lit->bytes.begin(), c.lit->bytes.end());
        while (out.litpool.size() % 4) out.litpool.push_back(0xFF);
        out.litpool.insert(out.litpool.end(), c.lit->mask.begin(), c.lit->mask.end());
        while (out.litpool.size() % 4) out.litpool.push_back(0xFF);
        for (int k = 0; k < granule; k++) {
            Probe pr;
            pr.pattern = (uint32_t)c.pattern;
            pr.lit_off = lit_off;
            pr.lit_len = (uint16_t)c.lit->bytes.size();
            pr.gram_off = (uint16_t)(c.o + k);
            by_gram[gram_at(*c.lit, c.o + k, fold)].push_back(pr);
        }
    }
    out.n_grams = (uint32_t)by_gram.size();
    out.log2_bits = 0;
    out.log2_cap = 4;
    if (out.n_grams) {
        out.log2_bits = 15;
        while (out.log2_bits < 20 && (1ull << out.log2_bits) < 1024ull * out.n_grams) out.log2_bits++;
        while ((1u << out.log2_cap) < 2 * out.n_grams) out.log2_cap++;
    }
    out.log2_bits2 = out.log2_bits ? std::min<uint32_t>(out.log2_bits, 18) : 0;
    out.bitmap.assign(out.log2_bits ? (1u << out.log2_bits) / 32 : 4, 0);
    out.bitmap2.assign(out.log2_bits2 ? (1u << out.log2_bits2) / 32 : 4, 0);
    out.gkeys.assign(1u << out.log2_cap, kGramEmpty);
    out.gfirst.assign(1u << out.log2_cap, 0);
    out.gcount.assign(1u << out.log2_cap, 0);
    for (auto &kv : by_gram) {
        uint32_t bit = gram_filter_bit(kv.first, out.log2_bits);
        out.bitmap[bit >> 5] |= 1u << (bit & 31);
        uint32_t bit2 = gram_filter_bit2(kv.first, out.log2_bits2);
        out.bitmap2[bit2 >> 5] |= 1u << (bit2 & 31);
        uint32_t slot = gram_table_slot(kv.first, out.log2_cap);
        while (out.gkeys[slot] != kGramEmpty) slot = (slot + 1) & ((1u << out.log2_cap) - 1);
        out.gkeys[slot] = kv.first;
        out.gfirst[slot] = out.n_probes;
        out.gcount[slot] = (uint32_t)kv.second.size();
        for (auto &pr : kv.second) out.probes.push_back(pr), out.n_probes++;
    }
}

void collect_all_grams(const std::vector<std::vector<Literal>> &literals, bool fold, std::vector<uint32_t> &grams) {
    grams.clear();
    for (auto &ls : literals)
        for (auto &l : ls)
            for (size_t o = 0; o + kGramLen <= l.bytes.size(); o++) grams.push_back(gram_at(l, o, fold));
    std::sort(grams.begin(), grams.end());
    grams.erase(std::unique(grams.begin(), grams.end()), grams.end());
}

int compile_library(const char *spec, size_t n, CompiledLibrary &out, pm_compile_report &rep) {
    memset(&rep, 0, sizeof rep);
    rep.pattern_index = rep.secondary_index = rep.regex_offset = -1;
    out = CompiledLibrary();
    try {
        JsonValue root;
        std::string jerr;
        size_t joff = 0;
        if (!json_parse(spec, n, root, jerr, &joff))
            throw CompileFail{PM_E_SYNTAX, -1, -1, (int)joff, "", "library JSON: " + jerr};
        std::vector<const JsonValue *> libs;
        if (root.type == JsonValue::Array)
            for (auto &v : root.arr) libs.push_back(&v);
        else
            libs.push_back(&root);

        std::vector<PatternDesc> pdesc;
        std::vector<SecDesc> sdesc;
        std::vector<Dfa> &units = out.unit_dfas;
        std::vector<NodePtr> asts;
        std::map<std::string, uint32_t> sec_units;   // secondary regex source -> unit
        std::vector<uint32_t> context_units;         // units walked by the resolve stage (secondaries, STACK)

        for (const JsonValue *lib : libs) {
            if (lib->type != JsonValue::Object) throw CompileFail{PM_E_SYNTAX, -1, -1, -1, "", "library must be a JSON object"};
            const JsonValue *pats = lib->get("patterns");
            if (!pats) continue;
            if (pats->type != JsonValue::Array) throw CompileFail{PM_E_SYNTAX, -1, -1, -1, "", "\"patterns\" must be an array"};
            for (auto &p : pats->arr) {
                int idx = (int)out.patterns.size();
                if (p.type != JsonValue::Object) throw CompileFail{PM_E_SYNTAX, idx, -1, -1, "", "pattern must be an object"};
                PatternInfo pi;
                pi.id = str_of(p.get("id"));
                pi.name = str_of(p.get("name"));
                pi.category = str_of(p.get("category"));
                std::string sev = str_of(p.get("severity"), "INFO");
                for (auto &c : sev) c = (char)toupper((unsigned char)c);
                int rank = -1;
                for (int i = 0; i < 5; i++)
                    if (sev == kSeverities[i]) rank = i;
                if (rank < 0) throw CompileFail{PM_E_SYNTAX, idx, -1, -1, pi.id, "unknown severity '" + sev + "'"};
                pi.severity = sev;
                const JsonValue *prim = p.get("primary_pattern");
                const JsonValue *rx = prim ? prim->get("regex") : nullptr;
                if (!rx || rx->type != JsonValue::String)
                    throw CompileFail{PM_E_SYNTAX, idx, -1, -1, pi.id, "primary_pattern.regex missing"};
                pi.regex = rx->str;

                PatternDesc d;
                memset(&d, 0, sizeof d);
                d.confidence = 1.0;
                num_of(prim->get("confidence"), d.confidence);
                d.mult = kSevMult[rank];
                d.severity = (uint32_t)rank;
                if (const JsonValue *ctx = p.get("context_extraction")) {
                    double v;
                    if (num_of(ctx->get("lines_before"), v)) d.lines_before = (int32_t)v;
                    if (num_of(ctx->get("lines_after"), v)) d.lines_after = (int32_t)v;
                    const JsonValue *st = ctx->get("include_stack_trace");
                    if (st && st->type == JsonValue::Bool) d.include_stack = st->b;
                }
                if (d.lines_before < 0 || d.lines_after < 0 || d.lines_before > 1000 || d.lines_after > 1000)
                    throw CompileFail{PM_E_SYNTAX, idx, -1, -1, pi.id, "context_extraction out of range"};

                // primary regex -> verify DFA (+ AST kept for literal analysis / grouping)
                d.unit = (uint32_t)units.size();
                units.emplace_back();
                NodePtr ast;
                compile_unit(pi.regex, idx, -1, pi.id, units.back(), &ast);
                asts.push_back(std::move(ast));

                d.sec_first = (uint32_t)sdesc.size();
                if (const JsonValue *secs = p.get("secondary_patterns")) {
                    if (secs->type != JsonValue::Array && secs->type != JsonValue::Null)
                        throw CompileFail{PM_E_SYNTAX, idx, -1, -1, pi.id, "secondary_patterns must be an array"};
                    if (secs->arr.size() > (size_t)kMaxSecondary)
                        throw CompileFail{PM_E_SYNTAX, idx, -1, -1, pi.id, "more than 8 secondary patterns"};
                    int j = 0;
                    for (auto &s : secs->arr) {
                        const JsonValue *srx = s.get("regex");
                        if (!srx || srx->type != JsonValue::String)
                            throw CompileFail{PM_E_SYNTAX, idx, j, -1, pi.id, "secondary regex missing"};
                        SecDesc sd;
                        memset(&sd, 0, sizeof sd);
                        double v;
                        if (num_of(s.get("weight"), v)) sd.weight = v;
                        if (num_of(s.get("proximity_window"), v)) sd.window = (int32_t)v;
                        if (sd.window < 0 || sd.window > 1000)
                            throw CompileFail{PM_E_SYNTAX, idx, j, -1, pi.id, "proximity_window out of range"};
                        // identical secondary regexes share one unit (one table, hot in L1/L2)
                        auto cached = sec_units.find(srx->str);
                        if (cached != sec_units.end()) {
                            sd.unit = cached->second;
                        } else {
                            sd.unit = (uint32_t)units.size();
                            units.emplace_back();
                            compile_unit(srx->str, idx, j, pi.id, units.back(), nullptr);
                            sec_units.emplace(srx->str, sd.unit);
                            context_units.push_back(sd.unit);
                        }
                        sdesc.push_back(sd);
                        j++;
                    }
                }
                d.sec_count = (uint32_t)sdesc.size() - d.sec_first;

                std::string mj = "{\"id\":";
                json_escape(pi.id, mj);
                mj += ",\"name\":";
                json_escape(pi.name, mj);
                mj += ",\"severity\":";
                json_escape(pi.severity, mj);
                mj += ",\"category\":";
                json_escape(pi.category, mj);
                if (const JsonValue *rem = p.get("remediation")) {
                    mj += ",\"remediation\":";
                    mj.append(spec + rem->raw_begin, rem->raw_end - rem->raw_begin);
                }
                mj += "}";
                pi.meta_json = std::move(mj);
                out.patterns.push_back(std::move(pi));
                pdesc.push_back(d);
            }
        }
        const int P = (int)out.patterns.size();
        // built-in STACK unit (SPEC.md section 6)
        uint32_t stack_unit = (uint32_t)units.size();
        units.emplace_back();
        compile_unit(kStackRegex, -1, -1, "<stack>", units.back(), nullptr);
        context_units.push_back(stack_unit);

        // ---- tiering: literal-gram filter where a long enough literal is required, DFA otherwise.  "Long enough" is 7
        //      bytes with the filter testing aligned words; if testing every even offset as well (5 bytes are enough then,
        //      k_scan does twice the lookups) takes EVERY pattern off the combined DFA, the library is compiled that way: a
        //      second, table-walking pass over the text costs five times what the extra lookups do.
        out.literals.resize(P);
        std::vector<int> tier2;
        std::vector<std::vector<Literal>> lits_of(P);
        std::vector<int> quality_of(P, 0);
        int shortest = 1 << 30;
        for (int i = 0; i < P; i++) {
            lits_of[i] = required_literals(*asts[i], &quality_of[i]);
            shortest = std::min(shortest, quality_of[i]);
        }
        static const bool allow_half = getenv("PM_FILTER_GRANULE4") == nullptr;
        const int granule = (allow_half && shortest >= kMinLiteralHalf && shortest < kMinLiteral) ? 2 : 4;
        const int min_literal = granule == 2 ? kMinLiteralHalf : kMinLiteral;
        out.granule = granule;
        std::vector<char> unit_literal_is_match(units.size(), 0);
        static const bool allow_literal_match = getenv("PM_NO_LITERAL_MATCH") == nullptr;
        for (int i = 0; i < P; i++) {
            if (quality_of[i] >= min_literal) {
                out.literals[i] = std::move(lits_of[i]);
                pdesc[i].tier = 1;
                // a plain literal (or an alternation of them): finding it IS the match, k_verify has nothing to walk
                if (allow_literal_match && literal_hit_is_match(*asts[i], out.literals[i])) unit_literal_is_match[pdesc[i].unit] = 1;
            } else {
                pdesc[i].tier = 2;
                tier2.push_back(i);
            }
            out.patterns[i].tier = (int)pdesc[i].tier;
        }

        // ---- gram selection (static byte-rarity model; the runtime may re-select from measured
        //      gram frequencies, see build_filter / pm_runtime.cu)
        FilterBuild fb;
        build_filter(out.literals, /*fold_mode=*/-1, nullptr, granule, fb);
        const bool fold = fb.fold;
        const uint32_t n_grams = fb.n_grams, n_probes = fb.n_probes, log2_bits = fb.log2_bits, log2_cap = fb.log2_cap;
        std::vector<uint32_t> &bitmap = fb.bitmap, &gkeys = fb.gkeys, &gfirst = fb.gfirst, &gcount = fb.gcount;
        Section litpool, probes_sec;
        litpool.bytes = fb.litpool;
        probes_sec.push_raw(fb.probes.data(), fb.probes.size() * sizeof(Probe));
        out.fold = fold;

        // ---- combined DFAs for tier 2
        std::vector<GroupBuild> groups;
        build_groups(tier2, asts, out.patterns, groups);

        // ---- tables
        Section units_sec, classmaps, tables, groups_sec, gaccept;
        auto put_dfa = [&](const Dfa &d, uint32_t &table_off, uint32_t &cm_off) {
            cm_off = (uint32_t)classmaps.push_raw(d.classmap, 256);
            while (tables.bytes.size() % 16) tables.bytes.push_back(0);   // 16-byte aligned: vector copies into shared memory
            table_off = (uint32_t)(tables.bytes.size() / 2);
            tables.push_raw(d.next.data(), d.next.size() * 2);
        };
        std::vector<char> is_context(units.size(), 0);
        for (uint32_t cu : context_units) is_context[cu] = 1;
        size_t wide_budget = 16u << 20;
        for (size_t ui = 0; ui < units.size(); ui++) {
            const Dfa &d = units[ui];
            UnitDesc u;
            memset(&u, 0, sizeof u);
            put_dfa(d, u.table_off, u.classmap_off);
            u.wide_off = 0xFFFFFFFFu;
            if (is_context[ui] && (size_t)d.nstates * 512 <= wide_budget) {
                wide_budget -= (size_t)d.nstates * 512;
                while (tables.bytes.size() % 16) tables.bytes.push_back(0);
                u.wide_off = (uint32_t)(tables.bytes.size() / 2);
                std::vector<uint16_t> wide((size_t)d.nstates * 256);
                for (int q = 0; q < d.nstates; q++)
                    for (int b = 0; b < 256; b++) wide[(size_t)q * 256 + b] = d.next[(size_t)q * d.stride() + d.classmap[b]];
                tables.push_raw(wide.data(), wide.size() * 2);
            }
            u.nstates = (uint16_t)d.nstates;
            u.stride = (uint16_t)d.stride();
            u.literal_is_match = unit_literal_is_match[ui];
            u.accept_state = (uint16_t)d.first_accept;   // early-accept: one sticky state
            // mid-line idle state: where the DFA sits after one byte that belongs to no pattern;
            // the entry-state guess of the speculative long-line walk (k_check_long)
            u.mid_state = d.next[(size_t)d.classmap[0x01]];
            {   // dead state: cannot reach the accepting state (backward reachability on the minimised DFA,
                // where all such states have been merged into one)
                const int ns = d.nstates, st = d.stride();
                std::vector<char> live(ns, 0);
                for (int q = d.first_accept; q < ns; q++) live[q] = 1;
                for (bool changed = true; changed;) {
                    changed = false;
                    for (int q = 0; q < ns; q++)
                        if (!live[q])
                            for (int c = 0; c < st; c++)
                                if (live[d.next[(size_t)q * st + c]]) {
                                    live[q] = 1;
                                    changed = true;
                                    break;
                                }
                }
                u.dead_state = 0xFFFF;
                for (int q = 0; q < ns; q++)
                    if (!live[q]) u.dead_state = (uint16_t)q;
            }
            if (d.first_accept >= d.nstates)
                throw CompileFail{PM_E_UNSUPPORTED, -1, -1, -1, "", "regex can never match"};
            units_sec.push(u);
        }
        for (auto &g : groups) {
            GroupDesc gd;
            memset(&gd, 0, sizeof gd);
            gd.classmap_off = (uint32_t)classmaps.push_raw(g.scan.cmap, 256);
            while (tables.bytes.size() % 16) tables.bytes.push_back(0);   // 16-byte aligned for bulk copies
            gd.table_off = (uint32_t)(tables.bytes.size() / 2);
            tables.push_raw(g.scan.tab.data(), g.scan.tab.size() * 2);
            gd.nstates = (uint16_t)g.scan.nstates;
            gd.stride = (uint16_t)g.scan.stride;
            gd.first_accept = (uint16_t)g.scan.first_accept;
            gd.mid_state = (uint16_t)g.scan.mid_state;
            gd.wide = g.scan.wide;
            gd.accept_off = (uint32_t)(gaccept.bytes.size() / 4);
            uint32_t na = (uint32_t)g.scan.accept.size(), run = na + 1;
            for (uint32_t a = 0; a < na; a++) {
                gaccept.push(run);
                run += (uint32_t)g.scan.accept[a].size();
            }
            gaccept.push(run);
            for (auto &lst : g.scan.accept)
                for (int u : lst) gaccept.push((uint32_t)g.patterns[u]);
            groups_sec.push(gd);
            out.group_dfas.push_back(g.dfa);
            out.group_scan.push_back(g.scan);
            out.group_patterns.push_back(g.patterns);
        }

        // ---- assemble
        BlobHeader h;
        memset(&h, 0, sizeof h);
        h.magic = kBlobMagic;
        h.version = kBlobVersion;
        h.n_patterns = (uint32_t)P;
        h.n_units = (uint32_t)units.size();
        h.n_secondaries = (uint32_t)sdesc.size();
        h.stack_unit = stack_unit;
        h.filter_log2_bits = log2_bits;
        h.filter_log2_bits2 = fb.log2_bits2;
        h.filter_fold = fold;
        h.filter_granule = (uint32_t)granule;
        h.gram_log2_cap = log2_cap;
        h.n_grams = n_grams;
        h.n_probes = n_probes;
        h.n_groups = (uint32_t)groups.size();
        std::vector<uint8_t> &blob = out.blob;
        blob.assign(align16(sizeof h), 0);
        auto place = [&](const void *p, size_t nbytes) -> uint64_t {
            size_t at = align16(blob.size());
            blob.resize(at + nbytes, 0);
            if (nbytes) memcpy(blob.data() + at, p, nbytes);
            blob.resize(align16(blob.size()), 0);
            return at;
        };
        h.off_bitmap = place(bitmap.data(), bitmap.size() * 4);
        h.off_bitmap2 = place(fb.bitmap2.data(), fb.bitmap2.size() * 4);
        h.off_gram_keys = place(gkeys.data(), gkeys.size() * 4);
        h.off_gram_first = place(gfirst.data(), gfirst.size() * 4);
        h.off_gram_count = place(gcount.data(), gcount.size() * 4);
        h.off_probes = place(probes_sec.bytes.data(), probes_sec.bytes.size());
        h.off_litpool = place(litpool.bytes.data(), litpool.bytes.size());
        h.off_units = place(units_sec.bytes.data(), units_sec.bytes.size());
        h.off_classmaps = place(classmaps.bytes.data(), classmaps.bytes.size());
        h.off_tables = place(tables.bytes.data(), tables.bytes.size());
        h.off_patterns = place(pdesc.data(), pdesc.size() * sizeof(PatternDesc));
        h.off_secondaries = place(sdesc.data(), sdesc.size() * sizeof(SecDesc));
        h.off_groups = place(groups_sec.bytes.data(), groups_sec.bytes.size());
        h.off_group_accept = place(gaccept.bytes.data(), gaccept.bytes.size());
        h.total_bytes = blob.size();
        memcpy(blob.data(), &h, sizeof h);

        rep.status = PM_OK;
        rep.n_patterns = P;
        rep.n_regex_units = (int)units.size();
        rep.n_filtered = P - (int)tier2.size();
        rep.n_dfa_scanned = (int)tier2.size();
        rep.n_grams = (int)n_grams;
        rep.filter_log2_bits = (int)log2_bits;
        rep.filter_fold_case = fold;
        rep.dfa_groups = (int)groups.size();
        for (auto &g : groups) rep.dfa_states += g.dfa.nstates, rep.dfa_classes = std::max(rep.dfa_classes, g.dfa.ncls);
        rep.blob_bytes = (int64_t)blob.size();
        snprintf(rep.message, sizeof rep.message, "ok");
        return PM_OK;
    } catch (const CompileFail &f) {
        rep.status = f.status;
        rep.pattern_index = f.pattern;
        rep.secondary_index = f.secondary;
        rep.regex_offset = f.offset;
        snprintf(rep.pattern_id, sizeof rep.pattern_id, "%s", f.id.c_str());
        snprintf(rep.message, sizeof rep.message, "%s", f.msg.c_str());
        out = CompiledLibrary();
        return f.status;
    }
}

}  // namespace pm

// ---- compiled library <-> bytes ------------------------------------
// The operator's sync step compiles once and leaves the result beside the YAML; the service loads
// it instead of compiling again (20 000 patterns: ~10 s of subset construction against milliseconds
// of reading).  What travels: the device blob, the per-pattern metadata the marshaller prints, and
// the required literals (the runtime rebuilds the gram filter from them when it has measured gram
// frequencies).  The automata kept for the compiler's unit tests do not travel.
namespace pm {
namespace {

constexpr uint32_t kSavedMagic = 0x424C4D50u, kSavedFormat = 1;   // "PMLB"

struct Writer {
    std::vector<uint8_t> &o;
    void raw(const void *p, size_t n) {
        const uint8_t *b = static_cast<const uint8_t *>(p);
        o.insert(o.end(), b, b + n);
    }
    void u32(uint32_t v) { raw(&v, 4); }
    void u64(uint64_t v) { raw(&v, 8); }
    void str(const std::string &s) {
        u32((uint32_t)s.size());
        raw(s.data(), s.size());
    }
    void bytes(const std::vector<uint8_t> &v) {
        u32((uint32_t)v.size());
        raw(v.data(), v.size());
    }
};

struct Reader {
    const uint8_t *p;
    size_t n, at = 0;
    bool ok = true;
    bool raw(void *dst, size_t k) {
        if (!ok || k > n - at) return ok = false;
        memcpy(dst, p + at, k);
        at += k;
        return true;
    }
    uint32_t u32() {
        uint32_t v = 0;
        raw(&v, 4);
        return v;
    }
    uint64_t u64() {
        uint64_t v = 0;
        raw(&v, 8);
        return v;
    }
    std::string str() {
        const uint32_t k = u32();
        if (!ok || k > n - at) {
            ok = false;
            return std::string();
        }
        std::string s(reinterpret_cast<const char *>(p + at), k);
        at += k;
        return s;
    }
    std::vector<uint8_t> bytes() {
        const uint32_t k = u32();
        if (!ok || k > n - at) {
            ok = false;
            return {};
        }
        std::vector<uint8_t> v(p + at, p + at + k);
        at += k;
        return v;
    }
};

uint64_t fnv1a64(const uint8_t *p, size_t n) {
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < n; i++) h = (h ^ p[i]) * 1099511628211ull;
    return h;
}

}  // namespace

void serialize_library(const CompiledLibrary &lib, std::vector<uint8_t> &out) {
    out.clear();
    Writer w{out};
    w.u32(kSavedMagic);
    w.u32(kSavedFormat);
    w.u32(kBlobVersion);
    w.u32(lib.fold ? 1u : 0u);
    w.u64(lib.blob.size());
    w.raw(lib.blob.data(), lib.blob.size());
    w.u32((uint32_t)lib.patterns.size());
    for (const PatternInfo &p : lib.patterns) {
        w.str(p.id);
        w.str(p.name);
        w.str(p.severity);
        w.str(p.category);
        w.str(p.regex);
        w.str(p.meta_json);
        w.u32((uint32_t)p.tier);
    }
    w.u32((uint32_t)lib.literals.size());
    for (const auto &lits : lib.literals) {
        w.u32((uint32_t)lits.size());
        for (const Literal &l : lits) {
            w.bytes(l.bytes);
            w.bytes(l.mask);
        }
    }
    w.u64(fnv1a64(out.data(), out.size()));
}

int deserialize_library(const uint8_t *data, size_t n, CompiledLibrary &lib, std::string &err) {
    lib = CompiledLibrary();
    if (n < 40) {
        err = "not a compiled pattern library (too short)";
        return PM_E_INVALID;
    }
    uint64_t want = 0;
    memcpy(&want, data + n - 8, 8);
    if (fnv1a64(data, n - 8) != want) {
        err = "compiled pattern library is damaged (checksum)";
        return PM_E_INVALID;
    }
    Reader r{data, n - 8};
    if (r.u32() != kSavedMagic || r.u32() != kSavedFormat) {
        err = "not a compiled pattern library (magic/format)";
        return PM_E_INVALID;
    }
    if (r.u32() != kBlobVersion) {
        err = "compiled pattern library was built by another version of the compiler: compile it again";
        return PM_E_UNSUPPORTED;
    }
    lib.fold = r.u32() != 0;
    const uint64_t blob_len = r.u64();
    if (!r.ok || blob_len > r.n - r.at || blob_len < sizeof(BlobHeader)) {
        err = "compiled pattern library is truncated";
        return PM_E_INVALID;
    }
    lib.blob.assign(data + r.at, data + r.at + blob_len);
    r.at += blob_len;
    const BlobHeader *h = lib.header();
    if (h->version != kBlobVersion) {
        err = "device blob has another layout version";
        return PM_E_UNSUPPORTED;
    }
    lib.granule = h->filter_granule == 2 ? 2 : 4;
    const uint32_t np = r.u32();
    for (uint32_t i = 0; i < np && r.ok; i++) {
        PatternInfo p;
        p.id = r.str();
        p.name = r.str();
        p.severity = r.str();
        p.category = r.str();
        p.regex = r.str();
        p.meta_json = r.str();
        p.tier = (int)r.u32();
        lib.patterns.push_back(std::move(p));
    }
    const uint32_t nl = r.u32();
    for (uint32_t i = 0; i < nl && r.ok; i++) {
        std::vector<Literal> lits(r.u32());
        for (Literal &l : lits) {
            l.bytes = r.bytes();
            l.mask = r.bytes();
            if (!r.ok) break;
        }
        lib.literals.push_back(std::move(lits));
    }
    if (!r.ok || r.at != r.n || lib.patterns.size() != h->n_patterns) {
        err = "compiled pattern library is inconsistent";
        lib = CompiledLibrary();
        return PM_E_INVALID;
    }
    return PM_OK;
}

}  // namespace pm
