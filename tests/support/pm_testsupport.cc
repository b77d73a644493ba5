// TEST SUPPORT ONLY -- not part of libpodmortem_scan.so and never loaded by the product.
//
// Exposes the pattern compiler's intermediate results (verify DFAs, combined DFAs, literals,
// gram filter) to the CPU test-suite so that compiled tables can be diffed against the oracle
// without a GPU.  The table walks below are a unit-test harness for host code, not a scan path.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../operator_b200/csrc/pm_json.h"
#include "../../operator_b200/csrc/pm_library.h"
#include "../../operator_b200/csrc/pm_split.h"

using namespace pm;

namespace pm {
bool dfa_line_accepts(const Dfa &d, const uint8_t *line, size_t n, std::vector<int> *units) {
    uint32_t s = 0;
    const int stride = d.stride();
    bool any = false;
    auto visit = [&](uint32_t st) {
        if ((int)st >= d.first_accept) {
            any = true;
            if (units)
                for (int u : d.accept[st - d.first_accept]) units->push_back(u);
        }
    };
    for (size_t i = 0; i < n; i++) {
        s = d.next[(size_t)s * stride + d.classmap[line[i]]];
        visit(s);
    }
    s = d.next[(size_t)s * stride + d.ncls];
    visit(s);
    return any;
}
}  // namespace pm

extern "C" {

struct pmt_lib {
    CompiledLibrary lib;
    std::string scratch;
};

int pmt_compile(const char *spec, size_t n, pmt_lib **out, pm_compile_report *rep) {
    pmt_lib *l = new pmt_lib();
    int rc = compile_library(spec, n, l->lib, *rep);
    if (rc != PM_OK) {
        delete l;
        *out = nullptr;
        return rc;
    }
    *out = l;
    return PM_OK;
}

void pmt_free(pmt_lib *l) { delete l; }

int pmt_n_patterns(pmt_lib *l) { return (int)l->lib.patterns.size(); }
int pmt_pattern_tier(pmt_lib *l, int p) { return l->lib.patterns[p].tier; }

static const PatternDesc *pdesc(pmt_lib *l) {
    return reinterpret_cast<const PatternDesc *>(l->lib.blob.data() + l->lib.header()->off_patterns);
}
static const SecDesc *sdesc(pmt_lib *l) {
    return reinterpret_cast<const SecDesc *>(l->lib.blob.data() + l->lib.header()->off_secondaries);
}

int pmt_primary_match(pmt_lib *l, int p, const uint8_t *line, size_t n) {
    return dfa_line_accepts(l->lib.unit_dfas[pdesc(l)[p].unit], line, n, nullptr);
}
int pmt_secondary_match(pmt_lib *l, int p, int j, const uint8_t *line, size_t n) {
    return dfa_line_accepts(l->lib.unit_dfas[sdesc(l)[pdesc(l)[p].sec_first + j].unit], line, n, nullptr);
}
int pmt_stack_match(pmt_lib *l, const uint8_t *line, size_t n) {
    return dfa_line_accepts(l->lib.unit_dfas[l->lib.header()->stack_unit], line, n, nullptr);
}
// 1 when k_verify treats a confirmed literal of pattern p as the match itself (UnitDesc::literal_is_match)
int pmt_literal_is_match(pmt_lib *l, int p) {
    const UnitDesc *ud = reinterpret_cast<const UnitDesc *>(l->lib.blob.data() + l->lib.header()->off_units);
    return ud[pdesc(l)[p].unit].literal_is_match;
}
int pmt_unit_states(pmt_lib *l, int p) { return l->lib.unit_dfas[pdesc(l)[p].unit].nstates; }

// Combined DFAs: which tier-2 patterns does the line match?  Returns the count written.
int pmt_group_match(pmt_lib *l, const uint8_t *line, size_t n, int *out, int cap) {
    std::vector<int> pats;
    for (size_t g = 0; g < l->lib.group_dfas.size(); g++) {
        std::vector<int> units;
        dfa_line_accepts(l->lib.group_dfas[g], line, n, &units);
        for (int u : units) pats.push_back(l->lib.group_patterns[g][u]);
    }
    std::sort(pats.begin(), pats.end());
    pats.erase(std::unique(pats.begin(), pats.end()), pats.end());
    int k = 0;
    for (int p : pats)
        if (k < cap) out[k++] = p;
    return (int)pats.size();
}

// The same question answered by walking the DEVICE form of the combined DFAs (ScanTable: newline
// folded into the byte alphabet, premultiplied entries): bytes of the line, then '\n'.
int pmt_group_scan_match(pmt_lib *l, const uint8_t *line, size_t n, int *out, int cap) {
    std::vector<int> pats;
    for (size_t g = 0; g < l->lib.group_scan.size(); g++) {
        const ScanTable &t = l->lib.group_scan[g];
        uint32_t s = 0;
        auto feed = [&](uint8_t b) {
            s = t.tab[s + (t.wide ? b : t.cmap[b])];
            uint32_t id = s / t.stride;
            if ((int)id >= t.first_accept)
                for (int u : t.accept[id - t.first_accept]) pats.push_back(l->lib.group_patterns[g][u]);
        };
        for (size_t i = 0; i < n; i++) feed(line[i]);
        feed('\n');
    }
    std::sort(pats.begin(), pats.end());
    pats.erase(std::unique(pats.begin(), pats.end()), pats.end());
    int k = 0;
    for (int p : pats)
        if (k < cap) out[k++] = p;
    return (int)pats.size();
}
int pmt_group_is_wide(pmt_lib *l, int g) { return l->lib.group_scan[g].wide; }
int pmt_group_scan_states(pmt_lib *l, int g) { return l->lib.group_scan[g].nstates; }

// Emulates the gram filter + gram table + literal confirmation on a text whose first byte sits
// at alignment `phase` (0..3) of the device buffer.  Returns 1 if pattern p would be handed to
// its verify DFA for this text.
int pmt_filter_triggers(pmt_lib *l, int p, const uint8_t *text, size_t n, int phase) {
    const BlobHeader *h = l->lib.header();
    if (!h->filter_log2_bits) return 0;
    const uint8_t *b = l->lib.blob.data();
    const uint32_t *bitmap = (const uint32_t *)(b + h->off_bitmap);
    const uint32_t *gk = (const uint32_t *)(b + h->off_gram_keys);
    const uint32_t *gf = (const uint32_t *)(b + h->off_gram_first);
    const uint32_t *gc = (const uint32_t *)(b + h->off_gram_count);
    const Probe *pr = (const Probe *)(b + h->off_probes);
    const uint8_t *lp = b + h->off_litpool;
    // the kernel tests the word at every text offset that is a multiple of the filter's granule (4, or 2 for libraries
    // with 5-byte literals); `phase` is the offset of text[0] in that grid
    const size_t gran = h->filter_granule == 2 ? 2 : 4;
    size_t first = (gran - (size_t)phase % gran) % gran;
    for (size_t i = first; i + 4 <= n; i += gran) {
        uint32_t g;
        memcpy(&g, text + i, 4);
        if (h->filter_fold) g |= 0x20202020u;
        uint32_t bit = gram_filter_bit(g, h->filter_log2_bits);
        if (!((bitmap[bit >> 5] >> (bit & 31)) & 1)) continue;
        const uint32_t *bitmap2 = (const uint32_t *)(b + h->off_bitmap2);
        uint32_t bit2 = gram_filter_bit2(g, h->filter_log2_bits2);
        if (!((bitmap2[bit2 >> 5] >> (bit2 & 31)) & 1)) continue;
        uint32_t mask = (1u << h->gram_log2_cap) - 1, slot = gram_table_slot(g, h->gram_log2_cap);
        while (gk[slot] != kGramEmpty && gk[slot] != g) slot = (slot + 1) & mask;
        if (gk[slot] != g) continue;
        for (uint32_t k = gf[slot]; k < gf[slot] + gc[slot]; k++) {
            if ((int)pr[k].pattern != p) continue;
            if (i < pr[k].gram_off) continue;
            size_t st = i - pr[k].gram_off;
            if (st + pr[k].lit_len > n) continue;
            const uint8_t *lit = lp + pr[k].lit_off, *m = lit + ((pr[k].lit_len + 3u) & ~3u);
            bool eq = true;
            for (size_t j = 0; j < pr[k].lit_len && eq; j++) eq = ((text[st + j] | m[j]) == lit[j]);
            if (eq) return 1;
        }
    }
    return 0;
}

// JSON list of the required literals of pattern p (bytes shown as latin-1 escaped), for test diagnostics.
const char *pmt_literals(pmt_lib *l, int p) {
    std::string &s = l->scratch;
    s = "[";
    bool firstl = true;
    for (auto &lit : l->lib.literals[p]) {
        if (!firstl) s += ",";
        firstl = false;
        s += "\"";
        for (size_t i = 0; i < lit.bytes.size(); i++) {
            char buf[8];
            uint8_t c = lit.bytes[i];
            if (c >= 0x20 && c < 0x7F && c != '"' && c != '\\')
                s.push_back((char)c);
            else {
                snprintf(buf, sizeof buf, "\\\\x%02x", c);
                s += buf;
            }
        }
        s += "\"";
    }
    s += "]";
    return s.c_str();
}

}  // extern "C"

// Transition-table entries (states x columns) of a pattern's verify DFA.
extern "C" int pmt_unit_table_entries(pmt_lib *l, int p) {
    const pm::Dfa &d = l->lib.unit_dfas[pdesc(l)[p].unit];
    return d.nstates * d.stride();
}

// The streaming-ingest helpers of pm_parse: span of a top-level string member and its decoding.
// Returns the decoded length (out holds at least n bytes), -1 when the key is absent, -2 on a bad escape.
extern "C" long pmt_json_extract(const char *body, size_t n, const char *key, char *out) {
    size_t b = 0, e = 0;
    if (!pm::json_find_top_level_string(body, n, key, &b, &e)) return -1;
    size_t len = pm::json_decode_string(body + b, e - b, out);
    return len == (size_t)-1 ? -2 : (long)len;
}

// split_log (pm_split.cc): writes begin, end, lead, trail per unit into out[4*k ..]; returns the unit count.
extern "C" int pmt_split(const uint8_t *log, size_t n, int pieces, int halo, int64_t *out, int cap) {
    std::vector<pm::SplitUnit> u;
    pm::split_log(log, n, pieces, halo, u);
    for (size_t k = 0; k < u.size() && (int)k < cap; k++) {
        out[4 * k] = (int64_t)u[k].begin;
        out[4 * k + 1] = (int64_t)u[k].end;
        out[4 * k + 2] = u[k].lead;
        out[4 * k + 3] = u[k].trail;
    }
    return (int)u.size();
}

// serialize_library / deserialize_library (pm_library.cc).  pmt_saved: the bytes of a compiled library
// (valid until the next call); pmt_load_status: what loading `bytes` answers; pmt_roundtrip_equal: 1 when
// the loaded copy has the same blob, pattern metadata and literals.
static std::vector<uint8_t> g_saved;
extern "C" const uint8_t *pmt_saved(pmt_lib *l, size_t *n) {
    pm::serialize_library(l->lib, g_saved);
    *n = g_saved.size();
    return g_saved.data();
}
extern "C" int pmt_load_status(const uint8_t *bytes, size_t n) {
    pm::CompiledLibrary c;
    std::string err;
    return pm::deserialize_library(bytes, n, c, err);
}
extern "C" int pmt_roundtrip_equal(pmt_lib *l) {
    std::vector<uint8_t> bytes;
    pm::serialize_library(l->lib, bytes);
    pm::CompiledLibrary c;
    std::string err;
    if (pm::deserialize_library(bytes.data(), bytes.size(), c, err) != PM_OK) return 0;
    if (c.blob != l->lib.blob || c.fold != l->lib.fold || c.patterns.size() != l->lib.patterns.size() || c.literals.size() != l->lib.literals.size()) return 0;
    for (size_t i = 0; i < c.patterns.size(); i++) {
        const auto &a = c.patterns[i], &b = l->lib.patterns[i];
        if (a.id != b.id || a.name != b.name || a.severity != b.severity || a.category != b.category || a.regex != b.regex || a.meta_json != b.meta_json || a.tier != b.tier) return 0;
    }
    for (size_t i = 0; i < c.literals.size(); i++)
        if (!(c.literals[i] == l->lib.literals[i])) return 0;
    return 1;
}
