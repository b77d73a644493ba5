// Pattern-library compiler: library JSON (SPEC.md section 1) -> device blob (pm_blob.h).
// This is row a9 of SURVEY.md section 8: the step that hooks into
// PatternSyncService.validatePatterns (PatternSyncService.java:228-253), which today only
// counts YAML files.  Host-only.
#pragma once
#include <functional>
#include <string>
#include <vector>

#include "pm_automata.h"
#include "pm_blob.h"
#include "pm_literal.h"
#include "podmortem_scan.h"

namespace pm {

struct PatternInfo {
    std::string id, name, severity, category;
    std::string regex;
    std::string meta_json;   // {"id":..,"name":..,"severity":..,"category":..,"remediation":..}
    int tier = 0;
};

// The gram filter of a library (bitmap + exact gram table + probes + literal pool), built from
// the required literals of the tier-1 patterns.  Separate from the rest of the blob because the
// runtime rebuilds it when it has measured which grams are frequent in the actual logs.
struct FilterBuild {
    uint32_t log2_bits = 0, log2_bits2 = 0, log2_cap = 4, n_grams = 0, n_probes = 0;
    int granule = 4;
    bool fold = false;
    std::vector<uint32_t> bitmap, bitmap2, gkeys, gfirst, gcount;
    std::vector<Probe> probes;
    std::vector<uint8_t> litpool;
};
using GramCostFn = std::function<double(uint32_t gram)>;   // expected hits per aligned text word

// fold_mode: -1 decide (fold iff a chosen window has a case-insensitive position), 0/1 fixed.
// cost == nullptr: static byte-rarity model.  granule: 4 (7-byte windows, 4 grams each) or 2 (5-byte windows, 2 grams).
void build_filter(const std::vector<std::vector<Literal>> &literals, int fold_mode, const GramCostFn *cost, int granule, FilterBuild &out);
// Every distinct 4-byte gram of every literal (folded when `fold`), sorted.
void collect_all_grams(const std::vector<std::vector<Literal>> &literals, bool fold, std::vector<uint32_t> &grams);

// Device form of a combined DFA (see GroupDesc in pm_blob.h).
struct ScanTable {
    bool wide = false;
    int nstates = 0, first_accept = 0, mid_state = 0, stride = 0;
    std::vector<uint16_t> tab;               // premultiplied next-state entries
    uint8_t cmap[256] = {0};
    std::vector<std::vector<int>> accept;    // unit lists of states >= first_accept
};
// Returns false if the table does not fit max_bytes in either form.
bool build_scan_table(const Dfa &d, size_t max_bytes, ScanTable &out);

struct CompiledLibrary {
    std::vector<uint8_t> blob;
    std::vector<PatternInfo> patterns;
    // kept for the compiler's unit tests (tests/support): the automata behind the blob
    std::vector<Dfa> unit_dfas;                      // index = unit id
    std::vector<std::vector<Literal>> literals;      // per pattern (empty for tier 2)
    bool fold = false;
    int granule = 4;                                 // of the gram filter (pm_blob.h)
    std::vector<Dfa> group_dfas;
    std::vector<ScanTable> group_scan;
    std::vector<std::vector<int>> group_patterns;    // unit id in a group -> pattern index
    const BlobHeader *header() const { return reinterpret_cast<const BlobHeader *>(blob.data()); }
};

// Returns PM_OK or a negative pm_status; `rep` always describes the outcome.
int compile_library(const char *spec, size_t n, CompiledLibrary &out, pm_compile_report &rep);

// Compiled library <-> bytes (what the sync hook leaves beside the YAML; pm_library.cc).
void serialize_library(const CompiledLibrary &lib, std::vector<uint8_t> &out);
int deserialize_library(const uint8_t *data, size_t n, CompiledLibrary &lib, std::string &err);

}  // namespace pm
