// Pattern-library compiler: library JSON (SPEC.md section 1) -> device blob (pm_blob.h).
// This is row a9 of SURVEY.md section 8: the step that hooks into
// PatternSyncService.validatePatterns (PatternSyncService.java:228-253), which today only
// counts YAML files.  Host-only.
#pragma once
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

struct CompiledLibrary {
    std::vector<uint8_t> blob;
    std::vector<PatternInfo> patterns;
    // kept for the compiler's unit tests (tests/support): the automata behind the blob
    std::vector<Dfa> unit_dfas;                      // index = unit id
    std::vector<std::vector<Literal>> literals;      // per pattern (empty for tier 2)
    std::vector<Dfa> group_dfas;
    std::vector<std::vector<int>> group_patterns;    // unit id in a group -> pattern index
    const BlobHeader *header() const { return reinterpret_cast<const BlobHeader *>(blob.data()); }
};

// Returns PM_OK or a negative pm_status; `rep` always describes the outcome.
int compile_library(const char *spec, size_t n, CompiledLibrary &out, pm_compile_report &rep);

}  // namespace pm
