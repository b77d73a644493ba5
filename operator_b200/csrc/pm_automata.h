// Regex AST -> byte-level NFA (UTF-8) -> search DFA with per-line semantics.  Host-only.
//
// The DFAs answer "does Pattern.compile(rx).matcher(line).find() hold" (SPEC.md section 3) for
// one line at a time: the device feeds the line's bytes, then one END-OF-LINE symbol.  Zero-
// width assertions (^ \z \b \B) are resolved inside the subset construction with one symbol of
// look-ahead, so the tables the kernels walk are plain next-state tables.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "pm_regex.h"

namespace pm {

enum AssertKind : uint8_t { A_NONE = 0, A_BOL, A_END, A_WORDB, A_NWORDB };

struct NfaState {
    struct Tr {
        uint8_t lo, hi;
        int to;
    };
    struct Eps {
        int to;
        uint8_t cond;
    };
    std::vector<Tr> tr;
    std::vector<Eps> eps;
    int match = -1;   // unit id when accepting
};

struct Nfa {
    std::vector<NfaState> st;
    std::vector<int> starts;   // one entry state per unit (all are re-armed at every position)
    bool uses_word_boundary = false;

    // Adds one regex as unit `unit`.  Returns false when the state budget is exceeded.
    bool add(const Node &ast, int unit, int max_states);
};

struct Dfa {
    int nstates = 0;
    int ncls = 0;                 // byte classes; column `ncls` is END-OF-LINE
    int first_accept = 0;         // states >= first_accept carry a non-empty accept list
    uint8_t classmap[256] = {0};
    std::vector<uint16_t> next;   // nstates x (ncls + 1); state 0 is the start-of-line state
    // accept lists of states first_accept.. (unit ids, sorted); with early_accept there is one
    // sticky accepting state and its list is the unit(s) that can no longer be lost.
    std::vector<std::vector<int>> accept;

    int stride() const { return ncls + 1; }
};

// Subset construction.  early_accept: the first acceptance collapses into one sticky state
// (single-pattern verification).  Returns false if more than max_states states are needed.
bool build_search_dfa(const Nfa &nfa, bool early_accept, int max_states, Dfa &out);

// Host-side table walk, used by the compiler's self-check and by the unit tests.  NOT a scan
// path: the product only ever walks these tables on the device.
bool dfa_line_accepts(const Dfa &d, const uint8_t *line, size_t n, std::vector<int> *units = nullptr);

}  // namespace pm
