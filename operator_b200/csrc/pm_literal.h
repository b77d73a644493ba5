// Required-literal analysis of a regex AST.  Host-only.
//
// A *required set* is a set of byte strings such that every match of the regex contains at
// least one of them.  The scan kernel's gram filter (pm_kernels.cu, k_scan) is sound only
// because of that property: a line that matches must contain one of the strings, so one of the
// string's sampled 4-byte grams must show up at a 4-aligned text offset.
#pragma once
#include <cstdint>
#include <vector>

#include "pm_regex.h"

namespace pm {

struct Literal {
    std::vector<uint8_t> bytes;   // case-insensitive positions are stored lower-case
    std::vector<uint8_t> mask;    // 0x20 where the position is a case-insensitive ASCII letter
    bool operator<(const Literal &o) const { return bytes != o.bytes ? bytes < o.bytes : mask < o.mask; }
    bool operator==(const Literal &o) const { return bytes == o.bytes && mask == o.mask; }
};

// Best required set of `ast` (empty when none exists).  Quality = shortest member length,
// returned in *min_len.
std::vector<Literal> required_literals(const Node &ast, int *min_len);

// True when `chosen` (the set required_literals returned for `ast`) is EVERY string the regex can match and
// the regex asserts nothing about the text around the match: an ASCII literal, an alternation of literals,
// "(?i)..." over such.  An occurrence of any member in a line is then a match of the line (Matcher.find()),
// and the confirm stage's literal comparison needs no DFA walk behind it (UnitDesc::literal_is_match).
bool literal_hit_is_match(const Node &ast, const std::vector<Literal> &chosen);

// Estimated commonness of a byte in log text (static table), -log2 probability scaled.
int byte_rarity(uint8_t b);

}  // namespace pm
