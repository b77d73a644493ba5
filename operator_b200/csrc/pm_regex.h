// Java-regex subset parser (SPEC.md section 3).  Host-only.
//
// The reference delegates matching to java.util.regex inside the out-of-tree log-parser
// service (call site: LogParserRestClient.java:37-39); the grammar accepted here is the subset
// of that engine this build compiles to DFAs, and everything else is rejected with an error
// that names the byte offset (north_star: "rejects unsupported constructs with an error").
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <utility>
#include <vector>

namespace pm {

struct RegexError {
    int status = 0;        // pm_status value
    int offset = -1;       // byte offset in the regex source
    std::string message;
};

using CpRange = std::pair<uint32_t, uint32_t>;   // inclusive code point range
constexpr uint32_t kMaxCp = 0x10FFFF;
constexpr int kMaxRepeat = 256;
constexpr int kInf = 1 << 30;

enum class NodeKind : uint8_t { Lit, Set, Dot, Cat, Alt, Rep, Bol, EolJava, End, WordB, NotWordB, Empty };

struct Node {
    NodeKind kind = NodeKind::Empty;
    uint32_t cp = 0;                 // Lit
    bool ci = false;                 // Lit: ASCII case-insensitive letter
    bool dotall = false;             // Dot
    std::vector<CpRange> ranges;     // Set: final (negation and case folding applied), sorted
    std::vector<std::unique_ptr<Node>> kids;
    int lo = 0, hi = -1;             // Rep: hi < 0 means unbounded
    bool lazy = false;
    int pos = 0;
};
using NodePtr = std::unique_ptr<Node>;

// Parses `src` (UTF-8).  On failure returns nullptr and fills `err`.
NodePtr parse_regex(const std::string &src, RegexError &err);

// (min, max) match width in code points; max == kInf when unbounded.
std::pair<int, int> node_width(const Node &n);

std::vector<CpRange> normalize_ranges(std::vector<CpRange> r);
std::vector<CpRange> negate_ranges(const std::vector<CpRange> &r);

}  // namespace pm
