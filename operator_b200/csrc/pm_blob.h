// Layout of a compiled pattern library as the device sees it: one contiguous blob, copied to
// each GPU by pm_library_publish.  Shared by the host compiler and the kernels, POD only.
//
// Sections (all offsets are bytes from the start of the blob, 16-byte aligned):
//   bitmap      gram filter, 2^filter_log2_bits bits, staged into shared memory by k_scan
//   gram_keys   open-addressing table gram -> probe run (exact check of a filter hit)
//   probes      (pattern, literal, offset of the gram inside the literal)
//   litpool     literal bytes followed by their case masks (both padded to whole 32-bit words with 0xFF)
//   units       one verify-DFA descriptor per regex unit (primaries, secondaries, STACK)
//   classmaps   256-byte byte->class maps
//   tables      u16 next-state tables
//   patterns    scoring/context parameters per primary pattern
//   secondaries weight/window/unit per secondary pattern
//   groups      combined shared-memory DFAs for primaries that have no usable literal
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define PM_HD __host__ __device__ __forceinline__
#else
#define PM_HD inline
#endif

namespace pm {

constexpr uint32_t kBlobMagic = 0x424C4D50u;   // "PMLB"
constexpr uint32_t kBlobVersion = 10;
constexpr int kGramLen = 4;        // bytes per gram = one aligned 32-bit text word
constexpr int kMinLiteral = 7;     // any 7-byte window covers one aligned 4-byte word
// Granule of the gram filter: the text offsets whose 4-byte word k_scan tests are the multiples of it.
//   4: aligned words, one lookup per 4 bytes; a literal needs a 7-byte window (4 grams, one per alignment);
//   2: also the words that straddle two aligned ones, one lookup per 2 bytes; a 5-byte window is enough (2 grams).
// A library compiles to granule 2 exactly when that takes every pattern off the combined DFA (pm_library.cc).
constexpr int kMinLiteralHalf = 5;
inline int window_bytes(int granule) { return kGramLen + granule - 1; }
constexpr uint32_t kGramEmpty = 0xFFFFFFFFu;   // 0xFF never occurs in valid UTF-8
constexpr int kStackMax = 50;      // SPEC.md section 6
constexpr int kMaxSecondary = 8;

struct BlobHeader {
    uint32_t magic, version;
    uint32_t n_patterns, n_units, n_secondaries, stack_unit;
    uint32_t filter_log2_bits;   // 0: no pattern uses the gram filter
    uint32_t filter_log2_bits2;  // second, smaller bitmap (independent hash) checked on a first-level hit
    uint32_t filter_fold;        // 1: text words are OR-ed with 0x20202020 before hashing
    uint32_t gram_log2_cap, n_grams, n_probes, n_groups;
    uint64_t off_bitmap2;
    uint64_t off_bitmap, off_gram_keys, off_gram_first, off_gram_count, off_probes, off_litpool;
    uint64_t off_units, off_classmaps, off_tables, off_patterns, off_secondaries, off_groups;
    uint64_t off_group_accept;   // u32 pool: per-group accept index + lists
    uint64_t total_bytes;
    uint32_t filter_granule;     // 4 or 2 (see kMinLiteralHalf)
    uint32_t pad_;
};

struct Probe {
    uint32_t pattern;
    uint32_t lit_off;    // into litpool, 4-byte aligned: bytes[lit_len], 0xFF to a whole word, mask[lit_len], 0xFF to a whole word
    uint16_t lit_len;
    uint16_t gram_off;   // offset of this gram inside the literal
};

struct UnitDesc {
    uint32_t table_off;      // in u16 elements from off_tables
    uint32_t classmap_off;   // in bytes from off_classmaps
    uint16_t nstates, stride, accept_state;
    uint16_t mid_state;      // state after one byte that belongs to no pattern (speculative walk entry guess)
    uint16_t dead_state;     // state from which acceptance is unreachable, 0xFFFF if none (anchored regexes)
    uint16_t literal_is_match;   // 1: the pattern's required literals are all it can match (pm_literal.h): k_verify skips the walk
    uint32_t wide_off;       // u16 elements from off_tables: 256-column copy (index = state*256 + byte) for the
                             // units the resolve stage walks most (STACK, secondaries); 0xFFFFFFFF if absent
};

struct PatternDesc {
    double confidence, mult;
    uint32_t sec_first, sec_count;
    int32_t lines_before, lines_after;
    uint32_t include_stack, severity;   // severity rank 0..4 = INFO..CRITICAL
    uint32_t unit, tier;                // tier 1: gram filter, tier 2: combined DFA
};

struct SecDesc {
    double weight;
    int32_t window;
    uint32_t unit;
};

// Combined scan DFA of one group, in the form k_dfa_scan walks: END-OF-LINE is folded into the
// '\n' byte (its targets are "terminal twins" whose rows equal the start-of-line row, so the
// kernel needs no newline branch), entries are premultiplied by the row stride.
//   wide:       256 columns, index = state*256 + byte                (one lookup per byte)
//   compressed: stride columns, index = state*stride + classmap[byte]
struct GroupDesc {
    uint32_t table_off;      // u16 elements from off_tables
    uint32_t classmap_off;
    uint32_t accept_off;     // u32 elements from off_group_accept: (nstates-first_accept+1) starts, then lists
    uint16_t nstates, stride, first_accept, mid_state;
    uint32_t wide, pad;
};

constexpr uint32_t kFilterMult0 = 0x9E3779B1u;   // default multiplier; the runtime may reseed (pm_runtime.cu)
// First-level filter bit of a gram: the bitmap WORD comes from the top (log2_bits - 5) bits of the
// multiplicative hash, the bit inside the word from the hash's low 5 bits (k_scan tests it with
// one wrapping shift, no extra instruction to extract the bit index).
PM_HD uint32_t gram_filter_bit(uint32_t gram, uint32_t log2_bits, uint32_t mult = kFilterMult0) {
    const uint32_t h = gram * mult;
    return ((h >> (37u - log2_bits)) << 5) | (h & 31u);
}
PM_HD uint32_t gram_filter_bit2(uint32_t gram, uint32_t log2_bits2) { return (gram * 0xC2B2AE3Du) >> (32u - log2_bits2); }
PM_HD uint32_t gram_table_slot(uint32_t gram, uint32_t log2_cap) { return (gram * 0x85EBCA6Bu) >> (32u - log2_cap); }

}  // namespace pm
