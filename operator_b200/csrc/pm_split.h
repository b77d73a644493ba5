// Cutting one large log into units for several devices.  Host-only, no CUDA:
// also linked into the CPU test harness.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace pm {

// Length without trailing "\r?\n" separators (Java's split drops trailing empty strings).
size_t effective_len(const uint8_t *p, size_t n);

struct SplitUnit {
    size_t begin, end;   // bytes [begin, end) of the log are staged as this unit's text
    int lead, trail;     // lines of halo in front of / behind the lines this unit reports
};

// A log is cut at line starts into up to `pieces` units.  Every unit carries a halo of `halo` lines in
// front of and behind the lines it reports, so that context windows, STACK runs and secondary
// proximity windows of its own lines see exactly the text they would see in the whole log; events on
// halo lines are dropped at gather time (they belong to the neighbour).  A trailing halo ends on a
// non-empty line (or at the end of the log): the line count of a unit then follows from the device's
// count without Java's "trailing empty strings are dropped" interfering.  Lines reported by unit k:
// (lines of the unit's text) - lead - trail, in log order.
void split_log(const uint8_t *log, size_t len, int pieces, int halo, std::vector<SplitUnit> &out);

}  // namespace pm
