#include "pm_split.h"

#include <cstring>

namespace pm {

size_t effective_len(const uint8_t *p, size_t n) {
    while (n > 0 && p[n - 1] == '\n') {
        n--;
        if (n > 0 && p[n - 1] == '\r') n--;
    }
    return n;
}

void split_log(const uint8_t *s, size_t len, int pieces, int H, std::vector<SplitUnit> &out) {
    const size_t n = effective_len(s, len);
    std::vector<size_t> cuts{0};
    for (int k = 1; k < pieces; k++) {
        size_t pos = n / (size_t)pieces * (size_t)k;
        if (pos <= cuts.back()) pos = cuts.back();
        const void *r = pos < n ? memchr(s + pos, '\n', n - pos) : nullptr;
        if (!r) break;
        size_t at = (size_t)((const uint8_t *)r - s) + 1;
        if (at < n && at > cuts.back()) cuts.push_back(at);
    }
    cuts.push_back(len);
    const int np = (int)cuts.size() - 1;
    for (int k = 0; k < np; k++) {
        const size_t core_begin = cuts[k], core_end = cuts[k + 1];
        size_t begin = core_begin;
        int lead = 0;
        while (lead < H && begin > 0) {   // begin is a line start: step to the start of the line before
            const void *r = begin >= 2 ? memrchr(s, '\n', begin - 1) : nullptr;
            begin = r ? (size_t)((const uint8_t *)r - s) + 1 : 0;
            lead++;
        }
        size_t end = core_end;
        int trail = 0;
        if (k != np - 1) {
            while (end < n) {
                const void *r = memchr(s + end, '\n', n - end);
                const size_t line_end = r ? (size_t)((const uint8_t *)r - s) : n;
                const bool empty = line_end == end || (line_end == end + 1 && s[end] == '\r');
                trail++;
                if (!r) {   // the log's last line
                    end = len;
                    break;
                }
                end = line_end + 1;
                if (trail >= H && !empty) break;
            }
            if (end >= n) end = len;
        }
        out.push_back(SplitUnit{begin, end, lead, trail});
    }
}

}  // namespace pm
