// Synthetic crash-log corpus generator, host side: bench and test INPUT
// only.  The text of a line is defined in pm_synth_core.h (shared with the device generator,
// pm_synth_gpu.cu); this file is the multi-threaded driver that fills a host buffer.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "pm_synth_core.h"

using namespace pm_synth;

namespace {

struct StringSink : SinkBase<StringSink> {
    std::string &s;
    void ch(char c) { s.push_back(c); }
    size_t size() const { return s.size(); }
};

}  // namespace

static int64_t corpus_host(const Params &p, uint64_t first_line, int64_t target_bytes, uint8_t *out, int64_t *n_lines_out, int threads);

extern "C" int64_t pm_synth_corpus(uint64_t seed, uint64_t first_line, int64_t target_bytes, double p_hit,
                                   const char *const *witnesses, const char *const *sec_witnesses, int n_wit, int crlf,
                                   uint8_t *out, int64_t *n_lines_out, int threads) {
    if (!out || target_bytes < 0) return -1;
    Params p{seed, p_hit, witnesses, sec_witnesses, n_wit, crlf != 0};
    return corpus_host(p, first_line, target_bytes, out, n_lines_out, threads);
}

// The same with near-miss lines (pm_synth_core.h, Params::p_near).
extern "C" int64_t pm_synth_corpus_near(uint64_t seed, uint64_t first_line, int64_t target_bytes, double p_hit,
                                        const char *const *witnesses, const char *const *sec_witnesses, int n_wit, int crlf,
                                        double p_near, const char *const *near, int n_near,
                                        uint8_t *out, int64_t *n_lines_out, int threads) {
    if (!out || target_bytes < 0) return -1;
    Params p{seed, p_hit, witnesses, sec_witnesses, n_wit, crlf != 0};
    p.p_near = p_near, p.near = near, p.n_near = n_near;
    return corpus_host(p, first_line, target_bytes, out, n_lines_out, threads);
}

static int64_t corpus_host(const Params &p, uint64_t first_line, int64_t target_bytes, uint8_t *out, int64_t *n_lines_out, int threads) {
    if (threads <= 0) threads = (int)std::max(1u, std::thread::hardware_concurrency());
    const uint64_t chunk_lines = 32768;
    int64_t written = 0, lines = 0;
    uint64_t next_line = first_line;
    bool done = false;
    while (!done) {
        std::vector<std::string> bufs(threads);
        std::vector<std::vector<uint32_t>> ends(threads);
        std::vector<std::thread> th;
        for (int t = 0; t < threads; t++) {
            th.emplace_back([&, t]() {
                uint64_t a = next_line + (uint64_t)t * chunk_lines;
                std::string &s = bufs[t];
                s.reserve(chunk_lines * 140);
                ends[t].reserve(chunk_lines);
                for (uint64_t i = a; i < a + chunk_lines; i++) {
                    StringSink o{{}, s};
                    gen_line(p, i, o);
                    if (p.crlf) s.push_back('\r');
                    s.push_back('\n');
                    ends[t].push_back((uint32_t)s.size());
                }
            });
        }
        for (auto &x : th) x.join();
        for (int t = 0; t < threads && !done; t++) {
            int64_t room = target_bytes - written;
            if ((int64_t)bufs[t].size() <= room) {
                memcpy(out + written, bufs[t].data(), bufs[t].size());
                written += (int64_t)bufs[t].size();
                lines += (int64_t)ends[t].size();
            } else {
                // last whole line that fits
                auto it = std::upper_bound(ends[t].begin(), ends[t].end(), (uint32_t)std::max<int64_t>(room, 0));
                size_t k = (size_t)(it - ends[t].begin());
                if (k > 0) {
                    memcpy(out + written, bufs[t].data(), ends[t][k - 1]);
                    written += ends[t][k - 1];
                    lines += (int64_t)k;
                }
                done = true;
            }
        }
        next_line += (uint64_t)threads * chunk_lines;
    }
    if (n_lines_out) *n_lines_out = lines;
    return written;
}
