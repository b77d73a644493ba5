// Launch wrappers of the sm_100a kernels (pm_kernels.cu).  All launches are asynchronous on the
// given stream and return the CUDA error of the launch; none of them allocates.
//
// Device data layout of one staged batch (see DESIGN.md "Data layout in HBM"):
//   text        [16 x '\n'] pod0 bytes, '\n' padding to the next 16-byte boundary (>= 1 byte),
//               pod1 ..., '\n' padding up to a multiple of kTileBytes, one more tile of '\n'.
//               Every line of every pod is therefore delimited by '\n' on both sides.
//   tile_nl     u32 per 2 KiB tile: number of '\n' bytes in the tile          (k_scan)
//   tile_prefix u64 per tile: '\n' bytes before the tile                      (k_prefix_*)
//   cand        u64 per filter hit: (16-byte group index << 8) | mask of the even offsets whose word hit   (k_scan)
//   raw events  u64 key = (line_start << 24) | pattern                        (k_confirm, k_dfa_scan)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "pm_blob.h"

namespace pm {

// PM_DEBUG_TIMING: the runtime points this at a function that records a named CUDA event on the
// batch's main stream; the multi-kernel launch wrappers call it between their kernels.
extern thread_local void (*g_stage_mark)(void *arg, const char *name);
extern thread_local void *g_stage_mark_arg;
inline void stage_mark(const char *name) {
    if (g_stage_mark) g_stage_mark(g_stage_mark_arg, name);
}

constexpr int kTileBytes = 2048;        // one warp iteration of k_scan: 4 x (32 lanes x 16 B)
constexpr int kScanThreads = 1024;      // one persistent CTA per SM
constexpr int kHeaderBytes = 16;
constexpr int kPatternBits = 24;        // raw event key = line_start << 24 | pattern
constexpr int kMaxCtxLines = 1000;      // lines_before / lines_after / proximity_window limit

// The gram filter in use on one device: device pointers into the blob as compiled, or into a
// separately uploaded rebuild (profile-guided re-selection, pm_runtime.cu).
struct FilterView {
    const uint32_t *bitmap = nullptr, *bitmap2 = nullptr, *gkeys = nullptr, *gfirst = nullptr, *gcount = nullptr;
    const Probe *probes = nullptr;
    const uint8_t *litpool = nullptr;
    uint32_t log2_bits = 0, log2_bits2 = 0, log2_cap = 0, fold = 0, n_grams = 0;
    uint32_t granule = 4;   // 4: k_scan tests the aligned words; 2: every even offset (pm_blob.h)
};

// The bytes a line matching the built-in STACK regex can start with (bit b of the 256-bit set), derived from the unit's
// DFA at publish time: k_plan looks at one byte before it collects the lines of a possible stack trace.
struct StackFirst {
    uint32_t bits[8];
};

struct DeviceLibrary {
    const uint8_t *blob = nullptr;      // device copy of the compiled blob
    FilterView fv;
    int ctx_slots = 0;                  // newline slots k_plan needs per direction (largest window of the library + 2)
    StackFirst stack_first = {{0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}};
    uint32_t verify_table_cap = 0;      // bytes of shared memory k_verify gives each warp for its task's table (0: walk through L1)
    BlobHeader h;                       // host copy of the blob header
};

struct ScanCounters {                   // device-resident, zeroed before every scan
    unsigned long long n_cand;          // candidate entries written by k_scan
    unsigned long long n_raw;           // raw events written by k_confirm / k_dfa_scan
    unsigned long long n_verified;      // literal confirmations that reached a verify DFA
    unsigned int cand_overflow, raw_overflow;
    unsigned int merge_barrier;         // grid-wide barrier of k_merge_all (arrivals, all rounds)
    unsigned int pad;
};

struct PodDev {                         // per pod, device-resident
    unsigned long long start;           // offset of the pod's first byte in `text`
    unsigned long long eff_end;         // start + length without trailing separators (SPEC.md s.2)
    unsigned long long raw_end;         // start + length
    unsigned long long nl_before;       // '\n' bytes in text[0, start)            (k_pods)
    int n_lines;                        // lines of the pod (0 when eff_end == start)
    int pad;
};

// Pod table as the kernels see it.  any_trailing_cr: some pod ends in a lone '\r'; only then
// can a '\r' in front of the '\n' padding be mistaken for the '\r' of a "\r\n" separator, and
// only then do the kernels pay for a pod lookup when they strip it.
struct PodTable {
    const PodDev *pods;
    int n;
    int any_trailing_cr;
};

struct EventOut {                       // SoA result rows, device pointers
    int32_t *pod, *pattern, *line_no, *line_len, *ctx_first, *ctx_last;
    int64_t *line_start, *ctx_start, *ctx_end;
    double *score;
};

cudaError_t launch_pad(uint8_t *text, const PodDev *pods, int n_pods, unsigned long long total_bytes, cudaStream_t s);
cudaError_t launch_scan(const uint8_t *text, unsigned long long n_tiles, const DeviceLibrary &lib, uint32_t *tile_nl,
                        unsigned long long *cand, unsigned long long cand_cap, ScanCounters *ctr, int num_sms, cudaStream_t s);
// Exclusive prefix over tile_nl -> tile_prefix.  `scratch` holds ceil(n_tiles / 4096) + 1 u64.
cudaError_t launch_prefix(const uint32_t *tile_nl, unsigned long long *tile_prefix, unsigned long long n_tiles,
                          unsigned long long *scratch, cudaStream_t s);
cudaError_t launch_pods(const uint8_t *text, const unsigned long long *tile_prefix, PodDev *pods, int n_pods, cudaStream_t s);
// Two launches: k_confirm (candidate -> exact gram -> literal -> verify task) and k_verify (task ->
// line -> verify DFA -> raw event).  `tasks` holds raw_cap u64.
cudaError_t launch_confirm(const uint8_t *text, unsigned long long total_bytes, const DeviceLibrary &lib, PodTable pt,
                           const unsigned long long *cand, unsigned long long cand_cap, unsigned long long *tasks, unsigned long long *raw,
                           unsigned long long raw_cap, ScanCounters *ctr, int num_sms, cudaStream_t s, bool tables_in_smem = true);
// wide / table_bytes: from the group's descriptor (GroupDesc.wide, nstates * stride * 2).
cudaError_t launch_dfa_scan(const uint8_t *text, unsigned long long total_bytes, const DeviceLibrary &lib, PodTable pt, int group, bool wide,
                            size_t table_bytes, unsigned long long *raw, unsigned long long raw_cap, ScanCounters *ctr, int num_sms, cudaStream_t s);
// Sorts raw[0, n) and removes duplicates; *n_unique (device) receives the count.  temp/temp_bytes:
// scratch from sort_temp_bytes(n).
size_t sort_temp_bytes(unsigned long long n);
cudaError_t launch_sort_unique(unsigned long long *raw, unsigned long long *sorted, unsigned long long *uniq,
                               unsigned long long n, unsigned long long *n_unique, void *temp, size_t temp_bytes, cudaStream_t s);
// Resolve stage scratch (device memory owned by the caller).  items: item_cap x resolve_item_bytes(),
// results: item_cap bytes, long_list: long_cap u32, aux: n_events x resolve_aux_bytes(), groups: group_cap x resolve_group_bytes().
struct ResolveCounters {
    unsigned long long n_items, n_long;
    unsigned int overflow, n_groups;
};
struct ResolveScratch {
    void *items;
    unsigned long long item_cap;
    unsigned char *results;
    uint32_t *long_list;
    unsigned long long long_cap;
    void *aux;
    ResolveCounters *ctr;
    void *groups;                       // group_cap x resolve_group_bytes(): runs of items that share a unit (k_check: a warp per run)
    unsigned long long group_cap;
};
size_t resolve_group_bytes();
size_t resolve_item_bytes();
size_t resolve_aux_bytes();
// Four launches (k_plan, k_check, k_check_long, k_finalize).  If rs.ctr->overflow is set afterwards the
// scratch was too small and the outputs are incomplete: call again with larger capacities.
// keys[0, *n_keys_dev) is what gets planned and checked (rs.aux holds key_cap entries): either uniq itself,
// or the raw event queue while the sort that produces uniq is still running on another stream -
// then uniq_ready is the event recorded after that sort, waited for just before k_finalize, which
// places each key's row at its rank in uniq[0, *n_unique_dev) (rank: the ranks launch_sort_small left in its scratch).  All counts are read on the device;
// *n_keys_dev > key_cap or *n_unique_dev > ev_cap sets rs.ctr->overflow.
cudaError_t launch_resolve(const uint8_t *text, const DeviceLibrary &lib, const unsigned long long *tile_prefix, PodTable pt,
                           const unsigned long long *keys, const unsigned long long *n_keys_dev, unsigned long long key_cap,
                           const unsigned long long *uniq, const unsigned long long *n_unique_dev, unsigned long long ev_cap, EventOut out,
                           const ResolveScratch &rs, int num_sms, cudaStream_t s, cudaStream_t side = nullptr, cudaEvent_t fork_ev = nullptr,
                           cudaEvent_t join_ev = nullptr, cudaEvent_t uniq_ready = nullptr, const uint32_t *rank = nullptr);
// Ordering + de-duplication without host involvement for ctr->n_raw <= kMediumSort (otherwise *big_flag = 1 and
// the caller falls back to launch_sort_unique once it knows the count).  Up to kSmallSort keys take one rank-merge
// launch in shared memory, more take kMergePasses rounds of pairwise merges.
constexpr int kSmallSort = 16384, kMediumSort = 131072;
// Five launches (those of the path that does not apply return at once); `scratch` holds 2 * kMediumSort u64.  Afterwards the
// first kMediumSort u32 of `scratch` hold, for every raw key, its rank in uniq (what k_finalize needs to place the key's row).
cudaError_t launch_sort_small(const unsigned long long *raw, ScanCounters *ctr, unsigned long long raw_cap, unsigned long long *scratch,
                              unsigned long long *uniq, unsigned long long *n_unique, unsigned int *big_flag, int num_sms, cudaStream_t s);
// Gram profile: counts, for every aligned word of text[0, sample_bytes), the occurrences of the
// keys of an open-addressing table (keys: 2^log2_cap u32, kGramEmpty = free; counts zeroed by the
// caller).  Used once per library to learn which literal grams are frequent in the real logs.
cudaError_t launch_gram_profile(const uint8_t *text, unsigned long long sample_bytes, const uint32_t *keys, uint32_t log2_cap, uint32_t fold,
                                unsigned int *counts, int num_sms, cudaStream_t s);
// Result rows (first *n_rows_dev of each array, if <= cap) and the counter block to pinned host memory, by a kernel.
// dst_host / ctr_host: device-visible addresses of pinned host buffers laid out like the device ones.
cudaError_t launch_export(EventOut src, EventOut dst_host, const unsigned long long *n_rows_dev, unsigned long long cap, const void *ctr_dev, void *ctr_host,
                          size_t ctr_bytes, int num_sms, cudaStream_t s);
// Line table of one pod: out[i] = offset (relative to the pod) of line i, i < n_lines.
cudaError_t launch_line_starts(const uint8_t *text, const unsigned long long *tile_prefix, unsigned long long nl_before, int n_lines,
                               unsigned long long pod_start, unsigned long long pod_eff_end, long long *out, int num_sms, cudaStream_t s);

}  // namespace pm
