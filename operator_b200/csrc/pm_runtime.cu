// Host runtime behind the C ABI (include/podmortem_scan.h): context, device memory pool,
// library publication, pod batches with pod-level sharding over the GPUs of one node, and the
// marshalling of device results into pinned structure-of-arrays views.
//
// Reference side of this boundary: LogParserRestClient.parseLogs (LogParserRestClient.java:37-39)
// carries one pod per request (PodFailureWatcher.java:196-199); a batch is what the new /parse
// resource submits after coalescing concurrent requests.  Pods are
// independent, so the multi-GPU path is a partition of the pods (longest-processing-time greedy
// by bytes) with no collective; match lists are gathered on the host.
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "pm_kernels.h"
#include "pm_library.h"
#include "pm_split.h"
#include "podmortem_scan.h"

using namespace pm;

namespace {

thread_local std::string g_err;

int fail(int status, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return status;
}

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) return fail(PM_E_CUDA, "%s: %s", #call, cudaGetErrorString(e__)); \
    } while (0)

struct Pool {   // size-bucketed free list; blocks are handed back, never split
    std::mutex mu;
    std::multimap<size_t, void *> free_blocks;
    bool pinned = false;
    int device = 0;

    int get(size_t bytes, void **out, size_t *cap) {
        if (bytes == 0) bytes = 256;
        {
            std::lock_guard<std::mutex> g(mu);
            auto it = free_blocks.lower_bound(bytes);
            if (it != free_blocks.end() && it->first <= bytes * 4 + (1u << 20)) {
                *out = it->second;
                *cap = it->first;
                free_blocks.erase(it);
                return PM_OK;
            }
        }
        size_t want = (bytes + 255) & ~(size_t)255;
        cudaError_t e;
        if (pinned) {
            e = cudaHostAlloc(out, want, cudaHostAllocPortable | cudaHostAllocMapped);
        } else {
            cudaSetDevice(device);
            e = cudaMalloc(out, want);
        }
        if (e != cudaSuccess) {
            cudaGetLastError();
            trim();
            e = pinned ? cudaHostAlloc(out, want, cudaHostAllocPortable | cudaHostAllocMapped) : cudaMalloc(out, want);
            if (e != cudaSuccess) return fail(PM_E_NOMEM, "%s of %zu bytes: %s", pinned ? "cudaHostAlloc" : "cudaMalloc", want, cudaGetErrorString(e));
        }
        *cap = want;
        return PM_OK;
    }
    void put(void *p, size_t cap) {
        if (!p) return;
        std::lock_guard<std::mutex> g(mu);
        free_blocks.emplace(cap, p);
    }
    void trim() {
        std::lock_guard<std::mutex> g(mu);
        for (auto &kv : free_blocks) pinned ? cudaFreeHost(kv.second) : cudaFree(kv.second);
        free_blocks.clear();
    }
};

struct Buf {
    void *p = nullptr;
    size_t cap = 0;
    template <class T>
    T *as() const { return reinterpret_cast<T *>(p); }
};

struct Device {
    int id = 0, num_sms = 0;
    Pool pool;
    // Batches in flight on this device: at most `depth` (pm_set_pipeline; 1 unless asked otherwise).
    std::mutex mu;
    std::condition_variable cv;
    int holders = 0, depth = 1;
    // depth > 1: the SMs are split into two green contexts.  k_scan takes a whole SM per CTA (1024
    // threads x 64 registers) for the length of a pass, so nothing shares an SM with it; with the
    // split, the scan of one batch streams on sms_scan SMs while the latency-bound tail (confirm,
    // verify, sort, resolve) of the previous batch runs on the other sms_post.
    CUgreenCtx gc_scan = nullptr, gc_post = nullptr;
    int sms_scan = 0, sms_post = 0;
    // Streams and events of finished batches, kept for the next one: creating three streams and
    // sixteen events costs more host time than a 1 MB request spends on the GPU.
    struct StreamSet {
        cudaStream_t stream = nullptr, stream2 = nullptr, stream3 = nullptr, scan_stream = nullptr;
        cudaEvent_t scan_fork = nullptr, scan_join = nullptr, pods_copied = nullptr, fork_ev[3] = {nullptr, nullptr, nullptr}, join_ev[3] = {nullptr, nullptr, nullptr};
        cudaEvent_t e[10] = {nullptr};
        void destroy() {
            for (cudaStream_t *st : {&stream, &stream2, &stream3, &scan_stream})
                if (*st) cudaStreamDestroy(*st), *st = nullptr;
            for (cudaEvent_t *ev : {&scan_fork, &scan_join, &pods_copied})
                if (*ev) cudaEventDestroy(*ev), *ev = nullptr;
            for (int k = 0; k < 3; k++) {
                if (fork_ev[k]) cudaEventDestroy(fork_ev[k]), fork_ev[k] = nullptr;
                if (join_ev[k]) cudaEventDestroy(join_ev[k]), join_ev[k] = nullptr;
            }
            for (auto &ev : e)
                if (ev) cudaEventDestroy(ev), ev = nullptr;
        }
    };
    std::vector<StreamSet> idle_sets;   // guarded by mu; emptied when the pipeline mode changes
    void drop_idle_sets() {
        std::vector<StreamSet> sets;
        {
            std::lock_guard<std::mutex> g(mu);
            sets.swap(idle_sets);
        }
        for (auto &ss : sets) ss.destroy();
    }
    void acquire() {
        std::unique_lock<std::mutex> g(mu);
        cv.wait(g, [&] { return holders < depth; });
        holders++;
    }
    void release() {
        {
            std::lock_guard<std::mutex> g(mu);
            holders--;
        }
        cv.notify_all();
    }
};

// Green contexts are driver API; the entry points are taken from the loaded driver so that the
// library keeps linking against the runtime only.
struct DriverApi {
    CUresult (*DeviceGet)(CUdevice *, int) = nullptr;
    CUresult (*DeviceGetDevResource)(CUdevice, CUdevResource *, CUdevResourceType) = nullptr;
    CUresult (*DevSmResourceSplitByCount)(CUdevResource *, unsigned int *, const CUdevResource *, CUdevResource *, unsigned int, unsigned int) = nullptr;
    CUresult (*DevResourceGenerateDesc)(CUdevResourceDesc *, CUdevResource *, unsigned int) = nullptr;
    CUresult (*GreenCtxCreate)(CUgreenCtx *, CUdevResourceDesc, CUdevice, unsigned int) = nullptr;
    CUresult (*GreenCtxDestroy)(CUgreenCtx) = nullptr;
    CUresult (*GreenCtxStreamCreate)(CUstream *, CUgreenCtx, unsigned int, int) = nullptr;
    bool load() {
        auto one = [](const char *name, void **fn) {
            cudaDriverEntryPointQueryResult st;
            return cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &st) == cudaSuccess && st == cudaDriverEntryPointSuccess && *fn;
        };
        return one("cuDeviceGet", (void **)&DeviceGet) && one("cuDeviceGetDevResource", (void **)&DeviceGetDevResource) &&
               one("cuDevSmResourceSplitByCount", (void **)&DevSmResourceSplitByCount) &&
               one("cuDevResourceGenerateDesc", (void **)&DevResourceGenerateDesc) && one("cuGreenCtxCreate", (void **)&GreenCtxCreate) &&
               one("cuGreenCtxDestroy", (void **)&GreenCtxDestroy) && one("cuGreenCtxStreamCreate", (void **)&GreenCtxStreamCreate);
    }
};

}  // namespace

// Which literal grams are frequent in benign text depends on the workload ("thre" of
// "thread-pool" is in every Quarkus log line), and the compiler's static byte-rarity model
// cannot know.  When the candidate rate of a scan is far above the bitmap's fill rate, the
// runtime counts EVERY gram of every literal on a sample of the staged text (k_gram_profile),
// re-selects each literal's 7-byte window with the measured frequencies and uploads the rebuilt
// filter.  A perf adaptation only: every candidate is confirmed exactly, results never change.
struct AdaptState {
    bool settled = false;
    int rounds = 0;
    // Small batches (a service that gets one 1 MB pod log per request) say little each, so their candidate counts are
    // added up: the decision is taken once 8 MiB of text have been seen, on whatever batch is staged then.
    uint64_t seen_bytes = 0, seen_cand = 0;
    // After the filter has settled the rate keeps being watched, 64 MiB at a time: logs whose vocabulary has drifted
    // (a new release, another tenant) open the selection again, at most kMaxReopen times per library and device.
    double settled_rate = 0.0;
    uint64_t watch_bytes = 0, watch_cand = 0;
    int reopened = 0;
    void *dev_filter = nullptr;     // rebuilt filter sections (one allocation)
    std::vector<void *> retired;    // earlier ones: a batch in flight may still read them; freed with the library
};

struct pm_library {
    CompiledLibrary lib;
    std::atomic<int> refs{1};
    std::vector<DeviceLibrary> dev;   // per context device, after publish
    std::vector<void *> dev_blobs;
    std::vector<AdaptState> adapt;
    std::mutex adapt_mu;              // guards dev[].fv against a concurrent maybe_adapt
    std::vector<uint8_t> saved;       // pm_library_serialize
    pm_ctx *owner = nullptr;
    uint64_t generation = 0;   // unique per compiled library (result-cache key of pm_parse)
};

struct pm_ctx {
    std::vector<Device *> devs;
    Pool pinned;
    std::mutex mu;
    pm_library *current = nullptr;
    DriverApi drv;
    void *service_state = nullptr;             // owned by pm_service.cc
    void (*service_deleter)(void *) = nullptr;
    std::mutex user_pinned_mu;
    std::map<void *, size_t> user_pinned;
    std::string parse_response;   // see pm_service.cc
};

namespace {

void lib_unref(pm_library *l) {
    if (!l) return;
    if (l->refs.fetch_sub(1) == 1) {
        for (size_t d = 0; d < l->dev_blobs.size(); d++)
            if (l->dev_blobs[d]) {
                cudaSetDevice(l->owner->devs[d]->id);
                cudaFree(l->dev_blobs[d]);
                if (d < l->adapt.size()) {
                    if (l->adapt[d].dev_filter) cudaFree(l->adapt[d].dev_filter);
                    for (void *r : l->adapt[d].retired) cudaFree(r);
                }
            }
        delete l;
    }
}

struct PodRef {
    const uint8_t *ptr;
    size_t len, eff_len;
    uint64_t tag;
    int on_device = -1;       // >= 0: `ptr` is device memory on that context device (pm_batch_add_pod_device); staged there, never cut
    bool last_cr = false;     // the log ends in a lone '\r'
    int dev = 0, local = 0;   // device slot and index among that device's pods
    uint64_t start = 0;       // offset in the device text buffer
    // As a staged unit (pm_batch::units): which pod, which bytes of it, and how many lines of
    // halo lie in front of / behind the lines this unit reports (0 for a pod staged whole).
    int pod = 0, lead = 0, trail = 0;
    uint64_t begin = 0;
};

struct DevWork {
    int slot = -1;   // index into ctx->devs
    std::vector<int> pods;   // batch pod indices, ascending
    uint64_t total = 0, n_tiles = 0;
    uint64_t cand_cap = 0, raw_cap = 0;
    Buf text, tile_nl, tile_prefix, scratch, cand, raw, sorted, uniq, temp, ctr, pods_dev, ev_soa, lines_tmp;
    Buf rs_items, rs_results, rs_long, rs_aux, rs_groups;   // resolve-stage scratch
    Buf tasks, sort_scratch;
    uint64_t item_cap = 0, long_cap = 0, ev_cap = 0, group_cap = 0;
    int any_trailing_cr = 0;
    PodTable table() const { return PodTable{pods_dev.as<PodDev>(), (int)pods.size(), any_trailing_cr}; }
    Buf h_ctr, h_pods, h_soa;   // pinned mirrors
    cudaStream_t stream = nullptr, stream2 = nullptr, stream3 = nullptr;   // 2, 3: work that is independent of the main chain
    cudaStream_t scan_stream = nullptr;   // pipelined mode: k_scan runs in the scan partition of the device
    cudaEvent_t scan_fork = nullptr, scan_join = nullptr, pods_copied = nullptr;
    int sms_scan = 0, sms_post = 0;       // SMs the scan / everything else is sized for
    DeviceLibrary dl;                     // the library view this scan runs with (copied under adapt_mu)
    cudaEvent_t fork_ev[3] = {nullptr, nullptr, nullptr}, join_ev[3] = {nullptr, nullptr, nullptr};
    uint64_t key_cap = 0;   // entries of rs_aux (keys the resolve stage can plan in one go)
    bool post_restarted = false;
    cudaEvent_t e[10] = {nullptr};   // 0 h2d0, 1 h2d1, 2 scan0, 3 scan1, 4 mid, 5 post0, 6 post1, 7 d2h1, 8 dfa0, 9 dfa1
    uint64_t n_events = 0, n_cand = 0, n_verified = 0;
    bool locked = false, staged = false, scanned = false;
    int launches = 0;
    // PM_DEBUG_TIMING=1: an event after every stage of the main stream, printed by finish()
    std::vector<std::pair<const char *, cudaEvent_t>> marks;
    void mark(const char *name) {
        static const bool on = getenv("PM_DEBUG_TIMING") != nullptr;
        if (!on) return;
        cudaEvent_t ev;
        cudaEventCreate(&ev);
        cudaEventRecord(ev, stream);
        marks.emplace_back(name, ev);
    }
    static void mark_hook(void *self, const char *name) { reinterpret_cast<DevWork *>(self)->mark(name); }
    void hook_marks() {   // lets the multi-kernel launch wrappers of pm_kernels.cu mark their stages as well
        g_stage_mark = mark_hook;
        g_stage_mark_arg = this;
    }
};

// Device-side layout of the small counters block `w.ctr`: ScanCounters, then the unique-event
// count (u64) and the "too many raw events for the one-CTA sort" flag (u32).
constexpr size_t kOffUnique = sizeof(ScanCounters);
constexpr size_t kOffBig = sizeof(ScanCounters) + 8;
constexpr size_t kOffResolve = sizeof(ScanCounters) + 16;   // ResolveCounters live in the same block: one copy back
constexpr size_t kCtrBytes = sizeof(ScanCounters) + 16 + sizeof(ResolveCounters);

size_t round_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// PM_VERIFY_SMEM_ALWAYS=1: k_verify stages tables in shared memory also with several batches in flight (measurement switch)
bool verify_smem_always() {
    static const bool v = getenv("PM_VERIFY_SMEM_ALWAYS") != nullptr;
    return v;
}

// PM_TEST_TINY_QUEUES=1 starts every internal queue (filter candidates, raw events, resolve work
// items) at a few entries, so that the overflow -> grow -> rerun paths run on small test inputs.
bool tiny_queues() {
    static const bool v = getenv("PM_TEST_TINY_QUEUES") != nullptr;
    return v;
}

}  // namespace

struct pm_batch {
    pm_ctx *ctx = nullptr;
    pm_library *lib = nullptr;
    std::vector<PodRef> pods;
    // What is actually staged: one unit per pod, or several for a log cut across devices
    // (split_pod).  DevWork::pods and PodDev tables index units.
    std::vector<PodRef> units;
    std::vector<std::pair<int, int>> pod_units;   // per pod: first unit, unit count
    std::vector<DevWork> work;
    bool staged = false, submitted = false;
    // final result (pinned, batch order)
    Buf r_soa, r_pod;
    pm_result_view view;
    std::vector<int32_t> pod_n_lines;
};

namespace {

int get(Pool &p, size_t bytes, Buf &b) {
    if (b.p && b.cap >= bytes) return PM_OK;
    if (b.p) p.put(b.p, b.cap);
    b.p = nullptr;
    b.cap = 0;
    return p.get(bytes, &b.p, &b.cap);
}
void put(Pool &p, Buf &b) {
    p.put(b.p, b.cap);
    b.p = nullptr;
    b.cap = 0;
}

void release_work(pm_batch *b) {
    for (auto &w : b->work) {
        Device *d = b->ctx->devs[w.slot];
        cudaSetDevice(d->id);
        if (w.stream) cudaStreamSynchronize(w.stream);
        for (Buf *x : {&w.text, &w.tile_nl, &w.tile_prefix, &w.scratch, &w.cand, &w.raw, &w.sorted, &w.uniq, &w.temp, &w.ctr, &w.pods_dev, &w.ev_soa, &w.lines_tmp, &w.rs_items, &w.rs_results, &w.rs_long, &w.rs_aux, &w.rs_groups, &w.tasks, &w.sort_scratch})
            put(d->pool, *x);
        for (Buf *x : {&w.h_ctr, &w.h_pods, &w.h_soa}) put(b->ctx->pinned, *x);
        if (w.stream) {
            for (cudaStream_t st : {w.stream2, w.stream3, w.scan_stream})
                if (st) cudaStreamSynchronize(st);
            Device::StreamSet ss;
            ss.stream = w.stream, ss.stream2 = w.stream2, ss.stream3 = w.stream3, ss.scan_stream = w.scan_stream;
            ss.scan_fork = w.scan_fork, ss.scan_join = w.scan_join, ss.pods_copied = w.pods_copied;
            for (int k = 0; k < 3; k++) ss.fork_ev[k] = w.fork_ev[k], ss.join_ev[k] = w.join_ev[k];
            for (int k = 0; k < 10; k++) ss.e[k] = w.e[k];
            w.stream = w.stream2 = w.stream3 = w.scan_stream = nullptr;
            w.scan_fork = w.scan_join = w.pods_copied = nullptr;
            for (int k = 0; k < 3; k++) w.fork_ev[k] = w.join_ev[k] = nullptr;
            for (auto &ev : w.e) ev = nullptr;
            bool keep = false;
            if (w.locked) {   // still holding the device: the pipeline mode cannot have changed under us
                std::lock_guard<std::mutex> g(d->mu);
                if (d->idle_sets.size() < 16) d->idle_sets.push_back(ss), keep = true;
            }
            if (!keep) ss.destroy();
        }
        if (w.locked) d->release(), w.locked = false;
    }
    b->work.clear();
}

// SoA layout of n events inside one allocation (8-byte fields first)
struct SoaLayout {
    size_t off_line_start, off_ctx_start, off_ctx_end, off_score, off_pod, off_pattern, off_line_no, off_line_len, off_ctx_first, off_ctx_last, bytes;
    explicit SoaLayout(uint64_t n) {
        size_t n8 = round_up(n * 8, 64), n4 = round_up(n * 4, 64);
        off_line_start = 0;
        off_ctx_start = n8;
        off_ctx_end = 2 * n8;
        off_score = 3 * n8;
        off_pod = 4 * n8;
        off_pattern = off_pod + n4;
        off_line_no = off_pattern + n4;
        off_line_len = off_line_no + n4;
        off_ctx_first = off_line_len + n4;
        off_ctx_last = off_ctx_first + n4;
        bytes = off_ctx_last + n4 + 64;
    }
    EventOut on(void *base) const {
        char *b = reinterpret_cast<char *>(base);
        EventOut o;
        o.line_start = reinterpret_cast<int64_t *>(b + off_line_start);
        o.ctx_start = reinterpret_cast<int64_t *>(b + off_ctx_start);
        o.ctx_end = reinterpret_cast<int64_t *>(b + off_ctx_end);
        o.score = reinterpret_cast<double *>(b + off_score);
        o.pod = reinterpret_cast<int32_t *>(b + off_pod);
        o.pattern = reinterpret_cast<int32_t *>(b + off_pattern);
        o.line_no = reinterpret_cast<int32_t *>(b + off_line_no);
        o.line_len = reinterpret_cast<int32_t *>(b + off_line_len);
        o.ctx_first = reinterpret_cast<int32_t *>(b + off_ctx_first);
        o.ctx_last = reinterpret_cast<int32_t *>(b + off_ctx_last);
        return o;
    }
};

int stage_impl(pm_batch *b) {
    pm_ctx *ctx = b->ctx;
    if (b->staged) return PM_OK;
    {
        std::lock_guard<std::mutex> g(ctx->mu);
        if (!ctx->current) return fail(PM_E_NO_LIBRARY, "no pattern library published");
        b->lib = ctx->current;
        b->lib->refs.fetch_add(1);
    }
    const int nd = (int)ctx->devs.size();
    // ---- pods -> units: a log of more than two minimum pieces is cut across the devices
    //      (PM_SPLIT_MIN_BYTES, default 64 MiB; PM_SPLIT_FORCE=n cuts into n pieces even on one device: tests)
    static const size_t split_min = getenv("PM_SPLIT_MIN_BYTES") ? (size_t)strtoull(getenv("PM_SPLIT_MIN_BYTES"), nullptr, 10) : (size_t)64 << 20;
    static const int split_force = getenv("PM_SPLIT_FORCE") ? atoi(getenv("PM_SPLIT_FORCE")) : 0;
    b->units.clear();
    b->pod_units.clear();
    for (size_t i = 0; i < b->pods.size(); i++) {
        const PodRef &p = b->pods[i];
        const int first = (int)b->units.size();
        int pieces = split_force > 1 ? split_force : nd;
        if (split_min > 0) pieces = (int)std::min<size_t>((size_t)pieces, p.eff_len / split_min);
        if (p.on_device >= 0) pieces = 1;   // the cut is made on host bytes
        if (pieces > 1) {
            const int halo = (b->lib->dev.empty() || b->lib->dev[0].ctx_slots <= 0) ? 64 : b->lib->dev[0].ctx_slots;
            std::vector<SplitUnit> cut;   // pm_split.cc
            split_log(p.ptr, p.len, pieces, halo, cut);
            for (const SplitUnit &c : cut) {
                PodRef u = p;
                u.ptr = p.ptr + c.begin;
                u.len = c.end - c.begin;
                u.eff_len = effective_len(u.ptr, u.len);
                u.last_cr = u.len && u.ptr[u.len - 1] == '\r';
                u.pod = (int)i;
                u.begin = c.begin;
                u.lead = c.lead;
                u.trail = c.trail;
                b->units.push_back(u);
            }
        } else {
            PodRef u = p;
            u.pod = (int)i;
            b->units.push_back(u);
        }
        b->pod_units.emplace_back(first, (int)b->units.size() - first);
    }
    // ---- unit -> device: longest-processing-time greedy by bytes
    std::vector<int> order(b->units.size());
    for (size_t i = 0; i < order.size(); i++) order[i] = (int)i;
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return b->units[x].len > b->units[y].len; });
    std::vector<uint64_t> load(nd, 0);
    for (const PodRef &u : b->units)
        if (u.on_device >= 0) load[u.on_device] += u.len + 64;   // device-resident pods stay where they are
    for (int i : order) {
        if (b->units[i].on_device >= 0) {
            b->units[i].dev = b->units[i].on_device;
            continue;
        }
        int best = 0;
        for (int d = 1; d < nd; d++)
            if (load[d] < load[best]) best = d;
        b->units[i].dev = best;
        load[best] += b->units[i].len + 64;
    }
    // Work entries in ascending device order, whatever the pods' order: the loop below takes one
    // in-flight slot per device and holds it while it waits for the next, so two threads staging
    // multi-device batches must ask for the devices in the same order or they can deadlock.
    b->work.clear();
    std::vector<int> slot_of(nd, -1);
    for (size_t i = 0; i < b->units.size(); i++) slot_of[b->units[i].dev] = 0;
    for (int d = 0; d < nd; d++)
        if (slot_of[d] == 0) {
            slot_of[d] = (int)b->work.size();
            b->work.emplace_back();
            b->work.back().slot = d;
        }
    for (size_t i = 0; i < b->units.size(); i++) {
        int d = b->units[i].dev;
        DevWork &w = b->work[slot_of[d]];
        b->units[i].local = (int)w.pods.size();
        w.pods.push_back((int)i);
    }
    // ---- per device: layout, buffers, copies
    for (auto &w : b->work) {
        Device *d = ctx->devs[w.slot];
        d->acquire();
        w.locked = true;
        CU(cudaSetDevice(d->id));
        bool reused = false;
        {
            std::lock_guard<std::mutex> g(d->mu);
            if (!d->idle_sets.empty()) {
                Device::StreamSet ss = d->idle_sets.back();
                d->idle_sets.pop_back();
                w.stream = ss.stream, w.stream2 = ss.stream2, w.stream3 = ss.stream3, w.scan_stream = ss.scan_stream;
                w.scan_fork = ss.scan_fork, w.scan_join = ss.scan_join, w.pods_copied = ss.pods_copied;
                for (int k = 0; k < 3; k++) w.fork_ev[k] = ss.fork_ev[k], w.join_ev[k] = ss.join_ev[k];
                for (int k = 0; k < 10; k++) w.e[k] = ss.e[k];
                reused = true;
            }
        }
        w.sms_scan = d->gc_scan ? d->sms_scan : d->num_sms;
        w.sms_post = d->gc_scan ? d->sms_post : d->num_sms;
        if (!reused) {
            if (d->gc_scan) {
                auto mk = [&](cudaStream_t *st, CUgreenCtx gc) {
                    CUresult r = ctx->drv.GreenCtxStreamCreate(st, gc, CU_STREAM_NON_BLOCKING, 0);
                    return r == CUDA_SUCCESS ? (int)PM_OK : fail(PM_E_CUDA, "cuGreenCtxStreamCreate failed (%d)", (int)r);
                };
                int rc;
                if ((rc = mk(&w.stream, d->gc_post)) || (rc = mk(&w.stream2, d->gc_post)) || (rc = mk(&w.stream3, d->gc_post)) ||
                    (rc = mk(&w.scan_stream, d->gc_scan)))
                    return rc;
                CU(cudaEventCreateWithFlags(&w.scan_fork, cudaEventDisableTiming));
                CU(cudaEventCreateWithFlags(&w.scan_join, cudaEventDisableTiming));
                w.sms_scan = d->sms_scan;
                w.sms_post = d->sms_post;
            } else if (d->depth > 1 && getenv("PM_PIPELINE_PRIORITY")) {
                // opt-in: the short kernels of the batches that are nearly done go ahead of the next
                // batch's scan CTAs whenever an SM frees up.  Measured on B200 (1 GiB batches, depth 5):
                // per-batch latency 1.07 -> 0.70 ms, throughput 3.96 -> 3.70 TB/s; off by default.
                int lo = 0, hi = 0;
                CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
                CU(cudaStreamCreateWithPriority(&w.stream, cudaStreamNonBlocking, hi));
                CU(cudaStreamCreateWithPriority(&w.stream2, cudaStreamNonBlocking, hi));
                CU(cudaStreamCreateWithPriority(&w.stream3, cudaStreamNonBlocking, hi));
                CU(cudaStreamCreateWithPriority(&w.scan_stream, cudaStreamNonBlocking, lo));
                CU(cudaEventCreateWithFlags(&w.scan_fork, cudaEventDisableTiming));
                CU(cudaEventCreateWithFlags(&w.scan_join, cudaEventDisableTiming));
                w.sms_scan = w.sms_post = d->num_sms;
            } else {
                CU(cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking));
                CU(cudaStreamCreateWithFlags(&w.stream2, cudaStreamNonBlocking));
                CU(cudaStreamCreateWithFlags(&w.stream3, cudaStreamNonBlocking));
                w.sms_scan = w.sms_post = d->num_sms;
            }
            CU(cudaEventCreateWithFlags(&w.pods_copied, cudaEventDisableTiming));
            for (int k = 0; k < 3; k++) {
                CU(cudaEventCreateWithFlags(&w.fork_ev[k], cudaEventDisableTiming));
                CU(cudaEventCreateWithFlags(&w.join_ev[k], cudaEventDisableTiming));
            }
            for (auto &ev : w.e) CU(cudaEventCreate(&ev));
        }
        const int np = (int)w.pods.size();
        uint64_t off = kHeaderBytes;
        int rc;
        if ((rc = get(ctx->pinned, sizeof(PodDev) * (np + 1), w.h_pods))) return rc;
        PodDev *hp = w.h_pods.as<PodDev>();
        w.any_trailing_cr = 0;
        for (int k = 0; k < np; k++) {
            PodRef &p = b->units[w.pods[k]];
            p.start = off;
            hp[k].start = off;
            hp[k].eff_end = off + p.eff_len;
            hp[k].raw_end = off + p.len;
            hp[k].nl_before = 0;
            hp[k].n_lines = 0;
            hp[k].pad = 0;
            if (p.last_cr) w.any_trailing_cr = 1;
            off = round_up(off + p.len + 1, 16);   // >= 1 byte of '\n' padding
        }
        // content end, rounded to a tile, plus one full tile of '\n' so that 512-byte look-ahead
        // reads of the context collector stay inside the buffer
        w.total = round_up(off, kTileBytes) + kTileBytes;
        w.n_tiles = w.total / kTileBytes;
        if (w.total >= (1ull << 40)) return fail(PM_E_INVALID, "more than 1 TiB of log text on one device");
        w.cand_cap = std::max<uint64_t>(1u << 20, w.total / 256);
        w.raw_cap = std::max<uint64_t>(1u << 18, w.total / 1024);
        if (tiny_queues()) w.cand_cap = 64, w.raw_cap = 32;
        if ((rc = get(d->pool, w.total, w.text))) return rc;
        if ((rc = get(d->pool, w.n_tiles * 4, w.tile_nl))) return rc;
        if ((rc = get(d->pool, w.n_tiles * 8, w.tile_prefix))) return rc;
        if ((rc = get(d->pool, (w.n_tiles / 4096 + 2) * 8, w.scratch))) return rc;
        if ((rc = get(d->pool, w.cand_cap * 8, w.cand))) return rc;
        if ((rc = get(d->pool, w.raw_cap * 8, w.raw))) return rc;
        if ((rc = get(d->pool, kCtrBytes + 64, w.ctr))) return rc;
        if ((rc = get(d->pool, sizeof(PodDev) * (np + 1), w.pods_dev))) return rc;
        if ((rc = get(ctx->pinned, kCtrBytes + 64, w.h_ctr))) return rc;
        memset(w.h_ctr.p, 0, kCtrBytes + 64);
        CU(cudaMemcpyAsync(w.pods_dev.p, hp, sizeof(PodDev) * np, cudaMemcpyHostToDevice, w.stream));
        CU(cudaEventRecord(w.e[0], w.stream));
        for (int k = 0; k < np; k++) {
            PodRef &p = b->units[w.pods[k]];
            if (p.len) CU(cudaMemcpyAsync(w.text.as<uint8_t>() + p.start, p.ptr, p.len, p.on_device >= 0 ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, w.stream));
        }
        CU(cudaEventRecord(w.e[1], w.stream));
        CU(launch_pad(w.text.as<uint8_t>(), w.pods_dev.as<PodDev>(), np, w.total, w.stream));
        w.launches = 1;
        w.staged = true;
    }
    b->staged = true;
    return PM_OK;
}

// A staging that fails half-way (pool out of memory, stream creation) gives back everything it took:
// the in-flight slots of the devices it had already acquired, their buffers and streams, and its
// reference to the library - otherwise a retry would leak them and, at depth 1, every later batch on
// those devices would wait for a slot that is never released.
int stage(pm_batch *b) {
    const int rc = stage_impl(b);
    if (rc != PM_OK && !b->staged) {
        const std::string msg = g_err;
        release_work(b);
        if (b->lib) lib_unref(b->lib), b->lib = nullptr;
        g_err = msg;
    }
    return rc;
}

// Called with the batch's own streams synchronised.  Other batches may be in flight on the device
// (pm_set_pipeline): they keep the filter view they copied at submit time, and a replaced filter
// is retired, not freed.
int maybe_adapt(pm_library *l, DevWork &w) {
    const int slot = w.slot;
    if (slot >= (int)l->adapt.size()) return PM_OK;
    std::unique_lock<std::mutex> adapt_lock(l->adapt_mu, std::try_to_lock);
    if (!adapt_lock.owns_lock()) return PM_OK;   // another batch is at it
    AdaptState &st = l->adapt[slot];
    DeviceLibrary &dl = l->dev[slot];
    if (!dl.fv.log2_bits) return PM_OK;
    const bool debug = getenv("PM_DEBUG") != nullptr;
    constexpr int kMaxReopen = 4;
    if (st.settled) {   // keep watching the rate
        st.watch_bytes += w.total;
        st.watch_cand += w.n_cand;
        if (st.watch_bytes < (64ull << 20)) return PM_OK;
        const double now = (double)st.watch_cand / ((double)st.watch_bytes / 4.0);
        st.watch_bytes = st.watch_cand = 0;
        if (now <= 4.0 * st.settled_rate + 1e-4 || st.reopened >= kMaxReopen) return PM_OK;
        if (debug) fprintf(stderr, "[pm] filter: rate %.5f against %.5f when it settled: selecting windows again\n", now, st.settled_rate);
        st.settled = false;
        st.rounds = 0;
        st.reopened++;
        st.seen_bytes = st.seen_cand = 0;
        return PM_OK;   // the next batches decide
    }
    st.seen_bytes += w.total;
    st.seen_cand += w.n_cand;
    if (st.seen_bytes < (8ull << 20) || w.total < (256u << 10)) return PM_OK;   // not enough text yet / too small a sample to profile
    const double words = (double)st.seen_bytes / 4.0;
    // what a filter whose grams are absent from the text would still let through: a word must collide in both bitmaps
    const double fill = (double)dl.fv.n_grams / (double)(1ull << dl.fv.log2_bits) * (double)dl.fv.n_grams / (double)(1ull << dl.fv.log2_bits2);
    const double rate = (double)st.seen_cand / words;
    if (debug) fprintf(stderr, "[pm] filter: %llu candidates in %llu bytes, rate %.5f, fill %.5f, round %d\n", (unsigned long long)st.seen_cand,
                       (unsigned long long)st.seen_bytes, rate, fill, st.rounds);
    st.seen_bytes = st.seen_cand = 0;
    if (rate <= 3.0 * fill + 2e-5 || st.rounds >= 2) {
        st.settled = true;
        st.settled_rate = rate;
        return PM_OK;
    }
    st.rounds++;
    // ---- 1. count every literal gram on a sample of the staged text
    const bool fold = dl.fv.fold != 0;
    std::vector<uint32_t> grams;
    collect_all_grams(l->lib.literals, fold, grams);
    uint32_t log2_cap = 4;
    while ((1ull << log2_cap) < 2ull * grams.size()) log2_cap++;
    const size_t cap = (size_t)1 << log2_cap;
    std::vector<uint32_t> keys(cap, kGramEmpty);
    for (uint32_t g : grams) {
        uint32_t sl = gram_table_slot(g, log2_cap);
        while (keys[sl] != kGramEmpty) sl = (sl + 1) & (uint32_t)(cap - 1);
        keys[sl] = g;
    }
    Device *d = l->owner->devs[slot];
    CU(cudaSetDevice(d->id));
    void *dkeys = nullptr, *dcounts = nullptr;
    CU(cudaMalloc(&dkeys, cap * 4));
    CU(cudaMalloc(&dcounts, cap * 4));
    CU(cudaMemcpyAsync(dkeys, keys.data(), cap * 4, cudaMemcpyHostToDevice, w.stream));
    CU(cudaMemsetAsync(dcounts, 0, cap * 4, w.stream));
    const unsigned long long sample = std::min<unsigned long long>(w.total, 256ull << 20) & ~15ull;
    CU(launch_gram_profile(w.text.as<uint8_t>(), sample, reinterpret_cast<const uint32_t *>(dkeys), log2_cap, fold, reinterpret_cast<unsigned int *>(dcounts),
                           w.sms_post, w.stream));
    std::vector<uint32_t> counts(cap);
    CU(cudaMemcpyAsync(counts.data(), dcounts, cap * 4, cudaMemcpyDeviceToHost, w.stream));
    CU(cudaStreamSynchronize(w.stream));
    cudaFree(dkeys);
    cudaFree(dcounts);
    // ---- 2. re-select windows with measured frequencies
    const double sample_words = (double)sample / 4.0;
    GramCostFn cost = [&](uint32_t g) -> double {
        uint32_t sl = gram_table_slot(g, log2_cap);
        while (keys[sl] != g && keys[sl] != kGramEmpty) sl = (sl + 1) & (uint32_t)(cap - 1);
        double c = keys[sl] == g ? (double)counts[sl] : 0.0;
        return (c + 0.25) / sample_words;
    };
    FilterBuild fb;
    build_filter(l->lib.literals, fold ? 1 : 0, &cost, l->lib.granule, fb);
    // ---- 3. upload the rebuilt sections and switch the view
    auto a16 = [](size_t x) { return (x + 15) & ~(size_t)15; };
    size_t o_bm = 0, o_b2 = a16(o_bm + fb.bitmap.size() * 4), o_k = a16(o_b2 + fb.bitmap2.size() * 4), o_f = a16(o_k + fb.gkeys.size() * 4), o_c = a16(o_f + fb.gfirst.size() * 4);
    size_t o_p = a16(o_c + fb.gcount.size() * 4), o_l = a16(o_p + fb.probes.size() * sizeof(Probe)), total = a16(o_l + fb.litpool.size()) + 16;
    std::vector<uint8_t> host(total, 0);
    memcpy(host.data() + o_bm, fb.bitmap.data(), fb.bitmap.size() * 4);
    memcpy(host.data() + o_b2, fb.bitmap2.data(), fb.bitmap2.size() * 4);
    memcpy(host.data() + o_k, fb.gkeys.data(), fb.gkeys.size() * 4);
    memcpy(host.data() + o_f, fb.gfirst.data(), fb.gfirst.size() * 4);
    memcpy(host.data() + o_c, fb.gcount.data(), fb.gcount.size() * 4);
    memcpy(host.data() + o_p, fb.probes.data(), fb.probes.size() * sizeof(Probe));
    if (!fb.litpool.empty()) memcpy(host.data() + o_l, fb.litpool.data(), fb.litpool.size());
    void *dev = nullptr;
    CU(cudaMalloc(&dev, total));
    CU(cudaMemcpy(dev, host.data(), total, cudaMemcpyHostToDevice));
    if (st.dev_filter) st.retired.push_back(st.dev_filter);
    st.dev_filter = dev;
    const uint8_t *db = reinterpret_cast<const uint8_t *>(dev);
    FilterView &fv = dl.fv;
    fv.bitmap = reinterpret_cast<const uint32_t *>(db + o_bm);
    fv.bitmap2 = reinterpret_cast<const uint32_t *>(db + o_b2);
    fv.log2_bits2 = fb.log2_bits2;
    fv.gkeys = reinterpret_cast<const uint32_t *>(db + o_k);
    fv.gfirst = reinterpret_cast<const uint32_t *>(db + o_f);
    fv.gcount = reinterpret_cast<const uint32_t *>(db + o_c);
    fv.probes = reinterpret_cast<const Probe *>(db + o_p);
    fv.litpool = db + o_l;
    fv.log2_bits = fb.log2_bits;
    fv.log2_cap = fb.log2_cap;
    fv.n_grams = fb.n_grams;
    if (debug) {
        double expect = 0;
        for (uint32_t i = 0; i < fb.gkeys.size(); i++)
            if (fb.gkeys[i] != kGramEmpty) expect += cost(fb.gkeys[i]);
        fprintf(stderr, "[pm] filter rebuilt from %zu profiled grams: %u grams, 2^%u bits, expected true-gram rate %.6f\n", grams.size(), fb.n_grams,
                fb.log2_bits, expect);
    }
    return PM_OK;
}

// Scratch of the resolve stage for up to ev_cap events (grow-only).
int ensure_resolve_scratch(pm_batch *b, DevWork &w, uint64_t ev_cap) {
    pm_ctx *ctx = b->ctx;
    Device *d = ctx->devs[w.slot];
    w.ev_cap = ev_cap;   // the result arrays are copied back at this capacity: keep it near the event count
    w.item_cap = std::min<uint64_t>(std::max<uint64_t>(w.item_cap, tiny_queues() ? 16 : ev_cap * 96 + (1u << 16)), 0xFFFFFFF0ull);
    w.long_cap = std::max<uint64_t>(w.long_cap, w.item_cap / 8 + (1u << 12));
    SoaLayout L(ev_cap);
    int rc;
    if ((rc = get(d->pool, L.bytes, w.ev_soa))) return rc;
    if ((rc = get(ctx->pinned, L.bytes, w.h_soa))) return rc;
    if ((rc = get(d->pool, w.item_cap * resolve_item_bytes(), w.rs_items))) return rc;
    if ((rc = get(d->pool, w.item_cap, w.rs_results))) return rc;
    if ((rc = get(d->pool, w.long_cap * 4, w.rs_long))) return rc;
    // runs of items that share a unit (one STACK run and one window per secondary and event): a run has at least one item,
    // most have 3-25; too few is an overflow like any other and grows with item_cap
    w.group_cap = tiny_queues() ? std::max<uint64_t>(w.group_cap, 4) : std::max<uint64_t>(w.group_cap, w.item_cap / 4 + (1u << 12));
    if ((rc = get(d->pool, w.group_cap * resolve_group_bytes(), w.rs_groups))) return rc;
    // the resolve stage plans either the sorted events (<= ev_cap) or the raw queue (<= kMediumSort)
    w.key_cap = std::max<uint64_t>(ev_cap, std::min<uint64_t>(w.raw_cap, kMediumSort));
    if ((rc = get(d->pool, w.key_cap * resolve_aux_bytes(), w.rs_aux))) return rc;
    if ((rc = get(d->pool, w.key_cap * 8 + 64, w.uniq))) return rc;   // the on-device sort may deliver up to key_cap unique keys, whatever ev_cap is
    return PM_OK;
}

// Enqueues the resolve stage and the result copies; the event count is read on the device.
// from_raw: plan and check the raw queue while the sort is still running on stream3 (join_ev[2]).
int enqueue_resolve(pm_batch *b, DevWork &w, bool from_raw = false) {
    const DeviceLibrary &dl = w.dl;
    w.hook_marks();
    SoaLayout L(w.ev_cap);
    ResolveScratch rs{w.rs_items.p, w.item_cap, w.rs_results.as<unsigned char>(), w.rs_long.as<uint32_t>(), w.long_cap, w.rs_aux.p,
                      reinterpret_cast<ResolveCounters *>(w.ctr.as<char>() + kOffResolve), w.rs_groups.p, w.group_cap};
    const unsigned long long *n_unique = reinterpret_cast<const unsigned long long *>(w.ctr.as<char>() + kOffUnique);
    const unsigned long long *keys = from_raw ? w.raw.as<unsigned long long>() : w.uniq.as<unsigned long long>();
    const unsigned long long *n_keys = from_raw ? &w.ctr.as<ScanCounters>()->n_raw : n_unique;
    const uint64_t key_cap = from_raw ? std::min<uint64_t>(w.raw_cap, kMediumSort) : w.ev_cap;
    CU(launch_resolve(w.text.as<uint8_t>(), dl, w.tile_prefix.as<unsigned long long>(), w.table(), keys, n_keys, key_cap,
                      w.uniq.as<unsigned long long>(), n_unique, w.ev_cap, L.on(w.ev_soa.p), rs, w.sms_post, w.stream, w.stream2, w.fork_ev[1],
                      w.join_ev[1], from_raw ? w.join_ev[2] : nullptr, from_raw ? w.sort_scratch.as<uint32_t>() : nullptr));
    w.launches += 4;
    w.mark("k_finalize");
    CU(cudaEventRecord(w.e[6], w.stream));
    // One batch at a time (latency mode): a small kernel writes exactly the rows there are, and the
    // counters, straight into pinned host memory - 26 instead of 35 us at the end of the 1 GiB step,
    // 0.25 -> 0.23 ms for a 1.4 MB request.  With several batches in flight (throughput mode) the copy
    // engines do it instead: they cost no SM, and an SM is what the other batches are waiting for
    // (measured: 1 % more throughput than with the kernel).  PM_RESULT_MEMCPY=1 forces the copy engines.
    static const bool force_copy_engine = getenv("PM_RESULT_MEMCPY") != nullptr;
    const bool copy_engine = force_copy_engine || (b->ctx->devs[w.slot]->depth > 1 && w.ev_cap <= kSmallSort);   // a large capacity is mostly empty: copy what there is
    if (copy_engine) {
        CU(cudaMemcpyAsync(w.h_soa.p, w.ev_soa.p, L.bytes, cudaMemcpyDeviceToHost, w.stream));
        CU(cudaMemcpyAsync(w.h_ctr.p, w.ctr.p, kCtrBytes, cudaMemcpyDeviceToHost, w.stream));
    } else {
        // unified addressing: pinned (portable, mapped) host memory has the same address on the device
        CU(launch_export(L.on(w.ev_soa.p), L.on(w.h_soa.p), n_unique, w.ev_cap, w.ctr.p, w.h_ctr.p, kCtrBytes, w.sms_post, w.stream));
        w.launches += 1;
    }
    CU(cudaEventRecord(w.e[7], w.stream));
    w.mark("result copies");
    return PM_OK;
}

// The whole pipeline of one scan, enqueued without a host round trip: scan, newline prefix,
// pod bases, confirm (+ combined DFAs), one-CTA sort/unique, resolve, result copies.
int scan_async(pm_batch *b) {
    pm_ctx *ctx = b->ctx;
    for (auto &w : b->work) {
        Device *d = ctx->devs[w.slot];
        CU(cudaSetDevice(d->id));
        {
            std::lock_guard<std::mutex> g(b->lib->adapt_mu);
            w.dl = b->lib->dev[w.slot];
        }
        const DeviceLibrary &dl = w.dl;
        w.hook_marks();
        const int np = (int)w.pods.size();
        // capacity of the result arrays: everything the small sort can deliver on the first scan of a
        // batch, the last event count plus 12 % on a re-scan (less to copy back)
        // (first scan: what the on-device sort can deliver for a batch of this size at a hit rate of 1e-2 of its lines;
        // a batch with more events than that comes back once with the count and repeats the resolve stage)
        uint64_t want = std::min<uint64_t>(kMediumSort, std::max<uint64_t>(kSmallSort, round_up(w.total / 8192, 1024)));
        if (w.scanned) want = std::min<uint64_t>(kMediumSort, round_up(w.n_events + w.n_events / 8 + 64, 256));
        int rc = ensure_resolve_scratch(b, w, want);
        if (rc) return rc;
        CU(cudaMemsetAsync(w.ctr.p, 0, kCtrBytes, w.stream));
        w.mark("start");
        cudaStream_t ss = w.scan_stream ? w.scan_stream : w.stream;
        if (w.scan_stream) {
            CU(cudaEventRecord(w.scan_fork, w.stream));
            CU(cudaStreamWaitEvent(ss, w.scan_fork, 0));
        }
        if ((rc = get(d->pool, w.raw_cap * 8, w.tasks))) return rc;
        if ((rc = get(d->pool, 2 * (size_t)kMediumSort * 8, w.sort_scratch))) return rc;
        CU(cudaEventRecord(w.e[2], ss));
        CU(launch_scan(w.text.as<uint8_t>(), w.n_tiles, dl, w.tile_nl.as<uint32_t>(), w.cand.as<unsigned long long>(), w.cand_cap,
                       w.ctr.as<ScanCounters>(), w.sms_scan, ss));
        CU(cudaEventRecord(w.e[3], ss));
        if (w.scan_stream) {
            CU(cudaEventRecord(w.scan_join, ss));
            CU(cudaStreamWaitEvent(w.stream, w.scan_join, 0));
        }
        w.mark("k_scan");
        // the newline prefix and the pod bases only feed the resolve stage: side stream, joined before k_plan
        CU(cudaEventRecord(w.fork_ev[0], w.stream));
        CU(cudaStreamWaitEvent(w.stream2, w.fork_ev[0], 0));
        CU(launch_prefix(w.tile_nl.as<uint32_t>(), w.tile_prefix.as<unsigned long long>(), w.n_tiles, w.scratch.as<unsigned long long>(), w.stream2));
        CU(launch_pods(w.text.as<uint8_t>(), w.tile_prefix.as<unsigned long long>(), w.pods_dev.as<PodDev>(), np, w.stream2));
        CU(cudaEventRecord(w.join_ev[0], w.stream2));
        // the host's copy of the pod table is only read at gather time: it leaves the device behind the
        // event k_plan waits for, and the main stream picks it up at the very end (pods_copied)
        CU(cudaMemcpyAsync(w.h_pods.p, w.pods_dev.p, sizeof(PodDev) * np, cudaMemcpyDeviceToHost, w.stream2));
        CU(cudaEventRecord(w.pods_copied, w.stream2));
        CU(launch_confirm(w.text.as<uint8_t>(), w.total, dl, w.table(), w.cand.as<unsigned long long>(), w.cand_cap, w.tasks.as<unsigned long long>(),
                          w.raw.as<unsigned long long>(), w.raw_cap, w.ctr.as<ScanCounters>(), w.sms_post, w.stream, /*tables_in_smem=*/d->depth == 1 || verify_smem_always()));
        w.launches += 1 + 2 + (np ? 1 : 0) + (dl.fv.log2_bits ? 2 : 0);
        w.mark("k_verify");
        // (a timing event between two kernels keeps the second from starting until the first has drained and the stamp is
        // written, 1-2 us each: none where there is nothing to time)
        if (dl.h.n_groups) CU(cudaEventRecord(w.e[8], w.stream));
        for (uint32_t g = 0; g < dl.h.n_groups; g++) {
            const GroupDesc &gd = reinterpret_cast<const GroupDesc *>(b->lib->lib.blob.data() + dl.h.off_groups)[g];
            CU(launch_dfa_scan(w.text.as<uint8_t>(), w.total, dl, w.table(), (int)g, gd.wide != 0, (size_t)gd.nstates * gd.stride * 2,
                               w.raw.as<unsigned long long>(), w.raw_cap, w.ctr.as<ScanCounters>(), w.sms_post, w.stream));
            w.launches++;
        }
        if (dl.h.n_groups) CU(cudaEventRecord(w.e[9], w.stream));
        // ordering and de-duplication run beside plan/check (which take the raw queue as it is);
        // k_finalize waits for them and places each row at its key's rank
        CU(cudaEventRecord(w.fork_ev[2], w.stream));
        CU(cudaStreamWaitEvent(w.stream3, w.fork_ev[2], 0));
        CU(launch_sort_small(w.raw.as<unsigned long long>(), w.ctr.as<ScanCounters>(), w.raw_cap, w.sort_scratch.as<unsigned long long>(),
                             w.uniq.as<unsigned long long>(),
                             reinterpret_cast<unsigned long long *>(w.ctr.as<char>() + kOffUnique),
                             reinterpret_cast<unsigned int *>(w.ctr.as<char>() + kOffBig), w.sms_post, w.stream3));
        CU(cudaEventRecord(w.join_ev[2], w.stream3));
        w.launches += 5;
        CU(cudaStreamWaitEvent(w.stream, w.join_ev[0], 0));
        w.mark("join prefix/pods");
        w.post_restarted = false;   // e[4] / e[5] bracket the host's part when the resolve stage has to be run again (finish)
        if ((rc = enqueue_resolve(b, w, true))) return rc;
        CU(cudaStreamWaitEvent(w.stream, w.pods_copied, 0));
    }
    return PM_OK;
}

int finish(pm_batch *b, pm_result_view *out) {
    pm_ctx *ctx = b->ctx;
    // ---- phase 1: wait; rerun with larger queues if the candidate / raw-event queues overflowed
    for (;;) {
        bool again = false;
        for (auto &w : b->work) {
            Device *d = ctx->devs[w.slot];
            CU(cudaSetDevice(d->id));
            CU(cudaStreamSynchronize(w.stream));
            ScanCounters *c = w.h_ctr.as<ScanCounters>();
            if (c->cand_overflow || c->raw_overflow) {
                int rc;
                if (c->cand_overflow) {
                    w.cand_cap = std::min<uint64_t>(w.cand_cap * 8, w.total / 16 + 64);
                    if ((rc = get(d->pool, w.cand_cap * 8, w.cand))) return rc;
                }
                if (c->raw_overflow) {
                    w.raw_cap *= 8;
                    if ((rc = get(d->pool, w.raw_cap * 8, w.raw))) return rc;
                }
                again = true;
            }
        }
        if (!again) break;
        int rc = scan_async(b);   // all devices: keeps the code simple, and overflow is rare
        if (rc) return rc;
    }
    // ---- phase 1b: batches with more raw events than the one-CTA sort takes go through the radix
    //      sort, which needs the count on the host; a too-small work-item queue is grown and
    //      the resolve stage repeated
    for (auto &w : b->work) {
        Device *d = ctx->devs[w.slot];
        CU(cudaSetDevice(d->id));
        ScanCounters *c = w.h_ctr.as<ScanCounters>();
        w.n_cand = c->n_cand;
        w.n_verified = c->n_verified;
        const bool big = *reinterpret_cast<unsigned int *>(w.h_ctr.as<char>() + kOffBig) != 0;
        if (big) {
            const uint64_t n_raw = c->n_raw;
            int rc;
            CU(cudaEventRecord(w.e[4], w.stream));   // (the stream is idle: the host has just read the counters)
            CU(cudaEventRecord(w.e[5], w.stream));
            w.post_restarted = true;
            size_t tb = sort_temp_bytes(n_raw);
            if ((rc = get(d->pool, n_raw * 8, w.sorted))) return rc;
            if ((rc = get(d->pool, tb, w.temp))) return rc;
            if ((rc = ensure_resolve_scratch(b, w, n_raw))) return rc;   // uniq holds up to n_raw keys
            unsigned long long *n_unique = reinterpret_cast<unsigned long long *>(w.ctr.as<char>() + kOffUnique);
            CU(launch_sort_unique(w.raw.as<unsigned long long>(), w.sorted.as<unsigned long long>(), w.uniq.as<unsigned long long>(), n_raw, n_unique,
                                  w.temp.p, tb, w.stream));
            w.launches += 1;
            if ((rc = enqueue_resolve(b, w))) return rc;
            CU(cudaStreamSynchronize(w.stream));
        }
        while (reinterpret_cast<ResolveCounters *>(w.h_ctr.as<char>() + kOffResolve)->overflow) {
            const uint64_t n_unique = *reinterpret_cast<unsigned long long *>(w.h_ctr.as<char>() + kOffUnique);
            int rc;
            if (n_unique > w.ev_cap) {   // more events than the result arrays were sized for
                uint64_t cap = w.ev_cap;
                while (cap < n_unique) cap <<= 1;
                if ((rc = ensure_resolve_scratch(b, w, cap))) return rc;
            } else {
                w.item_cap = std::min<uint64_t>(w.item_cap * 4, 0xFFFFFFF0ull);
                w.long_cap *= 4;
                w.group_cap *= 4;
                if ((rc = ensure_resolve_scratch(b, w, w.ev_cap))) return rc;
            }
            if ((rc = enqueue_resolve(b, w))) return rc;
            CU(cudaStreamSynchronize(w.stream));
        }
        w.n_events = *reinterpret_cast<unsigned long long *>(w.h_ctr.as<char>() + kOffUnique);
    }
    // ---- phase 2: gather in batch pod order
    uint64_t total_events = 0, bytes = 0, n_cand = 0, n_ver = 0;
    int launches = 0;
    double ms_h2d = 0, ms_scan = 0, ms_post = 0, ms_d2h = 0, ms_total = 0, ms_dfa = 0;
    for (auto &w : b->work) {
        CU(cudaSetDevice(ctx->devs[w.slot]->id));
        CU(cudaStreamSynchronize(w.stream));
        total_events += w.n_events;
        bytes += w.total;
        n_cand += w.n_cand;
        n_ver += w.n_verified;
        launches += w.launches;
        w.launches = 0;
        float a = 0, s = 0, p1 = 0, p2 = 0, dd = 0, tt = 0;
        cudaEventElapsedTime(&a, w.e[0], w.e[1]);
        cudaEventElapsedTime(&s, w.e[2], w.e[3]);
        if (w.post_restarted) {
            cudaEventElapsedTime(&p1, w.e[3], w.e[4]);
            cudaEventElapsedTime(&p2, w.e[5], w.e[6]);
        } else {
            cudaEventElapsedTime(&p1, w.e[3], w.e[6]);
        }
        cudaEventElapsedTime(&dd, w.e[6], w.e[7]);
        cudaEventElapsedTime(&tt, w.e[2], w.e[7]);
        ms_total = std::max<double>(ms_total, tt);
        float df = 0;
        if (w.dl.h.n_groups) cudaEventElapsedTime(&df, w.e[8], w.e[9]);
        if (w.dl.h.n_groups) ms_dfa = std::max<double>(ms_dfa, df);
        ms_h2d = std::max<double>(ms_h2d, a);
        ms_scan = std::max<double>(ms_scan, s);
        ms_post = std::max<double>(ms_post, p1 + p2);
        ms_d2h = std::max<double>(ms_d2h, dd);
        w.scanned = true;
        if (!w.marks.empty()) {
            for (size_t k = 1; k < w.marks.size(); k++) {
                float ms = 0;
                cudaEventElapsedTime(&ms, w.marks[k - 1].second, w.marks[k].second);
                fprintf(stderr, "[pm timing] %-32s %8.1f us\n", w.marks[k].first, ms * 1e3);
            }
            for (auto &mk : w.marks) cudaEventDestroy(mk.second);
            w.marks.clear();
        }
        int rs = maybe_adapt(b->lib, w);
        if (rs) return rs;
    }
    const size_t np = b->pods.size();
    SoaLayout L(total_events);
    int rc;
    if ((rc = get(ctx->pinned, L.bytes, b->r_soa))) return rc;
    if ((rc = get(ctx->pinned, np * 8 + (np + 1) * 8 + np * 4 + 64, b->r_pod))) return rc;
    EventOut R = L.on(b->r_soa.p);
    uint64_t *tags = b->r_pod.as<uint64_t>();
    int64_t *first = reinterpret_cast<int64_t *>(tags + np);
    int32_t *nlines = reinterpret_cast<int32_t *>(first + np + 1);
    // per device: event ranges per local pod (events are sorted by pod on each device)
    std::vector<std::vector<uint64_t>> dev_first(b->work.size());
    for (size_t k = 0; k < b->work.size(); k++) {
        DevWork &w = b->work[k];
        dev_first[k].assign(w.pods.size() + 1, w.n_events);
        if (w.n_events) {
            EventOut S = SoaLayout(w.ev_cap).on(w.h_soa.p);
            int prev = -1;
            for (uint64_t i = 0; i < w.n_events; i++)
                while (prev < S.pod[i]) dev_first[k][++prev] = i;
            for (size_t j = (size_t)prev + 1; j <= w.pods.size(); j++) dev_first[k][j] = w.n_events;
        }
        dev_first[k][w.pods.size()] = w.n_events;
    }
    uint64_t at = 0;
    std::vector<int> slot_index(ctx->devs.size(), -1);
    for (size_t k = 0; k < b->work.size(); k++) slot_index[b->work[k].slot] = (int)k;
    for (size_t i = 0; i < np; i++) {
        tags[i] = b->pods[i].tag;
        first[i] = (int64_t)at;
        int64_t line_base = 0;   // lines of this pod reported by the units before the current one
        for (int ui = b->pod_units[i].first; ui < b->pod_units[i].first + b->pod_units[i].second; ui++) {
            const PodRef &p = b->units[ui];
            DevWork &w = b->work[slot_index[p.dev]];
            const PodDev *hp = w.h_pods.as<PodDev>();
            const int core_lines = hp[p.local].n_lines - p.lead - p.trail;
            uint64_t a = dev_first[slot_index[p.dev]][p.local], e = dev_first[slot_index[p.dev]][p.local + 1];
            if (e > a) {
                EventOut S = SoaLayout(w.ev_cap).on(w.h_soa.p);
                if (b->pod_units[i].second == 1) {
                    size_t cnt = e - a;
                    memcpy(R.line_start + at, S.line_start + a, cnt * 8);
                    memcpy(R.ctx_start + at, S.ctx_start + a, cnt * 8);
                    memcpy(R.ctx_end + at, S.ctx_end + a, cnt * 8);
                    memcpy(R.score + at, S.score + a, cnt * 8);
                    memcpy(R.pattern + at, S.pattern + a, cnt * 4);
                    memcpy(R.line_no + at, S.line_no + a, cnt * 4);
                    memcpy(R.line_len + at, S.line_len + a, cnt * 4);
                    memcpy(R.ctx_first + at, S.ctx_first + a, cnt * 4);
                    memcpy(R.ctx_last + at, S.ctx_last + a, cnt * 4);
                    for (size_t j = 0; j < cnt; j++) R.pod[at + j] = (int32_t)i;
                    at += cnt;
                } else {   // a piece of a split log: own lines only, renumbered and re-based to the pod
                    const int64_t shift = line_base - p.lead;
                    for (uint64_t j = a; j < e; j++) {
                        const int ln = S.line_no[j];
                        if (ln <= p.lead || ln > p.lead + core_lines) continue;
                        R.line_start[at] = S.line_start[j] + (int64_t)p.begin;
                        R.ctx_start[at] = S.ctx_start[j] + (int64_t)p.begin;
                        R.ctx_end[at] = S.ctx_end[j] + (int64_t)p.begin;
                        R.score[at] = S.score[j];
                        R.pattern[at] = S.pattern[j];
                        R.line_no[at] = (int32_t)(ln + shift);
                        R.line_len[at] = S.line_len[j];
                        R.ctx_first[at] = (int32_t)(S.ctx_first[j] + shift);
                        R.ctx_last[at] = (int32_t)(S.ctx_last[j] + shift);
                        R.pod[at] = (int32_t)i;
                        at++;
                    }
                }
            }
            line_base += core_lines;
        }
        nlines[i] = b->pods[i].len == 0 ? 1 : (int32_t)line_base;   // "" splits into one empty line (SPEC.md section 2)
    }
    total_events = at;
    first[np] = (int64_t)at;
    pm_result_view &v = b->view;
    memset(&v, 0, sizeof v);
    v.n_events = (int64_t)total_events;
    v.n_pods = (int64_t)np;
    v.pod_index = R.pod;
    v.pattern_index = R.pattern;
    v.line_no = R.line_no;
    v.line_start = R.line_start;
    v.line_len = R.line_len;
    v.ctx_first_line = R.ctx_first;
    v.ctx_last_line = R.ctx_last;
    v.ctx_start = R.ctx_start;
    v.ctx_end = R.ctx_end;
    v.score = R.score;
    v.pod_tag = tags;
    v.pod_first_event = first;
    v.pod_n_lines = nlines;
    v.ms_h2d = ms_h2d;
    v.ms_scan = ms_scan;
    v.ms_post = ms_post;
    v.ms_d2h = ms_d2h;
    v.ms_total = ms_total;
    v.ms_dfa = ms_dfa;
    v.bytes_scanned = (int64_t)bytes;
    v.n_candidates = (int64_t)n_cand;
    v.n_verified = (int64_t)n_ver;
    v.gpu_launches = launches;
    if (out) *out = v;
    return PM_OK;
}

}  // namespace

// ============================================================================ C ABI
extern "C" {

int pm_abi_version(void) { return PM_ABI_VERSION; }

const char *pm_last_error(void) { return g_err.c_str(); }

int pm_init_devices(const int *device_ids, int n, pm_ctx **out) {
    if (!out || !device_ids || n <= 0) return fail(PM_E_INVALID, "bad arguments");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        cudaGetLastError();
        return fail(PM_E_NO_DEVICE, "no CUDA device (%s); this engine has no CPU path", e == cudaSuccess ? "count is 0" : cudaGetErrorString(e));
    }
    pm_ctx *ctx = new pm_ctx();
    ctx->pinned.pinned = true;
    for (int k = 0; k < n; k++) {
        int i = device_ids[k];
        cudaDeviceProp prop;
        if (i < 0 || i >= count || cudaGetDeviceProperties(&prop, i) != cudaSuccess || prop.major < 10) {
            cudaGetLastError();
            for (auto *d : ctx->devs) delete d;
            delete ctx;
            return fail(PM_E_NO_DEVICE, "device %d is missing or not sm_100-class: the kernels are built for sm_100a only", i);
        }
        Device *d = new Device();
        d->id = i;
        d->num_sms = prop.multiProcessorCount;
        d->pool.device = i;
        ctx->devs.push_back(d);
    }
    *out = ctx;
    return PM_OK;
}

int pm_init(int n_devices, pm_ctx **out) {
    if (!out) return fail(PM_E_INVALID, "out is null");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        cudaGetLastError();
        return fail(PM_E_NO_DEVICE, "no CUDA device (%s); this engine has no CPU path", e == cudaSuccess ? "count is 0" : cudaGetErrorString(e));
    }
    if (n_devices <= 0 || n_devices > count) n_devices = count;
    std::vector<int> ids(n_devices);
    for (int i = 0; i < n_devices; i++) ids[i] = i;
    return pm_init_devices(ids.data(), n_devices, out);
}

void pm_shutdown(pm_ctx *ctx) {
    if (!ctx) return;
    if (ctx->service_state && ctx->service_deleter) ctx->service_deleter(ctx->service_state);
    if (ctx->current) lib_unref(ctx->current);
    for (auto *d : ctx->devs) {
        cudaSetDevice(d->id);
        cudaDeviceSynchronize();
        d->pool.trim();
        d->drop_idle_sets();
        if (d->gc_scan) ctx->drv.GreenCtxDestroy(d->gc_scan);
        if (d->gc_post) ctx->drv.GreenCtxDestroy(d->gc_post);
        delete d;
    }
    ctx->pinned.trim();
    for (auto &kv : ctx->user_pinned) cudaFreeHost(kv.first);
    delete ctx;
}

int pm_device_count(const pm_ctx *ctx) { return ctx ? (int)ctx->devs.size() : 0; }

int pm_set_pipeline(pm_ctx *ctx, int depth, int post_sms, int *scan_sms_out, int *post_sms_out) {
    if (!ctx || depth < 1 || depth > 8) return fail(PM_E_INVALID, "depth must be 1..8");
    for (auto *d : ctx->devs) {
        std::lock_guard<std::mutex> g(d->mu);
        if (d->holders) return fail(PM_E_INVALID, "pm_set_pipeline with batches in flight on device %d", d->id);
    }
    if (depth > 1 && post_sms != 0 && !ctx->drv.GreenCtxCreate && !ctx->drv.load()) return fail(PM_E_CUDA, "this driver has no green-context entry points");
    for (auto *d : ctx->devs) {
        CU(cudaSetDevice(d->id));
        CU(cudaFree(nullptr));   // primary context up
        CU(cudaDeviceSynchronize());
        d->drop_idle_sets();   // streams of the old mode (green-context streams must go before their context)
        if (d->gc_scan) ctx->drv.GreenCtxDestroy(d->gc_scan), d->gc_scan = nullptr;
        if (d->gc_post) ctx->drv.GreenCtxDestroy(d->gc_post), d->gc_post = nullptr;
        d->depth = depth;
        d->sms_scan = d->sms_post = d->num_sms;
        if (depth == 1 || post_sms == 0) continue;   // no partition: the batches share all SMs
        if (post_sms < 1 || post_sms >= d->num_sms) return fail(PM_E_INVALID, "post_sms must be 1..%d", d->num_sms - 1);
        const DriverApi &A = ctx->drv;
        CUdevice cud;
        CUdevResource all, post, rest;
        CUdevResourceDesc desc_post, desc_scan;
        unsigned int groups = 1;
        CUresult r;
#define DRV(call)                                                    \
    if ((r = (call)) != CUDA_SUCCESS) {                              \
        d->depth = 1;                                                \
        return fail(PM_E_CUDA, "%s failed (%d)", #call, (int)r);     \
    }
        DRV(A.DeviceGet(&cud, d->id));
        DRV(A.DeviceGetDevResource(cud, &all, CU_DEV_RESOURCE_TYPE_SM));
        DRV(A.DevSmResourceSplitByCount(&post, &groups, &all, &rest, 0, (unsigned)post_sms));
        if (groups != 1 || rest.sm.smCount == 0) {
            d->depth = 1;
            return fail(PM_E_CUDA, "SM split of %d for the post partition left nothing for the scan", post_sms);
        }
        DRV(A.DevResourceGenerateDesc(&desc_post, &post, 1));
        DRV(A.DevResourceGenerateDesc(&desc_scan, &rest, 1));
        DRV(A.GreenCtxCreate(&d->gc_post, desc_post, cud, CU_GREEN_CTX_DEFAULT_STREAM));
        DRV(A.GreenCtxCreate(&d->gc_scan, desc_scan, cud, CU_GREEN_CTX_DEFAULT_STREAM));
#undef DRV
        d->sms_post = (int)post.sm.smCount;
        d->sms_scan = (int)rest.sm.smCount;
    }
    if (scan_sms_out) *scan_sms_out = ctx->devs.empty() ? 0 : ctx->devs[0]->sms_scan;
    if (post_sms_out) *post_sms_out = ctx->devs.empty() ? 0 : ctx->devs[0]->sms_post;
    return PM_OK;
}

int pm_library_compile(pm_ctx *ctx, const uint8_t *spec, size_t n, pm_library **out, pm_compile_report *rep) {
    pm_compile_report local;
    if (!rep) rep = &local;
    if (!out || !spec) return fail(PM_E_INVALID, "null argument");
    *out = nullptr;
    static std::atomic<uint64_t> next_generation{1};
    pm_library *l = new pm_library();
    l->owner = ctx;
    l->generation = next_generation.fetch_add(1);
    int rc = compile_library(reinterpret_cast<const char *>(spec), n, l->lib, *rep);
    if (rc != PM_OK) {
        delete l;
        return fail(rc, "pattern %d (%s): %s%s", rep->pattern_index, rep->pattern_id, rep->message,
                    rep->secondary_index >= 0 ? " [secondary pattern]" : "");
    }
    if (l->lib.patterns.size() >= (1u << kPatternBits)) {
        delete l;
        return fail(PM_E_TOO_COMPLEX, "more than %u patterns", 1u << kPatternBits);
    }
    *out = l;
    return PM_OK;
}

int pm_library_serialize(pm_library *l, const uint8_t **bytes, size_t *n) {
    if (!l || !bytes || !n) return fail(PM_E_INVALID, "null argument");
    if (l->saved.empty()) serialize_library(l->lib, l->saved);
    *bytes = l->saved.data();
    *n = l->saved.size();
    return PM_OK;
}

int pm_library_load(pm_ctx *ctx, const uint8_t *bytes, size_t n, pm_library **out) {
    if (!out || !bytes) return fail(PM_E_INVALID, "null argument");
    *out = nullptr;
    static std::atomic<uint64_t> next_generation{1ull << 32};
    pm_library *l = new pm_library();
    l->owner = ctx;
    l->generation = next_generation.fetch_add(1);
    std::string err;
    int rc = deserialize_library(bytes, n, l->lib, err);
    if (rc == PM_OK && l->lib.patterns.size() >= (1u << kPatternBits)) rc = PM_E_TOO_COMPLEX, err = "too many patterns";
    if (rc != PM_OK) {
        delete l;
        return fail(rc, "%s", err.c_str());
    }
    *out = l;
    return PM_OK;
}

int pm_library_publish(pm_ctx *ctx, pm_library *l) {
    if (!ctx || !l) return fail(PM_E_INVALID, "null argument");
    if (l->dev.empty()) {
        l->owner = ctx;
        l->dev.resize(ctx->devs.size());
        l->dev_blobs.assign(ctx->devs.size(), nullptr);
        l->adapt.assign(ctx->devs.size(), AdaptState());
        // an upload that fails on device d takes the copies on devices 0..d-1 back, so that a second
        // publish starts from nothing instead of skipping the upload and scanning with null tables
        auto upload = [&](size_t d) -> cudaError_t {
            cudaError_t e = cudaSetDevice(ctx->devs[d]->id);
            if (e == cudaSuccess) e = cudaMalloc(&l->dev_blobs[d], l->lib.blob.size());
            if (e == cudaSuccess) e = cudaMemcpy(l->dev_blobs[d], l->lib.blob.data(), l->lib.blob.size(), cudaMemcpyHostToDevice);
            return e;
        };
        for (size_t d = 0; d < ctx->devs.size(); d++) {
            const cudaError_t e = upload(d);
            if (e != cudaSuccess) {
                cudaGetLastError();
                for (size_t k = 0; k <= d; k++)
                    if (l->dev_blobs[k]) cudaSetDevice(ctx->devs[k]->id), cudaFree(l->dev_blobs[k]);
                l->dev.clear();
                l->dev_blobs.clear();
                l->adapt.clear();
                return fail(PM_E_CUDA, "library upload to device %d: %s", ctx->devs[d]->id, cudaGetErrorString(e));
            }
            l->dev[d].blob = reinterpret_cast<const uint8_t *>(l->dev_blobs[d]);
            l->dev[d].h = *l->lib.header();
            const BlobHeader &bh = *l->lib.header();
            FilterView &fv = l->dev[d].fv;
            const uint8_t *db = l->dev[d].blob;
            fv.bitmap = reinterpret_cast<const uint32_t *>(db + bh.off_bitmap);
            fv.bitmap2 = reinterpret_cast<const uint32_t *>(db + bh.off_bitmap2);
            fv.log2_bits2 = bh.filter_log2_bits2;
            fv.gkeys = reinterpret_cast<const uint32_t *>(db + bh.off_gram_keys);
            fv.gfirst = reinterpret_cast<const uint32_t *>(db + bh.off_gram_first);
            fv.gcount = reinterpret_cast<const uint32_t *>(db + bh.off_gram_count);
            fv.probes = reinterpret_cast<const Probe *>(db + bh.off_probes);
            fv.litpool = db + bh.off_litpool;
            fv.log2_bits = bh.filter_log2_bits;
            fv.log2_cap = bh.gram_log2_cap;
            fv.fold = bh.filter_fold;
            fv.granule = bh.filter_granule == 2 ? 2u : 4u;
            fv.n_grams = bh.n_grams;
            {   // largest context / proximity window of the library -> shared-memory slots of k_plan
                const uint8_t *hb = l->lib.blob.data();
                const PatternDesc *pdv = reinterpret_cast<const PatternDesc *>(hb + bh.off_patterns);
                const SecDesc *sdv = reinterpret_cast<const SecDesc *>(hb + bh.off_secondaries);
                int need = 1;
                for (uint32_t i = 0; i < bh.n_patterns; i++) {
                    need = std::max(need, std::max(pdv[i].lines_before, pdv[i].lines_after));
                    if (pdv[i].include_stack) need = std::max(need, kStackMax);
                }
                for (uint32_t i = 0; i < bh.n_secondaries; i++) need = std::max(need, sdv[i].window);
                l->dev[d].ctx_slots = need + 2;
                {   // first bytes of STACK lines: from the start state, every byte that does not lead to the dead state
                    const UnitDesc &su = reinterpret_cast<const UnitDesc *>(hb + bh.off_units)[bh.stack_unit];
                    const uint16_t *tab = reinterpret_cast<const uint16_t *>(hb + bh.off_tables) + su.table_off;
                    const uint8_t *cm = hb + bh.off_classmaps + su.classmap_off;
                    StackFirst sf;
                    memset(&sf, 0, sizeof sf);
                    for (int b = 0; b < 256; b++) {
                        const bool alive = su.dead_state == 0xFFFF || tab[cm[b]] != su.dead_state;
                        if (alive) sf.bits[b >> 5] |= 1u << (b & 31);
                    }
                    // an empty line presents its '\n' (or the '\r' of "\r\n") as first byte: allowed only if the unit accepts the empty line
                    const bool empty_ok = tab[su.stride - 1] == su.accept_state;
                    sf.bits['\n' >> 5] &= ~(1u << ('\n' & 31));
                    if (empty_ok) sf.bits['\n' >> 5] |= 1u << ('\n' & 31);
                    sf.bits['\r' >> 5] |= 1u << ('\r' & 31);   // could be the separator's '\r' or text: let the full path decide
                    l->dev[d].stack_first = sf;
                }
                // k_verify copies a task's table into shared memory when it fits the per-warp budget: the largest
                // table among the filtered primaries that still walk (one byte per entry up to 256 states), capped at 8 KiB: 4 warps x
                // budget + 3 KiB then still lets 6 CTAs share an SM, and since the literal-only patterns stopped walking the kernel no
                // longer needs more (4 KiB left 8 % of the walks on the L1 path, 15 us each: they set the kernel's time)
                const UnitDesc *udv = reinterpret_cast<const UnitDesc *>(hb + bh.off_units);
                uint32_t cap = 0;
                for (uint32_t i = 0; i < bh.n_patterns; i++)
                    if (pdv[i].tier == 1 && !udv[pdv[i].unit].literal_is_match) {   // one byte per entry up to 256 states (k_verify)
                        const UnitDesc &u = udv[pdv[i].unit];
                        cap = std::max(cap, (uint32_t)u.nstates * u.stride * (u.nstates <= 256 ? 1u : 2u));
                    }
                static const uint32_t cap_max = getenv("PM_VERIFY_TABLE_CAP") ? (uint32_t)atoi(getenv("PM_VERIFY_TABLE_CAP")) : 8192u;
                l->dev[d].verify_table_cap = std::min((cap + 255u) & ~255u, cap_max);
            }
        }
    }
    pm_library *old;
    {
        std::lock_guard<std::mutex> g(ctx->mu);
        old = ctx->current;
        l->refs.fetch_add(1);
        ctx->current = l;
    }
    lib_unref(old);
    return PM_OK;
}

void pm_library_free(pm_library *l) { lib_unref(l); }

int pm_library_pattern_count(const pm_library *l) { return l ? (int)l->lib.patterns.size() : 0; }

const char *pm_library_pattern_meta_json(const pm_library *l, int i) {
    if (!l || i < 0 || i >= (int)l->lib.patterns.size()) return nullptr;
    return l->lib.patterns[i].meta_json.c_str();
}

void *pm_host_alloc(pm_ctx *ctx, size_t bytes) {
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        fail(PM_E_NOMEM, "cudaHostAlloc(%zu) failed", bytes);
        return nullptr;
    }
    std::lock_guard<std::mutex> g(ctx->user_pinned_mu);
    ctx->user_pinned[p] = bytes;
    return p;
}

void pm_host_free(pm_ctx *ctx, void *p) {
    if (!p) return;
    std::lock_guard<std::mutex> g(ctx->user_pinned_mu);
    if (ctx->user_pinned.erase(p)) cudaFreeHost(p);
}

int pm_batch_begin(pm_ctx *ctx, pm_batch **out) {
    if (!ctx || !out) return fail(PM_E_INVALID, "null argument");
    pm_batch *b = new pm_batch();
    b->ctx = ctx;
    memset(&b->view, 0, sizeof b->view);
    *out = b;
    return PM_OK;
}

int pm_batch_add_pod(pm_batch *b, const uint8_t *log, size_t n, uint64_t tag) {
    if (!b || (!log && n)) return fail(PM_E_INVALID, "null argument");
    if (b->staged) return fail(PM_E_INVALID, "batch already staged");
    PodRef p;
    p.ptr = log;
    p.len = n;
    p.eff_len = effective_len(log, n);
    p.last_cr = n && log[n - 1] == '\r';
    p.tag = tag;
    b->pods.push_back(p);
    return PM_OK;
}

int pm_batch_add_pod_device(pm_batch *b, const void *d_log, size_t n, uint64_t tag) {
    if (!b || (!d_log && n)) return fail(PM_E_INVALID, "null argument");
    if (b->staged) return fail(PM_E_INVALID, "batch already staged");
    PodRef p;
    p.ptr = static_cast<const uint8_t *>(d_log);
    p.len = n;
    p.tag = tag;
    // which of the context's devices holds the bytes
    int ordinal = -1;
    if (n) {
        cudaPointerAttributes at;
        CU(cudaPointerGetAttributes(&at, d_log));
        if (at.type != cudaMemoryTypeDevice) return fail(PM_E_INVALID, "pm_batch_add_pod_device: not a device pointer (use pm_batch_add_pod for host memory)");
        ordinal = at.device;
    } else {
        CU(cudaGetDevice(&ordinal));
    }
    for (size_t k = 0; k < b->ctx->devs.size(); k++)
        if (b->ctx->devs[k]->id == ordinal) p.on_device = (int)k;
    if (p.on_device < 0) {
        if (n) return fail(PM_E_INVALID, "pm_batch_add_pod_device: the pointer lives on device %d, which this context does not own", ordinal);
        p.on_device = 0;
    }
    // Java's split drops trailing "\r?\n" separators: look at the end of the log, 4 KiB at a time.  A byte is kept
    // when it is neither '\r' nor '\n', or a '\r' that is not followed by '\n'; the effective end is behind the last kept byte.
    int previous = -1;
    cudaGetDevice(&previous);
    struct Restore {   // leave the caller's current device as it was
        int dev;
        ~Restore() {
            if (dev >= 0) cudaSetDevice(dev);
        }
    } restore{previous};
    CU(cudaSetDevice(ordinal));
    std::vector<uint8_t> tail(4097);
    size_t eff = 0, hi = n;
    bool found = false;
    while (hi > 0 && !found) {
        const size_t lo = hi > 4096 ? hi - 4096 : 0, upto = std::min(n, hi + 1);
        CU(cudaMemcpy(tail.data(), p.ptr + lo, upto - lo, cudaMemcpyDeviceToHost));
        for (size_t i = hi; i-- > lo;) {
            const uint8_t c = tail[i - lo];
            const bool keep = (c != '\r' && c != '\n') || (c == '\r' && (i + 1 == n || tail[i + 1 - lo] != '\n'));
            if (keep) {
                eff = i + 1;
                found = true;
                break;
            }
        }
        if (hi == n) p.last_cr = tail[n - 1 - lo] == '\r';
        hi = lo;
    }
    p.eff_len = eff;
    b->pods.push_back(p);
    return PM_OK;
}

int pm_batch_stage(pm_batch *b) {
    if (!b) return fail(PM_E_INVALID, "null argument");
    int rc = stage(b);
    if (rc) return rc;
    for (auto &w : b->work) {
        CU(cudaSetDevice(b->ctx->devs[w.slot]->id));
        CU(cudaStreamSynchronize(w.stream));
    }
    return PM_OK;
}

int pm_batch_submit(pm_batch *b) {
    if (!b) return fail(PM_E_INVALID, "null argument");
    int rc = stage(b);
    if (rc) return rc;
    rc = scan_async(b);
    if (rc) return rc;
    b->submitted = true;
    return PM_OK;
}

int pm_batch_wait(pm_batch *b, pm_result_view *out) {
    if (!b || !b->submitted) return fail(PM_E_INVALID, "batch was not submitted");
    b->submitted = false;
    return finish(b, out);
}

int pm_batch_rescan(pm_batch *b, pm_result_view *out) {
    if (!b || !b->staged) return fail(PM_E_INVALID, "batch is not staged");
    int rc = scan_async(b);
    if (rc) return rc;
    return finish(b, out);
}

void pm_batch_release(pm_batch *b) {
    if (!b) return;
    release_work(b);
    put(b->ctx->pinned, b->r_soa);
    put(b->ctx->pinned, b->r_pod);
    lib_unref(b->lib);
    delete b;
}

int64_t pm_batch_line_index(pm_batch *b, int pod, int64_t *out, int64_t cap) {
    if (!b || pod < 0 || pod >= (int)b->pods.size()) return fail(PM_E_INVALID, "bad pod index");
    if (!b->staged || b->pod_units[pod].second != 1)
        return fail(PM_E_INVALID, b->staged ? "the log was cut across devices: no single line table" : "batch is not staged");
    const PodRef &p = b->units[b->pod_units[pod].first];
    DevWork *w = nullptr;
    for (auto &x : b->work)
        if (x.slot == p.dev) w = &x;
    if (!w || !w->scanned) return fail(PM_E_INVALID, "batch has not been scanned");
    Device *d = b->ctx->devs[w->slot];
    CU(cudaSetDevice(d->id));
    const PodDev *hp = w->h_pods.as<PodDev>();
    int n_lines = hp[p.local].n_lines;
    if (p.len == 0) {
        if (cap >= 1 && out) out[0] = 0;
        return 1;
    }
    if (n_lines == 0) return 0;
    if (!out || cap < n_lines) return fail(PM_E_INVALID, "line table needs %d entries", n_lines);
    int rc;
    if ((rc = get(d->pool, (size_t)n_lines * 8, w->lines_tmp))) return rc;
    CU(launch_line_starts(w->text.as<uint8_t>(), w->tile_prefix.as<unsigned long long>(), hp[p.local].nl_before, n_lines, hp[p.local].start,
                          hp[p.local].eff_end, w->lines_tmp.as<long long>(), w->sms_post, w->stream));
    CU(cudaMemcpyAsync(out, w->lines_tmp.p, (size_t)n_lines * 8, cudaMemcpyDeviceToHost, w->stream));
    CU(cudaStreamSynchronize(w->stream));
    return n_lines;
}

}  // extern "C"

// Accessors for pm_service.cc (same shared object, not part of the C ABI)
namespace pm {
std::string &ctx_response_buffer(pm_ctx *ctx) { return ctx->parse_response; }
bool ctx_current_generation(pm_ctx *ctx, uint64_t *generation) {   // no pointer leaves the lock: a publish may free the library
    std::lock_guard<std::mutex> g(ctx->mu);
    if (!ctx->current) return false;
    *generation = ctx->current->generation;
    return true;
}
const CompiledLibrary &library_compiled(const pm_library *l) { return l->lib; }
uint64_t library_generation(const pm_library *l) { return l->generation; }
void *ctx_pinned_get(pm_ctx *ctx, size_t bytes, size_t *cap) {
    void *p = nullptr;
    if (ctx->pinned.get(bytes, &p, cap) != PM_OK) return nullptr;
    return p;
}
void ctx_pinned_put(pm_ctx *ctx, void *p, size_t cap) { ctx->pinned.put(p, cap); }
// pm_service.cc keeps its request batcher here; the deleter runs in pm_shutdown.
void *&ctx_service_state(pm_ctx *ctx) { return ctx->service_state; }
void (*&ctx_service_deleter(pm_ctx *ctx))(void *) { return ctx->service_deleter; }
std::mutex &ctx_mutex(pm_ctx *ctx) { return ctx->mu; }
int ctx_pipeline_depth(pm_ctx *ctx) { return ctx->devs.empty() ? 1 : ctx->devs[0]->depth; }
pm_library *batch_library_addref(pm_batch *b) {
    if (b->lib) b->lib->refs.fetch_add(1);
    return b->lib;
}
void library_release(pm_library *l) { lib_unref(l); }
void set_thread_error(const char *msg) { g_err = msg ? msg : ""; }
}  // namespace pm
