// The /parse contract as one call: PodFailureData JSON in, AnalysisResult JSON out (rows a3, a4,
// a14 of SURVEY.md section 8).
//
// Reference: the operator serialises `new PodFailureData(pod, podLogs, events)`
// (PodFailureWatcher.java:334, PodmortemReconciler.java:180-196) with Jackson and POSTs it to
// /parse (LogParserRestClient.java:37-39).  Of the response it reads
//   summary.highestSeverity, summary.significantEvents            EventService.java:75-83
//   events.size(), events[0..5).matchedPattern.{name,severity}, .score
//                                                       AnalysisStorageService.java:145-160, :298-325
// and forwards the whole object to the AI interface (AIInterfaceClient.java:51).  The JSON
// property names of the model jar (com.redhat.podmortem:common, pom.xml:83-87) are not in the
// reference tree; the names below follow SPEC.md section 7 (recollection, unpinned).
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <list>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "pm_json.h"
#include "pm_library.h"
#include "podmortem_scan.h"

namespace pm {
bool ctx_current_generation(pm_ctx *ctx, uint64_t *generation);
const CompiledLibrary &library_compiled(const pm_library *l);
uint64_t library_generation(const pm_library *l);
void *ctx_pinned_get(pm_ctx *ctx, size_t bytes, size_t *cap);
void ctx_pinned_put(pm_ctx *ctx, void *p, size_t cap);
void *&ctx_service_state(pm_ctx *ctx);
void (*&ctx_service_deleter(pm_ctx *ctx))(void *);
std::mutex &ctx_mutex(pm_ctx *ctx);
int ctx_pipeline_depth(pm_ctx *ctx);
pm_library *batch_library_addref(pm_batch *b);
void library_release(pm_library *l);
void set_thread_error(const char *msg);
}  // namespace pm

using namespace pm;

namespace {

thread_local std::string t_response;

const char *kSevNames[5] = {"INFO", "LOW", "MEDIUM", "HIGH", "CRITICAL"};

// Result cache.  The operator posts the same failure more than once: once
// per matching Podmortem CR (PodFailureWatcher.java:196-199), again on every reconcile
// (PodmortemReconciler.java:82-86), and again after a restart because its de-duplication map is
// memory-only (PodFailureWatcher.java:50).  Key: pod UID + latest finishedAt + library generation
// + a hash of the log; LRU of 256 responses per process.
struct ResultCache {
    std::mutex mu;
    std::list<std::pair<std::string, std::string>> lru;
    std::unordered_map<std::string, std::list<std::pair<std::string, std::string>>::iterator> index;
    bool get(const std::string &key, std::string &out) {
        std::lock_guard<std::mutex> g(mu);
        auto it = index.find(key);
        if (it == index.end()) return false;
        lru.splice(lru.begin(), lru, it->second);
        out = it->second->second;
        return true;
    }
    void put(const std::string &key, const std::string &val) {
        std::lock_guard<std::mutex> g(mu);
        if (index.count(key)) return;
        lru.emplace_front(key, val);
        index[key] = lru.begin();
        while (lru.size() > 256) {
            index.erase(lru.back().first);
            lru.pop_back();
        }
    }
};
ResultCache g_cache;

void collect_finished_at(const JsonValue &v, std::string &latest) {
    if (v.type == JsonValue::Object) {
        for (auto &kv : v.obj) {
            if (kv.first == "finishedAt" && kv.second.type == JsonValue::String) {
                if (kv.second.str > latest) latest = kv.second.str;   // RFC 3339 timestamps order lexicographically
            } else {
                collect_finished_at(kv.second, latest);
            }
        }
    } else if (v.type == JsonValue::Array) {
        for (auto &x : v.arr) collect_finished_at(x, latest);
    }
}

uint64_t fnv1a(const char *p, size_t n) {
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < n; i++) h = (h ^ (unsigned char)p[i]) * 1099511628211ull;
    return h;
}

// Appends the lines of text[a, b) (separators \r?\n) as JSON strings; returns how many.
void split_append(const char *text, int64_t a, int64_t b, std::vector<std::string> &out) {
    int64_t s = a;
    for (int64_t i = a; i <= b; i++) {
        if (i == b || text[i] == '\n') {
            int64_t e = i;
            if (i < b && e > s && text[e - 1] == '\r') e--;
            out.emplace_back(text + s, (size_t)(e - s));
            s = i + 1;
        }
    }
}

// ---- request batcher -------------------------------------------------
// The reference sends one pod per request and one request per matching CR
// (PodFailureWatcher.java:196-199), so a busy cluster produces many small concurrent /parse calls
// - and a 1 MB log is launch-bound on this GPU (~0.4 ms of fixed cost for ~20 us of scanning).
// Concurrent pm_parse calls are therefore coalesced: every caller queues its decoded log (already
// in pinned memory); whoever finds no batch being formed becomes the leader, takes everything that
// is queued as the pods of ONE batch, runs it and hands each caller its rows back; callers that
// arrive meanwhile form the next batch.  Nobody waits for a timer: a lone request is a batch of
// one and pays nothing.  JSON decoding before and marshalling after run on the callers' own
// threads.  Up to pipeline-depth leaders run at once (pm_set_pipeline), so the copy of one batch
// overlaps the kernels of another.  PM_PARSE_COALESCE=0 turns it off (one batch per request).
struct ParseJob {
    const uint8_t *log = nullptr;
    size_t len = 0;
    bool done = false;
    int rc = PM_OK;
    std::string err;
    pm_library *lib = nullptr;   // the library the batch ran with (one reference, released by the caller)
    int32_t n_lines = 0;
    std::vector<int32_t> pattern, line_no, ctx_first;
    std::vector<int64_t> ctx_start, ctx_end;
    std::vector<double> score;
};

struct Batcher {
    std::mutex mu;
    std::condition_variable cv;
    std::deque<ParseJob *> queue;
    int leaders = 0;
    int64_t n_requests = 0, n_batches = 0, max_batch = 0;
    static constexpr size_t kMaxJobs = 1024, kMaxBytes = 4ull << 30;
};

Batcher &batcher_of(pm_ctx *ctx) {
    std::lock_guard<std::mutex> g(ctx_mutex(ctx));
    void *&slot = ctx_service_state(ctx);
    if (!slot) {
        slot = new Batcher();
        ctx_service_deleter(ctx) = [](void *p) { delete static_cast<Batcher *>(p); };
    }
    return *static_cast<Batcher *>(slot);
}

// One batch for `jobs`; fills every job (rows or error).
void run_jobs(pm_ctx *ctx, const std::vector<ParseJob *> &jobs) {
    pm_batch *b = nullptr;
    pm_result_view v;
    int rc = pm_batch_begin(ctx, &b);
    for (size_t i = 0; !rc && i < jobs.size(); i++) rc = pm_batch_add_pod(b, jobs[i]->log, jobs[i]->len, i);
    if (!rc) rc = pm_batch_submit(b);
    if (!rc) rc = pm_batch_wait(b, &v);
    if (rc) {
        const char *msg = pm_last_error();
        for (ParseJob *j : jobs) {
            j->rc = rc;
            j->err = msg ? msg : "";
        }
        if (b) pm_batch_release(b);
        return;
    }
    for (size_t i = 0; i < jobs.size(); i++) {
        ParseJob *j = jobs[i];
        const int64_t a = v.pod_first_event[i], e = v.pod_first_event[i + 1];
        j->pattern.assign(v.pattern_index + a, v.pattern_index + e);
        j->line_no.assign(v.line_no + a, v.line_no + e);
        j->ctx_first.assign(v.ctx_first_line + a, v.ctx_first_line + e);
        j->ctx_start.assign(v.ctx_start + a, v.ctx_start + e);
        j->ctx_end.assign(v.ctx_end + a, v.ctx_end + e);
        j->score.assign(v.score + a, v.score + e);
        j->n_lines = v.pod_n_lines[i];
        j->lib = batch_library_addref(b);
        j->rc = PM_OK;
    }
    pm_batch_release(b);
}

void run_coalesced(pm_ctx *ctx, ParseJob &me) {
    static const bool off = getenv("PM_PARSE_COALESCE") && getenv("PM_PARSE_COALESCE")[0] == '0';
    if (off) {
        run_jobs(ctx, {&me});
        return;
    }
    Batcher &B = batcher_of(ctx);
    const int max_leaders = ctx_pipeline_depth(ctx);
    std::unique_lock<std::mutex> lk(B.mu);
    B.queue.push_back(&me);
    while (!me.done) {
        if (B.leaders >= max_leaders || B.queue.empty()) {
            B.cv.wait(lk);
            continue;
        }
        // lead: everything queued so far (bounded), in arrival order
        B.leaders++;
        std::vector<ParseJob *> take;
        size_t bytes = 0;
        while (!B.queue.empty() && take.size() < Batcher::kMaxJobs && (take.empty() || bytes + B.queue.front()->len <= Batcher::kMaxBytes)) {
            bytes += B.queue.front()->len;
            take.push_back(B.queue.front());
            B.queue.pop_front();
        }
        B.n_requests += (int64_t)take.size();
        B.n_batches++;
        if ((int64_t)take.size() > B.max_batch) B.max_batch = (int64_t)take.size();
        lk.unlock();
        run_jobs(ctx, take);
        lk.lock();
        for (ParseJob *j : take) j->done = true;
        B.leaders--;
        B.cv.notify_all();
    }
}

}  // namespace

extern "C" int pm_parse_stats(pm_ctx *ctx, int64_t *requests, int64_t *batches, int64_t *largest_batch) {
    if (!ctx) return PM_E_INVALID;
    Batcher &B = batcher_of(ctx);
    std::lock_guard<std::mutex> g(B.mu);
    if (requests) *requests = B.n_requests;
    if (batches) *batches = B.n_batches;
    if (largest_batch) *largest_batch = B.max_batch;
    return PM_OK;
}

extern "C" int pm_parse(pm_ctx *ctx, const uint8_t *request_json, size_t n, const char **response, size_t *response_len) {
    if (!ctx || !request_json || !response) return PM_E_INVALID;
    auto t0 = std::chrono::steady_clock::now();
    // Streaming ingest: the "logs" member is located without building a DOM for
    // it and decoded straight into pinned memory (asynchronous H2D), instead of JSON -> String ->
    // byte[] -> pageable copy.  The rest of the request (a few KB of pod metadata) goes through the DOM.
    const char *body = reinterpret_cast<const char *>(request_json);
    size_t lb = 0, le = 0;
    const bool has_logs = json_find_top_level_string(body, n, "logs", &lb, &le);
    std::string rest;
    if (has_logs) {
        rest.reserve(n - (le - lb));
        rest.append(body, lb);
        rest.append(body + le, n - le);
    }
    JsonValue req;
    std::string err;
    if (!json_parse(has_logs ? rest.data() : body, has_logs ? rest.size() : n, req, err)) return PM_E_SYNTAX;
    size_t pinned_cap = 0, log_len = 0;
    char *pinned = nullptr;
    struct PinnedGuard {
        pm_ctx *ctx;
        char *&p;
        size_t &cap;
        ~PinnedGuard() {
            if (p) ctx_pinned_put(ctx, p, cap);
        }
    } guard{ctx, pinned, pinned_cap};
    if (has_logs && le > lb) {
        pinned = static_cast<char *>(ctx_pinned_get(ctx, le - lb, &pinned_cap));
        if (!pinned) return PM_E_NOMEM;
        log_len = json_decode_string(body + lb, le - lb, pinned);
        if (log_len == (size_t)-1) return PM_E_SYNTAX;
    }
    struct LogView {
        const char *p;
        size_t n;
        const char *data() const { return p; }
        size_t size() const { return n; }
    } log{pinned ? pinned : "", log_len};
    std::string pod_name, pod_ns, pod_uid, finished;
    if (const JsonValue *pod = req.get("pod")) {
        if (const JsonValue *md = pod->get("metadata")) {
            if (const JsonValue *v = md->get("name")) pod_name = v->str;
            if (const JsonValue *v = md->get("namespace")) pod_ns = v->str;
            if (const JsonValue *v = md->get("uid")) pod_uid = v->str;
        }
        if (const JsonValue *st = pod->get("status")) collect_finished_at(*st, finished);
    }
    // Only the generation number of the current library is read here, under the context lock: a
    // concurrent pm_library_publish may drop the last reference to that library at any moment, so no
    // pointer to it is kept.  The scan below pins the library it actually runs with (job.lib).
    uint64_t generation = 0;
    if (!ctx_current_generation(ctx, &generation)) return PM_E_NO_LIBRARY;
    std::string cache_prefix, cache_tail;
    if (!pod_uid.empty()) {
        char buf[48];
        snprintf(buf, sizeof buf, "|%016llx|%zu", (unsigned long long)fnv1a(log.data(), log.size()), log.size());
        cache_prefix = pod_uid + "|" + finished + "|";
        cache_tail = buf;
        if (g_cache.get(cache_prefix + std::to_string((unsigned long long)generation) + cache_tail, t_response)) {
            *response = t_response.c_str();
            if (response_len) *response_len = t_response.size();
            return PM_OK;
        }
    }

    ParseJob job;
    job.log = reinterpret_cast<const uint8_t *>(log.data());
    job.len = log.size();
    run_coalesced(ctx, job);
    struct LibGuard {
        pm_library *l;
        ~LibGuard() {
            if (l) library_release(l);
        }
    } lib_guard{job.lib};
    if (job.rc) {
        set_thread_error(job.err.c_str());
        return job.rc;
    }
    pm_library *lib = job.lib;   // the library this request was scanned with (a publish may have happened since)
    const int64_t n_events = (int64_t)job.pattern.size();
    const CompiledLibrary &cl = library_compiled(lib);
    const PatternDesc *pd = reinterpret_cast<const PatternDesc *>(cl.blob.data() + cl.header()->off_patterns);

    std::string &o = t_response;
    o.clear();
    char id[64];
    snprintf(id, sizeof id, "%016llx", (unsigned long long)std::chrono::system_clock::now().time_since_epoch().count());
    o += "{\"analysisId\":\"";
    o += id;
    o += "\",\"events\":[";
    long dist[5] = {0, 0, 0, 0, 0};
    int top = -1;
    long significant = 0;
    std::vector<std::string> lines;
    for (int64_t i = 0; i < n_events; i++) {
        int p = job.pattern[i];
        int sev = (int)pd[p].severity;
        dist[sev]++;
        if (sev > top) top = sev;
        if (job.score[i] >= 0.5) significant++;
        if (i) o += ",";
        o += "{\"lineNumber\":" + std::to_string(job.line_no[i]);
        o += ",\"matchedPattern\":";
        o += cl.patterns[p].meta_json;
        lines.clear();
        split_append(log.data(), job.ctx_start[i], job.ctx_end[i], lines);
        int before = job.line_no[i] - job.ctx_first[i];
        o += ",\"context\":{\"linesBefore\":[";
        for (int k = 0; k < before && k < (int)lines.size(); k++) {
            if (k) o += ",";
            json_escape(lines[k], o);
        }
        o += "],\"matchedLine\":";
        json_escape(before < (int)lines.size() ? lines[before] : std::string(), o);
        o += ",\"linesAfter\":[";
        for (int k = before + 1; k < (int)lines.size(); k++) {
            if (k > before + 1) o += ",";
            json_escape(lines[k], o);
        }
        o += "]},\"score\":" + json_number(job.score[i]) + "}";
    }
    o += "],\"summary\":{\"significantEvents\":" + std::to_string(significant);
    o += ",\"highestSeverity\":\"";
    o += top >= 0 ? kSevNames[top] : "NONE";
    o += "\",\"totalEvents\":" + std::to_string((long long)n_events);
    o += ",\"severityDistribution\":{";
    for (int s = 4; s >= 0; s--) {
        o += "\"";
        o += kSevNames[s];
        o += "\":" + std::to_string(dist[s]);
        if (s) o += ",";
    }
    o += "}},\"metadata\":{\"podName\":";
    json_escape(pod_name, o);
    o += ",\"podNamespace\":";
    json_escape(pod_ns, o);
    o += ",\"totalLines\":" + std::to_string(job.n_lines);
    double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    o += ",\"processingTimeMs\":" + json_number(ms);
    o += ",\"patternsUsed\":" + std::to_string((long long)cl.patterns.size());
    o += ",\"engine\":\"podmortem-scan-b200\"}}";
    if (!cache_prefix.empty()) g_cache.put(cache_prefix + std::to_string((unsigned long long)library_generation(lib)) + cache_tail, o);
    *response = o.c_str();
    if (response_len) *response_len = o.size();
    return PM_OK;
}
