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
#include <cstdio>
#include <cstring>
#include <list>
#include <mutex>
#include <string>
#include <unordered_map>

#include "pm_json.h"
#include "pm_library.h"
#include "podmortem_scan.h"

namespace pm {
pm_library *ctx_current_library(pm_ctx *ctx);
const CompiledLibrary &library_compiled(const pm_library *l);
uint64_t library_generation(const pm_library *l);
void *ctx_pinned_get(pm_ctx *ctx, size_t bytes, size_t *cap);
void ctx_pinned_put(pm_ctx *ctx, void *p, size_t cap);
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

}  // namespace

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
    pm_library *lib = ctx_current_library(ctx);
    if (!lib) return PM_E_NO_LIBRARY;
    std::string cache_key;
    if (!pod_uid.empty()) {
        char buf[64];
        snprintf(buf, sizeof buf, "|%llu|%016llx|%zu", (unsigned long long)library_generation(lib), (unsigned long long)fnv1a(log.data(), log.size()), log.size());
        cache_key = pod_uid + "|" + finished + buf;
        if (g_cache.get(cache_key, t_response)) {
            *response = t_response.c_str();
            if (response_len) *response_len = t_response.size();
            return PM_OK;
        }
    }

    pm_batch *b = nullptr;
    int rc = pm_batch_begin(ctx, &b);
    if (rc) return rc;
    pm_result_view v;
    rc = pm_batch_add_pod(b, reinterpret_cast<const uint8_t *>(log.data()), log.size(), 0);
    if (!rc) rc = pm_batch_submit(b);
    if (!rc) rc = pm_batch_wait(b, &v);
    if (rc) {
        pm_batch_release(b);
        return rc;
    }
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
    for (int64_t i = 0; i < v.n_events; i++) {
        int p = v.pattern_index[i];
        int sev = (int)pd[p].severity;
        dist[sev]++;
        if (sev > top) top = sev;
        if (v.score[i] >= 0.5) significant++;
        if (i) o += ",";
        o += "{\"lineNumber\":" + std::to_string(v.line_no[i]);
        o += ",\"matchedPattern\":";
        o += cl.patterns[p].meta_json;
        lines.clear();
        split_append(log.data(), v.ctx_start[i], v.ctx_end[i], lines);
        int before = v.line_no[i] - v.ctx_first_line[i];
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
        o += "]},\"score\":" + json_number(v.score[i]) + "}";
    }
    o += "],\"summary\":{\"significantEvents\":" + std::to_string(significant);
    o += ",\"highestSeverity\":\"";
    o += top >= 0 ? kSevNames[top] : "NONE";
    o += "\",\"totalEvents\":" + std::to_string((long long)v.n_events);
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
    o += ",\"totalLines\":" + std::to_string(v.n_pods ? v.pod_n_lines[0] : 0);
    double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    o += ",\"processingTimeMs\":" + json_number(ms);
    o += ",\"patternsUsed\":" + std::to_string((long long)cl.patterns.size());
    o += ",\"engine\":\"podmortem-scan-b200\"}}";
    pm_batch_release(b);
    if (!cache_key.empty()) g_cache.put(cache_key, o);
    *response = o.c_str();
    if (response_len) *response_len = o.size();
    return PM_OK;
}
