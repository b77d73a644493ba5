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
#include <string>

#include "pm_json.h"
#include "pm_library.h"
#include "podmortem_scan.h"

namespace pm {
pm_library *ctx_current_library(pm_ctx *ctx);
const CompiledLibrary &library_compiled(const pm_library *l);
}  // namespace pm

using namespace pm;

namespace {

thread_local std::string t_response;

const char *kSevNames[5] = {"INFO", "LOW", "MEDIUM", "HIGH", "CRITICAL"};

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
    JsonValue req;
    std::string err;
    if (!json_parse(reinterpret_cast<const char *>(request_json), n, req, err)) return PM_E_SYNTAX;
    const JsonValue *logs = req.get("logs");
    static const std::string kEmpty;
    const std::string &log = (logs && logs->type == JsonValue::String) ? logs->str : kEmpty;
    std::string pod_name, pod_ns;
    if (const JsonValue *pod = req.get("pod"))
        if (const JsonValue *md = pod->get("metadata")) {
            if (const JsonValue *v = md->get("name")) pod_name = v->str;
            if (const JsonValue *v = md->get("namespace")) pod_ns = v->str;
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
    pm_library *lib = ctx_current_library(ctx);
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
    *response = o.c_str();
    if (response_len) *response_len = o.size();
    return PM_OK;
}
