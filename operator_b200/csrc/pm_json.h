// Minimal JSON DOM (parse + serialise helpers).  Host-only plumbing for the library spec
// (SPEC.md section 1) and for the /parse bodies: request = Jackson JSON of PodFailureData
// (PodFailureWatcher.java:334), response = AnalysisResult (LogParserRestClient.java:39).
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <utility>
#include <vector>

namespace pm {

struct JsonValue {
    enum Type { Null, Bool, Number, String, Array, Object } type = Null;
    bool b = false;
    double num = 0.0;
    bool is_int = false;
    int64_t inum = 0;
    std::string str;
    std::vector<JsonValue> arr;
    std::vector<std::pair<std::string, JsonValue>> obj;

    const JsonValue *get(const char *key) const;
    // Raw text span of this value inside the parsed buffer (used to echo opaque sub-trees).
    size_t raw_begin = 0, raw_end = 0;
};

// Returns false and fills err (with byte offset) on malformed input.
bool json_parse(const char *data, size_t n, JsonValue &out, std::string &err, size_t *err_off = nullptr);

// Span of the string VALUE of a top-level key (the bytes between its quotes, still escaped).
// Lets a large member (the pod log) be taken out of a request without building a DOM for it.
bool json_find_top_level_string(const char *data, size_t n, const char *key, size_t *begin, size_t *end);
// Decodes the escaped string body [data, data+n) to UTF-8 at `out` (capacity >= n); returns the
// decoded length or (size_t)-1 on a malformed escape.
size_t json_decode_string(const char *data, size_t n, char *out);

void json_escape(const std::string &s, std::string &out);   // appends a quoted JSON string
std::string json_number(double v);                          // shortest round-trip repr

}  // namespace pm
