/*
 * podmortem_scan.h -- C ABI of the B200 log-scan engine.
 *
 * This is the device boundary behind the operator's log-parser call.  The operator reaches the
 * scan through ONE interface, `Uni<AnalysisResult> parseLogs(PodFailureData)` =
 * `POST /parse` (reference: src/main/java/com/redhat/podmortem/operator/service/
 * LogParserRestClient.java:21-39, wrapped by LogParserClient.java:36-55).  The service behind
 * it is not in the reference tree, so the entry points below are what a Java 21 `/parse`
 * resource binds through the Panama FFM API (see INTEGRATION.md for the MethodHandle stubs);
 * the same symbols are bound with ctypes by the tests and the benchmark.
 *
 * Conventions
 *   - every function returns 0 (PM_OK) or a negative pm_status; no exceptions cross the ABI;
 *   - the message for the last failure on the calling thread: pm_last_error();
 *   - buffers that cross the ABI are owned by the C side (pm_host_alloc / result views);
 *     the caller never frees anything but through the matching pm_* call;
 *   - pm_ctx and a published pm_library are thread-safe; a pm_batch belongs to one thread;
 *   - a device holds one batch at a time unless pm_set_pipeline raised that: pm_batch_stage/submit
 *     block while the device is full and until a batch holding it is released (so release a batch
 *     before the same thread stages one more than the depth on the same devices);
 *   - there is no CPU fallback: without a usable CUDA device pm_init fails.
 */
#ifndef PODMORTEM_SCAN_H
#define PODMORTEM_SCAN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* 2: pm_set_pipeline, pm_parse_stats, pm_library_serialize/load, pm_line_index_noalloc, pm_noalloc_last_error
 * 3: pm_batch_add_pod_device */
#define PM_ABI_VERSION 3

typedef enum pm_status {
    PM_OK = 0,
    PM_E_INVALID = -1,       /* bad argument / bad call order */
    PM_E_SYNTAX = -2,        /* regex or library JSON does not parse */
    PM_E_UNSUPPORTED = -3,   /* regex construct outside the supported subset (SPEC.md s.3) */
    PM_E_TOO_COMPLEX = -4,   /* repetition bound or DFA state limit exceeded */
    PM_E_CUDA = -5,          /* CUDA runtime failure (message carries cudaGetErrorString) */
    PM_E_NO_DEVICE = -6,     /* no usable GPU: there is no CPU path */
    PM_E_NOMEM = -7,
    PM_E_NO_LIBRARY = -8,    /* scan submitted before pm_library_publish */
    PM_E_OVERFLOW = -9       /* internal queue overflow that could not be retried */
} pm_status;

typedef struct pm_ctx pm_ctx;
typedef struct pm_library pm_library;
typedef struct pm_batch pm_batch;

/* Where and why a library was rejected.  Surfaces in PatternLibrary status: the hook is
 * PatternSyncService.validatePatterns (PatternSyncService.java:228-253), the user-visible error
 * path is PatternLibraryReconciler.java:97-101. */
typedef struct pm_compile_report {
    int32_t status;            /* pm_status */
    int32_t pattern_index;     /* index over all patterns of the spec, -1 if not pattern-specific */
    int32_t secondary_index;   /* -1: primary regex, else index into secondary_patterns */
    int32_t regex_offset;      /* byte offset into the regex source, -1 if n/a */
    char pattern_id[64];
    char message[256];
    /* statistics of a successful compile */
    int32_t n_patterns, n_regex_units;
    int32_t n_filtered;        /* primaries served by the literal-gram filter */
    int32_t n_dfa_scanned;     /* primaries served by the combined shared-memory DFA */
    int32_t n_grams, filter_log2_bits, filter_fold_case;
    int32_t dfa_groups, dfa_states, dfa_classes;
    int64_t blob_bytes;
} pm_compile_report;

/* Result of a batch: structure-of-arrays in pinned host memory owned by the batch; valid until
 * pm_batch_release.  One row per event, ordered by (pod order, line_no, pattern_index).  The
 * rows carry everything AnalysisResult needs besides the text itself, which the caller slices
 * from its own copy of the log (fields consumed by the operator: EventService.java:75-83,
 * AnalysisStorageService.java:145-160 and :298-325). */
typedef struct pm_result_view {
    int64_t n_events;
    int64_t n_pods;
    const int32_t *pod_index;      /* order of pm_batch_add_pod */
    const int32_t *pattern_index;
    const int32_t *line_no;        /* 1-based */
    const int64_t *line_start;     /* byte offset in the pod's log */
    const int32_t *line_len;
    const int32_t *ctx_first_line; /* 1-based, inclusive */
    const int32_t *ctx_last_line;
    const int64_t *ctx_start;      /* byte range [ctx_start, ctx_end) in the pod's log */
    const int64_t *ctx_end;
    const double *score;
    /* per pod */
    const uint64_t *pod_tag;
    const int64_t *pod_first_event; /* n_pods + 1 entries: events of pod i are [first[i], first[i+1]) */
    const int32_t *pod_n_lines;
    /* timings of the last run, milliseconds, measured with CUDA events on the batch's streams
     * (max over devices) */
    double ms_h2d, ms_scan, ms_post, ms_d2h;
    /* first kernel of the scan to the results on the host, one CUDA-event pair per device, max over
     * devices (no host round trip inside unless a queue overflowed or the batch has more than 131072
     * raw events; with several batches in flight it includes waiting for SMs other batches hold) */
    double ms_total;
    /* the combined-DFA kernel(s) alone (0 when the library has no DFA-scanned pattern) */
    double ms_dfa;
    /* counters */
    int64_t bytes_scanned, n_candidates, n_verified, gpu_launches;
} pm_result_view;

/* ---- context -------------------------------------------------------------------------- */
int pm_abi_version(void);
/* n_devices <= 0: use every visible device. */
int pm_init(int n_devices, pm_ctx **out);
/* Explicit device ordinals (one process per GPU: pass the rank's local device). */
int pm_init_devices(const int *device_ids, int n, pm_ctx **out);
void pm_shutdown(pm_ctx *);
const char *pm_last_error(void);
int pm_device_count(const pm_ctx *);
/* Batches in flight per device.  depth 1 (the default): a device runs one batch at a time and every
 * kernel may use all of its SMs - the lowest latency for a single request.  depth 2..8: up to
 * `depth` batches hold a device at once and its SMs are split into two green contexts (driver API,
 * CUDA >= 12.4): the filter kernel of one batch streams on all but ~post_sms SMs while the
 * latency-bound confirm/sort/resolve tail of the batches before it runs on the rest - the
 * throughput mode of a service with requests queued (pm_batch_submit on several batches, then
 * pm_batch_wait in submission order).  The driver rounds post_sms up to its partition granularity;
 * the SM counts actually given to the two partitions of the first device are returned through
 * scan_sms / post_sms_actual (either may be NULL).  Call it with no batch staged.  No counterpart in
 * the reference (its /parse service handles one request per thread). */
int pm_set_pipeline(pm_ctx *, int depth, int post_sms, int *scan_sms, int *post_sms_actual);

/* ---- pattern library (a9; PatternSyncService.java:42-77 hands the files over) ---------- */
/* spec: UTF-8 JSON, one library object or an array of them (SPEC.md s.1). */
int pm_library_compile(pm_ctx *, const uint8_t *spec, size_t n, pm_library **out, pm_compile_report *rep);
/* Replicates the tables to every device of the context and makes the library current. */
int pm_library_publish(pm_ctx *, pm_library *);
/* A compiled library as bytes and back (the sync step,
 * PatternSyncService.java:228-253, compiles once and leaves the result beside the YAML; the /parse
 * service loads it instead of compiling again).  The bytes belong to the library and live as long
 * as it does.  pm_library_load rejects bytes of another compiler version (PM_E_UNSUPPORTED) and
 * damaged ones (PM_E_INVALID); a loaded library is published like a compiled one. */
int pm_library_serialize(pm_library *, const uint8_t **bytes, size_t *n);
int pm_library_load(pm_ctx *, const uint8_t *bytes, size_t n, pm_library **out);
void pm_library_free(pm_library *);
int pm_library_pattern_count(const pm_library *);
/* JSON text with per-pattern id/name/severity/category, for marshalling AnalysisResult. */
const char *pm_library_pattern_meta_json(const pm_library *, int pattern_index);

/* ---- pinned host memory (Java wraps the address as a MemorySegment) ------------------- */
void *pm_host_alloc(pm_ctx *, size_t bytes);
void pm_host_free(pm_ctx *, void *);

/* ---- batches: the unit the /parse resource submits (a1-a3, a10-a14) ------------------- */
int pm_batch_begin(pm_ctx *, pm_batch **out);
/* log: UTF-8 bytes; pinned (pm_host_alloc) memory gives asynchronous copies, pageable works.
 * The bytes must stay valid until pm_batch_wait returns. */
int pm_batch_add_pod(pm_batch *, const uint8_t *log, size_t n, uint64_t pod_tag);
int pm_batch_submit(pm_batch *);                      /* async: H2D + kernels on side streams */
int pm_batch_wait(pm_batch *, pm_result_view *out);   /* blocks the calling thread */
void pm_batch_release(pm_batch *);

/* ---- device-resident corpora (benchmarks; the roofline leg of bench.py) ---------------- */
/* A pod whose bytes already live in device memory on one of the context's GPUs (BASELINE configs 3 and 5:
 * corpora of 17-64 GB per GPU are generated where they are scanned).  The pod is staged on that GPU with a
 * device-to-device copy and is never cut across devices; everything else is as for pm_batch_add_pod.  No
 * counterpart in the reference (its logs arrive over HTTP, PodFailureWatcher.java:319-335). */
int pm_batch_add_pod_device(pm_batch *, const void *d_log, size_t n, uint64_t pod_tag);
/* Stages the batch's pods on their devices without scanning. */
int pm_batch_stage(pm_batch *);
/* Scans the staged bytes again (no H2D); result as in pm_batch_wait. */
int pm_batch_rescan(pm_batch *, pm_result_view *out);

/* ---- newline index (a10), materialised line table of one staged pod -------------------- */
/* Writes line_start offsets (bytes, relative to the pod) of the pod's lines into `out`
 * (host memory, capacity `cap` entries); returns the number of lines or a negative status.  A log that was cut across
 * devices (larger than 2 x PM_SPLIT_MIN_BYTES on a multi-GPU context) has no single table:
 * PM_E_INVALID. */
int64_t pm_batch_line_index(pm_batch *, int pod, int64_t *out, int64_t cap);

/* ---- line table without allocation ------------------------------------------------------ */
/* No counterpart in the reference: this entry point exists for callers of THIS library that are
 * given device buffers but may not create any themselves -- a plugin inside a host process that
 * hands out all GPU memory from its own pool, a latency-critical path that cannot afford
 * cudaMalloc's device-wide synchronisation, a job whose memory budget was fixed at start-up.
 * All other entry points draw working memory from the context's pool, which grows with
 * cudaMalloc; this one calls no allocator at all.  Its scratch is a fixed array inside the
 * loaded module (about 257 KiB on every GPU that loads the library, held until exit), hence:
 *   - n_bytes may not exceed PM_NOALLOC_MAX_BYTES; larger inputs are refused with PM_E_INVALID
 *     (message via pm_noalloc_last_error) and nothing is launched;
 *   - all calls on one GPU share that array: two calls whose kernels can overlap on the device
 *     (different streams, different host threads) race on it and corrupt each other's result.
 *     Serialise them: same stream, or an event between them;
 *   - the first call on each GPU looks up the array's device address (with that GPU current);
 *     make that first call outside any stream capture.  The address is kept for the life of the
 *     process: after a cudaDeviceReset() on that GPU it is stale, so do not call this entry point
 *     again in a process that reset the device;
 *   - a CUDA error left pending on the calling thread by an earlier, unrelated call is neither
 *     cleared nor blamed on this call: the entry point returns PM_E_CUDA and launches nothing.
 * d_text, d_line_start and d_n_lines are DEVICE pointers supplied by the caller; the current
 * device must be the one they live on.  d_line_start[i] receives the byte offset of line i for
 * i < min(*d_n_lines, cap), with the line model of SPEC.md section 2 (the same table
 * pm_batch_line_index returns); *d_n_lines receives the line count.  The work is enqueued on
 * `stream` (a cudaStream_t, NULL = default stream); results are valid once it has run. */
#define PM_NOALLOC_MAX_BYTES (64ull << 20)
int pm_line_index_noalloc(const uint8_t *d_text, uint64_t n_bytes, int64_t *d_line_start, uint64_t cap, uint64_t *d_n_lines, void *stream);
const char *pm_noalloc_last_error(void);

/* ---- the /parse contract as one call (a3, a4, a14) -------------------------------------- */
/* request_json: the PodFailureData body the operator posts (PodFailureWatcher.java:334;
 * fields used: "logs", and pod.metadata.name/namespace for the echo).  On success *response is
 * an AnalysisResult JSON document owned by the context, valid until the next pm_parse on the
 * same thread. */
int pm_parse(pm_ctx *, const uint8_t *request_json, size_t n, const char **response, size_t *response_len);
/* Concurrent pm_parse calls are coalesced into shared GPU batches (one pod per request; SURVEY.md
 * section 8(f).1 - the reference sends one pod per request, PodFailureWatcher.java:196-199).  Requests
 * served so far, the batches they were served in, and the largest batch; any pointer may be NULL. */
int pm_parse_stats(pm_ctx *, int64_t *requests, int64_t *batches, int64_t *largest_batch);

#ifdef __cplusplus
}
#endif
#endif /* PODMORTEM_SCAN_H */
