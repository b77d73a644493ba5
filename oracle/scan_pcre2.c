/* TEST INFRASTRUCTURE (oracle) -- never linked into or called by the product path.
 *
 * C restatement of the log-scan loop the operator reaches through POST /parse
 * (LogParserRestClient.java:37-39): split the log into lines (Java split("\\r?\\n")), then for
 * every line and every pattern run an unanchored regex search.  The matching engine of the
 * reference, java.util.regex of the JDK 21 runtime (pom.xml:10), is absent from this image
 * (no JVM); PCRE2 10.x (libpcre2-8.so.0, loaded with dlopen because the image has no headers)
 * stands in as a backtracking engine of the same family.  Patterns arrive already translated
 * to PCRE2 syntax by oracle/javaregex.py (to_pcre2).  Parity unpinned: the reference holds no
 * tests or vectors for this path.
 *
 * Used (a) as a second opinion on the Python oracle's regex verdicts and (b) as the timed CPU
 * baseline of bench.py ("port"), multi-threaded over line ranges.
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef struct pcre2_code pcre2_code;
typedef struct pcre2_match_data pcre2_match_data;
typedef int (*fn_match)(const pcre2_code *, const uint8_t *, size_t, size_t, uint32_t, pcre2_match_data *, void *);

static void *g_lib;
static pcre2_code *(*p_compile)(const uint8_t *, size_t, uint32_t, int *, size_t *, void *);
static int (*p_jit_compile)(pcre2_code *, uint32_t);
static pcre2_match_data *(*p_md_create)(const pcre2_code *, void *);
static void (*p_md_free)(pcre2_match_data *);
static void (*p_code_free)(pcre2_code *);
static fn_match p_match;

#define PCRE2_UTF 0x00080000u
#define PCRE2_NO_UTF_CHECK 0x40000000u
#define PCRE2_JIT_COMPLETE 0x00000001u

static int load(void) {
    if (g_lib) return 0;
    g_lib = dlopen("libpcre2-8.so.0", RTLD_NOW);
    if (!g_lib) return -1;
    p_compile = dlsym(g_lib, "pcre2_compile_8");
    p_jit_compile = dlsym(g_lib, "pcre2_jit_compile_8");
    p_md_create = dlsym(g_lib, "pcre2_match_data_create_from_pattern_8");
    p_md_free = dlsym(g_lib, "pcre2_match_data_free_8");
    p_code_free = dlsym(g_lib, "pcre2_code_free_8");
    p_match = dlsym(g_lib, "pcre2_match_8");
    return (p_compile && p_md_create && p_md_free && p_code_free && p_match) ? 0 : -1;
}

typedef struct {
    int n;
    pcre2_code **code;
} orc_lib;

/* patterns: n NUL-terminated PCRE2 sources.  Returns NULL and the failing index in *bad. */
orc_lib *orc_compile(const char *const *patterns, int n, int *bad) {
    if (load()) {
        if (bad) *bad = -2;
        return NULL;
    }
    orc_lib *l = calloc(1, sizeof *l);
    l->n = n;
    l->code = calloc((size_t)n, sizeof *l->code);
    for (int i = 0; i < n; i++) {
        int err;
        size_t off;
        l->code[i] = p_compile((const uint8_t *)patterns[i], strlen(patterns[i]), PCRE2_UTF, &err, &off, NULL);
        if (!l->code[i]) {
            if (bad) *bad = i;
            return NULL;
        }
        if (p_jit_compile) p_jit_compile(l->code[i], PCRE2_JIT_COMPLETE);
    }
    return l;
}

void orc_free(orc_lib *l) {
    if (!l) return;
    for (int i = 0; i < l->n; i++)
        if (l->code[i]) p_code_free(l->code[i]);
    free(l->code);
    free(l);
}

typedef struct {
    const orc_lib *lib;
    const uint8_t *log;
    const int64_t *starts;
    const int32_t *lens;
    int64_t a, b;
    int64_t *out_line;
    int32_t *out_pat;
    int64_t cap, n;
} job;

typedef struct {
    int64_t line;
    int32_t pat;
} hit;

static int hit_cmp(const void *a, const void *b) {
    const hit *x = a, *y = b;
    if (x->line != y->line) return x->line < y->line ? -1 : 1;
    return x->pat < y->pat ? -1 : x->pat > y->pat;
}

static void emit(job *j, int64_t line, int32_t pat) {
    if (j->n < j->cap) {
        j->out_line[j->n] = line;
        j->out_pat[j->n] = pat;
    }
    j->n++;
}

static void *worker(void *arg) {
    job *j = arg;
    pcre2_match_data **md = calloc((size_t)j->lib->n, sizeof *md);
    for (int p = 0; p < j->lib->n; p++) md[p] = p_md_create(j->lib->code[p], NULL);
    if (j->lib->n <= 2000) {
        /* the reference's shape: for every line, every pattern */
        for (int64_t i = j->a; i < j->b; i++)
            for (int p = 0; p < j->lib->n; p++) {
                int rc = p_match(j->lib->code[p], j->log + j->starts[i], (size_t)j->lens[i], 0, PCRE2_NO_UTF_CHECK, md[p], NULL);
                if (rc >= 0) emit(j, i, p);
            }
    } else {
        /* Large libraries (config 4: 20 000 patterns, ~20 MB of JIT code): the same (line, pattern) searches in blocks of
         * 1024 lines with the pattern loop outside, so that a pattern's code is fetched once per block instead of once
         * per line.  The hits of a block are put back in (line, pattern) order before they are emitted. */
        enum { BLOCK = 1024 };
        size_t hcap = 4096, hn;
        hit *hits = malloc(hcap * sizeof *hits);
        for (int64_t b0 = j->a; b0 < j->b; b0 += BLOCK) {
            const int64_t b1 = b0 + BLOCK < j->b ? b0 + BLOCK : j->b;
            hn = 0;
            for (int p = 0; p < j->lib->n; p++)
                for (int64_t i = b0; i < b1; i++) {
                    int rc = p_match(j->lib->code[p], j->log + j->starts[i], (size_t)j->lens[i], 0, PCRE2_NO_UTF_CHECK, md[p], NULL);
                    if (rc >= 0) {
                        if (hn == hcap) hits = realloc(hits, (hcap *= 2) * sizeof *hits);
                        hits[hn].line = i, hits[hn].pat = p, hn++;
                    }
                }
            qsort(hits, hn, sizeof *hits, hit_cmp);
            for (size_t k = 0; k < hn; k++) emit(j, hits[k].line, hits[k].pat);
        }
        free(hits);
    }
    for (int p = 0; p < j->lib->n; p++) p_md_free(md[p]);
    free(md);
    return NULL;
}

/* Java split("\\r?\\n") line table (SPEC.md section 2).  Returns the number of lines. */
int64_t orc_split(const uint8_t *log, int64_t n, int64_t *starts, int32_t *lens, int64_t cap) {
    int64_t k = 0, s = 0;
    if (n == 0) {
        if (cap > 0) starts[0] = 0, lens[0] = 0;
        return 1;
    }
    int found = 0;
    for (;;) {
        const uint8_t *q = memchr(log + s, '\n', (size_t)(n - s));
        int64_t e = q ? (int64_t)(q - log) : n, len = e - s;
        if (q) found = 1;
        if (q && len > 0 && log[e - 1] == '\r') len--;
        if (k < cap) starts[k] = s, lens[k] = (int32_t)len;
        k++;
        if (!q) break;
        s = e + 1;
    }
    if (found)
        while (k > 0 && (k > cap || lens[k - 1] == 0)) k--;
    return k;
}

/* Every (line, pattern) with a match, ordered by line then pattern within each thread's
 * range; ranges are consecutive, so the concatenation is globally ordered.  Returns the total
 * number of matches (which may exceed cap; then only cap were written). */
int64_t orc_scan(const orc_lib *lib, const uint8_t *log, const int64_t *starts, const int32_t *lens, int64_t n_lines, int threads,
                 int64_t *out_line, int32_t *out_pat, int64_t cap) {
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    job jobs[256];
    pthread_t th[256];
    int64_t per = (n_lines + threads - 1) / threads;
    int64_t *tl[256];
    int32_t *tp[256];
    for (int t = 0; t < threads; t++) {
        job *j = &jobs[t];
        memset(j, 0, sizeof *j);
        j->lib = lib, j->log = log, j->starts = starts, j->lens = lens;
        j->a = t * per < n_lines ? t * per : n_lines;
        j->b = (t + 1) * per < n_lines ? (t + 1) * per : n_lines;
        j->cap = cap;
        tl[t] = j->out_line = malloc((size_t)(cap > 0 ? cap : 1) * sizeof(int64_t));
        tp[t] = j->out_pat = malloc((size_t)(cap > 0 ? cap : 1) * sizeof(int32_t));
        pthread_create(&th[t], NULL, worker, j);
    }
    int64_t total = 0;
    for (int t = 0; t < threads; t++) {
        pthread_join(th[t], NULL);
        job *j = &jobs[t];
        int64_t w = j->n < cap ? j->n : cap;
        for (int64_t i = 0; i < w && total + i < cap; i++) out_line[total + i] = tl[t][i], out_pat[total + i] = tp[t][i];
        total += j->n;
        free(tl[t]);
        free(tp[t]);
    }
    return total;
}
