// Core of the synthetic crash-log generator: bench and test INPUT only.
//
// One definition of "the text of line i", compiled twice: for the host (pm_synth.cc, threads) and
// for the device (pm_synth_gpu.cu, one thread per line), so that a corpus generated on the GPU is
// byte-identical to the host's by construction (configs 3 and 5 of BASELINE.json need corpora of
// 17-64 GB per GPU, which neither fit the host nor should cross PCIe).
//
// Counter-based: the text of line i depends on (seed, i) alone, so any range of lines can be
// produced by any thread, and a rank of a multi-GPU run gets its own range through first_line.
// Shape: "<ts> <LEVEL> [<logger>] (<thread>) <message>" lines in the Quarkus console format the
// operator's pods emit, crash blocks (exception line, "\tat" frames, "Caused by:"), ~0.5 % long
// lines, and with probability p_hit a line carrying a witness string of one library pattern
// (optionally followed by a witness of that pattern's first secondary pattern).
//
// The generator writes through a sink `O` with str / ch / num / hex / size: a std::string on the
// host, a byte counter and a raw pointer on the device.
#pragma once
#include <stddef.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define PM_SYN_HD __host__ __device__
#else
#define PM_SYN_HD
#endif

namespace pm_synth {

PM_SYN_HD inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

struct Rng {
    uint64_t s;
    PM_SYN_HD uint64_t next() {
        s += 0x9E3779B97F4A7C15ull;
        return splitmix64(s);
    }
    PM_SYN_HD uint32_t below(uint32_t n) { return (uint32_t)(next() % n); }
    PM_SYN_HD double unit() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
};

// String tables as functions: a table of pointers to host string literals would not exist on the device.
constexpr uint32_t kLoggers_n = 32;
PM_SYN_HD inline const char *kLoggers(uint32_t i) {
    const char *const t[] = {"io.quarkus", "io.quarkus.runtime", "io.quarkus.arc.processor", "io.quarkus.vertx.http.runtime",
    "io.vertx.core.http.impl", "io.vertx.core.net.impl", "io.smallrye.mutiny", "io.smallrye.config",
    "org.hibernate.SQL", "org.hibernate.engine.jdbc", "org.jboss.resteasy.reactive", "org.jboss.threads",
    "org.acme.orders.OrderResource", "org.acme.orders.OrderService", "org.acme.billing.InvoiceJob",
    "org.acme.inventory.StockSync", "org.acme.catalog.SearchIndexer", "org.acme.auth.TokenFilter",
    "io.netty.channel.nio", "io.netty.handler.ssl", "org.apache.kafka.clients.consumer",
    "org.apache.kafka.clients.producer", "io.micrometer.core", "io.opentelemetry.sdk",
    "com.zaxxer.pool", "io.agroal.pool", "org.infinispan.client", "org.eclipse.microprofile.health",
    "io.fabric8.kubernetes.client", "org.quartz.core", "org.flywaydb.core", "io.grpc.internal",};
    return t[i];
}

constexpr uint32_t kLevels_n = 4;
PM_SYN_HD inline const char *kLevels(uint32_t i) {
    const char *const t[] = {"INFO ", "DEBUG", "WARN ", "ERROR"};
    return t[i];
}

constexpr uint32_t kVerbs_n = 16;
PM_SYN_HD inline const char *kVerbs(uint32_t i) {
    const char *const t[] = {"Processed", "Accepted", "Dispatched", "Completed", "Scheduled", "Flushed", "Loaded", "Cached",
                        "Published", "Consumed", "Committed", "Acquired", "Released", "Refreshed", "Indexed", "Rendered"};
    return t[i];
}

constexpr uint32_t kNouns_n = 16;
PM_SYN_HD inline const char *kNouns(uint32_t i) {
    const char *const t[] = {"request", "order", "invoice", "batch", "message", "record", "session", "snapshot",
                        "segment", "partition", "page", "document", "event", "task", "job", "span"};
    return t[i];
}

constexpr uint32_t kPaths_n = 12;
PM_SYN_HD inline const char *kPaths(uint32_t i) {
    const char *const t[] = {"/api/v1/orders", "/api/v1/orders/items", "/api/v2/invoices", "/q/health/live", "/q/health/ready",
                        "/q/metrics", "/api/v1/catalog/search", "/api/v1/stock/levels", "/auth/realms/acme/token",
                        "/api/v1/customers", "/api/v1/shipments/track", "/graphql"};
    return t[i];
}

constexpr uint32_t kExc_n = 10;
PM_SYN_HD inline const char *kExc(uint32_t i) {
    const char *const t[] = {"java.lang.IllegalStateException", "java.lang.IllegalArgumentException", "java.lang.NullPointerException",
                      "java.util.concurrent.CompletionException", "java.util.NoSuchElementException",
                      "jakarta.ws.rs.ProcessingException", "jakarta.persistence.PersistenceException",
                      "org.acme.orders.OrderNotFoundException", "java.lang.UnsupportedOperationException",
                      "java.util.concurrent.RejectedExecutionException"};
    return t[i];
}

constexpr uint32_t kFramePkgs_n = 12;
PM_SYN_HD inline const char *kFramePkgs(uint32_t i) {
    const char *const t[] = {"org.acme.orders", "org.acme.billing", "org.acme.inventory", "io.quarkus.arc.impl", "io.quarkus.vertx.core.runtime",
                            "io.vertx.core.impl", "org.jboss.resteasy.reactive.server.handlers", "io.smallrye.mutiny.operators",
                            "org.hibernate.internal", "java.base/java.util.concurrent", "java.base/java.lang", "io.netty.util.concurrent"};
    return t[i];
}

constexpr uint32_t kFrameCls_n = 16;
PM_SYN_HD inline const char *kFrameCls(uint32_t i) {
    const char *const t[] = {"OrderService", "OrderResource", "InvoiceJob", "StockSync", "ClientProxy", "VertxCoreRecorder", "ContextImpl",
                           "InvocationHandler", "UniOnItemTransform", "SessionImpl", "ThreadPoolExecutor", "Thread", "FastThreadLocalRunnable",
                           "AbstractResteasyReactiveContext", "DuplicatedContext", "EventLoopContext"};
    return t[i];
}

constexpr uint32_t kFrameMeth_n = 12;
PM_SYN_HD inline const char *kFrameMeth(uint32_t i) {
    const char *const t[] = {"handle", "run", "invoke", "apply", "accept", "call", "execute", "process", "onItem", "subscribe", "dispatch", "lambda$start$0"};
    return t[i];
}

PM_SYN_HD inline bool starts_with(const char *s, const char *prefix) {
    for (; *prefix; s++, prefix++)
        if (*s != *prefix) return false;
    return true;
}

// Shared formatting of the sinks: decimal with optional zero padding, fixed-width hex.
template <class Self>
struct SinkBase {
    PM_SYN_HD void num(uint64_t v, int width = 0) {
        char b[24];
        int n = 0;
        do {
            b[n++] = (char)('0' + v % 10);
            v /= 10;
        } while (v);
        while (n < width) b[n++] = '0';
        while (n) static_cast<Self *>(this)->ch(b[--n]);
    }
    PM_SYN_HD void hex(uint64_t v, int digits) {
        for (int i = digits - 1; i >= 0; i--) {
            const unsigned d = (unsigned)((v >> (4 * i)) & 15);
            static_cast<Self *>(this)->ch((char)(d < 10 ? '0' + d : 'a' + d - 10));
        }
    }
    PM_SYN_HD void str(const char *p) {
        for (; *p; p++) static_cast<Self *>(this)->ch(*p);
    }
};

struct CountSink : SinkBase<CountSink> {   // length only
    size_t n = 0;
    PM_SYN_HD void ch(char) { n++; }
    PM_SYN_HD size_t size() const { return n; }
};

struct PtrSink : SinkBase<PtrSink> {       // writes through a raw pointer
    uint8_t *base, *p;
    PM_SYN_HD explicit PtrSink(uint8_t *at) : base(at), p(at) {}
    PM_SYN_HD void ch(char c) { *p++ = (uint8_t)c; }
    PM_SYN_HD size_t size() const { return (size_t)(p - base); }
};

struct Params {
    uint64_t seed;
    double p_hit;
    const char *const *wit;
    const char *const *swit;
    int n_wit;
    bool crlf;
    // near-miss lines (hit-density sweeps): with probability p_near a line carries one of `near`, strings that share
    // vocabulary and literals with the library's witnesses without (in general) matching a pattern.  Decided on a
    // separate random stream: with p_near = 0 every line is what it was before this field existed.
    double p_near = 0.0;
    const char *const *near = nullptr;
    int n_near = 0;
};

constexpr uint64_t kBlock = 64;   // crash blocks are laid out per 64-line block

PM_SYN_HD inline Rng line_rng(const Params &p, uint64_t i) { return Rng{splitmix64(p.seed ^ (i * 0xD1342543DE82EF95ull))}; }
PM_SYN_HD inline Rng block_rng(const Params &p, uint64_t b) { return Rng{splitmix64(~p.seed ^ (b * 0xA24BAED4963EE407ull))}; }

// Is line i a hit line, and for which pattern?  (first draws of the line's stream)
PM_SYN_HD inline int hit_pattern(const Params &p, uint64_t i) {
    if (p.n_wit <= 0 || p.p_hit <= 0) return -1;
    Rng r = line_rng(p, i);
    if (r.unit() >= p.p_hit) return -1;
    return (int)r.below((uint32_t)p.n_wit);
}

template <class O>
PM_SYN_HD void prefix(O &o, Rng &r, uint64_t i, int level) {
    // 2026-03-14 09:26:53,589 INFO  [org.acm.Foo] (executor-thread-12)
    uint64_t ms = 1773480413000ull + i * 37 + r.below(37);
    uint64_t s = ms / 1000, day = s / 86400, sod = s % 86400;
    o.str("2026-03-");
    o.num(1 + day % 28, 2);
    o.ch(' ');
    o.num(sod / 3600, 2);
    o.ch(':');
    o.num(sod / 60 % 60, 2);
    o.ch(':');
    o.num(sod % 60, 2);
    o.ch(',');
    o.num(ms % 1000, 3);
    o.ch(' ');
    o.str(kLevels(level));
    o.str(" [");
    o.str(kLoggers(r.below(kLoggers_n)));
    o.str("] (");
    uint32_t t = r.below(10);
    if (t < 6) {
        o.str("executor-thread-");
        o.num(r.below(64));
    } else if (t < 9) {
        o.str("vert.x-eventloop-thread-");
        o.num(r.below(16));
    } else {
        o.str("main");
    }
    o.str(") ");
}

template <class O>
PM_SYN_HD void benign_message(O &o, Rng &r) {
    switch (r.below(12)) {
        case 0: case 1: case 2:
            o.str(kVerbs(r.below(kVerbs_n)));
            o.ch(' ');
            o.str(kNouns(r.below(kNouns_n)));
            o.str(" id=");
            o.hex(r.next(), 16);
            o.str(" in ");
            o.num(r.below(900));
            o.str("ms");
            break;
        case 3: case 4:
            o.str("HTTP ");
            o.str(r.below(4) ? "GET " : "POST ");
            o.str(kPaths(r.below(kPaths_n)));
            o.str(" -> ");
            o.num(r.below(20) ? 200 : (r.below(2) ? 404 : 503));
            o.str(" (");
            o.num(r.below(2000));
            o.str(" us) trace=");
            o.hex(r.next(), 16);
            o.hex(r.next(), 16);
            break;
        case 5:
            o.str("select o1_0.id,o1_0.customer_id,o1_0.status,o1_0.total from orders o1_0 where o1_0.id=");
            o.num(r.below(1000000));
            break;
        case 6:
            o.str("Heartbeat ok peers=");
            o.num(r.below(9));
            o.str(" lag=");
            o.num(r.below(5000));
            o.str(" epoch=");
            o.num(r.below(100000));
            break;
        case 7:
            o.str("GC pause (G1 Evacuation Pause) ");
            o.num(r.below(4096));
            o.str("M->");
            o.num(r.below(2048));
            o.str("M(");
            o.num(4096);
            o.str("M) ");
            o.num(r.below(80));
            o.ch('.');
            o.num(r.below(1000), 3);
            o.str("ms");
            break;
        case 8:
            o.str("Partition assignment topic=acme.");
            o.str(kNouns(r.below(kNouns_n)));
            o.str("s partition=");
            o.num(r.below(48));
            o.str(" offset=");
            o.num(r.next() % 100000000ull);
            break;
        case 9:
            o.str("Cache ");
            o.str(r.below(2) ? "hit" : "miss");
            o.str(" key=");
            o.str(kNouns(r.below(kNouns_n)));
            o.ch(':');
            o.hex(r.next(), 8);
            o.ch('-');
            o.hex(r.next(), 4);
            o.ch('-');
            o.hex(r.next(), 4);
            o.ch('-');
            o.hex(r.next(), 12);
            break;
        case 10:
            o.str("Client 10.");
            o.num(r.below(256));
            o.ch('.');
            o.num(r.below(256));
            o.ch('.');
            o.num(r.below(256));
            o.ch(':');
            o.num(1024 + r.below(60000));
            o.str(" authenticated as user");
            o.num(r.below(5000));
            o.str(" roles=[");
            o.str(r.below(2) ? "viewer" : "editor");
            o.str("]");
            break;
        default:
            o.str(kVerbs(r.below(kVerbs_n)));
            o.ch(' ');
            o.num(1 + r.below(500));
            o.ch(' ');
            o.str(kNouns(r.below(kNouns_n)));
            o.str("s, ");
            o.num(r.below(50));
            o.str(" skipped, queue depth ");
            o.num(r.below(10000));
            break;
    }
}

template <class O>
PM_SYN_HD void long_payload(O &o, Rng &r) {
    uint32_t target = 400 + r.below(3600);
    size_t start = o.size();
    o.str("Payload dump: {");
    while (o.size() - start < target) {
        o.ch('"');
        o.str(kNouns(r.below(kNouns_n)));
        o.str("\":\"");
        o.hex(r.next(), 16);
        o.str("\",");
    }
    o.str("\"end\":true}");
}

template <class O>
PM_SYN_HD void frame(O &o, Rng &r) {
    o.str("\tat ");
    o.str(kFramePkgs(r.below(kFramePkgs_n)));
    o.ch('.');
    const char *cls = kFrameCls(r.below(kFrameCls_n));
    o.str(cls);
    o.ch('.');
    o.str(kFrameMeth(r.below(kFrameMeth_n)));
    o.ch('(');
    o.str(cls);
    o.str(".java:");
    o.num(1 + r.below(900));
    o.ch(')');
}

template <class O>
PM_SYN_HD void gen_line(const Params &p, uint64_t i, O &o) {
    Rng r = line_rng(p, i);
    // -- hit line?
    bool hit = p.n_wit > 0 && p.p_hit > 0 && r.unit() < p.p_hit;
    int hp = hit ? (int)r.below((uint32_t)p.n_wit) : -1;
    if (hit) {
        prefix(o, r, i, 3);
        o.str(p.wit[hp]);
        return;
    }
    if (p.n_near > 0 && p.p_near > 0) {
        Rng q{splitmix64(p.seed ^ 0x6E6561726D697373ull ^ (i * 0x9FB21C651E98DF25ull))};
        if (q.unit() < p.p_near) {
            prefix(o, r, i, 2 + (int)q.below(2));
            o.str(p.near[q.below((uint32_t)p.n_near)]);
            return;
        }
    }
    // -- secondary witness right after a hit line (50 %)
    if (i > 0 && p.swit) {
        int prev = hit_pattern(p, i - 1);
        if (prev >= 0 && p.swit[prev] && r.below(2)) {
            if (p.swit[prev][0] == '\t' || starts_with(p.swit[prev], "Caused by")) {
                o.str(p.swit[prev]);
            } else {
                prefix(o, r, i, 2);
                o.str(p.swit[prev]);
            }
            return;
        }
    }
    // -- crash block layout of this 64-line block
    uint64_t b = i / kBlock, pos = i % kBlock;
    Rng br = block_rng(p, b);
    bool has_crash = br.below(6) == 0;
    uint32_t c_start = br.below(12), c_frames = 10 + br.below(31), c_cause = br.below(2), c_exc = br.below(kExc_n);
    if (has_crash && pos >= c_start) {
        uint64_t k = pos - c_start;
        if (k == 0) {
            prefix(o, r, i, 3);
            o.str("Request failed: ");
            o.str(kExc(c_exc));
            o.str(": ");
            o.str(kNouns(r.below(kNouns_n)));
            o.ch(' ');
            o.num(r.below(100000));
            o.str(" in unexpected state");
            return;
        }
        if (k <= c_frames) {
            frame(o, r);
            return;
        }
        if (c_cause) {
            uint64_t k2 = k - c_frames - 1;
            if (k2 == 0) {
                o.str("Caused by: ");
                o.str(kExc(r.below(kExc_n)));
                o.str(": value ");
                o.num(r.below(1000));
                o.str(" rejected");
                return;
            }
            if (k2 <= 6) {
                frame(o, r);
                return;
            }
            if (k2 == 7) {
                o.str("\t... ");
                o.num(5 + r.below(40));
                o.str(" more");
                return;
            }
        }
    }
    // -- ordinary line
    uint32_t lv = r.below(100);
    int level = lv < 70 ? 0 : lv < 85 ? 1 : lv < 95 ? 2 : 3;
    prefix(o, r, i, level);
    if (r.below(200) == 0)
        long_payload(o, r);
    else
        benign_message(o, r);
}


}  // namespace pm_synth
