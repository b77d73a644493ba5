// Synthetic crash-log corpus generator: bench and test INPUT only.
//
// Counter-based: the text of line i depends on (seed, i) alone, so any range of lines can be
// produced by any thread, and a rank of a multi-GPU run gets its own range through first_line.
// Shape: "<ts> <LEVEL> [<logger>] (<thread>) <message>" lines in the Quarkus console format the
// operator's pods emit, crash blocks (exception line, "\tat" frames, "Caused by:"), ~0.5 % long
// lines, and with probability p_hit a line carrying a witness string of one library pattern
// (optionally followed by a witness of that pattern's first secondary pattern).
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace {

inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

struct Rng {
    uint64_t s;
    uint64_t next() {
        s += 0x9E3779B97F4A7C15ull;
        return splitmix64(s);
    }
    uint32_t below(uint32_t n) { return (uint32_t)(next() % n); }
    double unit() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
};

const char *kLoggers[] = {
    "io.quarkus", "io.quarkus.runtime", "io.quarkus.arc.processor", "io.quarkus.vertx.http.runtime",
    "io.vertx.core.http.impl", "io.vertx.core.net.impl", "io.smallrye.mutiny", "io.smallrye.config",
    "org.hibernate.SQL", "org.hibernate.engine.jdbc", "org.jboss.resteasy.reactive", "org.jboss.threads",
    "org.acme.orders.OrderResource", "org.acme.orders.OrderService", "org.acme.billing.InvoiceJob",
    "org.acme.inventory.StockSync", "org.acme.catalog.SearchIndexer", "org.acme.auth.TokenFilter",
    "io.netty.channel.nio", "io.netty.handler.ssl", "org.apache.kafka.clients.consumer",
    "org.apache.kafka.clients.producer", "io.micrometer.core", "io.opentelemetry.sdk",
    "com.zaxxer.pool", "io.agroal.pool", "org.infinispan.client", "org.eclipse.microprofile.health",
    "io.fabric8.kubernetes.client", "org.quartz.core", "org.flywaydb.core", "io.grpc.internal",
};
const char *kLevels[] = {"INFO ", "DEBUG", "WARN ", "ERROR"};
const char *kVerbs[] = {"Processed", "Accepted", "Dispatched", "Completed", "Scheduled", "Flushed", "Loaded", "Cached",
                        "Published", "Consumed", "Committed", "Acquired", "Released", "Refreshed", "Indexed", "Rendered"};
const char *kNouns[] = {"request", "order", "invoice", "batch", "message", "record", "session", "snapshot",
                        "segment", "partition", "page", "document", "event", "task", "job", "span"};
const char *kPaths[] = {"/api/v1/orders", "/api/v1/orders/items", "/api/v2/invoices", "/q/health/live", "/q/health/ready",
                        "/q/metrics", "/api/v1/catalog/search", "/api/v1/stock/levels", "/auth/realms/acme/token",
                        "/api/v1/customers", "/api/v1/shipments/track", "/graphql"};
const char *kExc[] = {"java.lang.IllegalStateException", "java.lang.IllegalArgumentException", "java.lang.NullPointerException",
                      "java.util.concurrent.CompletionException", "java.util.NoSuchElementException",
                      "jakarta.ws.rs.ProcessingException", "jakarta.persistence.PersistenceException",
                      "org.acme.orders.OrderNotFoundException", "java.lang.UnsupportedOperationException",
                      "java.util.concurrent.RejectedExecutionException"};
const char *kFramePkgs[] = {"org.acme.orders", "org.acme.billing", "org.acme.inventory", "io.quarkus.arc.impl", "io.quarkus.vertx.core.runtime",
                            "io.vertx.core.impl", "org.jboss.resteasy.reactive.server.handlers", "io.smallrye.mutiny.operators",
                            "org.hibernate.internal", "java.base/java.util.concurrent", "java.base/java.lang", "io.netty.util.concurrent"};
const char *kFrameCls[] = {"OrderService", "OrderResource", "InvoiceJob", "StockSync", "ClientProxy", "VertxCoreRecorder", "ContextImpl",
                           "InvocationHandler", "UniOnItemTransform", "SessionImpl", "ThreadPoolExecutor", "Thread", "FastThreadLocalRunnable",
                           "AbstractResteasyReactiveContext", "DuplicatedContext", "EventLoopContext"};
const char *kFrameMeth[] = {"handle", "run", "invoke", "apply", "accept", "call", "execute", "process", "onItem", "subscribe", "dispatch", "lambda$start$0"};

template <size_t N>
constexpr uint32_t count_of(const char *(&)[N]) {
    return (uint32_t)N;
}

struct Out {
    std::string &s;
    void str(const char *p) { s.append(p); }
    void ch(char c) { s.push_back(c); }
    void num(uint64_t v, int width = 0) {
        char b[24];
        int n = 0;
        do {
            b[n++] = (char)('0' + v % 10);
            v /= 10;
        } while (v);
        while (n < width) b[n++] = '0';
        while (n) s.push_back(b[--n]);
    }
    void hex(uint64_t v, int digits) {
        static const char *d = "0123456789abcdef";
        for (int i = digits - 1; i >= 0; i--) s.push_back(d[(v >> (4 * i)) & 15]);
    }
};

struct Params {
    uint64_t seed;
    double p_hit;
    const char *const *wit;
    const char *const *swit;
    int n_wit;
    bool crlf;
};

constexpr uint64_t kBlock = 64;   // crash blocks are laid out per 64-line block

inline Rng line_rng(const Params &p, uint64_t i) { return Rng{splitmix64(p.seed ^ (i * 0xD1342543DE82EF95ull))}; }
inline Rng block_rng(const Params &p, uint64_t b) { return Rng{splitmix64(~p.seed ^ (b * 0xA24BAED4963EE407ull))}; }

// Is line i a hit line, and for which pattern?  (first draws of the line's stream)
inline int hit_pattern(const Params &p, uint64_t i) {
    if (p.n_wit <= 0 || p.p_hit <= 0) return -1;
    Rng r = line_rng(p, i);
    if (r.unit() >= p.p_hit) return -1;
    return (int)r.below((uint32_t)p.n_wit);
}

void prefix(Out &o, Rng &r, uint64_t i, int level) {
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
    o.str(kLevels[level]);
    o.str(" [");
    o.str(kLoggers[r.below(count_of(kLoggers))]);
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

void benign_message(Out &o, Rng &r) {
    switch (r.below(12)) {
        case 0: case 1: case 2:
            o.str(kVerbs[r.below(count_of(kVerbs))]);
            o.ch(' ');
            o.str(kNouns[r.below(count_of(kNouns))]);
            o.str(" id=");
            o.hex(r.next(), 16);
            o.str(" in ");
            o.num(r.below(900));
            o.str("ms");
            break;
        case 3: case 4:
            o.str("HTTP ");
            o.str(r.below(4) ? "GET " : "POST ");
            o.str(kPaths[r.below(count_of(kPaths))]);
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
            o.str(kNouns[r.below(count_of(kNouns))]);
            o.str("s partition=");
            o.num(r.below(48));
            o.str(" offset=");
            o.num(r.next() % 100000000ull);
            break;
        case 9:
            o.str("Cache ");
            o.str(r.below(2) ? "hit" : "miss");
            o.str(" key=");
            o.str(kNouns[r.below(count_of(kNouns))]);
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
            o.str(kVerbs[r.below(count_of(kVerbs))]);
            o.ch(' ');
            o.num(1 + r.below(500));
            o.ch(' ');
            o.str(kNouns[r.below(count_of(kNouns))]);
            o.str("s, ");
            o.num(r.below(50));
            o.str(" skipped, queue depth ");
            o.num(r.below(10000));
            break;
    }
}

void long_payload(Out &o, Rng &r) {
    uint32_t target = 400 + r.below(3600);
    size_t start = o.s.size();
    o.str("Payload dump: {");
    while (o.s.size() - start < target) {
        o.ch('"');
        o.str(kNouns[r.below(count_of(kNouns))]);
        o.str("\":\"");
        o.hex(r.next(), 16);
        o.str("\",");
    }
    o.str("\"end\":true}");
}

void frame(Out &o, Rng &r) {
    o.str("\tat ");
    o.str(kFramePkgs[r.below(count_of(kFramePkgs))]);
    o.ch('.');
    const char *cls = kFrameCls[r.below(count_of(kFrameCls))];
    o.str(cls);
    o.ch('.');
    o.str(kFrameMeth[r.below(count_of(kFrameMeth))]);
    o.ch('(');
    o.str(cls);
    o.str(".java:");
    o.num(1 + r.below(900));
    o.ch(')');
}

void gen_line(const Params &p, uint64_t i, std::string &s) {
    Out o{s};
    Rng r = line_rng(p, i);
    // -- hit line?
    bool hit = p.n_wit > 0 && p.p_hit > 0 && r.unit() < p.p_hit;
    int hp = hit ? (int)r.below((uint32_t)p.n_wit) : -1;
    if (hit) {
        prefix(o, r, i, 3);
        o.str(p.wit[hp]);
        return;
    }
    // -- secondary witness right after a hit line (50 %)
    if (i > 0 && p.swit) {
        int prev = hit_pattern(p, i - 1);
        if (prev >= 0 && p.swit[prev] && r.below(2)) {
            if (p.swit[prev][0] == '\t' || !strncmp(p.swit[prev], "Caused by", 9)) {
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
    uint32_t c_start = br.below(12), c_frames = 10 + br.below(31), c_cause = br.below(2), c_exc = br.below(count_of(kExc));
    if (has_crash && pos >= c_start) {
        uint64_t k = pos - c_start;
        if (k == 0) {
            prefix(o, r, i, 3);
            o.str("Request failed: ");
            o.str(kExc[c_exc]);
            o.str(": ");
            o.str(kNouns[r.below(count_of(kNouns))]);
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
                o.str(kExc[r.below(count_of(kExc))]);
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

}  // namespace

extern "C" int64_t pm_synth_corpus(uint64_t seed, uint64_t first_line, int64_t target_bytes, double p_hit,
                                   const char *const *witnesses, const char *const *sec_witnesses, int n_wit, int crlf,
                                   uint8_t *out, int64_t *n_lines_out, int threads) {
    if (!out || target_bytes < 0) return -1;
    Params p{seed, p_hit, witnesses, sec_witnesses, n_wit, crlf != 0};
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
                    gen_line(p, i, s);
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
