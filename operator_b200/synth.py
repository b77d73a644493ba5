"""Synthetic pattern libraries and crash-log corpora.

The reference ships no sample log, pattern file or benchmark, so the inputs
of every test and benchmark come from these seeded, counter-based generators.  The library
generator is pure Python; the corpus generator is native (csrc/synth/pm_synth.cc, bound with
ctypes) because BASELINE config 2 needs 1 GiB of text.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, List, Optional, Tuple

import numpy as np

LIB_SEED = 0x504D504154
LOG_SEED = 0x504D4C4F47
_M64 = (1 << 64) - 1


def splitmix64(x: int) -> int:
    x = (x + 0x9E3779B97F4A7C15) & _M64
    z = x
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return z ^ (z >> 31)


class _Rng:
    def __init__(self, seed: int):
        self.s = seed & _M64

    def next(self) -> int:
        self.s = (self.s + 0x9E3779B97F4A7C15) & _M64
        return splitmix64(self.s)

    def below(self, n: int) -> int:
        return self.next() % n

    def pick(self, seq):
        return seq[self.below(len(seq))]


SUBSYSTEMS = [
    "orders-db", "billing-db", "inventory-cache", "session-store", "kafka-broker", "amqp-channel",
    "payment-gateway", "auth-realm", "keycloak-adapter", "ldap-directory", "s3-bucket", "blob-store",
    "grpc-channel", "rest-client", "scheduler-quartz", "flyway-migration", "liquibase-changelog",
    "hibernate-orm", "agroal-pool", "vertx-eventbus", "infinispan-cluster", "redis-sentinel",
    "mongo-replica", "elastic-index", "smtp-relay", "otel-exporter", "vault-secret", "tls-keystore",
    "config-map", "service-mesh", "istio-sidecar", "envoy-listener", "ingress-route", "dns-resolver",
    "jwt-verifier", "oidc-tenant", "rate-limiter", "circuit-breaker", "bulkhead-queue", "saga-coordinator",
    "outbox-relay", "debezium-connector", "schema-registry", "cassandra-ring", "zookeeper-quorum",
    "etcd-lease", "nfs-mount", "pvc-volume", "cgroup-memory", "native-image", "jit-compiler",
    "gc-heap", "metaspace-region", "direct-buffer", "thread-pool", "netty-eventloop", "websocket-session",
    "graphql-resolver", "reactive-pg", "mysql-driver",
]
FAILURES = [
    "connection refused", "handshake failed", "pool exhausted", "lease expired", "quota exceeded",
    "deadlock detected", "checksum mismatch", "schema drift detected", "offset out of range",
    "leader not available", "replica lagging", "split brain suspected", "certificate expired",
    "certificate not yet valid", "hostname verification failed", "token signature invalid",
    "credentials rejected", "permission denied", "read timed out", "write timed out",
    "circuit opened", "retry budget exhausted", "backpressure overflow", "queue overflow",
    "segment corrupted", "index unavailable", "migration failed", "rollback failed",
    "lock wait timeout exceeded", "serialization failure", "constraint violated", "duplicate key rejected",
    "cannot allocate memory", "heap exhausted", "metaspace exhausted", "too many open files",
    "no space left on device", "volume not mounted", "read-only file system", "stale file handle",
    "liveness probe failed", "readiness probe failed", "startup probe failed", "image pull failed",
    "container OOMKilled", "evicted by kubelet", "signal SIGSEGV received", "signal SIGABRT received",
    "native library missing", "class definition not found", "bean resolution ambiguous",
    "unsatisfied dependency", "configuration property missing", "port already in use",
    "address not available", "route not found", "upstream reset before headers", "dns lookup failed",
    "protocol violation", "frame size exceeded",
]
CONTEXTS = [
    "during startup", "during shutdown", "while serving checkout", "while draining requests",
    "while replaying the journal", "while rebalancing partitions", "while rotating credentials",
    "while warming the cache", "after leader election", "after failover", "after config reload",
    "before first request", "on scheduled tick", "on health check", "on transaction commit",
    "on transaction rollback", "in background worker", "in reactive pipeline", "in blocking call",
    "in native mode", "in dev services", "under memory pressure", "under cpu throttling",
    "under network partition", "with stale snapshot", "with expired lease", "with empty response",
    "with partial write", "with malformed payload", "with unexpected status",
    "for tenant default", "for tenant premium", "for region eu-west", "for region us-east",
    "for shard primary", "for shard replica", "at retry limit", "at capacity limit",
    "at rate limit", "at deadline",
]
SEVERITIES = ["CRITICAL", "HIGH", "MEDIUM", "LOW", "INFO"]
SECONDARIES = [
    (r"Caused by: ", "Caused by: java.io.IOException: broken pipe"),
    (r"retrying in \d+ ?ms", "retrying in 250ms"),
    (r"giving up after \d+ attempts", "giving up after 5 attempts"),
    (r"(?i)shutting down", "Shutting down gracefully"),
    (r"exit code \d+", "process finished with exit code 137"),
    (r"\bat io\.quarkus\.", "\tat io.quarkus.runtime.Application.start(Application.java:101)"),
]
_ESC = set(".^$*+?()[]{}|\\")


def _q(s: str) -> str:
    return "".join("\\" + c if c in _ESC else c for c in s)


def make_library(n_patterns: int, seed: int = LIB_SEED, library_id: str = "synthetic") -> Tuple[dict, List[str], List[Optional[str]]]:
    """Returns (library dict per SPEC.md section 1, witness line per pattern, witness of the first
    secondary pattern or None).  Regex mix per SURVEY.md section 8(d): 60 % literal phrases, 30 % literals
    joined by \\d+ \\w+ \\S+ .*, 10 % alternation / classes / (?i)."""
    rng = _Rng(seed)
    total = len(SUBSYSTEMS) * len(FAILURES) * len(CONTEXTS)
    if n_patterns > total:
        raise ValueError("at most %d distinct synthetic patterns" % total)
    stride = 1000003            # coprime with total
    pats, wit, swit = [], [], []
    for i in range(n_patterns):
        k = (i * stride + 12345) % total
        sub = SUBSYSTEMS[k % len(SUBSYSTEMS)]
        fail = FAILURES[(k // len(SUBSYSTEMS)) % len(FAILURES)]
        ctx = CONTEXTS[k // (len(SUBSYSTEMS) * len(FAILURES))]
        kind = rng.below(100)
        if kind < 60:
            phrase = "%s %s %s" % (sub, fail, ctx)
            rx, w = _q(phrase), phrase
        elif kind < 90:
            j = rng.below(4)
            if j == 0:
                rx = "%s %s after \\d+ ?ms %s" % (_q(sub), _q(fail), _q(ctx))
                w = "%s %s after %d ms %s" % (sub, fail, 10 + rng.below(9000), ctx)
            elif j == 1:
                rx = "%s node \\w+ %s %s" % (_q(sub), _q(fail), _q(ctx))
                w = "%s node n%d %s %s" % (sub, rng.below(64), fail, ctx)
            elif j == 2:
                rx = "%s %s for \\S+ %s" % (_q(sub), _q(fail), _q(ctx))
                w = "%s %s for svc-%d.cluster.local:8443 %s" % (sub, fail, rng.below(500), ctx)
            else:
                rx = "%s .*%s %s" % (_q(sub), _q(fail), _q(ctx))
                w = "%s [attempt %d] %s %s" % (sub, 1 + rng.below(9), fail, ctx)
        else:
            j = rng.below(3)
            if j == 0:
                rx = "(?i)%s %s %s" % (_q(sub), _q(fail), _q(ctx))
                w = ("%s %s %s" % (sub, fail, ctx)).upper()
            elif j == 1:
                head, tail = fail.split(" ", 1) if " " in fail else (fail, "")
                rx = "%s [%s%s]%s %s %s" % (_q(sub), head[0].upper(), head[0].lower(), _q(head[1:]), _q(tail), _q(ctx))
                w = "%s %s%s %s %s" % (sub, head[0].upper(), head[1:], tail, ctx)
            else:
                rx = "(ERROR|FATAL|SEVERE) .*%s %s (%s|%s)" % (_q(sub), _q(fail), _q(ctx), _q(CONTEXTS[(k + 7) % len(CONTEXTS)]))
                w = "FATAL [%d] %s %s %s" % (rng.below(1000), sub, fail, ctx)
        secs = []
        sw = None
        if rng.below(100) < 30:
            for t in range(1 + rng.below(2)):
                srx, sexample = SECONDARIES[rng.below(len(SECONDARIES))]
                secs.append({"regex": srx, "weight": round(0.1 + 0.05 * rng.below(9), 2),
                             "proximity_window": 2 + rng.below(19)})
                if t == 0:
                    sw = sexample
        pats.append({
            "id": "syn-%05d" % i,
            "name": "%s: %s" % (sub, fail),
            "severity": SEVERITIES[rng.below(5)],
            "category": sub.split("-")[-1],
            "primary_pattern": {"regex": rx, "confidence": round(0.5 + 0.05 * rng.below(11), 2)},
            "secondary_patterns": secs,
            "context_extraction": {"lines_before": rng.below(11), "lines_after": rng.below(21),
                                   "include_stack_trace": bool(rng.below(2))},
            "remediation": {"description": "check %s" % sub},
        })
        wit.append(w)
        swit.append(sw)
    lib = {"metadata": {"library_id": library_id, "version": "1.0.0"}, "patterns": pats}
    return lib, wit, swit


# Short-literal and class-led primaries: what a hand-written library adds to the long phrases above
# ("\\bOOM\\b", "exit code \\d+", "(?i)error \\d+", "[Cc]ache (evicted|flushed)").  (regex, witness)
SHORT_CODES = ["OOM", "NPE", "EOF", "ANR", "SIGV", "ENOSPC", "EPIPE", "ECONNR", "EAGAIN", "E5XX", "OOMK", "DLQ", "GCX", "SSLX", "TLSX"]
SHORT_PRIMARIES = (
    [("\\b%s\\b" % c, "worker hit %s again" % c) for c in SHORT_CODES] +
    [("exit code %d\\b" % n, "process finished with exit code %d" % n) for n in (1, 2, 3, 126, 127, 134, 137, 139, 143, 255)] +
    [("(?i)%s \\d+" % w, "%s %d" % (w.upper(), 17 + i)) for i, w in enumerate(("errno", "fault", "abort", "crash", "fatal"))] +
    [("[%s%s]%s (evicted|flushed)" % (w[0].upper(), w[0], w[1:]), "%s evicted" % w) for w in ("cache", "buffer", "queue", "journal", "ledger")] +
    [("\\b%s\\d{3,4}\\b" % p_, "code %s%d seen" % (p_, 4000 + i)) for i, p_ in enumerate(("ERR-", "QK", "HHH", "RESTEASY", "SRMSG"))] +
    [("killed by signal %d" % n, "child killed by signal %d" % n) for n in (6, 9, 11, 15)] +
    [("%s: " % w, "goroutine 1 %s: nil map" % w) for w in ("panic", "fatal", "abort")] +
    [("status=5\\d\\d", "upstream status=503 retry"), ("\\btimed? ?out\\b", "request timed out"), ("(?i)segfault", "SEGFAULT at 0x0")]
)


def make_mixed_library(n_patterns: int, fraction: float = 0.1, seed: int = LIB_SEED, min_literal: int = 0) -> Tuple[dict, List[str], List[Optional[str]]]:
    """make_library with every (1/fraction)-th primary replaced by a short-literal / class-led regex.
    min_literal: only templates whose required literal has at least that many bytes (5: what the gram filter can
    still serve by testing every even offset; 0: all, including the 3-byte codes that need the combined DFA)."""
    lib, wit, swit = make_library(n_patterns, seed, "synthetic-mixed")
    step = max(1, int(round(1.0 / fraction)))
    templates = SHORT_PRIMARIES
    if min_literal >= 5:
        short = set("\\b%s\\b" % c for c in SHORT_CODES if len(c) < 5) | {"\\bERR-\\d{3,4}\\b", "\\bQK\\d{3,4}\\b", "\\bHHH\\d{3,4}\\b", "\\btimed? ?out\\b"}
        templates = [t for t in SHORT_PRIMARIES if t[0] not in short]
    k = 0
    for i in range(0, n_patterns, step):
        rx, w = templates[k % len(templates)]
        if k >= len(templates):            # more slots than templates: make the regex distinct
            rx, w = rx + " r%d" % k, w + " r%d" % k
        lib["patterns"][i]["primary_pattern"]["regex"] = rx
        lib["patterns"][i]["name"] = "short: " + rx
        wit[i] = w
        k += 1
    return lib, wit, swit


# --------------------------------------------------------------------------- native corpus generator
_lib = None


def _native():
    global _lib
    if _lib is None:
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libpm_synth.so")
        if not os.path.exists(path):
            raise RuntimeError("%s missing: run `make` (or __graft_entry__.build())" % path)
        _lib = ctypes.CDLL(path)
        _lib.pm_synth_corpus.restype = ctypes.c_int64
        _lib.pm_synth_corpus.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int64, ctypes.c_double,
                                         ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(ctypes.c_char_p),
                                         ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                         ctypes.POINTER(ctypes.c_int64), ctypes.c_int]
    return _lib


def near_misses(witnesses: List[str]) -> List[str]:
    """Strings that share vocabulary and literals with the witnesses without (in general) matching their
    patterns: every witness with its tail cut off, and with its head cut off."""
    out = []
    for w in witnesses:
        out.append(w[:max(8, int(len(w) * 0.7))])
        out.append(w[int(len(w) * 0.35):])
    return out


def make_corpus(target_bytes: int, witnesses: List[str], sec_witnesses: Optional[List[Optional[str]]] = None,
                p_hit: float = 1e-3, seed: int = LOG_SEED, first_line: int = 0, crlf: bool = False,
                out: Optional[np.ndarray] = None, threads: int = 0, p_near: float = 0.0,
                near: Optional[List[str]] = None) -> Tuple[np.ndarray, int]:
    """Generates <= target_bytes of Quarkus/JVM-style log text ending in a newline.
    Returns (uint8 array view of the bytes written, number of lines).  p_near / near: near-miss lines."""
    lib = _native()
    if p_near > 0 and near:
        return _make_corpus_near(lib, target_bytes, witnesses, sec_witnesses, p_hit, seed, first_line, crlf, out, threads, p_near, near)
    n = len(witnesses)
    W = (ctypes.c_char_p * max(n, 1))(*[w.encode() for w in witnesses])
    sw = sec_witnesses or [None] * n
    S = (ctypes.c_char_p * max(n, 1))(*[(s.encode() if s else None) for s in sw])
    if out is None:
        out = np.empty(target_bytes, dtype=np.uint8)
    assert out.dtype == np.uint8 and out.size >= target_bytes and out.flags["C_CONTIGUOUS"]
    nl = ctypes.c_int64(0)
    got = lib.pm_synth_corpus(seed, first_line, target_bytes, p_hit, W, S, n, int(crlf),
                              out.ctypes.data, ctypes.byref(nl), threads)
    if got < 0:
        raise RuntimeError("pm_synth_corpus failed")
    return out[:got], nl.value


def _make_corpus_near(lib, target_bytes, witnesses, sec_witnesses, p_hit, seed, first_line, crlf, out, threads, p_near, near):
    if not hasattr(lib, "_near_ready"):
        lib.pm_synth_corpus_near.restype = ctypes.c_int64
        lib.pm_synth_corpus_near.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int64, ctypes.c_double,
                                             ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(ctypes.c_char_p), ctypes.c_int, ctypes.c_int,
                                             ctypes.c_double, ctypes.POINTER(ctypes.c_char_p), ctypes.c_int,
                                             ctypes.c_void_p, ctypes.POINTER(ctypes.c_int64), ctypes.c_int]
        lib._near_ready = True
    n = len(witnesses)
    W = (ctypes.c_char_p * max(n, 1))(*[w.encode() for w in witnesses])
    sw = sec_witnesses or [None] * n
    S = (ctypes.c_char_p * max(n, 1))(*[(x.encode() if x else None) for x in sw])
    N = (ctypes.c_char_p * len(near))(*[x.encode() for x in near])
    if out is None:
        out = np.empty(target_bytes, dtype=np.uint8)
    nl = ctypes.c_int64(0)
    got = lib.pm_synth_corpus_near(seed, first_line, target_bytes, p_hit, W, S, n, int(crlf), p_near, N, len(near),
                                   out.ctypes.data, ctypes.byref(nl), threads)
    if got < 0:
        raise RuntimeError("pm_synth_corpus_near failed")
    return out[:got], nl.value


def make_corpus_device(device_ptr: int, target_bytes: int, witnesses: List[str], sec_witnesses: Optional[List[Optional[str]]] = None,
                       p_hit: float = 1e-3, seed: int = LOG_SEED, first_line: int = 0, crlf: bool = False, p_near: float = 0.0,
                       near: Optional[List[str]] = None) -> Tuple[int, int]:
    """The same corpus as make_corpus, written by the GPU into device memory at `device_ptr` (at least
    target_bytes bytes on the current CUDA device).  Returns (bytes written, number of lines)."""
    lib = _native()
    if not hasattr(lib, "_dev_ready"):
        lib.pm_synth_corpus_device_near.restype = ctypes.c_int64
        lib.pm_synth_corpus_device_near.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int64, ctypes.c_double,
                                                    ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(ctypes.c_char_p), ctypes.c_int, ctypes.c_int,
                                                    ctypes.c_double, ctypes.POINTER(ctypes.c_char_p), ctypes.c_int,
                                                    ctypes.c_void_p, ctypes.POINTER(ctypes.c_int64)]
        lib._dev_ready = True
    n = len(witnesses)
    w = (ctypes.c_char_p * max(n, 1))(*[x.encode("utf-8") for x in witnesses])
    sec = sec_witnesses or [None] * n
    sw = (ctypes.c_char_p * max(n, 1))(*[(x.encode("utf-8") if x else None) for x in sec])
    n_lines = ctypes.c_int64(0)
    near = near or []
    nm = (ctypes.c_char_p * max(len(near), 1))(*[x.encode("utf-8") for x in near])
    got = lib.pm_synth_corpus_device_near(seed, first_line, target_bytes, p_hit, w, sw, n, 1 if crlf else 0, p_near if near else 0.0, nm, len(near),
                                          device_ptr, ctypes.byref(n_lines))
    if got < 0:
        raise RuntimeError("pm_synth_corpus_device failed (%d)" % got)
    return int(got), int(n_lines.value)
