"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI, against
the oracle on the same seeded inputs.  Bar: bit-exact on every integer field; scores are
compared with == (SPEC.md section 5: every fp64 operation is a single IEEE operation in a fixed
order, so the stated tolerance is 0 ulp; a JVM comparison would use 1e-12 relative)."""
import json
import os

import numpy as np
import pytest

from oracle import oracle
from operator_b200 import synth
from tests import cases

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SCORE_TOL = 0.0


def _tuple(e):
    return (e.line_no, e.pattern, e.line_start, e.line_len, e.ctx_first, e.ctx_last, e.ctx_start, e.ctx_end, e.score)


def _got(res, pod):
    return [(g["line_no"], g["pattern"], g["line_start"], g["line_len"], g["ctx_first"], g["ctx_last"], g["ctx_start"], g["ctx_end"], g["score"])
            for g in res.events_of(pod)]


def check_parity(engine, lib_json, logs, expect_events=None):
    engine.compile_library(lib_json).publish()
    olib = oracle.load_library(lib_json)
    res = engine.scan(logs)
    assert res.n_pods == len(logs)
    total = 0
    for i, log in enumerate(logs):
        raw = bytes(log)
        want = [_tuple(e) for e in oracle.analyze(raw, olib)]
        got = _got(res, i)
        assert len(got) == len(want), (i, len(got), len(want), got[:3], want[:3])
        for g, w in zip(got, want):
            assert g[:8] == w[:8], (i, g, w)
            assert abs(g[8] - w[8]) <= SCORE_TOL, (i, g, w)
        assert int(res.pod_n_lines[i]) == len(oracle.split_lines(raw)), (i, raw[:40])
        total += len(want)
    if expect_events is not None:
        assert total >= expect_events
    return res


def lib_of(regexes, **kw):
    pats = []
    for i, r in enumerate(regexes):
        p = {"id": "p%d" % i, "name": "n%d" % i, "severity": ["INFO", "LOW", "MEDIUM", "HIGH", "CRITICAL"][i % 5],
             "primary_pattern": {"regex": r, "confidence": 0.5 + 0.05 * (i % 11)}}
        p.update(kw)
        pats.append(p)
    return {"patterns": pats}


def test_config1_single_pod_core_library(engine):
    """BASELINE config 1: one ~10k-line pod log against a ~50-pattern library."""
    lib_json, wit, swit = synth.make_library(50)
    corpus, n_lines = synth.make_corpus(1400 << 10, wit, swit, p_hit=5e-3)
    assert n_lines > 9000
    res = check_parity(engine, lib_json, [corpus], expect_events=20)
    assert res.gpu_launches >= 6


def test_500_patterns_sample(engine):
    """BASELINE config 2's library at a size the oracle finishes in seconds."""
    lib_json, wit, swit = synth.make_library(500)
    corpus, _ = synth.make_corpus(384 << 10, wit, swit, p_hit=2e-2)
    check_parity(engine, lib_json, [corpus], expect_events=30)


def test_handwritten_regexes_all_tiers(engine):
    """Every hand-written regex (filter tier and combined-DFA tier) on every hand-written line,
    LF and CRLF, with context windows and the STACK extension."""
    lib_json = lib_of(cases.REGEXES, context_extraction={"lines_before": 2, "lines_after": 3, "include_stack_trace": True})
    body = "\n".join(cases.LINES)
    logs = [body.encode("utf-8"), (body + "\n").encode("utf-8"), "\r\n".join(cases.LINES).encode("utf-8") + b"\r\n"]
    check_parity(engine, lib_json, logs, expect_events=100)


def test_edge_logs(engine):
    lib_json = lib_of([r"OutOfMemoryError", r"\bOOM\b", r"^x+$"], context_extraction={"lines_before": 5, "lines_after": 5, "include_stack_trace": True})
    T = chr(9)
    logs = [b"", b"\n", b"\n\n\n", b"\r\n", b"OutOfMemoryError", b"OutOfMemoryError\n", b"\nOutOfMemoryError\n\n", b"OOM\r\n\r\n",
            b"a\rOOM\rb", b"x", b"xx\r", b"xx\r\n", b"xx\r\r\n", b"\n\n\nOOM", ("OOM\n" * 200).encode(),
            ("java.lang.OutOfMemoryError\n" + "".join(T + "at a.B.c(B.java:%d)\n" % i for i in range(70)) + "tail\n").encode(),
            ("OutOfMemoryError " + "y" * 10000 + "\n" + "z" * 9000 + " OOM").encode(), b"OutOfMemoryErro", b"utOfMemoryError",
            ("é" * 1000 + "OutOfMemoryError").encode("utf-8")]
    check_parity(engine, lib_json, logs, expect_events=200)


def test_alignment_sweep(engine):
    """The filter samples 4-aligned words: a literal must be found at every alignment and right at
    pod boundaries (pods start on 16-byte boundaries; prefix lengths move the literal)."""
    lib_json = lib_of([r"kafka-broker leader not available", r"(?i)certificate expired", r"node \w+ quota exceeded at deadline"])
    logs = []
    for k in range(0, 40):
        logs.append(("a" * k + "kafka-broker leader not available").encode())
        logs.append(("b" * k + "Certificate EXPIRED" + "c" * (k % 5) + "\n" + "d" * k + "node n7 quota exceeded at deadline").encode())
    check_parity(engine, lib_json, logs, expect_events=120)


def test_short_literals_stay_on_the_filter(engine):
    """Required literals of 5 and 6 bytes (FATAL, panic:, Killed, (?i)error ...): the library compiles to filter
    granule 2 (k_scan_half tests the word at every even offset) instead of sending these patterns through the combined
    DFA.  A literal must be found at every alignment, across 16-byte, 512-byte and 2 KiB boundaries and at pod ends."""
    regs = [r"FATAL: \d+", r"panic: ", r"Killed process \d+", r"(?i)error \d+", r"status=5\d\d", r"OutOfMemoryError", r"\bEPIPE\b", r"[Cc]ache (evicted|flushed)"]
    lib_json = lib_of(regs, context_extraction={"lines_before": 2, "lines_after": 2})
    rep = engine.compile_library(lib_json).publish().report
    assert rep["n_dfa_scanned"] == 0 and rep["n_filtered"] == len(regs)
    logs = []
    for k in range(0, 70):
        logs.append(("a" * k + "FATAL: 12 " + "b" * (k % 7) + "\n" + "c" * k + " go panic: nil map\n" + "x" * (2040 - k) + "Killed process 77 at the tile edge").encode())
        logs.append(("d" * (500 + k) + "eRRoR 404 " + "e" * k + " status=503\n" + "f" * k + "EPIPE\n" + "g" * (k % 3) + "Cache evicted\nxEPIPEx FATAL:\n").encode())
        logs.append(("h" * k + "java.lang.OutOfMemoryError").encode())
    check_parity(engine, lib_json, logs, expect_events=450)
    # and on generated text with these patterns mixed into a synthetic library (still no DFA pass)
    lib2, wit, swit = synth.make_library(60)
    for i, rx in enumerate(regs):
        lib2["patterns"][i * 7]["primary_pattern"]["regex"] = rx
    wit2 = list(wit)
    for i, w in enumerate(["FATAL: 7", "panic: x", "Killed process 1", "ERROR 9", "status=500", "OutOfMemoryError", "EPIPE", "cache flushed"]):
        wit2[i * 7] = w
    rep = engine.compile_library(lib2).publish().report
    assert rep["n_dfa_scanned"] == 0
    corpus, _ = synth.make_corpus(3 << 20, wit2, swit, p_hit=5e-3)
    check_parity(engine, lib2, [corpus], expect_events=80)


def test_secondary_scoring_and_windows(engine):
    T = chr(9)
    lib_json = {"patterns": [
        {"id": "a", "severity": "CRITICAL", "primary_pattern": {"regex": "primary failure alpha", "confidence": 0.9},
         "secondary_patterns": [{"regex": "Caused by: ", "weight": 0.3, "proximity_window": 4},
                                {"regex": r"retrying in \d+ms", "weight": 0.25, "proximity_window": 10},
                                {"regex": "never present", "weight": 0.5, "proximity_window": 50}],
         "context_extraction": {"lines_before": 3, "lines_after": 2, "include_stack_trace": True}},
        {"id": "b", "severity": "LOW", "primary_pattern": {"regex": "primary failure beta", "confidence": 0.55},
         "secondary_patterns": [{"regex": "primary failure beta", "weight": 0.1, "proximity_window": 0},
                                {"regex": "primary failure alpha", "weight": 0.7, "proximity_window": 1000}],
         "context_extraction": {"lines_before": 1000, "lines_after": 1000}},
    ]}
    lines = []
    for i in range(400):
        if i % 37 == 5:
            lines.append("ERROR primary failure alpha #%d" % i)
            lines.append(T + "at x.Y.z(Y.java:1)")
            if i % 2:
                lines.append("Caused by: boom")
        elif i % 53 == 7:
            lines.append("WARN primary failure beta")
        elif i % 11 == 3:
            lines.append("INFO retrying in %dms" % i)
        else:
            lines.append("INFO filler line %d" % i)
    log = "\n".join(lines).encode()
    res = check_parity(engine, lib_json, [log, log[: len(log) // 3], b"primary failure beta"], expect_events=15)
    assert len(set(res.score.tolist())) > 4


def test_multi_pod_batch_equals_single_pods(engine):
    """Batching is transparent: a batch of pods gives each pod the events it gets alone."""
    lib_json, wit, swit = synth.make_library(50)
    engine.compile_library(lib_json).publish()
    pods = [synth.make_corpus(40000 + 7919 * k, wit, swit, p_hit=1e-2, first_line=k * 100000)[0].copy() for k in range(9)]
    pods.insert(3, np.zeros(0, dtype=np.uint8))
    batch = engine.scan(pods)
    for k, p in enumerate(pods):
        single = engine.scan([p])
        assert _got(batch, k) == _got(single, 0)
        assert int(batch.pod_n_lines[k]) == int(single.pod_n_lines[0])
    assert batch.pod_tag.tolist() == list(range(len(pods)))


def test_line_index_matches_oracle(engine):
    """a10: the materialised line table equals Java's split on the same bytes."""
    lib_json, wit, swit = synth.make_library(50)
    engine.compile_library(lib_json).publish()
    logs = [bytes(synth.make_corpus(300000, wit, swit)[0]), b"a\r\nb\n\nc", b"", b"\n\n", b"x\n\n\n", "é\nü\n".encode("utf-8"),
            bytes(synth.make_corpus(70000, wit, swit, crlf=True)[0])]
    b = engine.batch()
    try:
        for i, log in enumerate(logs):
            b.add_pod(log, i)
        b.submit()
        res = b.wait()
        for i, log in enumerate(logs):
            spans = oracle.split_lines(log)
            got = b.line_index(i, len(spans))
            assert got.tolist() == [s for s, _ in spans], (i, got[:5], spans[:5])
            assert int(res.pod_n_lines[i]) == len(spans)
    finally:
        b.release()


def test_rescan_is_idempotent_and_counts_injected_hits(engine):
    """256 MiB corpus, 500 patterns.  The Python oracle cannot run at this size, so check (1) two scans
    of the staged bytes give identical results, (2) every event line really contains a witness of its
    pattern, (3) the (line, pattern) pairs equal, exactly, those of the PCRE2 loop over the same bytes
    (~3.5 s on 16 host cores)."""
    lib_json, wit, swit = synth.make_library(500)
    engine.compile_library(lib_json).publish()
    corpus, n_lines = synth.make_corpus(256 << 20, wit, None, p_hit=1e-3)
    b = engine.batch()
    try:
        b.add_pod(corpus, 1)
        b.stage()
        r1 = b.rescan()
        r2 = b.rescan()
        assert r1.n_events == r2.n_events and r1.n_events > 1000
        for f in ("pattern_index", "line_no", "line_start", "line_len", "ctx_start", "ctx_end", "score"):
            assert np.array_equal(getattr(r1, f), getattr(r2, f)), f
        assert int(r1.pod_n_lines[0]) == n_lines
        blob = corpus.tobytes()
        for i in range(0, r1.n_events, max(1, r1.n_events // 400)):
            s, l, p = int(r1.line_start[i]), int(r1.line_len[i]), int(r1.pattern_index[i])
            line = blob[s:s + l].decode()
            assert wit[p] in line, (p, line)
            assert s == 0 or blob[s - 1] == 10
        # order: (line, pattern) strictly increasing
        key = r1.line_no.astype(np.int64) * 100000 + r1.pattern_index
        assert np.all(np.diff(key) > 0)
        # every (line, pattern) pair against the PCRE2 loop over the same 256 MiB: exact, no tolerance
        from oracle import pcre2_oracle
        if pcre2_oracle.available():
            pl = pcre2_oracle.Pcre2Library([p["primary_pattern"]["regex"] for p in lib_json["patterns"]])
            starts, lens = pcre2_oracle.split(corpus)
            ol, op, total = pcre2_oracle.scan(pl, corpus, starts, lens, threads=os.cpu_count() or 4)
            pl.close()
            assert total == ol.size
            want = np.stack([ol + 1, op.astype(np.int64)], axis=1)
            want = want[np.lexsort((want[:, 1], want[:, 0]))]
            got = np.stack([r1.line_no.astype(np.int64), r1.pattern_index.astype(np.int64)], axis=1)
            assert np.array_equal(got, want), (got.shape, want.shape)
    finally:
        b.release()


def test_parse_contract(engine):
    """a3/a4/a14: PodFailureData JSON in, AnalysisResult JSON out; the fields the operator reads
    (EventService.java:75-83, AnalysisStorageService.java:145-160, :298-325) equal the oracle's."""
    from operator_b200 import LogParserClient
    lib_json, wit, swit = synth.make_library(50)
    engine.compile_library(lib_json).publish()
    olib = oracle.load_library(lib_json)
    corpus, _ = synth.make_corpus(200000, wit, swit, p_hit=2e-2)
    text = bytes(corpus).decode()
    req = {"pod": {"metadata": {"name": "orders-7d9f", "namespace": "shop"}}, "logs": text, "events": []}
    out = LogParserClient(engine).analyze_log(req)
    ev = oracle.analyze(bytes(corpus), olib)
    summ = oracle.summarize(ev, olib)
    assert out["summary"]["highestSeverity"] == summ["highestSeverity"]
    assert out["summary"]["significantEvents"] == summ["significantEvents"]
    assert out["summary"]["severityDistribution"] == summ["severityDistribution"]
    assert len(out["events"]) == len(ev) > 5
    lines = text.split("\n")
    for o, e in zip(out["events"], ev):
        assert o["lineNumber"] == e.line_no
        assert o["matchedPattern"]["id"] == olib.patterns[e.pattern].id
        assert o["matchedPattern"]["severity"] == olib.patterns[e.pattern].severity
        assert o["score"] == e.score
        assert o["context"]["matchedLine"] == lines[e.line_no - 1]
        assert o["context"]["linesBefore"] == lines[e.ctx_first - 1:e.line_no - 1]
        assert o["context"]["linesAfter"] == lines[e.line_no:e.ctx_last]
    assert out["metadata"]["podName"] == "orders-7d9f"
    empty = LogParserClient(engine).analyze_log({"pod": {"metadata": {"name": "p"}}, "logs": "", "events": []})
    assert empty["events"] == [] and empty["summary"]["highestSeverity"] == "NONE"
    # escapes in the log (json.dumps emits \\uXXXX and surrogate pairs for non-ASCII), decoded straight into pinned memory
    uni = "caf\u00e9 \u20ac100 \U0001F600 \"quoted\" back\\slash\n" + wit[3] + " \u00e9\n" + chr(9) + "at x.Y(Z.java:1)\n"
    out_u = LogParserClient(engine).analyze_log({"pod": {"metadata": {"name": "u"}}, "logs": uni, "events": []})
    ev_u = oracle.analyze(uni.encode("utf-8"), olib)
    assert [o["lineNumber"] for o in out_u["events"]] == [e.line_no for e in ev_u] and len(ev_u) >= 1
    assert out_u["events"][0]["context"]["matchedLine"] == uni.split("\n")[1]
    # result cache: the operator re-posts the same failure (one request per matching CR, every reconcile)
    req2 = {"pod": {"metadata": {"name": "orders-7d9f", "namespace": "shop", "uid": "uid-123"},
                    "status": {"containerStatuses": [{"state": {"terminated": {"exitCode": 1, "finishedAt": "2026-03-03T09:27:00Z"}}}]}},
            "logs": text, "events": []}
    first = LogParserClient(engine).analyze_log(req2)
    again = LogParserClient(engine).analyze_log(req2)
    assert again == first and again["analysisId"] == first["analysisId"]       # served from the cache
    req2["pod"]["status"]["containerStatuses"][0]["state"]["terminated"]["finishedAt"] = "2026-03-03T09:30:00Z"
    later = LogParserClient(engine).analyze_log(req2)
    assert later["analysisId"] != first["analysisId"] and later["events"] == first["events"]


def test_golden_fixtures(engine):
    """Committed fixtures (tests/golden, generated by tests/golden/make_golden.py from the oracle):
    the device must reproduce them without the oracle in the loop."""
    with open(os.path.join(GOLDEN, "golden_events.json")) as f:
        gold = json.load(f)
    for case in gold["cases"]:
        engine.compile_library(case["library"]).publish()
        logs = [bytes.fromhex(h) for h in case["logs_hex"]]
        res = engine.scan(logs)
        for i, want in enumerate(case["events"]):
            got = [list(t) for t in _got(res, i)]
            assert got == want, (case["name"], i)
            assert int(res.pod_n_lines[i]) == case["n_lines"][i]


def test_unsupported_library_is_rejected(engine):
    from operator_b200 import ScanError
    with pytest.raises(ScanError) as ei:
        engine.compile_library(lib_of(["fine literal here", r"(?<=x)lookbehind"]))
    assert ei.value.status == -3 and ei.value.report["pattern_index"] == 1


def test_tier2_dfa_multi_tile(engine):
    """Patterns without a 7-byte literal go through the combined shared-memory DFA (k_dfa_scan).
    24 MiB gives every CTA several 64 KiB tiles, so tile carry, speculative entry states and
    range cuts at line starts are all exercised; LF and CRLF."""
    regexes = [r"\bGC\b", r"\d{3,}ms", r"^\s+at org\.acme", r"lag=\d+ epoch", r"[Cc]ache (hit|miss)", r"->\s5\d\d", r"id=[0-9a-f]{4}00", r"roles=\[editor\]$",
               r"(?i)caused by"]
    lib_json = lib_of(regexes, context_extraction={"lines_before": 1, "lines_after": 1})
    rep = engine.compile_library(lib_json).report
    assert rep["n_dfa_scanned"] >= 4 and rep["dfa_groups"] >= 1
    _, wit, swit = synth.make_library(50)
    lf = synth.make_corpus(24 << 20, wit, swit, p_hit=1e-3)[0]
    crlf = synth.make_corpus(3 << 20, wit, swit, p_hit=1e-3, crlf=True, first_line=10 ** 7)[0]
    check_parity(engine, lib_json, [lf, crlf], expect_events=5000)


def test_long_lines_through_verify_and_context(engine):
    """Lines of 0.2-9 KB with the literal at the start, in the middle, at the end and across the
    512-byte pieces of the warp walk; 'A.*B' regexes whose two halves are KBs apart (every
    speculative guess after A is wrong and is redone); long lines as context and secondary lines."""
    regs = [r"OutOfMemoryError: Java heap", r"fatal-alpha.*omega-done", r"^START-OF-LINE \w+ .*tail-marker$",
            r"(?i)connection refused by PEER \d+", r"begin-xyz\d+.*mid-xyz.*end-xyz"]
    lib = lib_of(regs, context_extraction={"lines_before": 2, "lines_after": 3, "include_stack_trace": True})
    lib["patterns"][0]["secondary_patterns"] = [{"regex": r"GC overhead", "weight": 0.4, "proximity_window": 4},
                                                 {"regex": r"payload-\w+-end", "weight": 0.3, "proximity_window": 6}]
    rng = np.random.default_rng(7)

    def filler(n):
        words = [b"alpha", b"beta", b"{\"k\":\"0123456789abcdef\"}", b"GC", b"overhead?", b"at", b"java.base/Thread.run", b"fatal", b"omega"]
        out = bytearray()
        while len(out) < n:
            out += words[int(rng.integers(len(words)))] + b" "
        return bytes(out[:n])

    lines = []
    for n in (10, 150, 200, 500, 511, 512, 513, 700, 1024, 2000, 4093, 9000):
        lines.append(b"OutOfMemoryError: Java heap " + filler(n))
        lines.append(filler(n) + b" OutOfMemoryError: Java heap")
        lines.append(filler(n // 2) + b" OutOfMemoryError: Java heap " + filler(n // 2))
        lines.append(b"fatal-alpha " + filler(n) + b" omega-done")
        lines.append(b"fatal-alpha " + filler(n))                       # first half only: no match
        lines.append(b"START-OF-LINE word " + filler(n) + b" tail-marker")
        lines.append(b"xSTART-OF-LINE word " + filler(n) + b" tail-marker")
        lines.append(filler(n) + b" Connection Refused by peer 42 " + filler(7))
        lines.append(b"begin-xyz7 " + filler(n) + b" mid-xyz " + filler(n) + b" end-xyz")
        lines.append(b"\tat " + filler(n))                                 # stack-like context line
        lines.append(b"GC overhead limit " + filler(n) + b" payload-abc-end")
        lines.append(filler(40))
    log = b"\n".join(lines) + b"\n"
    res = check_parity(engine, lib, [log, log[:len(log) // 2], b"\r\n".join(lines)])
    assert res.n_events > 100


def test_concurrent_parse_requests_share_batches(engine):
    """8(f).1: many /parse calls at once (one pod per request, as the operator sends them) are
    coalesced into shared GPU batches; every caller gets exactly the answer of a call made alone."""
    import json
    import threading
    lib_json, wit, swit = synth.make_library(50)
    engine.compile_library(lib_json).publish()
    olib = oracle.load_library(lib_json)
    n = 24
    logs = [bytes(synth.make_corpus(60000 + 7000 * k, wit, swit, p_hit=3e-2, seed=900 + k)[0]) for k in range(n)]
    logs[5] = b""                                        # an empty log rides along
    bodies = [json.dumps({"pod": {"metadata": {"name": "pod-%d" % k, "namespace": "ns"}}, "logs": l.decode(), "events": []}).encode()
              for k, l in enumerate(logs)]
    before = engine.parse_stats()
    out = [None] * n
    gate = threading.Barrier(n)

    def work(k):
        gate.wait()
        for _ in range(3):
            out[k] = json.loads(engine.parse_raw(bodies[k]))

    ts = [threading.Thread(target=work, args=(k,)) for k in range(n)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    after = engine.parse_stats()
    assert after["requests"] - before["requests"] == 3 * n
    assert after["batches"] - before["batches"] < 3 * n, after        # at least some requests shared a batch
    assert after["largest_batch"] >= 2
    for k in range(n):
        ev = oracle.analyze(logs[k], olib)
        assert out[k]["metadata"]["podName"] == "pod-%d" % k
        assert [(o["lineNumber"], o["matchedPattern"]["id"], o["score"]) for o in out[k]["events"]] == \
               [(e.line_no, olib.patterns[e.pattern].id, e.score) for e in ev], k
        lines = logs[k].decode().split("\n")
        for o, e in zip(out[k]["events"], ev):
            assert o["context"]["matchedLine"] == lines[e.line_no - 1]
            assert o["context"]["linesAfter"] == lines[e.line_no:e.ctx_last]
        assert out[k]["summary"]["totalEvents"] == len(ev)
    assert sum(len(o["events"]) for o in out) > 200


def test_http_front_end_speaks_the_parse_contract(engine):
    """a3: POST /parse with application/json, as LogParserRestClient.java:21-39 sends it; 404 elsewhere,
    400 for a malformed body, concurrent requests answered correctly."""
    import json
    import threading
    import urllib.error
    import urllib.request
    from http.server import ThreadingHTTPServer
    from operator_b200 import service
    lib_json, wit, swit = synth.make_library(50)
    engine.compile_library(lib_json).publish()
    olib = oracle.load_library(lib_json)
    srv = ThreadingHTTPServer(("127.0.0.1", 0), service.make_handler(engine))
    port = srv.server_address[1]
    t = threading.Thread(target=srv.serve_forever, daemon=True)
    t.start()
    try:
        logs = [bytes(synth.make_corpus(40000 + 9000 * k, wit, swit, p_hit=3e-2, seed=70 + k)[0]) for k in range(6)]

        def post(path, body):
            req = urllib.request.Request("http://127.0.0.1:%d%s" % (port, path), data=body, headers={"Content-Type": "application/json"}, method="POST")
            with urllib.request.urlopen(req, timeout=30) as r:
                return r.status, r.headers.get("Content-Type"), json.loads(r.read())

        out = [None] * len(logs)

        def work(k):
            body = json.dumps({"pod": {"metadata": {"name": "p%d" % k, "namespace": "n"}}, "logs": logs[k].decode(), "events": []}).encode()
            out[k] = post("/parse", body)

        ts = [threading.Thread(target=work, args=(k,)) for k in range(len(logs))]
        for x in ts:
            x.start()
        for x in ts:
            x.join()
        for k, (status, ctype, res) in enumerate(out):
            ev = oracle.analyze(logs[k], olib)
            assert status == 200 and ctype == "application/json"
            assert [(o["lineNumber"], o["matchedPattern"]["id"], o["score"]) for o in res["events"]] == [(e.line_no, olib.patterns[e.pattern].id, e.score) for e in ev]
            assert res["summary"]["highestSeverity"] == oracle.summarize(ev, olib)["highestSeverity"] and res["metadata"]["podName"] == "p%d" % k
        for path, body, code in (("/other", b"{}", 404), ("/parse", b"{\"logs\": \"unterminated", 400)):
            try:
                post(path, body)
                raise AssertionError("expected HTTP %d" % code)
            except urllib.error.HTTPError as e:
                assert e.code == code
    finally:
        srv.shutdown()
        srv.server_close()


@pytest.mark.gpu
def test_context_units_inside_and_outside_the_shared_memory_budget(engine):
    """k_check walks a run of context lines through the warp's shared-memory copy of the unit's table when
    the table fits (1536 one-byte entries) and through the 256-column table in global memory when it does
    not: one secondary of each kind (760 and 4680 entries), lines of every length up to the long-line
    threshold and beyond, hits at every alignment, windows that reach over the pod's ends."""
    import random
    rnd = random.Random(77)
    big = r"(?i)(alpha|bravo|charlie|delta|echo|foxtrot|golf|hotel|india|juliet)[ _-](kilo|lima|mike|november|oscar|papa|quebec|romeo)[0-9]{3}"
    small = r"connection (refused|reset|timed out) by peer [0-9]+"
    lib = {"patterns": [
        {"id": "a", "name": "a", "severity": "HIGH", "primary_pattern": {"regex": r"FATAL: worker crashed", "confidence": 0.9},
         "secondary_patterns": [{"regex": big, "weight": 0.4, "proximity_window": 9}, {"regex": small, "weight": 0.3, "proximity_window": 4}],
         "context_extraction": {"lines_before": 3, "lines_after": 4, "include_stack_trace": True}},
        {"id": "b", "name": "b", "severity": "LOW", "primary_pattern": {"regex": r"panic: runtime error", "confidence": 0.7},
         "secondary_patterns": [{"regex": big, "weight": 0.5, "proximity_window": 16}]},
    ]}
    fill = "abcdefghijklmnopqrstuvwxyz 0123456789 _-:/"
    def filler(n):
        return "".join(rnd.choice(fill) for _ in range(n))
    logs = []
    for pod in range(3):
        lines = []
        for i in range(1500):
            r = rnd.random()
            n = rnd.choice([0, 1, 7, 15, 16, 17, 31, 60, 100, 127, 128, 129, 200, 255, 256, 257, 300, 700])
            if r < 0.03:
                lines.append(filler(rnd.randint(0, 20)) + "FATAL: worker crashed" + filler(rnd.randint(0, 30)))
            elif r < 0.05:
                lines.append(filler(rnd.randint(0, 9)) + "panic: runtime error: index out of range")
            elif r < 0.12:
                lines.append(filler(n) + rnd.choice(["Echo-Mike123", "juliet_ROMEO007", "alpha papa999", "golf-lima12"]) + filler(rnd.randint(0, 40)))
            elif r < 0.18:
                lines.append(filler(n) + "connection %s by peer %d" % (rnd.choice(["refused", "reset", "timed out", "closed"]), rnd.randint(0, 99)))
            elif r < 0.24:
                lines.append("\tat com.example.Worker.run(Worker.java:%d)" % rnd.randint(1, 999))
            else:
                lines.append(filler(n))
        logs.append(("\n".join(lines) + "\n").encode())
    check_parity(engine, lib, logs, expect_events=100)
