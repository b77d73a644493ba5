"""CPU tests: the oracle against known answers and the committed golden vectors, the C-ABI
library (loads, exports every symbol of include/podmortem_scan.h, fails loudly without a GPU),
and the host-side sharding / reduction plumbing under gloo with world_size 2."""
import ctypes
import json
import os
import re
import subprocess
import sys

import pytest

from oracle import javaregex, oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "golden_events.json")


# ---------------------------------------------------------------- known answers: Java split
# java.lang.String#split(String) Javadoc: "boo:and:foo".split("o") -> {"b", "", ":and:f"}
# (trailing empty strings removed); the same rule with the separator \r?\n:
@pytest.mark.parametrize("log,want", [
    (b"", [b""]),                          # no separator found -> the input itself
    (b"abc", [b"abc"]),
    (b"abc\n", [b"abc"]),                  # trailing empty string removed
    (b"abc\n\n\n", [b"abc"]),
    (b"\n\n", []),                         # all pieces empty -> empty array
    (b"\nabc", [b"", b"abc"]),             # leading empty string kept
    (b"a\n\nb", [b"a", b"", b"b"]),
    (b"a\r\nb\r\n", [b"a", b"b"]),
    (b"a\rb\n", [b"a\rb"]),                # lone \r is not a separator
    (b"a\r\r\n", [b"a\r"]),
    (b"a\n\r", [b"a", b"\r"]),
    (b"\r\n", []),
])
def test_split_known_answers(log, want):
    got = [log[s:s + l] for s, l in oracle.split_lines(log)]
    assert got == want


# ---------------------------------------------------------------- known answers: java.util.regex
# Behaviours documented in the java.util.regex.Pattern Javadoc (JDK 21) that differ from Python's
# defaults and that the translator must restate.
@pytest.mark.parametrize("rx,line,want", [
    (r"a.c", "a\rc", False),               # '.' excludes line terminators \n \r 
    (r"a.c", "a\u0085c", False),
    (r"a.c", "a\u2028c", False),
    (r"(?s)a.c", "a\rc", True),            # DOTALL
    (r"a$", "a\r", True),                  # '$' matches before a final line terminator
    (r"a$", "a\u2029", True),
    (r"a$", "a\r\r", False),
    (r"a\z", "a\r", False),                # \z is the very end
    (r"a\Z", "a", True),
    (r"\w+", "é", False),                  # \w is ASCII-only without UNICODE_CHARACTER_CLASS
    (r"\W", "é", True),
    (r"(?i)é", "É", False),                # (?i) is ASCII-only without UNICODE_CASE
    (r"(?i)straße", "STRAßE", True),
    (r"\bé", "é", False),                  # \b follows \w
    (r"x(?i)y", "xY", True),               # a flag change lasts to the end of the group
    (r"x(?i)y", "XY", False),
    (r"(a(?i)b)c", "aBC", False),
    (r"(a(?i)b|c)d", "Cd", True),          # ... including later alternatives of that group
    (r"[^a]", "\r", True),                 # negated classes do match line terminators
    (r"\p{Punct}", "_", True),
    (r"\Qa+b\E", "a+b", True),
    (r"^\Qab\E?c$", "c", False),          # quotation is rewritten in place: the quantifier binds to 'b' alone
    (r"^\Qab\E?c$", "ac", True),
    (r"^\Qab\E{2}$", "abb", True),
    (r"^\Qab\E{2}$", "abab", False),
    (r"[\Q^-\E]x", "^x", True),           # ... and it works inside a character class
    (r"[\Q^-\E]x", "ax", False),
    (r"a\Q1\E2+", "a122", True),          # a digit that opens a quotation becomes \x31
    (r"a\Q\E+", "aa", True),              # an empty quotation vanishes: the quantifier binds to 'a'
    (r"a{2}", "aa", True),
    (r"a{2,}?b", "aaab", True),
])
def test_java_regex_known_answers(rx, line, want):
    assert (re.search(javaregex.to_python(rx), line, re.ASCII) is not None) == want


def test_oracle_scoring_and_context_by_hand():
    T = chr(9)
    lib = oracle.load_library({"patterns": [{
        "id": "oom", "severity": "CRITICAL", "primary_pattern": {"regex": "OutOfMemoryError", "confidence": 0.9},
        "secondary_patterns": [{"regex": "heap space", "weight": 0.3, "proximity_window": 2},
                               {"regex": "GC overhead", "weight": 0.5, "proximity_window": 3}],
        "context_extraction": {"lines_before": 1, "lines_after": 1, "include_stack_trace": True}}]})
    log = ("start\r\nfoo\njava.lang.OutOfMemoryError: Java heap space\n" + T + "at a.b.C(D.java:1)\n" + T + "at x.y.Z(Q.java:2)\n"
           "Caused by: boom\ndone\nGC overhead limit\n\n").encode()
    ev = oracle.analyze(log, lib)
    assert len(ev) == 1
    e = ev[0]
    # secondary 0 on the same line: d=0 -> 0.3 * (1 - 0/3); secondary 1 five lines away: outside the window
    assert e.score == (0.9 * 1.0) * (1.0 + 0.3 * (1.0 - 0.0 / 3.0))
    assert (e.line_no, e.ctx_first, e.ctx_last) == (3, 2, 6)      # 1 before; 1 after + 2 more stack lines
    assert log[e.ctx_start:e.ctx_end].split(b"\n")[0] == b"foo" and log[e.ctx_start:e.ctx_end].endswith(b"Caused by: boom")
    s = oracle.summarize(ev, lib)
    assert s == {"highestSeverity": "CRITICAL", "significantEvents": 1, "totalEvents": 1,
                 "severityDistribution": {"CRITICAL": 1, "HIGH": 0, "MEDIUM": 0, "LOW": 0, "INFO": 0}}


def test_oracle_reproduces_golden_vectors():
    with open(GOLDEN) as f:
        gold = json.load(f)
    n = 0
    for case in gold["cases"]:
        olib = oracle.load_library(case["library"])
        for log_hex, want, nl in zip(case["logs_hex"], case["events"], case["n_lines"]):
            log = bytes.fromhex(log_hex)
            got = [[e.line_no, e.pattern, e.line_start, e.line_len, e.ctx_first, e.ctx_last, e.ctx_start, e.ctx_end, e.score]
                   for e in oracle.analyze(log, olib)]
            assert got == want, case["name"]
            assert len(oracle.split_lines(log)) == nl
            n += len(got)
    assert n > 150


# ---------------------------------------------------------------- C ABI
def _header_symbols():
    src = open(os.path.join(ROOT, "include", "podmortem_scan.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pm_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from operator_b200 import engine
    dll = engine.load_library()
    syms = _header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(dll, s), s
    assert sorted(engine.ABI_SYMBOLS) == [s for s in syms if s in engine.ABI_SYMBOLS]
    assert set(engine.ABI_SYMBOLS) <= set(syms)
    assert dll.pm_abi_version() == 3


def test_signatures_carry_no_torch_types():
    src = open(os.path.join(ROOT, "include", "podmortem_scan.h")).read()
    assert "torch" not in src and "at::" not in src and "std::" not in src
    assert 'extern "C"' in src


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="a GPU is present")
def test_init_fails_loudly_without_gpu():
    from operator_b200 import Engine, ScanError
    with pytest.raises(ScanError) as ei:
        Engine(1)
    assert ei.value.status == -6 and "no CPU path" in str(ei.value)


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under operator_b200/ may import or link it."""
    bad = []
    for dirpath, _, files in os.walk(os.path.join(ROOT, "operator_b200")):
        for fn in files:
            if fn.endswith((".py", ".cc", ".cu", ".h", ".cuh")):
                text = open(os.path.join(dirpath, fn), errors="replace").read()
                if re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M) or "scan_pcre2" in text or "liboracle" in text:
                    bad.append(fn)
    assert not bad, bad


# ---------------------------------------------------------------- multi-process plumbing (gloo, world_size 2)
_WORKER = r"""
import os, sys
sys.path.insert(0, %r)
from operator_b200 import dist
rank, local, world = dist.init("gloo")
sizes = [900, 100, 500, 400, 50, 950, 20]
shards = dist.shard_pods(sizes, world)
mine = sum(sizes[i] for i in shards[rank])
dist.barrier()
total = dist.sum_over_ranks(mine)
slow = dist.max_over_ranks(1.0 + rank)
assert total == sum(sizes), total
assert slow == float(world), slow
assert sorted(i for s in shards for i in s) == list(range(len(sizes)))
os.write(1, ("rank " + str(rank) + " ok " + str(mine) + "\n").encode())   # one syscall: the two ranks share the pipe
"""


def test_sharding_and_reductions_under_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER % ROOT)
    import socket
    with socket.socket() as sk:          # a port that is free right now
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), str(script)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "rank 0 ok" in r.stdout and "rank 1 ok" in r.stdout


def test_lpt_sharding_balances_bytes():
    from operator_b200 import dist
    sizes = [12_000_000 + 1000 * (i % 17) for i in range(64)] + [90_000_000]
    shards = dist.shard_pods(sizes, 8)
    loads = [sum(sizes[i] for i in s) for s in shards]
    assert max(loads) - min(loads) <= max(sizes)


def test_streaming_ingest_extracts_logs_like_a_json_parser():
    """8(f).3: pm_parse takes "logs" out of the request without a DOM; the span finder and the
    decoder must agree with a real JSON parser on escapes, quotes and look-alike keys."""
    from tests.conftest import CompilerHarness
    L = CompilerHarness().L
    L.pmt_json_extract.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_char_p]
    L.pmt_json_extract.restype = ctypes.c_long
    logs = ["", "plain", "a\"b", "ends with backslash\\", "\\\"", "\\\\\"\\", "tab\tnl\nquote\" \\\" \\\\\" x", "café € \U0001F600",
            "{\"logs\":\"decoy\"}", "x" * 70000 + "\"" + "\\" * 7 + "\"" + "y" * 9000, "\\" * 64, "\"" * 33]
    for log in logs:
        for doc in ({"pod": {"logs": "nested decoy", "metadata": {"name": "p\"q"}}, "logs": log, "events": ["logs", "\"logs\":\"no\""]},
                    {"logs": log}, {"a": "\\", "b": ["\"", {"logs": "inner"}], "logs": log, "z": "\"logs\""}):
            for ascii_only in (True, False):
                body = json.dumps(doc, ensure_ascii=ascii_only).encode("utf-8")
                out = ctypes.create_string_buffer(len(body) + 1)
                n = L.pmt_json_extract(body, len(body), b"logs", out)
                assert n >= 0 and out.raw[:n].decode("utf-8") == log, (log[:40], ascii_only, n)
    body = json.dumps({"pod": {"logs": "only nested"}, "events": []}).encode()
    assert L.pmt_json_extract(body, len(body), b"logs", ctypes.create_string_buffer(len(body))) == -1
    bad = b'{"logs":"bad \\u12G4 escape"}'
    assert L.pmt_json_extract(bad, len(bad), b"logs", ctypes.create_string_buffer(len(bad))) == -2


def test_split_log_units_add_up():
    """8(e): a log cut into units with halos.  The lines a unit reports (its line count under the
    Java split model minus its halos) must tile the log's lines exactly, whatever the cut points
    land on: runs of empty lines, CRLF, no final newline, fewer lines than pieces."""
    import random
    from tests.conftest import CompilerHarness
    L = CompilerHarness().L
    L.pmt_split.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int64), ctypes.c_int]
    rnd = random.Random(5)

    def nlines(b):
        return len(oracle.split_lines(b)) if b else 0

    for trial in range(300):
        sep = b"\r\n" if trial % 3 == 0 else b"\n"
        n = rnd.choice([1, 2, 3, 7, 40, 300])
        lines = [b"" if rnd.random() < 0.35 else bytes(rnd.choice(b"abc xyz\r") for _ in range(rnd.randint(1, 30))).rstrip(b"\r") for _ in range(n)]
        log = sep.join(lines) + rnd.choice([b"", sep, sep * 3])
        pieces, halo = rnd.randint(2, 6), rnd.randint(0, 5)
        out = (ctypes.c_int64 * (4 * 16))()
        k = L.pmt_split(log, len(log), pieces, halo, out, 16)
        assert 1 <= k <= pieces
        units = [tuple(out[4 * i:4 * i + 4]) for i in range(k)]
        total, reported = nlines(log), 0
        prev_core_end = 0
        for i, (b, e, lead, trail) in enumerate(units):
            assert 0 <= b <= e <= len(log) and (b == 0 or log[b - 1:b] == b"\n")
            text = log[b:e]
            core = nlines(text) - lead - trail
            assert core >= 0 and lead <= halo, (trial, units)
            assert lead == halo or b == 0                        # a full halo unless the log starts first
            if i < k - 1:
                assert trail >= halo or e == len(log)
            # the first reported line starts where the previous unit's reported lines ended
            starts = [s for s, _ in oracle.split_lines(text)] if text else []
            if core > 0:
                assert b + starts[lead] == prev_core_end, (trial, i, units)
                last = lead + core - 1
                nxt = b + starts[last + 1] if last + 1 < len(starts) else None
                prev_core_end = nxt if nxt is not None else len(log)
            reported += core
        assert reported == total, (trial, total, reported, units, log)
