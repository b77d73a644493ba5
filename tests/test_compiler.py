"""CPU tests of the pattern compiler (a9): C++ parser/NFA/DFA/literal analysis vs the oracle's
independent Python parser + Python ``re`` (and PCRE2) on the same patterns and lines."""
import ctypes
import re

import numpy as np
import pytest

from oracle import javaregex, oracle, pcre2_oracle
from operator_b200 import synth
from tests import cases


def _py(rx):
    return re.compile(javaregex.to_python(rx), re.ASCII)


def test_verify_dfa_matches_python_re(harness):
    rc, h, rep = harness.compile_regexes(cases.REGEXES)
    assert rc == 0, rep.message
    pys = [_py(r) for r in cases.REGEXES]
    bad = []
    for line in cases.LINES:
        raw = line.encode("utf-8")
        for i, p in enumerate(pys):
            want = p.search(line) is not None
            got = bool(harness.L.pmt_primary_match(h, i, raw, len(raw)))
            if want != got:
                bad.append((cases.REGEXES[i], line[:60], want, got))
    assert not bad, bad[:10]
    harness.L.pmt_free(h)


@pytest.mark.skipif(not pcre2_oracle.available(), reason="libpcre2-8 not loadable")
def test_python_oracle_agrees_with_pcre2():
    lib = pcre2_oracle.Pcre2Library(cases.REGEXES)
    blob = "\n".join(cases.LINES).encode("utf-8") + b"\n"
    arr = np.frombuffer(blob, dtype=np.uint8)
    starts, lens = pcre2_oracle.split(arr)
    ol, op, total = pcre2_oracle.scan(lib, arr, starts, lens, threads=2)
    got = set(zip(ol.tolist(), op.tolist()))
    want = set()
    spans = oracle.split_lines(blob)
    assert [s for s, _ in spans] == starts.tolist() and [l for _, l in spans] == lens.tolist()
    for li, (s, l) in enumerate(spans):
        text = blob[s:s + l].decode("utf-8")
        for pi, rx in enumerate(cases.REGEXES):
            if _py(rx).search(text):
                want.add((li, pi))
    assert got == want, (sorted(got - want)[:5], sorted(want - got)[:5])


def test_rejections_match_oracle_parser(harness):
    for rx, status, off in cases.REJECTED:
        rc, h, rep = harness.compile_regexes([rx])
        assert rc == status, (rx, rc, rep.message)
        assert rep.pattern_index == 0
        with pytest.raises(javaregex.JavaRegexError) as ei:
            javaregex.parse(rx)
        kind = {-2: "syntax", -3: "unsupported", -4: "too_complex"}[status]
        assert ei.value.kind == kind, (rx, ei.value)
        if off is not None:
            assert rep.regex_offset == off, (rx, rep.regex_offset, rep.message)
            assert ei.value.offset == off, (rx, ei.value)


def test_report_names_secondary(harness):
    spec = {"patterns": [{"id": "ok", "primary_pattern": {"regex": "fine pattern"}},
                         {"id": "bad-one", "primary_pattern": {"regex": "also fine"},
                          "secondary_patterns": [{"regex": "good"}, {"regex": "(?=x)y"}]}]}
    rc, h, rep = harness.compile(spec)
    assert rc == -3 and rep.pattern_index == 1 and rep.secondary_index == 1 and rep.pattern_id == b"bad-one"


def test_library_json_errors(harness):
    assert harness.compile("{")[0] == -2
    assert harness.compile({"patterns": [{"id": "x"}]})[0] == -2
    assert harness.compile({"patterns": [{"id": "x", "severity": "BOGUS", "primary_pattern": {"regex": "abc"}}]})[0] == -2
    rc, h, rep = harness.compile({"patterns": []})
    assert rc == 0 and rep.n_patterns == 0


def test_tiers_and_literals(harness):
    rx = [r"OutOfMemoryError", r"(?i)out of memory", r"(ERROR|FATAL).*database", r"[Cc]onnection (refused|reset)", r"\bOOM\b", r"\d+ms"]
    rc, h, rep = harness.compile_regexes(rx)
    assert rc == 0
    tiers = [harness.L.pmt_pattern_tier(h, i) for i in range(len(rx))]
    assert tiers == [1, 1, 1, 1, 2, 2]
    assert rep.n_filtered == 4 and rep.n_dfa_scanned == 2 and rep.filter_fold_case == 1 and rep.dfa_groups == 1
    assert harness.L.pmt_literals(h, 0) == b'["OutOfMemoryError"]'
    assert harness.L.pmt_literals(h, 3) == b'["connection refused","connection reset"]'


@pytest.mark.parametrize("n_patterns", [50, 500])
def test_synthetic_library_vs_oracle(harness, n_patterns):
    """Every line of a synthetic corpus: verify DFA == oracle, and the gram filter is SOUND
    (a matching line always triggers its pattern, at every buffer alignment)."""
    lib_json, wit, swit = synth.make_library(n_patterns)
    rc, h, rep = harness.compile(lib_json)
    assert rc == 0, rep.message
    assert rep.n_patterns == n_patterns
    lib = oracle.load_library(lib_json)
    corpus, n_lines = synth.make_corpus((256 << 10) if n_patterns <= 50 else (96 << 10), wit, swit, p_hit=0.05)
    blob = bytes(corpus)
    hits = 0
    for s, l in oracle.split_lines(blob):
        raw = blob[s:s + l]
        text = raw.decode("utf-8")
        for p in lib.patterns:
            want = p.rx.search(text) is not None
            got = bool(harness.L.pmt_primary_match(h, p.index, raw, l))
            assert want == got, (p.regex, text)
            if want:
                hits += 1
                if harness.L.pmt_pattern_tier(h, p.index) == 1:
                    for phase in range(4):
                        assert harness.L.pmt_filter_triggers(h, p.index, raw, l, phase), (p.regex, text, phase)
        assert bool(harness.L.pmt_stack_match(h, raw, l)) == (lib.stack_rx.search(text) is not None)
    assert hits > 20
    # every witness is recognised by its own pattern (and by the filter)
    for i, w in enumerate(wit):
        raw = ("2026-01-01 ERROR [x] (main) " + w).encode()
        assert harness.L.pmt_primary_match(h, i, raw, len(raw)), (i, w)
        if harness.L.pmt_pattern_tier(h, i) == 1:
            assert all(harness.L.pmt_filter_triggers(h, i, raw, len(raw), ph) for ph in range(4)), (i, w)
    harness.L.pmt_free(h)


def test_combined_dfa_groups(harness):
    """Tier-2 patterns: the combined (multi-pattern) DFA reports exactly the patterns whose
    own verify DFA accepts."""
    rc, h, rep = harness.compile_regexes(cases.REGEXES)
    assert rc == 0
    tier2 = [i for i in range(len(cases.REGEXES)) if harness.L.pmt_pattern_tier(h, i) == 2]
    assert tier2 and rep.dfa_groups >= 1
    out = (ctypes.c_int * 128)()
    for line in cases.LINES:
        raw = line.encode("utf-8")
        n = harness.L.pmt_group_match(h, raw, len(raw), out, 128)
        got = sorted(out[i] for i in range(n))
        want = [i for i in tier2 if harness.L.pmt_primary_match(h, i, raw, len(raw))]
        assert got == want, (line[:60], got, want)
        # the device form of the same automaton (newline folded in, premultiplied rows)
        n2 = harness.L.pmt_group_scan_match(h, raw, len(raw), out, 128)
        assert sorted(out[i] for i in range(n2)) == want, (line[:60], want)
    harness.L.pmt_free(h)


def test_scan_table_line_sequences(harness):
    """The scan table is walked across line boundaries without a reset: feeding several lines
    back to back (separated by the newline byte) must give each line's own verdict."""
    rx = [r"^\s+at\s", r"\bOOM\b", r"a$", r"foo\z", r"^ERROR", r"x{3,5}y", r"\d+ms"]
    rc, h, rep = harness.compile_regexes(rx)
    assert rc == 0 and rep.n_dfa_scanned == len(rx)
    out = (ctypes.c_int * 64)()
    lines = [b"ERROR first", b"  at x", b"not ERROR", b"OOM", b"ZOOM", b"a", b"ab", b"foo", b"xxxy 12ms", b"", b"ERROR"]
    for i, line in enumerate(lines):
        prefix = b"\n".join(lines[:i]) + (b"\n" if i else b"")
        whole = prefix + line
        # verdict of the last line when walked after all previous ones = verdict of the line alone,
        # up to the patterns matched by earlier lines (which the walk also reports)
        n_all = harness.L.pmt_group_scan_match(h, whole, len(whole), out, 64)
        got_all = set(out[k] for k in range(n_all))
        n_alone = harness.L.pmt_group_match(h, line, len(line), out, 64)
        alone = set(out[k] for k in range(n_alone))
        earlier = set()
        for prev in lines[:i]:
            m = harness.L.pmt_group_match(h, prev, len(prev), out, 64)
            earlier |= set(out[k] for k in range(m))
        assert got_all == alone | earlier, (i, line, got_all, alone, earlier)
    harness.L.pmt_free(h)


def test_large_library_compiles(harness):
    lib_json, wit, _ = synth.make_library(20000)
    rc, h, rep = harness.compile(lib_json)
    assert rc == 0, rep.message
    assert rep.n_patterns == 20000 and rep.n_filtered == 20000 and 15 <= rep.filter_log2_bits <= 20
    harness.L.pmt_free(h)


def test_compiled_library_survives_a_round_trip_through_bytes(harness):
    """8(f).2: the sync step leaves the compiled library beside the YAML.  Same blob, metadata and
    literals after serialize -> deserialize; damaged or foreign bytes are refused."""
    import ctypes
    from operator_b200 import synth
    L = harness.L
    L.pmt_saved.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_size_t)]
    L.pmt_saved.restype = ctypes.POINTER(ctypes.c_uint8)
    L.pmt_load_status.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
    L.pmt_roundtrip_equal.argtypes = [ctypes.c_void_p]
    for spec in (synth.make_library(50)[0], synth.make_library(500)[0],
                 {"patterns": [{"id": "t", "primary_pattern": {"regex": r"\bOOM\b"}}, {"id": "u", "name": "café", "primary_pattern": {"regex": "(?i)connection refused"}}]}):
        rc, h, rep = harness.compile(spec)
        assert rc == 0
        assert L.pmt_roundtrip_equal(h) == 1
        n = ctypes.c_size_t()
        p = L.pmt_saved(h, ctypes.byref(n))
        data = ctypes.string_at(p, n.value)
        assert L.pmt_load_status(data, len(data)) == 0
        assert L.pmt_load_status(data[:-1], len(data) - 1) == -1                    # truncated
        assert L.pmt_load_status(data[:100], 100) == -1
        bad = bytearray(data)
        bad[len(bad) // 2] ^= 0x40
        assert L.pmt_load_status(bytes(bad), len(bad)) == -1                         # checksum
        assert L.pmt_load_status(b"", 0) == -1 and L.pmt_load_status(b"x" * 64, 64) == -1
        L.pmt_free(h)
