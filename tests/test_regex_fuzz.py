"""Three-way differential fuzz over the SPEC.md section 3 regex subset (CPU): random regexes from
tests/regex_fuzz.py, each against lines built from its own samples, decided three times --

  1. Python ``re`` on ``oracle.javaregex.to_python``   (the oracle every parity claim rests on),
  2. PCRE2 on ``oracle.javaregex.to_pcre2``            (a second engine, a second translation),
  3. the compiled verify DFA, walked on the host by tests/support (what the kernels walk).

A disagreement between 1 and 2 is a translator (or subset) problem in the oracle; between 1 and 3 a
compiler problem.  Fixed seed: the run is reproducible and its size is bounded.
"""
import random

import numpy as np
import pytest

from oracle import javaregex, pcre2_oracle
from tests import regex_fuzz

import os
SEED = int(os.environ.get("PM_FUZZ_SEED", "20261107"))
N_BATCHES = 8
REGEXES_PER_BATCH = 60
LINES_PER_REGEX = 8


def _batch(rnd):
    nodes, regexes = [], []
    while len(regexes) < REGEXES_PER_BATCH:
        node = regex_fuzz.random_regex(rnd)
        rx = node.regex()
        if rx in regexes:
            continue
        nodes.append(node)
        regexes.append(rx)
    return nodes, regexes


@pytest.mark.skipif(not pcre2_oracle.available(), reason="libpcre2-8 not loadable")
def test_three_engines_agree_on_random_regexes(harness):
    import re
    rnd = random.Random(SEED)
    n_pairs = n_match = n_rejected = 0
    for _ in range(N_BATCHES):
        nodes, regexes = _batch(rnd)
        # the compiler and the oracle's parser must reject the same regexes (for instance those that can
        # match the empty string); what both accept is tested
        keep = []
        for node, rx in zip(nodes, regexes):
            rc, h, rep = harness.compile_regexes([rx])
            try:
                javaregex.parse(rx)
                oracle_ok = True
            except javaregex.JavaRegexError:
                oracle_ok = False
            if rc != 0 and oracle_ok and b"never match" in rep.message:
                # the compiler proves emptiness (contradicting assertions such as \B.\b between word characters) and
                # refuses the pattern; the oracle has no such analysis.  Consistent iff nothing matches it.
                p = re.compile(javaregex.to_python(rx), re.ASCII)
                assert not any(p.search(l) for l in regex_fuzz.lines_for(node, rnd, 16)), rx
                n_rejected += 1
                continue
            if rc == -4 and oracle_ok and (b"state limit" in rep.message or b"shared-memory budget" in rep.message):
                n_rejected += 1   # too large a DFA: a limit of the compiler alone (PM_E_TOO_COMPLEX), nothing to compare
                continue
            assert (rc == 0) == oracle_ok, (rx, rc, rep.message)
            if rc == 0:
                harness.L.pmt_free(h)
                keep.append((node, rx))
            else:
                n_rejected += 1
        nodes, regexes = [k[0] for k in keep], [k[1] for k in keep]
        lines = []
        for node in nodes:
            lines += regex_fuzz.lines_for(node, rnd, LINES_PER_REGEX)
        lines.append("last line")   # Java's split drops trailing empty lines: keep the count stable
        rc, h, rep = harness.compile_regexes(regexes)
        assert rc == 0, rep.message
        pys = [re.compile(javaregex.to_python(rx), re.ASCII) for rx in regexes]
        blob = "\n".join(lines).encode("utf-8") + b"\n"
        arr = np.frombuffer(blob, dtype=np.uint8)
        starts, lens = pcre2_oracle.split(arr)
        assert starts.size == len(lines)
        plib = pcre2_oracle.Pcre2Library(regexes)
        ol, op, total = pcre2_oracle.scan(plib, arr, starts, lens, threads=2)
        plib.close()
        assert total == ol.size
        pcre = set(zip(ol.tolist(), op.tolist()))
        bad = []
        for li, line in enumerate(lines):
            raw = line.encode("utf-8")
            for pi, p in enumerate(pys):
                want = p.search(line) is not None
                got_dfa = bool(harness.L.pmt_primary_match(h, pi, raw, len(raw)))
                got_pcre = (li, pi) in pcre
                n_pairs += 1
                n_match += want
                if not (want == got_dfa == got_pcre):
                    bad.append((regexes[pi], line, {"python": want, "pcre2": got_pcre, "dfa": got_dfa}))
        harness.L.pmt_free(h)
        assert not bad, bad[:8]
    # the generator must produce matches, not only noise, and most regexes must be inside the subset
    print('fuzz: %d pairs, %d matches, %d rejected regexes' % (n_pairs, n_match, n_rejected))
    assert n_pairs > 100000 and n_match > 2000, (n_pairs, n_match)
    assert n_rejected < N_BATCHES * REGEXES_PER_BATCH // 3, n_rejected


# ---- patterns that are nothing but their literals: k_verify takes the confirmed literal as the match
# (UnitDesc::literal_is_match, pm_literal.cc literal_hit_is_match).  Sound iff "a member of the literal set
# occurs in the line" and "the regex finds a match in the line" are the same predicate.

_PLAIN_FLAGGED = [
    r"OutOfMemoryError", r"(?i)connection refused", r"(timeout exceeded|deadline exceeded)", r"Fail(ed|ure) to bind",
    r"exit code 13[79] seen", r"\QTypeError: x.y(z)\E", r"segfault at 0x0{4}", r"killed by signal 9!?",
]
_PLAIN_NOT_FLAGGED = [
    r"\bfatal error\b", r"^ERROR happened", r"failed to start$", r"error: .*failed", r"exit code [0-9]+ seen",
    "naïve résumé x", r"retrying \w+ connection", r"disk full\r",
]


def test_literal_is_match_flag_known_cases(harness):
    harness.L.pmt_literal_is_match.argtypes = [__import__("ctypes").c_void_p, __import__("ctypes").c_int]
    rc, h, rep = harness.compile_regexes(_PLAIN_FLAGGED + _PLAIN_NOT_FLAGGED)
    assert rc == 0, rep.message
    flags = [harness.L.pmt_literal_is_match(h, i) for i in range(len(_PLAIN_FLAGGED) + len(_PLAIN_NOT_FLAGGED))]
    harness.L.pmt_free(h)
    assert flags[:len(_PLAIN_FLAGGED)] == [1] * len(_PLAIN_FLAGGED), list(zip(_PLAIN_FLAGGED, flags))
    assert flags[len(_PLAIN_FLAGGED):] == [0] * len(_PLAIN_NOT_FLAGGED), list(zip(_PLAIN_NOT_FLAGGED, flags[len(_PLAIN_FLAGGED):]))


def _plain_regex(rnd):
    """(regex, sample strings): a literal, an alternation, small sets and optional pieces, maybe (?i)."""
    alphabet = "abcdeXYZ 019:_-./"
    def word(lo, hi):
        return "".join(rnd.choice(alphabet) for _ in range(rnd.randint(lo, hi)))
    def esc(s):
        return "".join("\\" + c if c in ".-/" and rnd.random() < 0.7 else c for c in s)
    kind = rnd.randrange(5)
    if kind == 0:
        w = word(7, 18)
        rx, samples = esc(w), [w]
    elif kind == 1:
        ws = [word(7, 14) for _ in range(rnd.randint(2, 4))]
        rx, samples = "(" + "|".join(esc(w) for w in ws) + ")", ws
    elif kind == 2:
        a, b, c = word(4, 8), word(1, 3), word(4, 8)
        rx, samples = "%s(%s)?%s" % (esc(a), esc(b), esc(c)), [a + c, a + b + c]
    elif kind == 3:
        a, c = word(4, 8), word(4, 8)
        rx, samples = "%s[ab]%s" % (esc(a), esc(c)), [a + "a" + c, a + "b" + c]
    else:
        a, c = word(3, 6), word(4, 8)
        rx, samples = "%s(XY|01){2}%s" % (esc(a), esc(c)), [a + "XY01" + c, a + "0101" + c]
    if rnd.random() < 0.4:
        rx = "(?i)" + rx
        samples = samples + [s.upper() for s in samples] + [s.swapcase() for s in samples]
    return rx, samples


def test_literal_is_match_is_the_regex(harness):
    import ctypes
    import re
    harness.L.pmt_literal_is_match.argtypes = [ctypes.c_void_p, ctypes.c_int]
    rnd = random.Random(SEED + 1)
    alphabet = "abcdeXYZ 019:_-./"
    n_flagged = n_pairs = n_hits = 0
    for _ in range(6):
        regexes, samples = [], []
        while len(regexes) < 40:
            rx, sm = _plain_regex(rnd)
            try:
                javaregex.parse(rx)
            except javaregex.JavaRegexError:
                continue
            regexes.append(rx)
            samples.append(sm)
        rc, h, rep = harness.compile_regexes(regexes)
        assert rc == 0, rep.message
        lines = []
        for sm in samples:
            for s in sm[:4]:
                pre = "".join(rnd.choice(alphabet) for _ in range(rnd.randint(0, 9)))
                lines.append(pre + s + " tail")
                k = rnd.randrange(len(s))
                lines.append(pre + s[:k] + ("#" if s[k] != "#" else "%") + s[k + 1:])    # one byte off
                lines.append(pre + s[:-1])                                               # cut short
                lines.append(pre + s.swapcase() + ".")                                    # case flipped
        pys = [re.compile(javaregex.to_python(rx), re.ASCII) for rx in regexes]
        for pi, rx in enumerate(regexes):
            if not harness.L.pmt_literal_is_match(h, pi):
                continue
            n_flagged += 1
            for line in lines:
                raw = line.encode()
                want = pys[pi].search(line) is not None
                n_pairs += 1
                n_hits += want
                for phase in range(4):
                    got = bool(harness.L.pmt_filter_triggers(h, pi, raw, len(raw), phase))
                    assert got == want, (rx, line, phase, got, want)
        harness.L.pmt_free(h)
    print("literal-is-match: %d flagged patterns, %d pairs, %d hits" % (n_flagged, n_pairs, n_hits))
    assert n_flagged > 150 and n_hits > 500, (n_flagged, n_hits)
