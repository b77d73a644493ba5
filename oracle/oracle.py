"""TEST INFRASTRUCTURE (oracle) -- never imported by the product path.

CPU restatement of the log-scan path that the operator reaches through
``POST /parse`` (LogParserRestClient.java:37-39, called from LogParserClient.java:40-41,
PodFailureWatcher.java:285-286 and PodmortemReconciler.java:146-147).  The service behind that
call (image ghcr.io/podmortem/podmortem-log-parser:latest, log-parser-deployment.yaml:27) and the
model jar com.redhat.podmortem:common:1.0-2112dab-SNAPSHOT (pom.xml:19, :83-87) are NOT in
/root/reference and the reference has no tests, so: **parity unpinned**.  The rules restated here
are the ones fixed in SPEC.md; the consumer-side anchors that do exist in the tree are cited on
the functions that honour them.

Per line x per pattern ``re.search`` (the shape of the JVM loop), scoring and context selection
in plain Python.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import this module.
"""
from __future__ import annotations

import json
import re
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

from . import javaregex

SEVERITIES = ["INFO", "LOW", "MEDIUM", "HIGH", "CRITICAL"]          # rank = index
SEV_MULT = {"CRITICAL": 1.0, "HIGH": 0.8, "MEDIUM": 0.6, "LOW": 0.4, "INFO": 0.2}
STACK_MAX = 50
STACK_RX = r"^(?:\s+at\s|Caused by:|\s+\.\.\.\s\d+\s(?:more|common frames omitted)|\s+Suppressed:)"
SIGNIFICANT_SCORE = 0.5
MAX_SECONDARY = 8


def compile_java(rx: str) -> "re.Pattern[str]":
    return re.compile(javaregex.to_python(rx), re.ASCII)


@dataclass
class Secondary:
    regex: str
    weight: float
    window: int
    rx: "re.Pattern[str]"


@dataclass
class Pattern:
    index: int
    id: str
    name: str
    severity: str
    confidence: float
    regex: str
    rx: "re.Pattern[str]"
    secondaries: List[Secondary]
    lines_before: int
    lines_after: int
    include_stack: bool
    category: str = ""


@dataclass
class Library:
    patterns: List[Pattern]
    stack_rx: "re.Pattern[str]" = field(default_factory=lambda: compile_java(STACK_RX))


class LibraryError(Exception):
    def __init__(self, kind, pattern_index, offset, msg):
        super().__init__(f"pattern {pattern_index}: {kind} at {offset}: {msg}")
        self.kind, self.pattern_index, self.offset = kind, pattern_index, offset


def load_library(spec) -> Library:
    """spec: dict, list of dicts, or JSON text of either (SPEC.md section 1)."""
    if isinstance(spec, (str, bytes)):
        spec = json.loads(spec)
    libs = spec if isinstance(spec, list) else [spec]
    out: List[Pattern] = []
    for lib in libs:
        for p in lib.get("patterns", []):
            idx = len(out)
            prim = p.get("primary_pattern") or {}
            sev = str(p.get("severity", "INFO")).upper()
            if sev not in SEV_MULT:
                raise LibraryError("spec", idx, 0, "unknown severity %r" % sev)
            try:
                rx = compile_java(prim["regex"])
                secs = []
                for s in p.get("secondary_patterns") or []:
                    secs.append(Secondary(s["regex"], float(s.get("weight", 0.0)),
                                          int(s.get("proximity_window", 0)), compile_java(s["regex"])))
            except javaregex.JavaRegexError as e:
                raise LibraryError(e.kind, idx, e.offset, e.msg) from None
            if len(secs) > MAX_SECONDARY:
                raise LibraryError("spec", idx, 0, "more than %d secondary patterns" % MAX_SECONDARY)
            ctx = p.get("context_extraction") or {}
            out.append(Pattern(idx, str(p.get("id", "")), str(p.get("name", "")), sev,
                               float(prim.get("confidence", 1.0)), prim["regex"], rx, secs,
                               int(ctx.get("lines_before", 0)), int(ctx.get("lines_after", 0)),
                               bool(ctx.get("include_stack_trace", False)), str(p.get("category", ""))))
    return Library(out)


def split_lines(log: bytes) -> List[Tuple[int, int]]:
    """Java ``logs.split("\\\\r?\\\\n")`` on the UTF-8 bytes -> [(start, length)] (SPEC.md section 2).

    \\r and \\n are ASCII, so splitting the bytes equals splitting the decoded String.
    """
    if not log:
        return [(0, 0)]
    lines: List[Tuple[int, int]] = []
    start = 0
    n = len(log)
    while True:
        nl = log.find(b"\n", start)
        end = n if nl < 0 else nl
        ln = end - start
        if nl >= 0 and ln > 0 and log[end - 1] == 0x0D:
            ln -= 1
        lines.append((start, ln))
        if nl < 0:
            break
        start = nl + 1
    # a separator was found at least once, or the single piece is the whole input
    if len(lines) > 1:
        while lines and lines[-1][1] == 0:
            lines.pop()
    return lines


@dataclass
class Event:
    line_no: int            # 1-based
    pattern: int
    line_start: int         # byte offset in the pod log
    line_len: int
    ctx_first: int          # first context line (1-based, inclusive)
    ctx_last: int
    ctx_start: int          # byte range of lines ctx_first..ctx_last, no final separator
    ctx_end: int
    score: float


def analyze(log: bytes, lib: Library) -> List[Event]:
    """All events of one pod log, ordered by (line, pattern index) -- SPEC.md sections 4-6."""
    spans = split_lines(log)
    texts = [log[s:s + l].decode("utf-8") for s, l in spans]
    n = len(spans)
    sec_cache: Dict[Tuple[int, int, int], bool] = {}
    stack_cache: Dict[int, bool] = {}

    def sec_match(p: Pattern, j: int, line: int) -> bool:
        key = (p.index, j, line)
        v = sec_cache.get(key)
        if v is None:
            v = p.secondaries[j].rx.search(texts[line - 1]) is not None
            sec_cache[key] = v
        return v

    def is_stack(line: int) -> bool:
        v = stack_cache.get(line)
        if v is None:
            v = lib.stack_rx.search(texts[line - 1]) is not None
            stack_cache[line] = v
        return v

    events: List[Event] = []
    for li, text in enumerate(texts):
        L = li + 1
        for p in lib.patterns:
            if p.rx.search(text) is None:
                continue
            acc = 0.0
            for j, s in enumerate(p.secondaries):
                best: Optional[int] = None
                for d in range(0, s.window + 1):
                    hit = False
                    for cand in ((L - d, L + d) if d else (L,)):
                        if 1 <= cand <= n and sec_match(p, j, cand):
                            hit = True
                    if hit:
                        best = d
                        break
                if best is not None:
                    acc = acc + s.weight * (1.0 - float(best) / float(s.window + 1))
            score = (p.confidence * SEV_MULT[p.severity]) * (1.0 + acc)
            first = max(1, L - p.lines_before)
            last = min(n, L + p.lines_after)
            if p.include_stack:
                while last < n and last - L < STACK_MAX and is_stack(last + 1):
                    last += 1
            cs = spans[first - 1][0]
            ce = spans[last - 1][0] + spans[last - 1][1]
            events.append(Event(L, p.index, spans[li][0], spans[li][1], first, last, cs, ce, score))
    return events


def summarize(events: Sequence[Event], lib: Library) -> dict:
    """The fields the operator reads back: EventService.java:75-83 (highestSeverity,
    significantEvents), AnalysisStorageService.java:145-160 and :298-325 (events.size(), first
    five events in list order, pattern name/severity, score)."""
    dist = {s: 0 for s in reversed(SEVERITIES)}
    top = -1
    sig = 0
    for e in events:
        sev = lib.patterns[e.pattern].severity
        dist[sev] += 1
        top = max(top, SEVERITIES.index(sev))
        if e.score >= SIGNIFICANT_SCORE:
            sig += 1
    return {"highestSeverity": SEVERITIES[top] if top >= 0 else "NONE",
            "significantEvents": sig, "totalEvents": len(events), "severityDistribution": dist}
