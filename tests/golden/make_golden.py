"""Generates tests/golden/golden_events.json from the oracle (run from the repo root:
`python tests/golden/make_golden.py`).  The reference has no fixtures for this path
(SURVEY.md section 4, 8(c)), so these vectors pin the ORACLE: CPU tests check that the oracle still
reproduces them, GPU tests check that the device reproduces them without the oracle."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import oracle  # noqa: E402
from operator_b200 import synth  # noqa: E402
from tests import cases  # noqa: E402


def case(name, lib_json, logs):
    olib = oracle.load_library(lib_json)
    evs, nl = [], []
    for log in logs:
        evs.append([[e.line_no, e.pattern, e.line_start, e.line_len, e.ctx_first, e.ctx_last, e.ctx_start, e.ctx_end, e.score]
                    for e in oracle.analyze(log, olib)])
        nl.append(len(oracle.split_lines(log)))
    return {"name": name, "library": lib_json, "logs_hex": [l.hex() for l in logs], "events": evs, "n_lines": nl}


def main():
    out = []
    pats = [{"id": "p%d" % i, "name": "n%d" % i, "severity": ["INFO", "LOW", "MEDIUM", "HIGH", "CRITICAL"][i % 5],
             "primary_pattern": {"regex": r, "confidence": 0.5 + 0.05 * (i % 11)},
             "context_extraction": {"lines_before": 2, "lines_after": 3, "include_stack_trace": True}}
            for i, r in enumerate(cases.REGEXES)]
    body = "\n".join(cases.LINES)
    out.append(case("handwritten", {"patterns": pats}, [body.encode("utf-8"), ("\r\n".join(cases.LINES) + "\r\n").encode("utf-8")]))
    lib_json, wit, swit = synth.make_library(40)
    logs = [bytes(synth.make_corpus(60000, wit, swit, p_hit=3e-2)[0]), bytes(synth.make_corpus(30000, wit, swit, p_hit=3e-2, first_line=10 ** 6, crlf=True)[0]),
            b"", b"\n\n"]
    out.append(case("synthetic40", lib_json, logs))
    with open(os.path.join(ROOT, "tests", "golden", "golden_events.json"), "w") as f:
        json.dump({"generator": "tests/golden/make_golden.py", "cases": out}, f, separators=(",", ":"))
    print("wrote", sum(len(e) for c in out for e in c["events"]), "events")


if __name__ == "__main__":
    main()
