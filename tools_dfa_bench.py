"""Throughput of the combined-DFA tier alone (k_dfa_scan): a library whose patterns have no
7-byte literal, 1 GiB corpus resident in HBM.  Prints one JSON line; not the headline metric."""
import json
import sys

sys.path.insert(0, ".")
from operator_b200 import Engine, synth

REGEXES = [r"\bOOM\b", r"sig(segv|abrt|kill)", r"\bNPE\b", r"pid=\d+ died", r"rc=[1-9]\d* fatal", r"\[crit\]", r"E\d{4}\b:", r"^panic:"]


def main(nbytes=1 << 30, steps=10):
    eng = Engine(1)
    lib = {"patterns": [{"id": "t%d" % i, "primary_pattern": {"regex": r}} for i, r in enumerate(REGEXES)]}
    rep = eng.compile_library(lib).publish().report
    _, wit, swit = synth.make_library(50)
    pinned = eng.pinned(nbytes)
    corpus, n_lines = synth.make_corpus(nbytes, wit, swit, p_hit=1e-3, out=pinned.array)
    b = eng.batch()
    b.add_pod(corpus, 0)
    b.stage()
    for _ in range(3):
        r = b.rescan(copy=False)
    ms = 0.0
    for _ in range(steps):
        r = b.rescan(copy=False)
        ms += r.ms_dfa
    ms /= steps
    print(json.dumps({"kernel": "k_dfa_scan", "patterns": len(REGEXES), "dfa_patterns": rep["n_dfa_scanned"], "groups": rep["dfa_groups"],
                      "states": rep["dfa_states"], "bytes": int(corpus.size), "ms": ms, "GB/s": corpus.size / ms / 1e6, "events": int(r.n_events),
                      "frac_of_measured_hbm": corpus.size / ms / 1e6 / 6532.5}))
    b.release()
    pinned.free()
    eng.close()


if __name__ == "__main__":
    main()
