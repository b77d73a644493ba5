#!/usr/bin/env python
"""k_scan_half (filter granule 2) against the combined-DFA route for a library with 5-6 byte literals:
    python tools_short_bench.py            # granule 2
    PM_FILTER_GRANULE4=1 python tools_short_bench.py   # same library forced to granule 4: the short literals go to k_dfa_scan
"""
import json
import os

import torch

from operator_b200 import Engine, synth


def main():
    torch.cuda.set_device(0)
    eng = Engine(devices=[0])
    lib, wit, swit = synth.make_mixed_library(500, min_literal=5)
    rep = eng.compile_library(lib).publish().report
    nbytes = 1 << 30
    t = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    n, lines = synth.make_corpus_device(t.data_ptr(), nbytes, wit, swit, p_hit=1e-3, first_line=79 << 36)
    b = eng.batch()
    b.add_pod_device(t.data_ptr(), n, 0, keep=t)
    b.stage()
    for _ in range(5):
        r = b.rescan(copy=True)
    fp = (int(r.n_events), int(r.line_no.astype("int64").sum()), int(r.pattern_index.astype("int64").sum()), float(r.score.sum()))
    tot = scan = dfa = 0.0
    k = 20
    for _ in range(k):
        r = b.rescan(copy=False)
        tot += r.ms_total
        scan += r.ms_scan
        dfa += r.ms_dfa
    print(json.dumps({"granule4_forced": "PM_FILTER_GRANULE4" in os.environ, "n_filtered": rep["n_filtered"], "n_dfa_scanned": rep["n_dfa_scanned"],
                      "ms_step": tot / k, "ms_scan": scan / k, "ms_dfa": dfa / k, "gbs": n / (tot / k) / 1e6, "k_scan_gbs": n / (scan / k) / 1e6,
                      "cand": int(r.n_candidates), "events": int(r.n_events), "fingerprint": fp}))
    b.release()
    eng.close()


if __name__ == "__main__":
    main()
