#!/usr/bin/env python
"""One 1 GiB batch at a time, device-resident: step / k_scan / post-scan times and a result
fingerprint, for comparing variants of the post-scan chain (environment switches of the runtime).

    python tools_tail_bench.py [--steps 30] [--corpus-mib 1024] [--patterns 500] [--p-hit 1e-3]
    PM_DEBUG_TIMING=1 python tools_tail_bench.py --steps 2      # per-stage times on stderr
"""
import argparse
import json
import os

import numpy as np
import torch

from operator_b200 import Engine, synth


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--corpus-mib", type=int, default=1024)
    ap.add_argument("--patterns", type=int, default=500)
    ap.add_argument("--p-hit", type=float, default=1e-3)
    ap.add_argument("--tag", default="")
    ap.add_argument("--stages", action="store_true", help="also print the per-stage device times of one more step (PM_DEBUG_TIMING marks)")
    args = ap.parse_args()
    if args.stages:
        os.environ["PM_DEBUG_TIMING"] = "1"   # read once, when the library records its first mark
    torch.cuda.set_device(0)
    eng = Engine(devices=[0])
    lib_json, wit, swit = synth.make_library(args.patterns)
    eng.compile_library(lib_json).publish()
    nbytes = (args.corpus_mib << 20) - 29
    pinned = eng.pinned(nbytes)
    corpus, n_lines = synth.make_corpus(nbytes, wit, swit, p_hit=args.p_hit, out=pinned.array)
    nbytes = int(corpus.size)
    b = eng.batch()
    b.add_pod(corpus, 0)
    b.stage()
    for _ in range(5):
        r = b.rescan(copy=True)
    fp = (int(r.n_events), int(r.line_no.astype(np.int64).sum()), int(r.pattern_index.astype(np.int64).sum()), int(r.ctx_end.sum()),
          float(r.score.sum()))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tot = scan = post = 0.0
    e0.record()
    for _ in range(args.steps):
        r = b.rescan(copy=False)
        tot += r.ms_total
        scan += r.ms_scan
        post += r.ms_post
    e1.record()
    torch.cuda.synchronize()
    k = args.steps
    print(json.dumps({"tag": args.tag, "env": {k_: v for k_, v in os.environ.items() if k_.startswith("PM_")}, "ms_step": tot / k, "ms_scan": scan / k,
                      "ms_post": post / k, "gbs": nbytes / (tot / k) / 1e6, "wall_ms_per_step": e0.elapsed_time(e1) / k, "events": int(r.n_events),
                      "cand": int(r.n_candidates), "verified": int(r.n_verified), "launches": int(r.gpu_launches), "fingerprint": fp}), flush=True)
    if args.stages:
        # the marks of a step are printed to stderr when it finishes: catch them for one step
        import re
        import sys
        import tempfile
        with tempfile.TemporaryFile(mode="w+") as tmp:
            sys.stderr.flush()
            saved = os.dup(2)
            os.dup2(tmp.fileno(), 2)
            try:
                b.rescan(copy=False)
            finally:
                os.dup2(saved, 2)
                os.close(saved)
            tmp.seek(0)
            stages = [(m.group(1).strip(), float(m.group(2))) for m in re.finditer(r"\[pm timing\] (.+?)\s+([0-9.]+) us", tmp.read())]
        print(json.dumps({"tag": args.tag, "stages_us": stages, "note": "every mark is an event record on the main stream and adds ~2 us"}), flush=True)
    b.release()
    pinned.free()
    eng.close()


if __name__ == "__main__":
    main()
