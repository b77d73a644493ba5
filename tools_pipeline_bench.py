#!/usr/bin/env python
"""Throughput of the device-resident scan with several batches in flight per GPU (pm_set_pipeline):
sweeps the pipeline depth and the SM split, checks every pipelined result against the
one-batch-at-a-time result of the same corpus.  Prints one JSON line per configuration.

    python tools_pipeline_bench.py [--steps 40] [--corpus-mib 1024]
"""
import argparse
import json
import time

import numpy as np
import torch

from operator_b200 import Engine, synth


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--corpus-mib", type=int, default=1024)
    ap.add_argument("--patterns", type=int, default=500)
    ap.add_argument("--splits", default="8,16,24,32")
    ap.add_argument("--depths", default="2,3")
    args = ap.parse_args()
    torch.cuda.set_device(0)
    eng = Engine(devices=[0])
    lib_json, wit, swit = synth.make_library(args.patterns)
    eng.compile_library(lib_json).publish()
    nbytes = (args.corpus_mib << 20) - 29
    pinned = eng.pinned(nbytes)
    corpus, n_lines = synth.make_corpus(nbytes, wit, swit, p_hit=1e-3, out=pinned.array)
    nbytes = int(corpus.size)

    def fingerprint(r):
        return (int(r.n_events), int(r.line_no.astype(np.int64).sum()), int(r.pattern_index.astype(np.int64).sum()),
                int(r.ctx_end.sum()), float(r.score.sum()))

    # reference: one batch at a time, all SMs
    b = eng.batch()
    b.add_pod(corpus, 0)
    b.stage()
    for _ in range(4):
        r = b.rescan(copy=True)
    want = fingerprint(r)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        r = b.rescan(copy=False)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    print(json.dumps({"depth": 1, "scan_sms": 148, "post_sms": 148, "ms_per_step": ms, "gbs": nbytes / ms / 1e6, "k_scan_ms": r.ms_scan,
                      "step_device_ms": r.ms_total}), flush=True)
    b.release()

    for depth in [int(x) for x in args.depths.split(",")]:
        for post in [int(x) for x in args.splits.split(",")]:
            scan_sms, post_sms = eng.set_pipeline(depth, post)
            batches = []
            for _ in range(depth):
                bb = eng.batch()
                bb.add_pod(corpus, 0)
                bb.stage()
                batches.append(bb)
            ok = True
            for bb in batches:
                for _ in range(2):
                    r = bb.rescan(copy=True)
                ok = ok and fingerprint(r) == want
            # pipelined: keep `depth` scans in flight, wait in submission order
            for bb in batches:
                bb.submit()
            for i in range(2 * depth):
                bb = batches[i % depth]
                r = bb.wait(copy=True)
                ok = ok and fingerprint(r) == want
                bb.submit()
            for bb in batches:
                bb.wait(copy=False)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            e0.record()
            for bb in batches:
                bb.submit()
            scan_ms = tot_ms = 0.0
            for i in range(args.steps):
                bb = batches[i % depth]
                r = bb.wait(copy=False)
                scan_ms += r.ms_scan
                tot_ms += r.ms_total
                if i + depth < args.steps:
                    bb.submit()
            e1.record()
            torch.cuda.synchronize()
            wall = (time.perf_counter() - t0) * 1e3 / args.steps
            ms = e0.elapsed_time(e1) / args.steps
            print(json.dumps({"depth": depth, "scan_sms": scan_sms, "post_sms": post_sms, "ms_per_step": ms, "wall_ms_per_step": wall,
                              "gbs": nbytes / ms / 1e6, "k_scan_ms": scan_ms / args.steps, "step_latency_ms": tot_ms / args.steps,
                              "results_identical": ok}), flush=True)
            for bb in batches:
                bb.release()
    eng.set_pipeline(1, 0)
    pinned.free()
    eng.close()


if __name__ == "__main__":
    main()
