#!/usr/bin/env python
"""Where the time of the small post-scan kernels goes: per-warp %globaltimer stamps of one step.

Needs a -DPM_TRACE build of the library (the stamps compile to nothing otherwise):

    make trace
    PM_SCAN_LIBRARY=$PWD/operator_b200/libpodmortem_scan_trace.so python tools_warp_trace.py [--p-hit 1e-3]

Prints, per kernel, the spread of the warps' start times, of their end times, and the time between
consecutive stamps (percentiles over the warps that had work), all in microseconds from the first
stamp of that kernel.
"""
import argparse
import ctypes
import json

import numpy as np
import torch

from operator_b200 import Engine, synth
from operator_b200 import engine as _engine

REGIONS = {0: ("k_confirm", ["enter", "candidate read", "exit"]),
           1: ("k_verify", ["enter", "task read", "line bytes here", "line bounds + tables", "walk done", "literal-only: emit"]),
           2: ("k_plan", ["enter", "loop", "key + pattern + pod", "line number", "forward lines", "backward lines"]),
           3: ("k_check", ["enter", "group read", "table staged", "lines walked"])}
WARPS = 16384


def pct(a):
    if a.size == 0:
        return None
    return [round(float(x), 2) for x in np.percentile(a, [0, 10, 50, 90, 99, 100])]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--corpus-mib", type=int, default=1024)
    ap.add_argument("--patterns", type=int, default=500)
    ap.add_argument("--p-hit", type=float, default=1e-3)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    eng = Engine(devices=[0])
    L = ctypes.CDLL(_engine.LIB_PATH)   # the same handle the engine holds
    L.pm_debug_trace_read.argtypes = [ctypes.c_void_p, ctypes.c_uint64]
    lib_json, wit, swit = synth.make_library(args.patterns)
    eng.compile_library(lib_json).publish()
    nbytes = (args.corpus_mib << 20) - 29
    pinned = eng.pinned(nbytes)
    corpus, _ = synth.make_corpus(nbytes, wit, swit, p_hit=args.p_hit, out=pinned.array)
    b = eng.batch()
    b.add_pod(corpus, 0)
    b.stage()
    for _ in range(5):
        r = b.rescan(copy=False)
    buf = np.zeros(4 * WARPS * 8, dtype=np.uint64)
    rc = L.pm_debug_trace_read(buf.ctypes.data, buf.size)   # clears what the warm-up left
    assert rc == 0, rc
    r = b.rescan(copy=False)
    rc = L.pm_debug_trace_read(buf.ctypes.data, buf.size)
    assert rc == 0, rc
    t = buf.reshape(4, WARPS, 8).astype(np.int64)
    print(json.dumps({"events": int(r.n_events), "candidates": int(r.n_candidates), "verify_tasks": int(r.n_verified), "ms_step": r.ms_total,
                      "percentiles": [0, 10, 50, 90, 99, 100], "unit": "us"}))
    for reg, (name, slots) in REGIONS.items():
        x = t[reg]
        seen = x[:, 0] > 0
        if not seen.any():
            any_stamp = (x > 0).any(axis=1)
            if not any_stamp.any():
                print(json.dumps({"kernel": name, "note": "no stamps"}))
                continue
        x = x.copy()
        vals = x[:, 6:].copy()
        x[:, 6:] = 0
        if name == "k_check":
            x[:, 4:6] = 0
        t0 = x[x > 0].min()
        out = {"kernel": name, "warps_entered": int(seen.sum()), "kernel_span_us": round(float(x.max() - t0) / 1e3, 2),
               "enter": pct((x[seen, 0] - t0) / 1e3)}
        for k in range(1, len(slots)):
            have = x[:, k] > 0
            out["n[%s]" % slots[k]] = int(have.sum())
            out["at[%s]" % slots[k]] = pct((x[have, k] - t0) / 1e3)
            prev = k - 1
            if name == "k_verify" and k == 5:
                prev = 3
            both = have & (x[:, prev] > 0)
            out["d[%s -> %s]" % (slots[prev], slots[k])] = pct((x[both, k] - x[both, prev]) / 1e3)
        if name == "k_verify":
            walked = x[:, 4] > 0
            ln, tb = vals[walked, 0], vals[walked, 1] >> 1
            staged = (vals[walked, 1] & 1) == 1
            dt = (x[walked, 4] - x[walked, 3]) / 1e3
            out["walked: line bytes"] = pct(ln)
            out["walked: table bytes"] = pct(tb)
            out["walked: table in shared memory"] = "%d of %d" % (int(staged.sum()), int(walked.sum()))
            out["walked: ns per line byte"] = pct(dt * 1e3 / np.maximum(ln, 1))
            for nm, sel in (("staged", staged), ("through L1", ~staged)):
                if sel.any():
                    out["walk us, " + nm] = pct(dt[sel])
        if name == "k_check":
            grp = vals[:, 1] > 0
            if grp.any():
                ent, st = vals[grp, 0] >> 1, (vals[grp, 0] & 1) == 1
                out["groups"] = int(grp.sum())
                out["group: items"] = pct(vals[grp, 1])
                out["group: table entries"] = pct(ent)
                out["group: table in shared memory"] = "%d of %d" % (int(st.sum()), int(grp.sum()))
                w = (t[reg][:, 5] > 0)
                out["lane 0: pieces"] = pct(t[reg][w, 5])
                out["lane 0: cycles per piece"] = pct(t[reg][w, 4] / np.maximum(t[reg][w, 5], 1))
        print(json.dumps(out))
    b.release()
    pinned.free()
    eng.close()


if __name__ == "__main__":
    main()
