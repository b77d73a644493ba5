#!/usr/bin/env python
"""Benchmark of the log-scan hot path (BASELINE.json metric: log-scan GB/s, % of HBM roofline).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # the CPU loop on host cores

A step is one pass of the scan over one batch: the synthetic 1 GiB crash-log corpus of BASELINE
config 2 against the 500-pattern synthetic library, per GPU (pods shard over ranks with no
collective, so N ranks scan N corpora: weak scaling).  One JSON line is printed by rank 0.

  value      corpus bytes / device time of a step with the corpus already resident in HBM
             (one CUDA-event pair from the first kernel to the last result copy; max over ranks)
  e2e        the same metric through the public C ABI with HOST buffers: begin/add_pod/submit/
             wait/release per step, H2D of the log and D2H of the results inside the timed region
  roofline   k_scan alone: corpus bytes / its CUDA-event duration against the measured HBM peak
  cpu_baseline  the reference-style loop (per line x per pattern regex search) on host cores over
             a bounded sample, timed beside it (N=1 only)
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CORPUS_BYTES = 1 << 30
N_PATTERNS = 500
P_HIT = 1e-3
CPU_SAMPLE_BYTES = 1 << 30      # the whole config-2 corpus: ~15 s on 16 cores, no extrapolation needed


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--corpus-bytes", type=int, default=CORPUS_BYTES)
    ap.add_argument("--patterns", type=int, default=N_PATTERNS)
    ap.add_argument("--depth", type=int, default=5, help="batches in flight per GPU in the throughput legs (pm_set_pipeline); 1 = one at a time")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--config5-gib", type=int, default=64, help="size of the fixed corpus of the strong-scaling leg (0: skip)")
    ap.add_argument("--config3-pods-per-gpu", type=int, default=1250, help="pods per GPU of the crash-storm leg (0: skip)")
    ap.add_argument("--ref-budget-s", type=float, default=150.0, help="--impl reference: wall-clock budget for warm-up + timed steps")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md)."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index: str):
        self.samples = []
        self.proc = None
        self.gpu = gpu_index
        self.marks = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", self.gpu, "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def mark(self):
        self.marks.append(time.time())

    def stop(self):
        if self.proc:
            self.proc.terminate()

    def summary(self):
        rows = []
        lo = self.marks[0] if self.marks else 0
        hi = self.marks[-1] if len(self.marks) > 1 else 1e18
        for t, line in self.samples:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                rows.append((t, float(parts[0]), float(parts[1]), parts[3:7]))
            except ValueError:
                continue
        sel = [r for r in rows if lo - 0.05 <= r[0] <= hi + 0.05] or rows
        if not sel:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        mhz = sorted(r[1] for r in sel)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in sel for i in range(4) if r[3][i].lower().startswith("active")})
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_max_mhz": sel[0][2], "reasons": reasons, "samples": len(sel)}


def peak_hbm():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def cpu_scan(corpus, lib_json, nbytes, threads):
    """The reference-style CPU loop on `nbytes` of the corpus; returns (seconds, bytes, matches,
    sorted (line_no, pattern_index) pairs or None)."""
    import numpy as np
    from oracle import pcre2_oracle as po
    sample = corpus[:nbytes]
    cut = int(np.flatnonzero(sample == 10)[-1]) + 1
    sample = sample[:cut]
    lib = po.Pcre2Library([p["primary_pattern"]["regex"] for p in lib_json["patterns"]])
    t0 = time.perf_counter()
    starts, lens = po.split(sample)
    lines, pats, total = po.scan(lib, sample, starts, lens, threads=threads)
    dt = time.perf_counter() - t0
    lib.close()
    pairs = None
    if total == lines.size:   # all (line, pattern) matches fit the output arrays: usable as the full-size parity reference
        pairs = np.stack([lines.astype(np.int64) + 1, pats.astype(np.int64)], axis=1)
        pairs = pairs[np.lexsort((pairs[:, 1], pairs[:, 0]))]
    return dt, cut, total, pairs


def workload_config(args, world):
    """The `config` object: identical in both arms (ours and --impl reference)."""
    return {"workload": "configs[1]: 1 GiB synthetic crash-log corpus vs 500-pattern library, per GPU", "corpus_bytes": args.corpus_bytes,
            "patterns": args.patterns, "p_hit": P_HIT, "parallelism": "pod-sharded x%d, no collective" % world,
            "l2": "input (1 GiB per GPU) is larger than the 126 MB L2; no flush between iterations"}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU shape for this path (split lines, per line x per
    pattern unanchored regex search) on all host cores, on the SAME corpus, pattern count, steps and
    warm-up as our arm.  The JVM engine (java.util.regex) is not in this image; PCRE2-JIT stands in
    (oracle/scan_pcre2.c) -- kind "port".  One step over 1 GiB takes ~14 s on 16 threads, so the run is
    capped by a wall-clock budget (--ref-budget-s, default 150 s): steps beyond it are not run and the
    line says so."""
    if rank != 0:
        return
    import numpy as np
    from operator_b200 import synth
    lib_json, wit, swit = synth.make_library(args.patterns)
    corpus, _ = synth.make_corpus(args.corpus_bytes, wit, swit, p_hit=P_HIT)
    cores = os.cpu_count() or 1
    t_begin = time.perf_counter()
    budget = args.ref_budget_s
    warm_done = 0
    t_step_est = None
    for _ in range(args.warmup):
        if t_step_est is not None and (time.perf_counter() - t_begin) + 2 * t_step_est > budget:   # keep room for one timed step
            break
        dt, _, _, _ = cpu_scan(corpus, lib_json, args.corpus_bytes, cores)
        t_step_est = dt
        warm_done += 1
    t_total, b_total, steps = 0.0, 0, 0
    for _ in range(max(1, args.steps)):
        if steps >= 1 and (time.perf_counter() - t_begin) + t_total / steps > budget:
            break
        dt, nb, _, _ = cpu_scan(corpus, lib_json, args.corpus_bytes, cores)
        t_total += dt
        b_total += nb
        steps += 1
    gbs = b_total / t_total / 1e9
    sample = "%d steps x the whole %d MiB config-2 corpus, %d patterns, PCRE2-JIT, %d threads" % (steps, args.corpus_bytes >> 20, args.patterns, cores)
    print(json.dumps({
        "impl": "reference", "metric": "log_scan_throughput", "value": gbs, "unit": "GB/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warm_done, "ms_per_step": 1e3 * t_total / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic", "config": workload_config(args, world),
        "steps_requested": args.steps, "warmup_requested": args.warmup,
        "budget": {"seconds": budget, "capped": steps < args.steps or warm_done < args.warmup,
                   "note": "a step scans the whole corpus; steps and warm-up beyond the wall-clock budget are not run"},
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))



# ------------------------------------------------------------------------------------------------
# BASELINE configs 3 and 5: corpora of 17-64 GB per GPU, generated on the device (the counter-based
# generator as a kernel, byte-identical to the host generator) and handed over as device pointers.

def device_corpus(torch, target_bytes, wit, swit, p_hit, first_line):
    from operator_b200 import synth
    t = torch.empty(target_bytes, dtype=torch.uint8, device="cuda")
    n, lines = synth.make_corpus_device(t.data_ptr(), target_bytes, wit, swit, p_hit=p_hit, first_line=first_line)
    return t, n, lines


def timed_rescans(b, dist, steps, warmup, sync_all):
    """Device time of `steps` one-batch-at-a-time scans of a staged batch (CUDA events inside the library)."""
    for _ in range(warmup):
        r = b.rescan(copy=False)
    sync_all()
    ms = scan = 0.0
    for _ in range(steps):
        r = b.rescan(copy=False)
        ms += r.ms_total
        scan += r.ms_scan
    sync_all()
    return ms / steps, scan / steps, r


def leg_config5(args, eng, torch, dist, dev, rank, world, wit, swit, sync_all, peak):
    """64 GiB corpus, fixed size, split by pod over the ranks: strong scaling (N = 1 holds all of it)."""
    n_pods, pod_bytes = args.config5_gib, 1 << 30
    b = eng.batch()
    my_bytes = my_lines = 0
    tensors = []
    for k in range(rank, n_pods, world):
        t, n, lines = device_corpus(torch, pod_bytes, wit, swit, P_HIT, (1000 + k) << 36)
        b.add_pod_device(t.data_ptr(), n, k, keep=t)
        tensors.append(t)
        my_bytes += n
        my_lines += lines
    b.stage()
    b._keep.clear()
    del tensors, t
    torch.cuda.empty_cache()          # the staged copy is what gets scanned
    steps = max(2, min(args.steps, 5))
    ms, ms_scan, r = timed_rescans(b, dist, steps, 2, sync_all)
    events = int(r.n_events)
    b.release()
    ms_max = dist.max_over_ranks(ms, dev)
    scan_max = dist.max_over_ranks(ms_scan, dev)
    total = dist.sum_over_ranks(my_bytes, dev)
    all_events = dist.sum_over_ranks(events, dev)
    gbs = total / (ms_max * 1e-3) / 1e9
    return {"workload": "configs[4]: %d GiB synthetic corpus (%d pods of 1 GiB, generated on the device), fixed size, pods split over %d GPU(s)" % (n_pods, n_pods, world),
            "scaling": "strong", "bytes": int(total), "n_gpus": world, "steps": steps, "ms_per_step": ms_max, "k_scan_ms": scan_max,
            "gbs": gbs, "frac_of_measured_hbm_x_gpus": gbs / (peak * world), "k_scan_gbs": total / (scan_max * 1e-3) / 1e9,
            "events": int(all_events), "one_batch_at_a_time": True}


def leg_config3(args, eng, torch, dist, dev, rank, world, lib_json, wit, swit, sync_all, peak):
    """Crash storm: 10 000 failed pods x 100 000 lines, pod-sharded over 8 GPUs = 1250 pods per GPU.  With
    fewer GPUs the pod count shrinks in proportion (1250 x N pods): per-GPU work is the configuration's."""
    import numpy as np
    from oracle import pcre2_oracle as po
    per_gpu = args.config3_pods_per_gpu
    pod_bytes = 100_000 * 143          # ~100 000 lines of ~143 bytes
    b = eng.batch()
    tensors, sizes, lines_of = [], [], []
    my_bytes = 0
    for j in range(per_gpu):
        pod_id = rank * per_gpu + j
        t, n, lines = device_corpus(torch, pod_bytes, wit, swit, P_HIT, (5000 + pod_id) << 36)
        b.add_pod_device(t.data_ptr(), n, pod_id, keep=t)
        tensors.append(t)
        sizes.append(n)
        lines_of.append(lines)
        my_bytes += n
    b.stage()
    b._keep.clear()
    del tensors, t
    torch.cuda.empty_cache()
    steps = max(2, min(args.steps, 5))
    for _ in range(2):
        r = b.rescan(copy=False)
    sync_all()
    ms = scan = 0.0
    for i in range(steps):
        r = b.rescan(copy=(i == steps - 1))
        ms += r.ms_total
        scan += r.ms_scan
    sync_all()
    ms, scan = ms / steps, scan / steps
    # parity on 64 seeded pods (over all ranks) regenerated on the host: the (line, pattern) pairs of each must equal
    # the PCRE2 loop's
    sample = [p for p in np.random.RandomState(3).permutation(per_gpu * world)[:64].tolist() if p // per_gpu == rank]
    ok, checked, pairs = True, 0, 0
    plib = po.Pcre2Library([p["primary_pattern"]["regex"] for p in lib_json["patterns"]])
    for pod_id in sample:
        from operator_b200 import synth
        j = pod_id - rank * per_gpu
        host, host_lines = synth.make_corpus(pod_bytes, wit, swit, p_hit=P_HIT, first_line=(5000 + pod_id) << 36)
        starts, lens = po.split(host)
        ol, op, total = po.scan(plib, host, starts, lens, threads=max(1, (os.cpu_count() or 1) // world))
        want = np.stack([ol.astype(np.int64) + 1, op.astype(np.int64)], axis=1)
        want = want[np.lexsort((want[:, 1], want[:, 0]))]
        a, e = int(r.pod_first_event[j]), int(r.pod_first_event[j + 1])
        got = np.stack([r.line_no[a:e].astype(np.int64), r.pattern_index[a:e].astype(np.int64)], axis=1)
        ok = ok and int(host.size) == sizes[j] and host_lines == lines_of[j] == int(r.pod_n_lines[j]) and want.shape == got.shape and bool(np.array_equal(want, got))
        checked += 1
        pairs += int(want.shape[0])
    plib.close()
    events = int(r.n_events)
    b.release()
    ms_max = dist.max_over_ranks(ms, dev)
    scan_max = dist.max_over_ranks(scan, dev)
    total = dist.sum_over_ranks(my_bytes, dev)
    gbs = total / (ms_max * 1e-3) / 1e9
    return {"workload": "configs[2]: crash storm, %d pods x ~100 000 lines (%d per GPU; the configuration is 10 000 pods on 8 GPUs, scaled in proportion "
                        "to the GPU count), generated on the device, pod-sharded" % (per_gpu * world, per_gpu),
            "scaling": "weak", "bytes": int(total), "pods": per_gpu * world, "n_gpus": world, "steps": steps, "ms_per_step": ms_max, "k_scan_ms": scan_max,
            "gbs": gbs, "frac_of_measured_hbm_x_gpus": gbs / (peak * world), "events": int(dist.sum_over_ranks(events, dev)),
            "parity": {"pods_checked": int(dist.sum_over_ranks(checked, dev)), "pairs": int(dist.sum_over_ranks(pairs, dev)),
                       "equal": dist.sum_over_ranks(0 if ok else 1, dev) == 0,
                       "what": "per sampled pod: bytes and line count of the host-regenerated text, and every (line_no, pattern_index) pair vs the PCRE2 loop"}}


def measure_corpus(eng, torch, dist, dev_tensor, nbytes, steps, depth, sync_all):
    """One device-resident corpus: step time one batch at a time (latency) and with `depth` batches in flight
    (throughput), plus the counters of a step."""
    b = eng.batch()
    b.add_pod_device(dev_tensor.data_ptr(), nbytes, 0, keep=dev_tensor)
    b.stage()
    ms, ms_scan, r = timed_rescans(b, dist, steps, 4, sync_all)
    one = {"ms_per_step": ms, "k_scan_ms": ms_scan, "post_ms": float(r.ms_post), "dfa_ms": float(r.ms_dfa), "gbs": nbytes / (ms * 1e-3) / 1e9}
    stats = {"filter_candidates": int(r.n_candidates), "literal_confirms": int(r.n_verified), "events": int(r.n_events)}
    want = int(r.n_events)
    b.release()
    piped = None
    if depth > 1:
        eng.set_pipeline(depth, 0)
        batches = []
        for _ in range(depth):
            bb = eng.batch()
            bb.add_pod_device(dev_tensor.data_ptr(), nbytes, 0, keep=dev_tensor)
            bb.stage()
            batches.append(bb)

        def run(n):
            ok = True
            for bb in batches[:min(depth, n)]:
                bb.submit()
            for i in range(n):
                rr = batches[i % depth].wait(copy=False)
                ok = ok and int(rr.n_events) == want
                if i + depth < n:
                    batches[i % depth].submit()
            return ok

        run(2 * depth)
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n_timed = max(2 * depth, steps)
        e0.record()
        ok = run(n_timed)
        e1.record()
        sync_all()
        ms_p = e0.elapsed_time(e1) / n_timed
        piped = {"batches_in_flight": depth, "ms_per_step": ms_p, "gbs": nbytes / (ms_p * 1e-3) / 1e9, "same_events_every_step": ok}
        for bb in batches:
            bb.release()
        eng.set_pipeline(1, 0)
    return one, piped, stats


def leg_hit_sweep(args, eng, torch, dist, lib_json, wit, swit, sync_all, peak):
    """Hit density and near-miss lines: p_hit in {0, 1e-4, 1e-3, 1e-2}; in every corpus 1e-3 of the lines
    are near misses (a witness with its head or tail cut off: library vocabulary, often the required literal, no match)."""
    from operator_b200 import synth
    near = synth.near_misses(wit)
    eng.compile_library(lib_json).publish()
    out = []
    for p_hit in (0.0, 1e-4, 1e-3, 1e-2):
        t = torch.empty(args.corpus_bytes, dtype=torch.uint8, device="cuda")
        n, lines = synth.make_corpus_device(t.data_ptr(), args.corpus_bytes, wit, swit, p_hit=p_hit, first_line=77 << 36, p_near=1e-3, near=near)
        one, piped, stats = measure_corpus(eng, torch, dist, t, n, max(3, min(args.steps, 10)), args.depth, sync_all)
        row = {"p_hit": p_hit, "p_near": 1e-3, "lines": lines}
        row.update(stats)
        row["one_batch_at_a_time"] = one
        row["pipelined"] = piped
        row["frac_of_measured_hbm"] = (piped or one)["gbs"] / peak
        out.append(row)
        del t
        torch.cuda.empty_cache()
    return out


def leg_short_literal_library(args, eng, torch, dist, sync_all, peak):
    """The 500-pattern library with every tenth primary replaced by a regex whose required literal has 5 or 6 bytes
    (\\bEPIPE\\b, (?i)errno \\d+, [Cc]ache (evicted|flushed), ...): the library compiles to filter granule 2 - k_scan_half
    tests the word at every even offset, twice the lookups - and no pattern needs the combined DFA."""
    from operator_b200 import synth
    lj, w, sw = synth.make_mixed_library(args.patterns, min_literal=5)
    rep = eng.compile_library(lj).publish().report
    t = torch.empty(args.corpus_bytes, dtype=torch.uint8, device="cuda")
    n, lines = synth.make_corpus_device(t.data_ptr(), args.corpus_bytes, w, sw, p_hit=P_HIT, first_line=79 << 36)
    one, piped, stats = measure_corpus(eng, torch, dist, t, n, max(3, min(args.steps, 10)), args.depth, sync_all)
    del t
    torch.cuda.empty_cache()
    row = {"workload": "config-2 corpus vs the 500-pattern library with 10 % primaries of 5-6 byte literals (filter granule 2, no DFA pass)",
           "library": {k: rep[k] for k in ("n_patterns", "n_filtered", "n_dfa_scanned", "dfa_groups", "n_grams", "filter_log2_bits", "blob_bytes")},
           "k_scan_half_gbs": n / (one["k_scan_ms"] * 1e-3) / 1e9, "k_scan_half_frac_of_measured_hbm": n / (one["k_scan_ms"] * 1e-3) / 1e9 / peak}
    row.update(stats)
    row["one_batch_at_a_time"] = one
    row["pipelined"] = piped
    row["frac_of_measured_hbm"] = (piped or one)["gbs"] / peak
    return row


def leg_mixed_library(args, eng, torch, dist, sync_all, peak):
    """The 500-pattern library with every tenth primary replaced by a short-literal / class-led regex
    (\\bOOM\\b, exit code 137\\b, (?i)errno \\d+, [Cc]ache (evicted|flushed), ...): what has no 7-byte literal goes to
    the combined shared-memory DFA (k_dfa_scan), a second pass over the text."""
    from operator_b200 import synth
    lj, w, sw = synth.make_mixed_library(args.patterns)
    rep = eng.compile_library(lj).publish().report
    t = torch.empty(args.corpus_bytes, dtype=torch.uint8, device="cuda")
    n, lines = synth.make_corpus_device(t.data_ptr(), args.corpus_bytes, w, sw, p_hit=P_HIT, first_line=78 << 36)
    one, piped, stats = measure_corpus(eng, torch, dist, t, n, max(3, min(args.steps, 10)), args.depth, sync_all)
    del t
    torch.cuda.empty_cache()
    groups = max(1, rep["dfa_groups"])
    row = {"workload": "config-2 corpus vs the mixed 500-pattern library (10 % short-literal / class-led primaries)",
           "library": {k: rep[k] for k in ("n_patterns", "n_filtered", "n_dfa_scanned", "dfa_groups", "dfa_states", "dfa_classes", "blob_bytes")},
           "k_dfa_scan_gbs_per_pass": n / (one["dfa_ms"] / groups * 1e-3) / 1e9 if one["dfa_ms"] > 0 else None,
           "k_dfa_scan_frac_of_measured_hbm": (n / (one["dfa_ms"] / groups * 1e-3) / 1e9 / peak) if one["dfa_ms"] > 0 else None}
    row.update(stats)
    row["one_batch_at_a_time"] = one
    row["pipelined"] = piped
    row["frac_of_measured_hbm"] = (piped or one)["gbs"] / peak
    return row


def leg_single_context(args, torch, world, local, lib_json, wit, swit, peak):
    """Rank 0 alone: ONE context over all N GPUs of the node, the way a single /parse service on a multi-GPU box runs
    (pods placed over the devices, per-device streams, host gather in batch order) - a crash-storm-shaped batch of
    device-resident pods, checked pod for pod against a one-GPU context; and one host-resident log large enough to be
    cut across the devices (line-aligned pieces with halos), checked against the uncut scan."""
    import numpy as np
    from operator_b200 import Engine, synth
    pods_per_dev, pod_bytes = 64, 100_000 * 143
    eng = Engine(devices=list(range(world)))
    one = Engine(devices=[0])
    out = {"n_gpus": world}
    try:
        eng.compile_library(lib_json).publish()
        one.compile_library(lib_json).publish()
        b = eng.batch()
        total = 0
        for d in range(world):
            torch.cuda.set_device(d)
            for j in range(pods_per_dev):
                pod_id = d * pods_per_dev + j
                t, n, _ = device_corpus(torch, pod_bytes, wit, swit, P_HIT, (9000 + pod_id) << 36)
                b.add_pod_device(t.data_ptr(), n, pod_id, keep=t)
                total += n
        b.stage()
        for _ in range(2):
            r = b.rescan(copy=False)
        ms = 0.0
        t0 = time.perf_counter()
        for i in range(3):
            r = b.rescan(copy=(i == 2))
            ms += r.ms_total
        wall = (time.perf_counter() - t0) / 3
        ms /= 3

        def fingerprints(res, n_pods):
            fp = []
            for k in range(n_pods):
                a, e = int(res.pod_first_event[k]), int(res.pod_first_event[k + 1])
                fp.append((e - a, int(res.line_no[a:e].astype(np.int64).sum()), int(res.pattern_index[a:e].astype(np.int64).sum()),
                           int(res.ctx_end[a:e].sum()), float(res.score[a:e].sum()), int(res.pod_n_lines[k])))
            return fp

        got = fingerprints(r, world * pods_per_dev)
        events = int(r.n_events)
        b.release()
        torch.cuda.empty_cache()
        # the same pods, 64 at a time, on a one-GPU context
        torch.cuda.set_device(0)
        same = True
        for d in range(world):
            rb = one.batch()
            for j in range(pods_per_dev):
                pod_id = d * pods_per_dev + j
                t, n, _ = device_corpus(torch, pod_bytes, wit, swit, P_HIT, (9000 + pod_id) << 36)
                rb.add_pod_device(t.data_ptr(), n, pod_id, keep=t)
            rb.submit()
            want = fingerprints(rb.wait(), pods_per_dev)
            rb.release()
            same = same and want == got[d * pods_per_dev:(d + 1) * pods_per_dev]
        gbs = total / (ms * 1e-3) / 1e9
        out["pod_sharded_batch"] = {"pods": world * pods_per_dev, "bytes": total, "events": events, "device_ms_per_step": ms, "host_wall_ms_per_step": wall * 1e3,
                                    "gbs": gbs, "frac_of_measured_hbm_x_gpus": gbs / (peak * world), "equals_one_gpu_context_pod_for_pod": same}
        # one large host-resident log, cut across the devices
        log_bytes = min(world + 1, 9) * (64 << 20)
        log, lines = synth.make_corpus(log_bytes, wit, swit, p_hit=P_HIT, first_line=9900 << 36)
        t0 = time.perf_counter()
        rc = eng.scan([log])
        wall_cut = time.perf_counter() - t0
        ru = one.scan([log])
        equal = rc.n_events == ru.n_events and all(np.array_equal(getattr(rc, f), getattr(ru, f)) for f in
                                                   ("pattern_index", "line_no", "line_start", "line_len", "ctx_first_line", "ctx_last_line", "ctx_start", "ctx_end", "score", "pod_n_lines"))
        out["large_log_cut_across_devices"] = {"bytes": int(log.size), "lines": lines, "events": int(rc.n_events), "pieces": min(world, int(log.size) // (64 << 20)),
                                               "host_wall_ms": wall_cut * 1e3, "equals_uncut_scan": bool(equal)}
    finally:
        eng.close()
        one.close()
        torch.cuda.set_device(local)
    return out


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch
    from operator_b200 import Engine, dist, synth

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: the scan has no CPU path (use --impl reference for the CPU loop)")
    torch.cuda.set_device(local)
    dev = "cuda:%d" % local
    if world > 1:
        dist.init("nccl")

    # ---- inputs: library + this rank's corpus in pinned host memory (a1: the ingest side)
    lib_json, wit, swit = synth.make_library(args.patterns)
    # one context per process, on this rank's GPU
    visible = os.environ.get("CUDA_VISIBLE_DEVICES")
    gpu_index = visible.split(",")[local] if visible else str(local)
    eng = Engine(devices=[local])
    lib_report = eng.compile_library(lib_json).publish().report
    pinned = eng.pinned(args.corpus_bytes)
    corpus, n_lines = synth.make_corpus(args.corpus_bytes, wit, swit, p_hit=P_HIT, first_line=rank << 36, out=pinned.array)
    nbytes = int(corpus.size)

    sampler = ClockSampler(gpu_index)
    sampler.start()

    def sync_all():
        torch.cuda.synchronize()
        dist.barrier()

    # ---- leg 1a: device-resident corpus, one batch at a time: the filter kernel alone on the GPU
    #      (roofline) and the latency of one step
    b = eng.batch()
    b.add_pod(corpus, rank)
    b.stage()
    for _ in range(max(args.warmup, 3)):
        r = b.rescan(copy=False)
    sync_all()
    sampler.mark()
    serial_steps = max(3, min(args.steps, 30))
    ms_total = ms_scan = ms_post = 0.0
    for _ in range(serial_steps):
        r = b.rescan(copy=False)
        ms_total += r.ms_total
        ms_scan += r.ms_scan
        ms_post += r.ms_post
    sync_all()
    n_events, n_cand, n_ver = r.n_events, r.n_candidates, r.n_verified
    want_events = int(r.n_events)
    gpu_pairs = np.stack([r.line_no.astype(np.int64), r.pattern_index.astype(np.int64)], axis=1)   # a copy: the views die with the batch
    d2h_bytes = int(r.n_events) * 64 + 64
    b.release()

    # ---- leg 1b (value): the same step with `depth` batches in flight per GPU.  A step is still one
    #      full pass over one device-resident 1 GiB batch with its results back on the host; the tail of
    #      one batch (confirm, sort, resolve, result copies) overlaps the next batches' scans.  Timed with
    #      a CUDA-event pair around exactly `steps` steps, device idle on both sides.
    depth = max(1, min(args.depth, args.steps))
    if depth > 1:
        eng.set_pipeline(depth, 0)
    batches = []
    for _ in range(depth):
        bb = eng.batch()
        bb.add_pod(corpus, rank)
        bb.stage()
        batches.append(bb)

    def run_steps(n):
        launches_ = 0
        ok = True
        for bb in batches[:min(depth, n)]:
            bb.submit()
        for i in range(n):
            bb = batches[i % depth]
            rr_ = bb.wait(copy=False)
            launches_ += rr_.gpu_launches
            ok = ok and int(rr_.n_events) == want_events
            if i + depth < n:
                bb.submit()
        return launches_, ok

    run_steps(max(args.warmup, 3, depth))   # every batch in flight has been through its first scan
    sync_all()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    ev0.record()
    launches, same = run_steps(args.steps)
    ev1.record()
    sync_all()
    wall_dev = time.perf_counter() - t0
    ms_bracket = ev0.elapsed_time(ev1)
    if not same:
        raise SystemExit("bench.py: a pipelined step returned a different event count than the serial step")
    for bb in batches:
        bb.release()

    # ---- leg 2: end to end through the C ABI with host buffers, two requests in flight (the copy of
    #      the next request's log overlaps the kernels of the current one)
    for _ in range(max(min(args.warmup, 3), 1)):
        eng.scan([corpus])
    sync_all()
    t0 = time.perf_counter()
    e2e_steps = max(1, min(args.steps, 20))
    e2e_depth = 2 if depth > 1 else 1
    ms_h2d = 0.0
    inflight = []
    for i in range(e2e_steps + e2e_depth - 1):
        if i < e2e_steps:
            bb = eng.batch()
            bb.add_pod(corpus, rank)
            bb.submit()
            inflight.append(bb)
        if i >= e2e_depth - 1:
            bb = inflight.pop(0)
            rr = bb.wait(copy=False)
            ms_h2d += rr.ms_h2d
            bb.release()
    sync_all()
    wall_e2e = time.perf_counter() - t0
    sampler.mark()
    sampler.stop()
    if depth > 1:
        eng.set_pipeline(1, 0)

    # ---- pattern-count sweep (the metric is "GB/s vs pattern count"): same corpus, nested libraries
    sweep = []
    if world == 1 and not args.no_sweep:
        for n_pat in (50, 5000, 20000):
            lj, _, _ = synth.make_library(n_pat)
            t0 = time.perf_counter()
            rep = eng.compile_library(lj).publish().report
            compile_s = time.perf_counter() - t0
            sb = eng.batch()
            sb.add_pod(corpus, 0)
            sb.stage()
            for _ in range(4):          # includes the one-off filter adaptation
                sr = sb.rescan(copy=False)
            ms_t = ms_s = 0.0
            for _ in range(5):
                sr = sb.rescan(copy=False)
                ms_t += sr.ms_total
                ms_s += sr.ms_scan
            sweep.append({"patterns": n_pat, "k_scan_gbs": nbytes / (ms_s / 5 * 1e-3) / 1e9, "step_gbs": nbytes / (ms_t / 5 * 1e-3) / 1e9,
                          "frac_of_measured_hbm": nbytes / (ms_s / 5 * 1e-3) / 1e9 / peak_hbm()[0], "events": int(sr.n_events),
                          "filter_candidates": int(sr.n_candidates), "library_blob_bytes": rep["blob_bytes"], "compile_s": round(compile_s, 2)})
            sb.release()
        # config 4 (20 000 patterns): where the post-scan time goes, stage by stage - a child process with the stage marks
        # switched on (every mark is an event record on the main stream and adds ~2 us, so the marked steps are not the timed ones)
        try:
            child = subprocess.run([sys.executable, os.path.join(ROOT, "tools_tail_bench.py"), "--patterns", "20000", "--steps", "3", "--stages"],
                                   capture_output=True, text=True, timeout=600)
            last = [l for l in child.stdout.splitlines() if l.startswith("{")][-1]
            sweep[-1]["post_scan_stages_us"] = json.loads(last)["stages_us"]
            sweep[-1]["post_scan_stages_note"] = "CUDA events between the kernels of one step, in stream order; ncu L2 hit rates of these kernels: profiles/r02_config4_ncu.md"
        except Exception as e:   # the detail is a report, not a gate
            sweep[-1]["post_scan_stages_us"] = None
            sweep[-1]["post_scan_stages_note"] = "unavailable: %r" % (e,)
        eng.compile_library(lib_json).publish()

    # ---- configs[4] shaped (one large device-resident batch: here 8 pods x the 1 GiB corpus): the fixed
    #      latency of the small kernels is amortised, one batch at a time
    large = None
    if world == 1 and not args.no_sweep:
        lb = eng.batch()
        n_big = 8
        for k in range(n_big):
            lb.add_pod(corpus, k)
        lb.stage()
        for _ in range(3):
            lr = lb.rescan(copy=False)
        ms_t = ms_s = 0.0
        for _ in range(5):
            lr = lb.rescan(copy=False)
            ms_t += lr.ms_total
            ms_s += lr.ms_scan
        large = {"workload": "%d pods x the 1 GiB corpus in one batch, device-resident, one batch at a time" % n_big, "bytes": n_big * nbytes,
                 "k_scan_gbs": n_big * nbytes / (ms_s / 5 * 1e-3) / 1e9, "k_scan_frac_of_measured_hbm": n_big * nbytes / (ms_s / 5 * 1e-3) / 1e9 / peak_hbm()[0],
                 "step_gbs": n_big * nbytes / (ms_t / 5 * 1e-3) / 1e9, "ms_per_step": ms_t / 5, "events": int(lr.n_events),
                 "events_match": int(lr.n_events) == n_big * want_events}
        lb.release()

    # ---- BASELINE config 1: one ~10k-line pod log vs the ~50-pattern library, end-to-end latency
    latency = None
    if world == 1 and not args.no_sweep:
        lj, w50, s50 = synth.make_library(50)
        eng.compile_library(lj).publish()
        small, small_lines = synth.make_corpus(1400 << 10, w50, s50, p_hit=1e-3)
        small_pinned = eng.pinned(int(small.size))
        small_pinned.array[:small.size] = small
        for _ in range(5):
            eng.scan([small_pinned.array[:small.size]])
        t0 = time.perf_counter()
        for _ in range(50):
            lr = eng.scan([small_pinned.array[:small.size]])
        latency = {"workload": "configs[0]: one pod log, %d lines, %d bytes, 50 patterns" % (small_lines, small.size),
                   "ms_end_to_end": (time.perf_counter() - t0) / 50 * 1e3, "ms_h2d": lr.ms_h2d, "ms_k_scan": lr.ms_scan, "ms_device_total": lr.ms_total,
                   "events": int(lr.n_events), "note": "launch/latency-bound at this size: a roofline fraction is not meaningful"}
        small_pinned.free()
        eng.compile_library(lib_json).publish()

    # ---- hit density / near misses, and the mixed library that reaches the combined DFA (N = 1)
    hit_sweep = mixed = short5 = None
    if world == 1 and not args.no_sweep:
        hit_sweep = leg_hit_sweep(args, eng, torch, dist, lib_json, wit, swit, sync_all, peak_hbm()[0])
        mixed = leg_mixed_library(args, eng, torch, dist, sync_all, peak_hbm()[0])
        short5 = leg_short_literal_library(args, eng, torch, dist, sync_all, peak_hbm()[0])
        eng.compile_library(lib_json).publish()

    # ---- BASELINE configs 3 and 5 (every rank takes part)
    cfg3 = cfg5 = None
    if not args.no_sweep:
        eng.compile_library(lib_json).publish()
        if args.config5_gib > 0:
            cfg5 = leg_config5(args, eng, torch, dist, dev, rank, world, wit, swit, sync_all, peak_hbm()[0])
        if args.config3_pods_per_gpu > 0:
            cfg3 = leg_config3(args, eng, torch, dist, dev, rank, world, lib_json, wit, swit, sync_all, peak_hbm()[0])

    # ---- reductions: max time over ranks, sum of bytes
    step_ms = dist.max_over_ranks(ms_bracket / args.steps, dev)
    serial_ms = dist.max_over_ranks(ms_total / serial_steps, dev)
    scan_ms = dist.max_over_ranks(ms_scan / serial_steps, dev)
    e2e_s = dist.max_over_ranks(wall_e2e / e2e_steps, dev)
    wall_dev_s = dist.max_over_ranks(wall_dev / args.steps, dev)
    total_bytes = dist.sum_over_ranks(nbytes, dev)
    h2d_ms = dist.max_over_ranks(ms_h2d / e2e_steps, dev)

    # ---- the product's own multi-GPU path: rank 0 alone opens ONE context over all N GPUs (the other ranks have
    #      closed theirs and wait at the final barrier)
    single = None
    if world > 1 and not args.no_sweep:
        eng.close()
        eng = None
        torch.cuda.empty_cache()
        dist.barrier()
        if rank == 0:
            single = leg_single_context(args, torch, world, local, lib_json, wit, swit, peak_hbm()[0])

    parity_failed = False
    if rank == 0:
        peak, peak_src = peak_hbm()
        achieved = nbytes / (scan_ms * 1e-3) / 1e9          # per GPU: the kernel's own launch
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "k_scan_traffic.json")) as f:
                traffic = json.load(f).get("dram_bytes_per_launch")
        except Exception:
            pass
        out = {
            "metric": "log_scan_throughput", "value": total_bytes / (step_ms * 1e-3) / 1e9, "unit": "GB/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": workload_config(args, world),
            "workload_stats": {"corpus_bytes_per_gpu": nbytes, "lines_per_gpu": n_lines, "batches_in_flight_per_gpu": depth,
                               "events_per_step": int(n_events), "filter_candidates_per_step": int(n_cand), "literal_confirms_per_step": int(n_ver),
                               "library": {k: lib_report[k] for k in ("n_patterns", "n_filtered", "n_dfa_scanned", "dfa_groups", "n_grams", "filter_log2_bits", "blob_bytes")}},
            "e2e": {"value": total_bytes / e2e_s / 1e9, "unit": "GB/s", "h2d_bytes_per_step": nbytes + 64, "d2h_bytes_per_step": d2h_bytes,
                    "h2d_ms_per_copy": h2d_ms, "requests_in_flight": e2e_depth, "steps": e2e_steps,
                    "note": "PCIe-bound: with 2 requests in flight their host-to-device copies share the link (each takes ~2x as long), "
                            "and the kernels of one request run under the copy of the next"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "k_scan", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "frac_of_nominal_8TBs": achieved / 8000.0, "peak_source": peak_src, "traffic": traffic,
                         "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of one k_scan launch from the ncu --set full capture under profiles/ "
                                           "(k_scan_traffic.json names it); not measured in this run",
                         "algorithmic_bytes_per_launch": nbytes, "avg_launch_ms": scan_ms,
                         "timed": "CUDA events around each k_scan launch, %d launches, kernel alone on the GPU (one batch at a time)" % serial_steps,
                         "whole_step_frac": (nbytes / (step_ms * 1e-3) / 1e9) / peak},
            "timing": {"device_ms_per_step": step_ms, "how": "CUDA-event pair around the %d timed steps, %d batches in flight" % (args.steps, depth),
                       "host_wall_ms_per_step": wall_dev_s * 1e3,
                       "one_batch_at_a_time": {"device_ms_per_step": serial_ms, "k_scan_ms": scan_ms, "post_ms": ms_post / serial_steps,
                                               "gbs": nbytes / (serial_ms * 1e-3) / 1e9, "steps": serial_steps}},
            "clocks": sampler.summary(),
            "pattern_sweep": sweep,
            "large_batch": large,
            "config1_latency": latency,
            "hit_sweep": hit_sweep,
            "mixed_library": mixed,
            "short_literal_library": short5,
            "config3_crash_storm": cfg3,
            "config5_fixed_corpus": cfg5,
            "single_context": single,
        }
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            dt, nb, n_match, cpu_pairs = cpu_scan(corpus, lib_json, min(CPU_SAMPLE_BYTES, nbytes), cores)
            out["cpu_baseline"] = {"value": nb / dt / 1e9, "unit": "GB/s", "cores": cores, "kind": "port",
                                   "sample": "first %d MiB of the same corpus, all %d patterns, per line x per pattern PCRE2-JIT search "
                                             "(JVM unavailable on this box), %d threads, %.1f s" % (nb >> 20, args.patterns, cores, dt)}
            # Full-size parity of the headline configuration: the CPU loop has just produced every (line, pattern)
            # match of the whole corpus; the GPU step's events must be exactly those pairs.
            if cpu_pairs is not None and nb == nbytes:
                equal = cpu_pairs.shape == gpu_pairs.shape and bool(np.array_equal(cpu_pairs, gpu_pairs))
                out["parity_check"] = {"pairs": int(cpu_pairs.shape[0]), "gpu_pairs": int(gpu_pairs.shape[0]), "equal": equal,
                                       "what": "(line_no, pattern_index) of every event of one timed step vs the PCRE2 loop over the same %d bytes" % nb}
                parity_failed = not equal
            else:
                out["parity_check"] = {"pairs": int(n_match), "equal": None, "what": "not comparable: the CPU leg did not cover the whole corpus"}
        print(json.dumps(out))
    if eng is not None:
        if pinned is not None:
            pinned.free()
        eng.close()
    if world > 1:
        dist.barrier()
        import torch.distributed as td
        td.destroy_process_group()
    if parity_failed:
        raise SystemExit("bench.py: the GPU events differ from the CPU loop's matches on the full corpus (parity_check.equal = false)")


if __name__ == "__main__":
    main()
