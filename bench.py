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

    run_steps(max(args.warmup, 3))
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

    # ---- reductions: max time over ranks, sum of bytes
    step_ms = dist.max_over_ranks(ms_bracket / args.steps, dev)
    serial_ms = dist.max_over_ranks(ms_total / serial_steps, dev)
    scan_ms = dist.max_over_ranks(ms_scan / serial_steps, dev)
    e2e_s = dist.max_over_ranks(wall_e2e / e2e_steps, dev)
    wall_dev_s = dist.max_over_ranks(wall_dev / args.steps, dev)
    total_bytes = dist.sum_over_ranks(nbytes, dev)
    h2d_ms = dist.max_over_ranks(ms_h2d / e2e_steps, dev)

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
    pinned.free()
    eng.close()
    if world > 1:
        import torch.distributed as td
        td.destroy_process_group()
    if parity_failed:
        raise SystemExit("bench.py: the GPU events differ from the CPU loop's matches on the full corpus (parity_check.equal = false)")


if __name__ == "__main__":
    main()
