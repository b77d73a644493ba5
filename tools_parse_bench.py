#!/usr/bin/env python
"""Requests/s of the /parse entry point (pm_parse) under concurrent callers: BASELINE config 1
shaped requests (one ~1.4 MB pod log, ~50 patterns), T caller threads, with and without request
coalescing (PM_PARSE_COALESCE=0 in a child process).  One JSON line per run.

    python tools_parse_bench.py [--threads 32] [--seconds 4]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def run(threads, seconds, depth, log_bytes):
    from operator_b200 import Engine, synth
    eng = Engine(devices=[0])
    lib_json, wit, swit = synth.make_library(50)
    eng.compile_library(lib_json).publish()
    if depth > 1:
        eng.set_pipeline(depth, 0)
    bodies = []
    for k in range(threads):
        corpus, n_lines = synth.make_corpus(log_bytes, wit, swit, p_hit=1e-3, seed=4000 + k)
        bodies.append(json.dumps({"pod": {"metadata": {"name": "pod-%d" % k, "namespace": "bench"}}, "logs": bytes(corpus).decode(), "events": []}).encode())
    for b in bodies[:4]:
        eng.parse_raw(b)
    s0 = eng.parse_stats()
    stop = time.perf_counter() + seconds
    counts = [0] * threads
    lat = [[] for _ in range(threads)]
    gate = threading.Barrier(threads + 1)

    def work(k):
        gate.wait()
        while time.perf_counter() < stop:
            t0 = time.perf_counter()
            eng.parse_raw(bodies[k])
            lat[k].append(time.perf_counter() - t0)
            counts[k] += 1

    ts = [threading.Thread(target=work, args=(k,)) for k in range(threads)]
    for t in ts:
        t.start()
    t0 = time.perf_counter()
    gate.wait()
    for t in ts:
        t.join()
    dt = time.perf_counter() - t0
    s1 = eng.parse_stats()
    total = sum(counts)
    all_lat = sorted(x for l in lat for x in l)
    req_bytes = sum(len(b) for b in bodies) / len(bodies)
    out = {"threads": threads, "pipeline_depth": depth, "coalesce": os.environ.get("PM_PARSE_COALESCE", "1") != "0",
           "request_bytes": int(req_bytes), "requests": total, "seconds": round(dt, 3), "requests_per_s": total / dt,
           "log_gbs": total * log_bytes / dt / 1e9, "batches": s1["batches"] - s0["batches"],
           "requests_per_batch": (s1["requests"] - s0["requests"]) / max(1, s1["batches"] - s0["batches"]),
           "latency_ms_p50": all_lat[len(all_lat) // 2] * 1e3, "latency_ms_p99": all_lat[int(len(all_lat) * 0.99)] * 1e3,
           "host_cores": os.cpu_count()}
    eng.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--threads", type=int, default=32)
    ap.add_argument("--seconds", type=float, default=4.0)
    ap.add_argument("--depth", type=int, default=2)
    ap.add_argument("--log-bytes", type=int, default=1400 << 10)
    ap.add_argument("--child", action="store_true")
    args = ap.parse_args()
    if args.child:
        print(json.dumps(run(args.threads, args.seconds, args.depth, args.log_bytes)), flush=True)
        return
    for threads, depth, coalesce in ((1, 1, "1"), (args.threads, 1, "0"), (args.threads, args.depth, "0"), (args.threads, 1, "1"), (args.threads, args.depth, "1")):
        env = dict(os.environ, PM_PARSE_COALESCE=coalesce)
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", "--threads", str(threads), "--seconds", str(args.seconds),
                            "--depth", str(depth), "--log-bytes", str(args.log_bytes)], env=env, capture_output=True, text=True, timeout=600)
        sys.stdout.write(r.stdout if r.returncode == 0 else json.dumps({"error": r.stderr[-500:]}) + "\n")
        sys.stdout.flush()


if __name__ == "__main__":
    main()
