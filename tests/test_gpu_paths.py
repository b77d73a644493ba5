"""GPU tests of the less-travelled paths: queue overflow -> rerun, the radix-sort path for many
raw events, the 20k-pattern library (BASELINE config 4 shape), many small pods (config 3 shape)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import oracle, pcre2_oracle
from operator_b200 import synth
from tests.test_gpu_parity import check_parity, lib_of, _got

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_TINY = r"""
import sys
sys.path.insert(0, %r)
from operator_b200 import Engine, synth
from oracle import oracle
lib_json, wit, swit = synth.make_library(50)
corpus, _ = synth.make_corpus(600000, wit, swit, p_hit=5e-2)
eng = Engine(1)
eng.compile_library(lib_json).publish()
for rounds in range(2):                      # second scan reuses grown queues from the pool
    res = eng.scan([corpus, corpus[:200000]])
    for pod, log in enumerate([bytes(corpus), bytes(corpus[:200000])]):
        want = [(e.line_no, e.pattern, e.ctx_first, e.ctx_last, e.score) for e in oracle.analyze(log, oracle.load_library(lib_json))]
        got = [(g["line_no"], g["pattern"], g["ctx_first"], g["ctx_last"], g["score"]) for g in res.events_of(pod)]
        assert got == want and len(want) > 50, (pod, len(got), len(want))
eng.close()
print("tiny-queues ok")
"""


def test_overflow_rerun_paths(tmp_path):
    """Every internal queue starts at a few entries (PM_TEST_TINY_QUEUES): candidate queue, raw
    events, verify tasks and resolve work items all overflow and are grown, results stay exact."""
    script = tmp_path / "tiny.py"
    script.write_text(_TINY % ROOT)
    env = dict(os.environ, PM_TEST_TINY_QUEUES="1")
    r = subprocess.run([sys.executable, str(script)], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "tiny-queues ok" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]


@pytest.mark.timeout(300)
def test_many_raw_events_take_the_merge_passes(engine):
    """More than 16384 raw events: the shared-memory rank merge declines, the seven merge rounds of k_merge_all order
    them on the device (up to 131072).  Every field against the oracle."""
    lib_json, wit, swit = synth.make_library(50)
    corpus, n_lines = synth.make_corpus(12 << 20, wit, swit, p_hit=0.25)
    res = check_parity(engine, lib_json, [corpus], expect_events=17000)
    assert res.n_events > 16384
    # the merge rounds wait for each other at a grid-wide barrier: with the post-scan kernels confined to a small
    # green-context partition (8 SMs) the grid must still fit it at once
    engine.set_pipeline(2, 8)
    try:
        res2 = check_parity(engine, lib_json, [corpus], expect_events=17000)
        assert res2.n_events == res.n_events
    finally:
        engine.set_pipeline(1, 0)


@pytest.mark.skipif(not pcre2_oracle.available(), reason="libpcre2-8 not loadable")
def test_more_raw_events_than_the_device_sort_takes(engine):
    """More than 131072 raw events: the host learns the count and the radix sort runs; the resolve stage is repeated
    with result arrays of that size.  (line, pattern) pairs against the PCRE2 loop, every field of the head of the
    log against the oracle."""
    lib_json, wit, swit = synth.make_library(50)
    engine.compile_library(lib_json).publish()
    corpus, n_lines = synth.make_corpus(64 << 20, wit, swit, p_hit=0.4)
    res = engine.scan([corpus])
    assert res.n_events > 131072
    pl = pcre2_oracle.Pcre2Library([p["primary_pattern"]["regex"] for p in lib_json["patterns"]])
    starts, lens = pcre2_oracle.split(corpus)
    ol, op, total = pcre2_oracle.scan(pl, corpus, starts, lens, threads=os.cpu_count() or 4)
    pl.close()
    assert total == ol.size
    want = np.stack([ol + 1, op.astype(np.int64)], axis=1)
    want = want[np.lexsort((want[:, 1], want[:, 0]))]
    got = np.stack([res.line_no.astype(np.int64), res.pattern_index.astype(np.int64)], axis=1)
    assert np.array_equal(got, want)
    head = bytes(corpus[:60000])
    head = head[:head.rfind(b"\n") + 1]
    ev = oracle.analyze(head, oracle.load_library(lib_json))
    n_head = len(oracle.split_lines(head))
    got_head = [t for t in _got(res, 0) if t[0] <= n_head - 80]      # lines whose context window (after + STACK run) lies inside the head
    want_head = [e for e in ev if e.line_no <= n_head - 80]
    assert len(got_head) == len(want_head) > 50
    for g, w in zip(got_head, want_head):
        assert g[:8] == (w.line_no, w.pattern, w.line_start, w.line_len, w.ctx_first, w.ctx_last, w.ctx_start, w.ctx_end) and g[8] == w.score, (g, w)


@pytest.mark.skipif(not pcre2_oracle.available(), reason="libpcre2-8 not loadable")
def test_large_library_20k_patterns(engine):
    """BASELINE config 4 shape: 20 000 patterns.  (line, pattern) pairs against PCRE2 over the whole
    sample, full event parity against the Python oracle on the head of it."""
    lib_json, wit, swit = synth.make_library(20000)
    rep = engine.compile_library(lib_json).publish().report
    assert rep["n_patterns"] == 20000
    corpus, n_lines = synth.make_corpus(300000, wit, swit, p_hit=5e-2)
    res = engine.scan([corpus])
    pl = pcre2_oracle.Pcre2Library([p["primary_pattern"]["regex"] for p in lib_json["patterns"]])
    starts, lens = pcre2_oracle.split(corpus)
    ol, op, total = pcre2_oracle.scan(pl, corpus, starts, lens, threads=os.cpu_count() or 4)
    want = sorted(zip((ol + 1).tolist(), op.tolist()))
    got = sorted(zip(res.line_no.tolist(), res.pattern_index.tolist()))
    assert got == want and len(want) > 50
    head = bytes(corpus[:20000])
    head = head[:head.rfind(b"\n") + 1]
    olib = oracle.load_library(lib_json)
    ev = oracle.analyze(head, olib)
    got_head = [t for t in _got(res, 0) if t[2] < len(head)]
    assert [t[:2] for t in got_head] == [(e.line_no, e.pattern) for e in ev]
    assert [t[8] for t in got_head] == [e.score for e in ev]
    # 64 MiB at the benchmark's hit density: every (line, pattern) pair against the PCRE2 loop (~9.4e9 regex searches
    # on the host cores, about half a minute)
    corpus, n_lines = synth.make_corpus(64 << 20, wit, swit, p_hit=1e-3, first_line=1 << 40)
    res = engine.scan([corpus])
    starts, lens = pcre2_oracle.split(corpus)
    ol, op, total = pcre2_oracle.scan(pl, corpus, starts, lens, threads=os.cpu_count() or 4)
    assert total == ol.size
    want = np.stack([ol + 1, op.astype(np.int64)], axis=1)
    want = want[np.lexsort((want[:, 1], want[:, 0]))]
    got = np.stack([res.line_no.astype(np.int64), res.pattern_index.astype(np.int64)], axis=1)
    assert want.shape[0] > 300 and np.array_equal(got, want)
    pl.close()


def test_crash_storm_many_small_pods(engine):
    """BASELINE config 3 shape, scaled down: 3000 pods in one batch; every pod keeps its own line
    numbering and context clipping.  Parity on a seeded sample of 64 pods."""
    lib_json, wit, swit = synth.make_library(50)
    engine.compile_library(lib_json).publish()
    olib = oracle.load_library(lib_json)
    big, _ = synth.make_corpus(24 << 20, wit, swit, p_hit=2e-2)
    nl = np.flatnonzero(big == 10)
    pods, pos = [], 0
    rng = np.random.default_rng(7)
    for k in range(3000):
        n = int(rng.integers(5, 120))
        hi = nl[min(len(nl) - 1, np.searchsorted(nl, pos) + n)] + 1
        pods.append(big[pos:hi] if k % 7 else big[pos:hi - 1])      # some pods lack the final newline
        pos = hi
    res = engine.scan(pods)
    assert res.n_pods == 3000 and res.n_events > 1000
    for k in rng.choice(3000, 64, replace=False):
        want = [(e.line_no, e.pattern, e.line_start, e.line_len, e.ctx_first, e.ctx_last, e.ctx_start, e.ctx_end, e.score)
                for e in oracle.analyze(bytes(pods[k]), olib)]
        assert _got(res, int(k)) == want, k
        assert int(res.pod_n_lines[k]) == len(oracle.split_lines(bytes(pods[k])))


_PIPE = r"""
import sys, threading
sys.path.insert(0, %r)
import numpy as np
from operator_b200 import Engine, ScanError, synth
from oracle import oracle
lib_json, wit, swit = synth.make_library(50)
logs = [synth.make_corpus(700000 + 90000 * k, wit, swit, p_hit=2e-2, seed=100 + k)[0] for k in range(4)]
olib = oracle.load_library(lib_json)
want = [[(e.line_no, e.pattern, e.ctx_first, e.ctx_last, e.score) for e in oracle.analyze(bytes(l), olib)] for l in logs]
def rows(res, pod=0):
    return [(g["line_no"], g["pattern"], g["ctx_first"], g["ctx_last"], g["score"]) for g in res.events_of(pod)]
eng = Engine(1)
eng.compile_library(lib_json).publish()
for depth, post in ((3, 0), (2, 16), (1, 0)):
    sms = eng.set_pipeline(depth, post)
    if post:
        assert 0 < sms[1] < sms[0] and sms[0] + sms[1] <= 160, sms
    # several batches submitted before the first wait, re-submitted while others are in flight
    batches = []
    for k in range(depth):
        b = eng.batch(); b.add_pod(logs[k], k); b.stage(); batches.append(b)
    for rounds in range(3):
        for b in batches: b.submit()
        for k, b in enumerate(batches):
            got = rows(b.wait())
            assert got == want[k] and len(got) > 20, (depth, post, rounds, k, len(got), len(want[k]))
    if depth > 1:
        try:
            eng.set_pipeline(1, 0)
            raise SystemExit("set_pipeline with batches staged must fail")
        except ScanError:
            pass
    for b in batches: b.release()
    # threads, one request each, all at once (the /parse service shape)
    out = [None] * 4
    def work(k):
        out[k] = rows(eng.scan([logs[k]]))
    ts = [threading.Thread(target=work, args=(k,)) for k in range(4)]
    for t in ts: t.start()
    for t in ts: t.join()
    assert out == want, (depth, post)
eng.close()
print("pipeline ok")
"""


def test_batches_in_flight_share_a_device(tmp_path):
    """pm_set_pipeline: depth 3 on shared SMs, depth 2 on a green-context SM split, back to depth 1;
    interleaved submits/waits and concurrent threads give the oracle's events every time."""
    script = tmp_path / "pipe.py"
    script.write_text(_PIPE % ROOT)
    r = subprocess.run([sys.executable, str(script)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "pipeline ok" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]
    # the same with every queue starting tiny: overflow -> grow -> rerun while other batches are in flight
    r = subprocess.run([sys.executable, str(script)], capture_output=True, text=True, timeout=600, env=dict(os.environ, PM_TEST_TINY_QUEUES="1"))
    assert r.returncode == 0 and "pipeline ok" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]


_SPLIT = r"""
import sys
sys.path.insert(0, %r)
import numpy as np
from operator_b200 import Engine, synth
from oracle import oracle
lib_json, wit, swit = synth.make_library(50)
olib = oracle.load_library(lib_json)
big, _ = synth.make_corpus(900000, wit, swit, p_hit=3e-2, seed=31)
crlf, _ = synth.make_corpus(400000, wit, swit, p_hit=3e-2, seed=32, crlf=True)
gaps = bytes(big[:300000]).replace(b"\n", b"\n\n\n", 400) + b"\n\n\n\n"      # runs of empty lines around the cuts, empty tail
small = bytes(big[:15000])
logs = [bytes(big), bytes(crlf), gaps, small, b"", bytes(big[:200000]) + b"no newline at the end " + wit[0].encode()]
eng = Engine(0)
eng.compile_library(lib_json).publish()
res = eng.scan(logs)
n_split = 0
for pod, log in enumerate(logs):
    want = [(e.line_no, e.pattern, e.ctx_first, e.ctx_last, e.ctx_start, e.ctx_end, e.score) for e in oracle.analyze(log, olib)]
    got = [(g["line_no"], g["pattern"], g["ctx_first"], g["ctx_last"], g["ctx_start"], g["ctx_end"], g["score"]) for g in res.events_of(pod)]
    assert got == want, (pod, len(got), len(want), [x for x in zip(got, want) if x[0] != x[1]][:3])
    assert int(res.pod_n_lines[pod]) == len(oracle.split_lines(log)), (pod, int(res.pod_n_lines[pod]))
assert res.n_events > 300 and res.n_events == sum(len(res.events_of(k)) for k in range(len(logs)))
b = eng.batch(); b.add_pod(logs[0], 0); b.add_pod(small, 1); b.stage(); b.rescan()
try:
    b.line_index(0, 10)
    raise SystemExit("line_index of a split log must fail")
except Exception as e:
    assert "cut across" in str(e), e
assert len(b.line_index(1, int(res.pod_n_lines[3]))) == int(res.pod_n_lines[3])
b.release()
eng.close()
print("split ok")
"""


def test_one_large_log_is_cut_across_devices(tmp_path):
    """8(e): a log above the split threshold is cut at line starts into units with context halos
    (here forced onto however many GPUs there are, 3 pieces at least); events, line numbers, byte
    offsets, contexts and scores equal the oracle's on the whole log."""
    script = tmp_path / "split.py"
    script.write_text(_SPLIT % ROOT)
    env = dict(os.environ, PM_SPLIT_FORCE="3", PM_SPLIT_MIN_BYTES="60000")
    r = subprocess.run([sys.executable, str(script)], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "split ok" in r.stdout, r.stdout[-1500:] + r.stderr[-2500:]


def test_device_generator_is_byte_identical_and_device_pods_scan_alike(engine):
    """The corpus generator as a kernel (configs 3 and 5 generate their text where it is scanned) writes
    exactly the host generator's bytes, and a pod handed over as a device pointer gives the events of
    the same pod handed over as host bytes."""
    import torch
    lib_json, wit, swit = synth.make_library(50)
    engine.compile_library(lib_json).publish()
    for target, kw in ((64 << 20, dict(p_hit=1e-3)), (5 << 20, dict(p_hit=0.02, crlf=True, first_line=(3 << 36) + 17)), (1000, dict(p_hit=0.5)), (5, dict())):
        host, host_lines = synth.make_corpus(target, wit, swit, **kw)
        dev = torch.empty(max(target, 1), dtype=torch.uint8, device="cuda")
        got, dev_lines = synth.make_corpus_device(dev.data_ptr(), target, wit, swit, **kw)
        assert (got, dev_lines) == (int(host.size), host_lines)
        assert np.array_equal(dev[:got].cpu().numpy(), host)
    host, _ = synth.make_corpus(6 << 20, wit, swit, p_hit=2e-3)
    raw = bytes(host).rstrip(b"\n") + b"\r\n\n\r\n"      # trailing separators are dropped on the device side too
    want = engine.scan([raw, b"", raw[:100000] + b"\r"])
    d0 = torch.frombuffer(bytearray(raw), dtype=torch.uint8).cuda()
    d2 = torch.frombuffer(bytearray(raw[:100000] + b"\r"), dtype=torch.uint8).cuda()
    b = engine.batch()
    try:
        b.add_pod_device(d0.data_ptr(), d0.numel(), 0, keep=d0)
        b.add_pod_device(0, 0, 1)
        b.add_pod_device(d2.data_ptr(), d2.numel(), 2, keep=d2)
        b.submit()
        got = b.wait()
    finally:
        b.release()
    assert got.n_events == want.n_events and want.n_events > 0
    assert got.pod_n_lines.tolist() == want.pod_n_lines.tolist()
    for name in ("pod_index", "pattern_index", "line_no", "line_start", "line_len", "ctx_first_line", "ctx_last_line", "ctx_start", "ctx_end", "score"):
        assert np.array_equal(getattr(got, name), getattr(want, name)), name


def test_filter_adapts_over_many_small_requests():
    """A service sees one ~1 MB pod log per request: no single batch says enough about gram frequencies, so the
    candidate counts of successive batches are added up and the profile-guided window selection runs once 8 MiB
    have gone by (on the batch staged then).  Results never change; the candidate count of a request drops."""
    from operator_b200 import Engine
    lib_json, wit, swit = synth.make_library(500)
    eng = Engine(1)
    try:
        eng.compile_library(lib_json).publish()
        pods = [synth.make_corpus(1400 << 10, wit, swit, p_hit=1e-3, first_line=k << 30)[0].copy() for k in range(16)]
        cand, events = [], []
        for p in pods:
            r = eng.scan([p])
            cand.append(r.n_candidates)
            events.append(r.n_events)
        again = [eng.scan([p]) for p in pods[:3]]
        assert [a.n_events for a in again] == events[:3]            # same answers before and after the re-selection
        assert min(cand[:4]) > 2000                                 # static model: "thre", "conn", ... are everywhere in this text
        assert max(cand[-4:]) * 5 < min(cand[:4]), cand             # after ~8 MiB the windows were re-selected
    finally:
        eng.close()
