"""pm_line_index_noalloc (collected last on purpose: if it fails under -x, every other test has
already run).

The entry point serves callers that receive device buffers but must not allocate: they run
inside a host that owns all GPU memory, or on a path where cudaMalloc's implicit device
synchronisation is not acceptable.  Its scratch is a fixed array inside the library image, so
what needs checking beyond "same answer as the pooled entry point" is exactly what a shared,
never-reinitialised scratch can break: a second call right after a first one (stale counts and
prefixes from the larger first input must not leak into the smaller second one), and the size
limit.  Device buffers here come from torch purely as the test's own allocator; the entry point
is called through the C ABI with raw pointers.
"""
import ctypes

import numpy as np
import pytest

from operator_b200 import synth
from operator_b200.engine import load_library
from oracle import oracle

pytestmark = pytest.mark.gpu

MAX_BYTES = 64 << 20


def _reference_table(engine, log: bytes):
    """Line table from the existing, pool-backed entry point (pm_batch_line_index)."""
    b = engine.batch()
    try:
        b.add_pod(log, 0)
        b.submit()
        res = b.wait()
        n = int(res.pod_n_lines[0])
        return b.line_index(0, n).copy(), n
    finally:
        b.release()


def _oracle_table(log: bytes):
    """Line starts straight from the oracle's restatement of Java's split (SPEC.md section 2)."""
    lines = oracle.split_lines(log)
    return np.array([a for a, _ in lines], dtype=np.int64), len(lines)


def _noalloc_table(log: bytes, cap=None):
    import torch
    dll = load_library()
    n = len(log)
    d_text = torch.frombuffer(bytearray(log), dtype=torch.uint8).cuda() if n else torch.empty(0, dtype=torch.uint8, device="cuda")
    cap = (log.count(b"\n") + 2) if cap is None else cap
    d_out = torch.full((max(cap, 1),), -7, dtype=torch.int64, device="cuda")
    d_n = torch.zeros(1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    rc = dll.pm_line_index_noalloc(d_text.data_ptr() if n else None, n, d_out.data_ptr(), cap, d_n.data_ptr(), None)
    assert rc == 0, dll.pm_noalloc_last_error()
    torch.cuda.synchronize()
    n_lines = int(d_n.item())
    return d_out[:min(n_lines, cap)].cpu().numpy(), n_lines


def test_two_calls_in_a_row_match_pooled_entry_point(engine):
    lib_json, wit, swit = synth.make_library(50)
    engine.compile_library(lib_json).publish()
    big = bytes(synth.make_corpus(3 << 20, wit, swit)[0]) + b"tail without newline"
    small = b"a\r\nbb\n\nccc\n\n\n"          # trailing separators dropped, inner empty line kept
    for log in (big, small):                   # second call runs on the first call's leftovers
        want, want_n = _reference_table(engine, log)
        got, got_n = _noalloc_table(log)
        assert got_n == want_n
        assert np.array_equal(got, want)       # byte offsets: exact, like every other index test here
        ref, ref_n = _oracle_table(log)        # and the oracle itself, not only the entry point that was tested against it
        assert got_n == ref_n and np.array_equal(got, ref)
    assert _noalloc_table(small)[1] == 4      # "a", "bb", "", "ccc"


def test_edge_inputs_and_limit(engine):
    lib_json, _, _ = synth.make_library(50)
    engine.compile_library(lib_json).publish()
    for log in (b"", b"\n", b"\n\n\n", b"x", b"x\n", b"\nx", b"\r\n", b"a\rb\n\r", b"q" * 5000 + b"\n" + b"r" * 4096,
                b"x\n\r\r\n", b"a\r\n\r\n", b"k\n" + b"\r\n" * 3000 + b"\n" * 9000, b"\n" * 70000):
        want, want_n = _reference_table(engine, log)
        got, got_n = _noalloc_table(log)
        assert (got_n, got.tolist()) == (want_n, want.tolist()), log[:20]
        ref, ref_n = _oracle_table(log)
        assert (got_n, got.tolist()) == (ref_n, ref.tolist()), log[:20]
    # capacity smaller than the table: count is still reported, only `cap` entries written
    got, n = _noalloc_table(b"1\n2\n3\n4\n5", cap=2)
    assert n == 5 and got.tolist() == [0, 2]
    # beyond the fixed scratch: refused before anything is launched
    dll = load_library()
    rc = dll.pm_line_index_noalloc(None, MAX_BYTES + 1, None, 0, ctypes.c_void_p(8), None)
    assert rc == -1 and b"exceed" in dll.pm_noalloc_last_error()


def test_input_of_exactly_the_limit():
    """The largest input the fixed scratch takes: 64 MiB = 32768 tiles, i.e. every slot of the
    in-module arrays and 32 rounds of the single-CTA prefix.  Checked against the oracle."""
    lib_json, wit, swit = synth.make_library(20)
    corpus = bytes(synth.make_corpus(MAX_BYTES - 4096, wit, swit)[0])
    log = corpus + b"z" * (MAX_BYTES - len(corpus) - 3) + b"\n\r\n"      # ends in separators that are dropped
    assert len(log) == MAX_BYTES
    for _ in range(2):                                                   # twice: the second run starts from a full scratch
        got, got_n = _noalloc_table(log)
        ref, ref_n = _oracle_table(log)
        assert got_n == ref_n
        assert np.array_equal(got, ref)
