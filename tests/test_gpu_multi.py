"""One context over several GPUs (runs only where >= 2 B200s are visible): pods are partitioned
over the devices (LPT by bytes), scanned concurrently, and gathered in batch order -- the result
must equal the single-device result pod for pod."""
import numpy as np
import pytest

from operator_b200 import Engine, synth

pytestmark = pytest.mark.gpu


def _device_count():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_device_count() < 2, reason="needs two GPUs")
def test_pod_sharding_over_two_devices_matches_one_device():
    lib_json, wit, swit = synth.make_library(50)
    pods = [synth.make_corpus(200000 + 50021 * (k % 5), wit, swit, p_hit=1e-2, first_line=k * 10 ** 6)[0].copy() for k in range(23)]
    pods[7] = np.zeros(0, dtype=np.uint8)
    one = Engine(devices=[0])
    two = Engine(2)
    try:
        assert two.device_count == 2
        one.compile_library(lib_json).publish()
        two.compile_library(lib_json).publish()
        r1 = one.scan(pods)
        for _ in range(2):                       # second round exercises pooled buffers on both devices
            r2 = two.scan(pods)
            assert r2.n_events == r1.n_events > 100
            for f in ("pod_index", "pattern_index", "line_no", "line_start", "line_len", "ctx_first_line", "ctx_last_line", "ctx_start", "ctx_end", "score",
                      "pod_first_event", "pod_n_lines", "pod_tag"):
                assert np.array_equal(getattr(r1, f), getattr(r2, f)), f
    finally:
        one.close()
        two.close()


@pytest.mark.skipif(_device_count() < 2, reason="needs two GPUs")
def test_two_threads_stage_multi_device_batches_in_opposite_orders():
    """Two host threads submit two-GPU batches whose pods arrive in opposite size orders ([small, big] and
    [big, small]) at depth 1.  Largest-first placement then puts the big pod of each batch on a different
    device first; the device slots must still be taken in one global order, or each thread ends up holding
    one device and waiting for the other (the deadlock ADVICE round 1 described).  A watchdog turns a hang
    into a failure."""
    import threading
    lib_json, wit, swit = synth.make_library(20)
    big = synth.make_corpus(3 << 20, wit, swit, p_hit=1e-2, first_line=1 << 20)[0].copy()
    small = synth.make_corpus(200000, wit, swit, p_hit=1e-2, first_line=2 << 20)[0].copy()
    eng = Engine(2)
    errors, done = [], []
    try:
        eng.compile_library(lib_json).publish()
        want = {0: eng.scan([small, big]).n_events, 1: eng.scan([big, small]).n_events}

        def worker(kind):
            try:
                for _ in range(40):
                    r = eng.scan([small, big] if kind == 0 else [big, small])
                    assert r.n_events == want[kind]
                done.append(kind)
            except Exception as e:   # noqa: BLE001
                errors.append(e)

        threads = [threading.Thread(target=worker, args=(k,), daemon=True) for k in (0, 1)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=120)
        assert not errors, errors
        assert sorted(done) == [0, 1], "a thread is still waiting for a device slot: lock-order deadlock"
    finally:
        if sorted(done) == [0, 1]:
            eng.close()
