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
