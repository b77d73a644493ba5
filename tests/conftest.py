import ctypes
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run on the GPU box with -m gpu)")


class Rep(ctypes.Structure):
    _fields_ = [("status", ctypes.c_int32), ("pattern_index", ctypes.c_int32), ("secondary_index", ctypes.c_int32),
                ("regex_offset", ctypes.c_int32), ("pattern_id", ctypes.c_char * 64), ("message", ctypes.c_char * 256),
                ("n_patterns", ctypes.c_int32), ("n_regex_units", ctypes.c_int32), ("n_filtered", ctypes.c_int32),
                ("n_dfa_scanned", ctypes.c_int32), ("n_grams", ctypes.c_int32), ("filter_log2_bits", ctypes.c_int32),
                ("filter_fold_case", ctypes.c_int32), ("dfa_groups", ctypes.c_int32), ("dfa_states", ctypes.c_int32),
                ("dfa_classes", ctypes.c_int32), ("blob_bytes", ctypes.c_int64)]


class CompilerHarness:
    """tests/support/libpm_testsupport.so: the pattern compiler's tables walked on the host."""

    def __init__(self):
        path = os.path.join(ROOT, "tests", "support", "libpm_testsupport.so")
        if not os.path.exists(path):
            import subprocess
            subprocess.run(["make", "-C", ROOT, "tests/support/libpm_testsupport.so"], check=True)
        self.L = ctypes.CDLL(path)
        vp = ctypes.c_void_p
        self.L.pmt_literals.restype = ctypes.c_char_p
        self.L.pmt_compile.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(vp), ctypes.POINTER(Rep)]
        self.L.pmt_free.argtypes = [vp]
        self.L.pmt_n_patterns.argtypes = [vp]
        self.L.pmt_pattern_tier.argtypes = [vp, ctypes.c_int]
        self.L.pmt_primary_match.argtypes = [vp, ctypes.c_int, ctypes.c_char_p, ctypes.c_size_t]
        self.L.pmt_secondary_match.argtypes = [vp, ctypes.c_int, ctypes.c_int, ctypes.c_char_p, ctypes.c_size_t]
        self.L.pmt_stack_match.argtypes = [vp, ctypes.c_char_p, ctypes.c_size_t]
        self.L.pmt_group_match.argtypes = [vp, ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_int), ctypes.c_int]
        self.L.pmt_group_scan_match.argtypes = [vp, ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_int), ctypes.c_int]
        self.L.pmt_filter_triggers.argtypes = [vp, ctypes.c_int, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_int]
        self.L.pmt_literals.argtypes = [vp, ctypes.c_int]
        self.L.pmt_unit_states.argtypes = [vp, ctypes.c_int]

    def compile(self, spec):
        if isinstance(spec, (dict, list)):
            spec = json.dumps(spec)
        if isinstance(spec, str):
            spec = spec.encode()
        h, rep = ctypes.c_void_p(), Rep()
        rc = self.L.pmt_compile(spec, len(spec), ctypes.byref(h), ctypes.byref(rep))
        return rc, h, rep

    def compile_regexes(self, regexes):
        return self.compile({"patterns": [{"id": "p%d" % i, "primary_pattern": {"regex": r}} for i, r in enumerate(regexes)]})


@pytest.fixture(scope="session")
def harness():
    return CompilerHarness()


@pytest.fixture(scope="session")
def engine():
    from operator_b200 import Engine
    e = Engine(1)
    yield e
    e.close()
