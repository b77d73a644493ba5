"""ctypes binding of the C ABI (include/podmortem_scan.h) and the host-side mirror of the
reference's client interface for this path.

Reference interface mirrored here: ``LogParserClient.analyzeLog(PodFailureData) ->
Uni<AnalysisResult>`` (LogParserClient.java:36-55) over ``LogParserRestClient.parseLogs``
(LogParserRestClient.java:37-39).  The production host is Java 21 over Panama FFM
(INTEGRATION.md); this module binds the very same symbols so that tests and bench.py drive the
library exactly as the JVM would.  There is no Python or CPU fallback: without the shared
object or without a GPU every entry point raises.
"""
from __future__ import annotations

import ctypes
import json
import os
from typing import Dict, Iterable, List, Optional, Sequence, Union

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# PM_SCAN_LIBRARY: another build of the same library (kernel experiments); the default is the in-tree build
LIB_PATH = os.environ.get("PM_SCAN_LIBRARY") or os.path.join(_HERE, "libpodmortem_scan.so")

PM_OK = 0
STATUS = {0: "PM_OK", -1: "PM_E_INVALID", -2: "PM_E_SYNTAX", -3: "PM_E_UNSUPPORTED", -4: "PM_E_TOO_COMPLEX",
          -5: "PM_E_CUDA", -6: "PM_E_NO_DEVICE", -7: "PM_E_NOMEM", -8: "PM_E_NO_LIBRARY", -9: "PM_E_OVERFLOW"}


class CompileReport(ctypes.Structure):
    _fields_ = [("status", ctypes.c_int32), ("pattern_index", ctypes.c_int32), ("secondary_index", ctypes.c_int32),
                ("regex_offset", ctypes.c_int32), ("pattern_id", ctypes.c_char * 64), ("message", ctypes.c_char * 256),
                ("n_patterns", ctypes.c_int32), ("n_regex_units", ctypes.c_int32), ("n_filtered", ctypes.c_int32),
                ("n_dfa_scanned", ctypes.c_int32), ("n_grams", ctypes.c_int32), ("filter_log2_bits", ctypes.c_int32),
                ("filter_fold_case", ctypes.c_int32), ("dfa_groups", ctypes.c_int32), ("dfa_states", ctypes.c_int32),
                ("dfa_classes", ctypes.c_int32), ("blob_bytes", ctypes.c_int64)]

    def as_dict(self) -> dict:
        d = {}
        for name, _ in self._fields_:
            v = getattr(self, name)
            d[name] = v.decode("utf-8", "replace") if isinstance(v, bytes) else v
        return d


_P = ctypes.POINTER


class ResultView(ctypes.Structure):
    _fields_ = [("n_events", ctypes.c_int64), ("n_pods", ctypes.c_int64),
                ("pod_index", _P(ctypes.c_int32)), ("pattern_index", _P(ctypes.c_int32)), ("line_no", _P(ctypes.c_int32)),
                ("line_start", _P(ctypes.c_int64)), ("line_len", _P(ctypes.c_int32)),
                ("ctx_first_line", _P(ctypes.c_int32)), ("ctx_last_line", _P(ctypes.c_int32)),
                ("ctx_start", _P(ctypes.c_int64)), ("ctx_end", _P(ctypes.c_int64)), ("score", _P(ctypes.c_double)),
                ("pod_tag", _P(ctypes.c_uint64)), ("pod_first_event", _P(ctypes.c_int64)), ("pod_n_lines", _P(ctypes.c_int32)),
                ("ms_h2d", ctypes.c_double), ("ms_scan", ctypes.c_double), ("ms_post", ctypes.c_double), ("ms_d2h", ctypes.c_double), ("ms_total", ctypes.c_double), ("ms_dfa", ctypes.c_double),
                ("bytes_scanned", ctypes.c_int64), ("n_candidates", ctypes.c_int64), ("n_verified", ctypes.c_int64),
                ("gpu_launches", ctypes.c_int64)]


class ScanError(RuntimeError):
    def __init__(self, status: int, message: str, report: Optional[dict] = None):
        super().__init__("%s: %s" % (STATUS.get(status, status), message))
        self.status, self.message, self.report = status, message, report


_dll = None

# every symbol include/podmortem_scan.h declares (tests check the library exports all of them)
ABI_SYMBOLS = ["pm_abi_version", "pm_init", "pm_init_devices", "pm_shutdown", "pm_last_error", "pm_device_count", "pm_library_compile",
               "pm_library_publish", "pm_library_free", "pm_library_pattern_count", "pm_library_pattern_meta_json",
               "pm_host_alloc", "pm_host_free", "pm_batch_begin", "pm_batch_add_pod", "pm_batch_submit", "pm_batch_wait",
               "pm_batch_release", "pm_batch_stage", "pm_batch_rescan", "pm_batch_line_index", "pm_parse",
               "pm_line_index_noalloc", "pm_noalloc_last_error", "pm_batch_add_pod_device", "pm_set_pipeline", "pm_parse_stats", "pm_library_serialize", "pm_library_load"]


def load_library() -> ctypes.CDLL:
    """Loads libpodmortem_scan.so; fails loudly when it has not been built."""
    global _dll
    if _dll is not None:
        return _dll
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("%s is missing: build it with `make` (or __graft_entry__.build()); "
                           "there is no fallback implementation" % LIB_PATH)
    d = ctypes.CDLL(LIB_PATH)
    vp, sz = ctypes.c_void_p, ctypes.c_size_t
    d.pm_abi_version.restype = ctypes.c_int
    d.pm_init.argtypes = [ctypes.c_int, _P(vp)]
    d.pm_init_devices.argtypes = [_P(ctypes.c_int), ctypes.c_int, _P(vp)]
    d.pm_shutdown.argtypes = [vp]
    d.pm_shutdown.restype = None
    d.pm_last_error.restype = ctypes.c_char_p
    d.pm_device_count.argtypes = [vp]
    d.pm_library_compile.argtypes = [vp, ctypes.c_char_p, sz, _P(vp), _P(CompileReport)]
    d.pm_library_publish.argtypes = [vp, vp]
    d.pm_library_serialize.argtypes = [vp, _P(vp), _P(sz)]
    d.pm_library_load.argtypes = [vp, ctypes.c_char_p, sz, _P(vp)]
    d.pm_library_free.argtypes = [vp]
    d.pm_library_free.restype = None
    d.pm_library_pattern_count.argtypes = [vp]
    d.pm_library_pattern_meta_json.argtypes = [vp, ctypes.c_int]
    d.pm_library_pattern_meta_json.restype = ctypes.c_char_p
    d.pm_host_alloc.argtypes = [vp, sz]
    d.pm_host_alloc.restype = vp
    d.pm_host_free.argtypes = [vp, vp]
    d.pm_host_free.restype = None
    d.pm_batch_begin.argtypes = [vp, _P(vp)]
    d.pm_batch_add_pod.argtypes = [vp, vp, sz, ctypes.c_uint64]
    d.pm_batch_add_pod_device.argtypes = [vp, vp, sz, ctypes.c_uint64]
    d.pm_batch_submit.argtypes = [vp]
    d.pm_batch_wait.argtypes = [vp, _P(ResultView)]
    d.pm_batch_release.argtypes = [vp]
    d.pm_batch_release.restype = None
    d.pm_batch_stage.argtypes = [vp]
    d.pm_batch_rescan.argtypes = [vp, _P(ResultView)]
    d.pm_parse_stats.argtypes = [vp, _P(ctypes.c_int64), _P(ctypes.c_int64), _P(ctypes.c_int64)]
    d.pm_set_pipeline.argtypes = [vp, ctypes.c_int, ctypes.c_int, _P(ctypes.c_int), _P(ctypes.c_int)]
    d.pm_batch_line_index.argtypes = [vp, ctypes.c_int, _P(ctypes.c_int64), ctypes.c_int64]
    d.pm_batch_line_index.restype = ctypes.c_int64
    d.pm_parse.argtypes = [vp, ctypes.c_char_p, sz, _P(ctypes.c_char_p), _P(sz)]
    d.pm_line_index_noalloc.argtypes = [vp, ctypes.c_uint64, vp, ctypes.c_uint64, vp, vp]
    d.pm_noalloc_last_error.restype = ctypes.c_char_p
    _dll = d
    return d


def _check(rc: int):
    if rc != PM_OK:
        raise ScanError(rc, (load_library().pm_last_error() or b"").decode("utf-8", "replace"))


def _arr(ptr, n, dtype):
    if n == 0 or not ptr:
        return np.empty(0, dtype=dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).view(dtype)


class Result:
    """Copy of a pm_result_view (numpy arrays), detached from the batch."""

    def __init__(self, v: ResultView, copy: bool = True):
        n, p = int(v.n_events), int(v.n_pods)
        c = (lambda a: a.copy()) if copy else (lambda a: a)
        self.n_events, self.n_pods = n, p
        self.pod_index = c(_arr(v.pod_index, n, np.int32))
        self.pattern_index = c(_arr(v.pattern_index, n, np.int32))
        self.line_no = c(_arr(v.line_no, n, np.int32))
        self.line_start = c(_arr(v.line_start, n, np.int64))
        self.line_len = c(_arr(v.line_len, n, np.int32))
        self.ctx_first_line = c(_arr(v.ctx_first_line, n, np.int32))
        self.ctx_last_line = c(_arr(v.ctx_last_line, n, np.int32))
        self.ctx_start = c(_arr(v.ctx_start, n, np.int64))
        self.ctx_end = c(_arr(v.ctx_end, n, np.int64))
        self.score = c(_arr(v.score, n, np.float64))
        self.pod_tag = c(_arr(v.pod_tag, p, np.uint64))
        self.pod_first_event = c(_arr(v.pod_first_event, p + 1, np.int64))
        self.pod_n_lines = c(_arr(v.pod_n_lines, p, np.int32))
        self.ms_h2d, self.ms_scan, self.ms_post, self.ms_d2h, self.ms_total = v.ms_h2d, v.ms_scan, v.ms_post, v.ms_d2h, v.ms_total
        self.ms_dfa = v.ms_dfa
        self.bytes_scanned, self.n_candidates, self.n_verified = int(v.bytes_scanned), int(v.n_candidates), int(v.n_verified)
        self.gpu_launches = int(v.gpu_launches)

    def events_of(self, pod: int):
        a, b = int(self.pod_first_event[pod]), int(self.pod_first_event[pod + 1])
        return [dict(line_no=int(self.line_no[i]), pattern=int(self.pattern_index[i]), line_start=int(self.line_start[i]),
                     line_len=int(self.line_len[i]), ctx_first=int(self.ctx_first_line[i]), ctx_last=int(self.ctx_last_line[i]),
                     ctx_start=int(self.ctx_start[i]), ctx_end=int(self.ctx_end[i]), score=float(self.score[i]))
                for i in range(a, b)]


class Library:
    def __init__(self, engine: "Engine", handle: int, report: CompileReport):
        self._engine, self._h, self.report = engine, handle, report.as_dict()

    @property
    def n_patterns(self) -> int:
        return load_library().pm_library_pattern_count(self._h)

    def pattern_meta(self, i: int) -> dict:
        s = load_library().pm_library_pattern_meta_json(self._h, i)
        return json.loads(s.decode()) if s else {}

    def publish(self) -> "Library":
        _check(load_library().pm_library_publish(self._engine._ctx, self._h))
        return self

    def to_bytes(self) -> bytes:
        """The compiled library as bytes (pm_library_serialize): what the sync hook leaves beside the YAML."""
        p, n = ctypes.c_void_p(), ctypes.c_size_t()
        _check(load_library().pm_library_serialize(self._h, ctypes.byref(p), ctypes.byref(n)))
        return ctypes.string_at(p, n.value)

    def free(self):
        if self._h:
            load_library().pm_library_free(self._h)
            self._h = None


class PinnedBuffer:
    """Host memory from pm_host_alloc viewed as a numpy uint8 array (what Java wraps as a MemorySegment)."""

    def __init__(self, engine: "Engine", nbytes: int):
        self._engine = engine
        self.ptr = load_library().pm_host_alloc(engine._ctx, nbytes)
        if not self.ptr:
            _check(-7)
        self.nbytes = nbytes
        self.array = np.ctypeslib.as_array((ctypes.c_uint8 * nbytes).from_address(self.ptr))

    def free(self):
        if self.ptr:
            load_library().pm_host_free(self._engine._ctx, self.ptr)
            self.ptr = None


class Batch:
    def __init__(self, engine: "Engine"):
        self._engine = engine
        self._keep: List[object] = []
        h = ctypes.c_void_p()
        _check(load_library().pm_batch_begin(engine._ctx, ctypes.byref(h)))
        self._h = h

    def add_pod(self, log: Union[bytes, np.ndarray], tag: int = 0):
        if isinstance(log, np.ndarray):
            assert log.dtype == np.uint8 and log.flags["C_CONTIGUOUS"]
            self._keep.append(log)
            ptr, n = log.ctypes.data, log.size
        else:
            buf = ctypes.create_string_buffer(log, len(log)) if len(log) else ctypes.create_string_buffer(1)
            self._keep.append(buf)
            ptr, n = ctypes.addressof(buf), len(log)
        _check(load_library().pm_batch_add_pod(self._h, ptr, n, tag))

    def add_pod_device(self, device_ptr: int, nbytes: int, tag: int = 0, keep: object = None):
        """A pod that already lives in device memory (pm_batch_add_pod_device); `keep` is whatever owns it."""
        if keep is not None:
            self._keep.append(keep)
        _check(load_library().pm_batch_add_pod_device(self._h, device_ptr, nbytes, tag))

    def submit(self):
        _check(load_library().pm_batch_submit(self._h))

    def wait(self, copy: bool = True) -> Result:
        v = ResultView()
        _check(load_library().pm_batch_wait(self._h, ctypes.byref(v)))
        return Result(v, copy)

    def stage(self):
        _check(load_library().pm_batch_stage(self._h))

    def rescan(self, copy: bool = True) -> Result:
        v = ResultView()
        _check(load_library().pm_batch_rescan(self._h, ctypes.byref(v)))
        return Result(v, copy)

    def line_index(self, pod: int, n_lines: int) -> np.ndarray:
        out = np.empty(max(n_lines, 1), dtype=np.int64)
        got = load_library().pm_batch_line_index(self._h, pod, out.ctypes.data_as(_P(ctypes.c_int64)), out.size)
        if got < 0:
            _check(int(got))
        return out[:got]

    def release(self):
        if self._h:
            load_library().pm_batch_release(self._h)
            self._h = None
            self._keep.clear()


class Engine:
    """One pm_ctx: the GPUs of this process (one process per GPU under torchrun, or all GPUs of
    the node when used stand-alone)."""

    def __init__(self, n_devices: int = 0, devices: Optional[Sequence[int]] = None):
        d = load_library()
        ctx = ctypes.c_void_p()
        if devices is not None:
            ids = (ctypes.c_int * len(devices))(*devices)
            _check(d.pm_init_devices(ids, len(devices), ctypes.byref(ctx)))
        else:
            _check(d.pm_init(n_devices, ctypes.byref(ctx)))
        self._ctx = ctx

    @property
    def device_count(self) -> int:
        return load_library().pm_device_count(self._ctx)

    def set_pipeline(self, depth: int, post_sms: int = 16):
        """Batches in flight per device (pm_set_pipeline).  Returns (scan_sms, post_sms) of the
        partition the driver made; depth 1 goes back to one batch per device on all SMs."""
        a, b = ctypes.c_int(0), ctypes.c_int(0)
        _check(load_library().pm_set_pipeline(self._ctx, depth, post_sms, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def compile_library(self, spec: Union[dict, list, str, bytes]) -> Library:
        """a9: library JSON (SPEC.md section 1) -> compiled tables.  Raises ScanError carrying the
        pm_compile_report when a pattern is rejected (this is what the sync hook at
        PatternSyncService.java:228-253 surfaces through PatternLibraryReconciler.java:97-101)."""
        if isinstance(spec, (dict, list)):
            spec = json.dumps(spec)
        if isinstance(spec, str):
            spec = spec.encode("utf-8")
        h, rep = ctypes.c_void_p(), CompileReport()
        rc = load_library().pm_library_compile(self._ctx, spec, len(spec), ctypes.byref(h), ctypes.byref(rep))
        if rc != PM_OK:
            raise ScanError(rc, rep.message.decode("utf-8", "replace"), rep.as_dict())
        return Library(self, h, rep)

    def load_library(self, data: bytes) -> Library:
        """A library compiled earlier (Library.to_bytes) without compiling again (pm_library_load)."""
        h = ctypes.c_void_p()
        _check(load_library().pm_library_load(self._ctx, data, len(data), ctypes.byref(h)))
        lib = Library(self, h, CompileReport())
        lib.report["n_patterns"] = lib.n_patterns
        return lib

    def batch(self) -> Batch:
        return Batch(self)

    def pinned(self, nbytes: int) -> PinnedBuffer:
        return PinnedBuffer(self, nbytes)

    def scan(self, logs: Sequence[Union[bytes, np.ndarray]]) -> Result:
        b = self.batch()
        try:
            for i, log in enumerate(logs):
                b.add_pod(log, i)
            b.submit()
            return b.wait()
        finally:
            b.release()

    def parse(self, failure_data: Union[dict, bytes, str]) -> dict:
        """The /parse call: PodFailureData JSON -> AnalysisResult (LogParserRestClient.java:37-39)."""
        if isinstance(failure_data, dict):
            failure_data = json.dumps(failure_data)
        if isinstance(failure_data, str):
            failure_data = failure_data.encode("utf-8")
        out, n = ctypes.c_char_p(), ctypes.c_size_t()
        _check(load_library().pm_parse(self._ctx, failure_data, len(failure_data), ctypes.byref(out), ctypes.byref(n)))
        return json.loads(ctypes.string_at(out, n.value).decode("utf-8"))

    def parse_raw(self, body: bytes) -> bytes:
        """pm_parse without the Python-side JSON decode (what an HTTP front end forwards)."""
        out, n = ctypes.c_char_p(), ctypes.c_size_t()
        _check(load_library().pm_parse(self._ctx, body, len(body), ctypes.byref(out), ctypes.byref(n)))
        return ctypes.string_at(out, n.value)

    def parse_stats(self) -> dict:
        a, b, c = ctypes.c_int64(), ctypes.c_int64(), ctypes.c_int64()
        _check(load_library().pm_parse_stats(self._ctx, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)))
        return {"requests": a.value, "batches": b.value, "largest_batch": c.value}

    def close(self):
        if self._ctx:
            load_library().pm_shutdown(self._ctx)
            self._ctx = None


class LogParserClient:
    """Mirror of LogParserClient (LogParserClient.java:36-55): ``analyze_log(failure_data)``
    returns the AnalysisResult; failures propagate as exceptions, which is how the reference's
    callers see a failed Uni (PodFailureWatcher.java:295-304)."""

    def __init__(self, engine: Engine):
        self._engine = engine

    def analyze_log(self, failure_data: dict) -> dict:
        return self._engine.parse(failure_data)
