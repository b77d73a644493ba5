"""B200-native log-scan engine behind the podmortem operator's log-parser call.

Only what the hot path needs lives here: `csrc/` (CUDA kernels, pattern compiler, C ABI),
`engine.py` (ctypes binding + mirror of the reference's LogParserClient), `patterns.py`
(pattern-file discovery and the sync-time compile hook), `synth.py` (seeded inputs) and
`dist.py` (one-process-per-GPU plumbing for the benchmark).
"""
from .engine import Engine, LogParserClient, ScanError, load_library, LIB_PATH, ABI_SYMBOLS  # noqa: F401
