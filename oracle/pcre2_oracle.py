"""TEST INFRASTRUCTURE (oracle) -- ctypes wrapper of oracle/scan_pcre2.c.

Second opinion on the Python oracle's regex verdicts (PCRE2 instead of Python ``re``) and the
timed CPU baseline of bench.py.  Never imported by the product path.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Sequence, Tuple

import numpy as np

from . import javaregex

_SO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_build", "liboracle_pcre2.so")
_dll = None


def _load():
    global _dll
    if _dll is None:
        if not os.path.exists(_SO):
            raise RuntimeError("%s missing: run `make`" % _SO)
        d = ctypes.CDLL(_SO)
        d.orc_compile.restype = ctypes.c_void_p
        d.orc_compile.argtypes = [ctypes.POINTER(ctypes.c_char_p), ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        d.orc_free.argtypes = [ctypes.c_void_p]
        d.orc_split.restype = ctypes.c_int64
        d.orc_split.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64]
        d.orc_scan.restype = ctypes.c_int64
        d.orc_scan.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int,
                               ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64]
        _dll = d
    return _dll


def available() -> bool:
    try:
        d = _load()
        bad = ctypes.c_int(0)
        arr = (ctypes.c_char_p * 1)(b"a")
        h = d.orc_compile(arr, 1, ctypes.byref(bad))
        if not h:
            return False
        d.orc_free(h)
        return True
    except Exception:
        return False


class Pcre2Library:
    def __init__(self, java_regexes: Sequence[str]):
        d = _load()
        srcs = [javaregex.to_pcre2(r).encode("utf-8") for r in java_regexes]
        arr = (ctypes.c_char_p * max(len(srcs), 1))(*srcs)
        bad = ctypes.c_int(-1)
        self._h = d.orc_compile(arr, len(srcs), ctypes.byref(bad))
        if not self._h:
            raise RuntimeError("PCRE2 rejected pattern %d" % bad.value)
        self.n = len(srcs)

    def close(self):
        if self._h:
            _load().orc_free(self._h)
            self._h = None


def split(log: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    d = _load()
    n = int(log.size)
    cap = int((log == 10).sum()) + 2
    starts = np.empty(cap, dtype=np.int64)
    lens = np.empty(cap, dtype=np.int32)
    k = d.orc_split(log.ctypes.data, n, starts.ctypes.data, lens.ctypes.data, cap)
    return starts[:k], lens[:k]


def scan(lib: Pcre2Library, log: np.ndarray, starts: np.ndarray, lens: np.ndarray, threads: int = 1,
         cap: int = 1 << 22) -> Tuple[np.ndarray, np.ndarray, int]:
    """Returns (line index 0-based, pattern index, total matches)."""
    d = _load()
    ol = np.empty(cap, dtype=np.int64)
    op = np.empty(cap, dtype=np.int32)
    total = d.orc_scan(lib._h, log.ctypes.data, starts.ctypes.data, lens.ctypes.data, int(starts.size), threads,
                       ol.ctypes.data, op.ctypes.data, cap)
    k = min(int(total), cap)
    return ol[:k], op[:k], int(total)
