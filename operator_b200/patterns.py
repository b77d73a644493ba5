"""Pattern hand-off: file discovery and the sync-time compile hook (rows a6-a8 of SURVEY.md section 8).

Reference behaviour mirrored here:

* ``PatternSyncService.syncRepository`` lays repositories out as
  ``/shared/patterns/<library>/<repo>/**`` (PatternSyncService.java:42-77, PATTERN_CACHE_DIR at
  :28) and its ``validatePatterns`` only COUNTS ``.yaml``/``.yml`` files (:228-253);
* ``getAvailableLibraries`` walks the library directory recursively and names each library by
  the file's basename without extension (:88-114);
* a failure that must reach the user surfaces through the reconciler's catch block as status
  ``Failed`` / ``"Failed to reconcile: " + message`` (PatternLibraryReconciler.java:97-101).

``validate_patterns`` is the drop-in for the count-only validation: it parses every discovered
file, compiles the whole set with the engine and raises ``PatternCompileError`` (whose message
names file, pattern id and regex offset) instead of logging a count.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Tuple

import yaml

from .engine import Engine, Library, ScanError

PATTERN_CACHE_DIR = "/shared/patterns"      # PatternSyncService.java:28


class PatternCompileError(Exception):
    """What PatternLibraryReconciler.reconcile would put into status.message (:97-101)."""

    def __init__(self, message: str, report: dict, source_file: str = ""):
        super().__init__(message)
        self.report, self.source_file = report, source_file

    def status_message(self) -> str:
        return "Failed to reconcile: " + str(self)


def discover_pattern_files(root: str) -> List[str]:
    """Recursive walk, regular files ending in .yaml or .yml (PatternSyncService.java:236-243)."""
    out = []
    for dirpath, dirnames, files in os.walk(root):
        dirnames.sort()
        for fn in sorted(files):
            if fn.endswith(".yaml") or fn.endswith(".yml"):
                p = os.path.join(dirpath, fn)
                if os.path.isfile(p):
                    out.append(p)
    return out


def available_libraries(library_path: str) -> List[str]:
    """Library id = file basename without .yaml/.yml (PatternSyncService.java:88-114)."""
    names = []
    for p in discover_pattern_files(library_path):
        base = os.path.basename(p)
        names.append(base[:-5] if base.endswith(".yaml") else base[:-4])
    return names


def load_pattern_files(paths: List[str]) -> Tuple[List[dict], List[Tuple[int, str]]]:
    """Parses the YAML files into library dicts (SPEC.md section 1).  Returns the libraries and, for
    error reporting, (first pattern index, file) per library."""
    libs, index = [], []
    n = 0
    for p in paths:
        with open(p, "r", encoding="utf-8") as f:
            doc = yaml.safe_load(f)
        if not isinstance(doc, dict) or "patterns" not in doc:
            continue        # not a pattern library (e.g. CI config living in the same repo)
        doc.setdefault("metadata", {}).setdefault("library_id", os.path.splitext(os.path.basename(p))[0])
        libs.append(doc)
        index.append((n, p))
        n += len(doc.get("patterns") or [])
    return libs, index


COMPILED_NAME = ".podmortem-compiled.pmlb"   # beside the YAML tree; not a .yaml, so the file count of the reference is unchanged


def _fingerprint(paths: List[str]) -> bytes:
    """SHA-256 over the pattern files (relative order, names and contents): a compiled library on the
    volume is only good for exactly the YAML it was compiled from."""
    import hashlib
    h = hashlib.sha256()
    for path in paths:
        with open(path, "rb") as f:
            data = f.read()
        h.update(os.path.basename(path).encode() + b"\0" + str(len(data)).encode() + b"\0" + data)
    return h.digest()


def load_compiled(engine: Engine, repo_path: str, publish: bool = True) -> Optional[Library]:
    """Service start-up: take the library the sync step compiled (8(f).2) if it matches the YAML on the
    volume; None when there is none, it is stale, or it was built by another compiler version."""
    path = os.path.join(repo_path, COMPILED_NAME)
    if not os.path.isfile(path):
        return None
    with open(path, "rb") as f:
        data = f.read()
    if len(data) < 32 or data[:32] != _fingerprint(discover_pattern_files(repo_path)):
        return None
    try:
        lib = engine.load_library(data[32:])
    except ScanError:
        return None
    if publish:
        lib.publish()
    return lib


def validate_patterns(engine: Engine, repo_path: str, publish: bool = True, save: bool = False) -> Library:
    """Sync-time hook (replaces the count in PatternSyncService.validatePatterns, :228-253):
    compile everything under repo_path; on rejection raise with file + pattern id + offset.
    save: leave the compiled library beside the YAML for the service to load (load_compiled)."""
    paths = discover_pattern_files(repo_path)
    libs, index = load_pattern_files(paths)
    try:
        lib = engine.compile_library(libs)
    except ScanError as e:
        rep = e.report or {}
        src = ""
        for first, path in index:
            if rep.get("pattern_index", -1) >= first:
                src = path
        where = "%s: pattern %r" % (os.path.relpath(src, repo_path) if src else "?", rep.get("pattern_id", ""))
        if rep.get("secondary_index", -1) >= 0:
            where += " secondary #%d" % rep["secondary_index"]
        if rep.get("regex_offset", -1) >= 0:
            where += " at regex offset %d" % rep["regex_offset"]
        raise PatternCompileError("%s: %s" % (where, e.message), rep, src) from None
    if save:
        tmp = os.path.join(repo_path, COMPILED_NAME + ".tmp")
        with open(tmp, "wb") as f:
            f.write(_fingerprint(paths) + lib.to_bytes())
        os.replace(tmp, os.path.join(repo_path, COMPILED_NAME))
    if publish:
        lib.publish()
    return lib
