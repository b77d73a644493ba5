"""Pattern hand-off (a6-a8): file discovery rule and the sync-time compile hook.  The CPU tests use
the compiler harness in place of a GPU context; the GPU test goes through pm_library_compile."""
import os

import pytest
import yaml

from operator_b200 import patterns, synth


def _tree(tmp_path):
    lib, _, _ = synth.make_library(12)
    root = tmp_path / "mylib" / "repo"
    (root / "quarkus").mkdir(parents=True)
    (root / "quarkus" / "core-patterns.yaml").write_text(yaml.safe_dump({"metadata": lib["metadata"], "patterns": lib["patterns"][:8]}))
    (root / "extra.yml").write_text(yaml.safe_dump({"patterns": lib["patterns"][8:]}))
    (root / "README.md").write_text("not yaml")
    (root / ".github").mkdir()
    (root / ".github" / "ci.yaml").write_text("on: push\n")
    return root


def test_discovery_rule(tmp_path):
    root = _tree(tmp_path)
    files = patterns.discover_pattern_files(str(root))
    assert [os.path.relpath(f, root) for f in files] == ["extra.yml", ".github/ci.yaml", "quarkus/core-patterns.yaml"]
    # library id = basename without extension (PatternSyncService.java:88-114)
    assert patterns.available_libraries(str(tmp_path / "mylib")) == ["extra", "ci", "core-patterns"]
    libs, index = patterns.load_pattern_files(files)
    assert len(libs) == 2 and [i for i, _ in index] == [0, 4]     # ci.yaml is not a pattern library
    assert sum(len(l["patterns"]) for l in libs) == 12


class _FakeEngine:
    """Routes compile_library through the CPU compiler harness (same C++ compiler, no device)."""

    def __init__(self, harness):
        self.h = harness

    def compile_library(self, spec):
        from operator_b200.engine import ScanError
        rc, h, rep = self.h.compile(spec)
        if rc != 0:
            d = {n: (getattr(rep, n).decode() if isinstance(getattr(rep, n), bytes) else getattr(rep, n)) for n, _ in rep._fields_}
            raise ScanError(rc, d["message"], d)

        class L:
            report = {"n_patterns": rep.n_patterns}

            def publish(self):
                return self
        return L()


def test_validate_hook_accepts_and_rejects(tmp_path, harness):
    root = _tree(tmp_path)
    lib = patterns.validate_patterns(_FakeEngine(harness), str(root))
    assert lib.report["n_patterns"] == 12
    bad = {"patterns": [{"id": "uses-lookahead", "primary_pattern": {"regex": r"timeout (?=after)"}}]}
    (root / "zz-bad.yaml").write_text(yaml.safe_dump(bad))
    with pytest.raises(patterns.PatternCompileError) as ei:
        patterns.validate_patterns(_FakeEngine(harness), str(root))
    msg = ei.value.status_message()
    assert msg.startswith("Failed to reconcile: zz-bad.yaml: pattern 'uses-lookahead' at regex offset 8") and "look-ahead" in msg
    assert ei.value.report["pattern_index"] == 4 and ei.value.report["status"] == -3


@pytest.mark.gpu
def test_validate_hook_on_device(tmp_path, engine):
    root = _tree(tmp_path)
    lib = patterns.validate_patterns(engine, str(root))
    assert lib.n_patterns == 12 and lib.pattern_meta(0)["id"] == "syn-00008"   # extra.yml sorts first


@pytest.mark.gpu
def test_compiled_library_on_the_volume(tmp_path, engine):
    """8(f).2: the sync hook leaves the compiled library beside the YAML; the service loads it without
    compiling, scans with it exactly as with a freshly compiled one, and ignores it once the YAML changed."""
    from oracle import oracle
    root = _tree(tmp_path)
    assert patterns.load_compiled(engine, str(root)) is None                       # nothing there yet
    lib = patterns.validate_patterns(engine, str(root), save=True)
    assert os.path.isfile(root / patterns.COMPILED_NAME)
    assert [os.path.relpath(f, root) for f in patterns.discover_pattern_files(str(root))] == ["extra.yml", ".github/ci.yaml", "quarkus/core-patterns.yaml"]
    full, wit, swit = synth.make_library(12)
    corpus, _ = synth.make_corpus(300000, wit, swit, p_hit=3e-2)
    want = engine.scan([corpus])
    loaded = patterns.load_compiled(engine, str(root))                              # publishes the loaded copy
    assert loaded is not None and loaded.n_patterns == lib.n_patterns == 12
    assert [loaded.pattern_meta(i) for i in range(12)] == [lib.pattern_meta(i) for i in range(12)]
    got = engine.scan([corpus])
    assert got.n_events == want.n_events > 20
    for name in ("pattern_index", "line_no", "ctx_first_line", "ctx_last_line", "ctx_start", "ctx_end", "score"):
        assert (getattr(got, name) == getattr(want, name)).all(), name
    # stale: a pattern file changed after the compile
    with open(root / "extra.yml", "a") as f:
        f.write("\n# touched\n")
    assert patterns.load_compiled(engine, str(root)) is None
    # damaged
    patterns.validate_patterns(engine, str(root), save=True)
    p = root / patterns.COMPILED_NAME
    data = bytearray(p.read_bytes())
    data[len(data) // 2] ^= 1
    p.write_bytes(bytes(data))
    assert patterns.load_compiled(engine, str(root)) is None
