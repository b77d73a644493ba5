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
