"""Docs stay honest: everything PARITY.md points at (source files, test names) must exist."""
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_parity_table_references_exist():
    txt = open(os.path.join(ROOT, "PARITY.md")).read()
    defined = set()
    for f in glob.glob(os.path.join(ROOT, "tests", "*.py")):
        defined |= set(re.findall(r"def (test_[a-z0-9_]+)", open(f).read()))
        defined.add(os.path.basename(f)[:-3])
    named = set(re.findall(r"test_[a-z0-9_]+", txt))
    # a trailing '_*' in the table means "every test with this prefix"
    missing = [t for t in named if t not in defined and not any(d.startswith(t.rstrip("_")) for d in defined)]
    assert not missing, missing
    for rel in set(re.findall(r"`(pkg/[a-z_/]+\.py|csrc/[a-z_/]+\.(?:cu|cc|cuh|h))", txt)):
        path = os.path.join(ROOT, rel.replace("pkg/", "distributedtensorflowexample_b200/"))
        assert os.path.exists(path), rel
    for rel in ("README.md", "DESIGN.md", "INTEGRATION.md", "BASELINE.md", "run_single_gpu.sh", "run_multi_gpu.sh",
                "main.py", "bench.py", "__graft_entry__.py", "profiles/README.md"):
        assert os.path.exists(os.path.join(ROOT, rel)), rel


def test_profiles_readme_lists_existing_files():
    txt = open(os.path.join(ROOT, "profiles", "README.md")).read()
    for name in set(re.findall(r"`([a-z0-9_]+\.(?:jsonl|txt|md))`", txt)):
        assert os.path.exists(os.path.join(ROOT, "profiles", name)), name
