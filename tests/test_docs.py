"""The binding a maintainer would paste (INTEGRATION.md's Rust `extern "C"` block) names every entry point include/rt3.h declares."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_integration_md_lists_every_entry_point():
    with open(os.path.join(ROOT, "include", "rt3.h")) as f:
        declared = set(re.findall(r"\b(rt3_[a-z0-9_]+)\s*\(", f.read()))
    with open(os.path.join(ROOT, "INTEGRATION.md")) as f:
        text = f.read()
    missing = sorted(d for d in declared if ("fn " + d + "(") not in text)
    assert not missing, missing
