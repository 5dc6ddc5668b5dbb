"""Every function declared in include/kdtree_b200.h must be exported and have a ctypes signature
(an undeclared signature silently truncates 64-bit pointers)."""
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_header_and_ctypes_table_agree(kd):
    text = (ROOT / "include" / "kdtree_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    declared = set(re.findall(r"\b(kd_[a-z0-9_]+)\s*\(", text))
    lib = kd.lib()
    assert declared, "no declarations parsed"
    missing = sorted(declared - lib._kd_declared)
    assert not missing, f"functions without a ctypes signature: {missing}"
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in the header but not exported"
