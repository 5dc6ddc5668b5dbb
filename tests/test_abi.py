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


def test_torch_ops_are_registered_and_agree_with_the_python_entry_points(kd):
    import torch
    pts, qs = kd.ops.generate_problem(2, 3, 2000, 7)
    perm, tp = kd.ops.build(pts)
    p2, t2 = torch.ops.kd.build(pts)
    assert torch.equal(perm, p2) and torch.equal(tp, t2)
    a = kd.ops.nearest(tp, perm, qs)
    b = torch.ops.kd.nearest(tp, perm, qs)
    assert all(torch.equal(x, y) for x, y in zip(a, b))
    assert all(torch.equal(x, y) for x, y in zip(kd.ops.materialize_nodes(perm), torch.ops.kd.materialize_nodes(perm)))
    assert all(torch.equal(x, y) for x, y in zip(kd.ops.bbox(pts), torch.ops.kd.bbox(pts)))
