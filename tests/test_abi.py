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


def test_distributed_level_schedule_posts_before_waits():
    # every wait phase of a distributed level has its post phase earlier in the list (virtual-rank drivers rely on it)
    from parallel_kd_tree_b200.parallel import dist_build as db
    steps = [ph for ph, _ in db.level_schedule()]
    pairs = [(db.PH_SAMPLE_POST, db.PH_PIVOT_WAIT), (db.PH_BRACKET_POST, db.PH_BRACKET_WAIT), (db.PH_HIST, db.PH_SCAN),
             (db.PH_CNT_POST, db.PH_CNT_WAIT), (db.PH_PART_POST, db.PH_PART_WAIT), (db.PH_MAX_POST, db.PH_MAX_WAIT),
             (db.PH_FIX_POST, db.PH_FIX_WAIT)]
    for post, wait in pairs:
        assert post in steps and wait in steps and steps.index(post) < steps.index(wait)
    assert steps[0] == db.PH_STATE and steps.count(db.PH_HIST) == 6 and steps.count(db.PH_PART) == 1
