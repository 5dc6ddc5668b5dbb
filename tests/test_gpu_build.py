"""sm_100a build kernels against the native CPU builder (which the CPU suite ties to the reference)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _compare(kd, pts, what):
    cpu_perm = kd.ops.build(pts)[0]
    dev = pts.cuda()
    perm, tree_pts = kd.ops.build(dev)
    torch.cuda.synchronize()
    got = perm.cpu()
    if not torch.equal(got, cpu_perm):
        bad = (got != cpu_perm).nonzero().flatten()
        viol = kd.ops.check_tree(pts, got) if got.min() >= 0 and got.max() < len(got) else -1
        pytest.fail(f"{what}: {bad.numel()} of {got.numel()} slots differ (first {bad[:8].tolist()}); "
                    f"rule violations {viol}; launches {kd.ops.last_launch_count()}")
    assert torch.equal(tree_pts.cpu(), pts[cpu_perm.long()]), f"{what}: tree_points != points[perm]"


@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 6, 7, 8, 9, 15, 16, 17, 31, 33, 63, 64, 65, 100, 127, 129, 255, 257, 511,
                               513, 1000, 1023, 1025, 2047, 2049, 3000, 4095, 4096])
@pytest.mark.parametrize("dim", [1, 2, 3, 4, 8])
def test_finisher_only_sizes(kd, n, dim):
    pts, _ = kd.ops.generate_problem(n + dim, dim, n, 0)
    _compare(kd, pts, f"n={n} dim={dim}")


@pytest.mark.parametrize("n,dim", [(4097, 3), (8191, 2), (8193, 3), (10007, 1), (65536, 3), (100003, 2),
                                   (500000, 3), (1 << 20, 2), (1 << 20, 3), ((1 << 21) + 5, 4)])
def test_select_partition_levels(kd, n, dim):
    pts, _ = kd.ops.generate_problem(n % 97 + 1, dim, n, 0)
    _compare(kd, pts, f"n={n} dim={dim}")


@pytest.mark.parametrize("n,dim", [(300, 5), (5000, 8), (20000, 16), (70000, 8), (30000, 128)])
def test_gather_mode_high_dim(kd, n, dim):
    pts, _ = kd.ops.generate_problem(3, dim, n, 0)
    _compare(kd, pts, f"n={n} dim={dim}")


def test_heavy_ties_and_signed_zero(kd):
    g = torch.Generator().manual_seed(0)
    for n, dim in [(3000, 3), (50000, 2), (300000, 3)]:
        pts = torch.randint(-3, 4, (n, dim), generator=g).float()
        pts[::17] = -0.0
        _compare(kd, pts, f"ties n={n} dim={dim}")
    pts = torch.zeros(20000, 3)           # every key equal: all six digit passes, pure index order
    _compare(kd, pts, "all-equal")
    for n in (7, 100, 1000, 4096):        # the same inside the shared-memory sub-tree builder (index digits only)
        _compare(kd, torch.zeros(n, 2), f"all-equal n={n}")
        pts = torch.randint(-2, 3, (n, 3), generator=g).float()
        _compare(kd, pts, f"ties n={n}")
    pts = torch.tensor([0.0, 1e-30, -1e-30, 1.0, -1.0, 3e38, -3e38, 1e-38]).repeat(500)[:, None].repeat(1, 2)
    pts = pts + torch.arange(pts.shape[0])[:, None] * 0.0   # wide exponent range across zero
    _compare(kd, pts[torch.randperm(pts.shape[0], generator=g)], "wide-range")


def test_clustered(kd):
    pts = kd.ops.generate_clustered(5, 3, 400000, n_clusters=8, sigma=2.0)
    _compare(kd, pts, "clustered")


def test_build_is_repeatable_and_workspace_reusable(kd):
    pts, _ = kd.ops.generate_problem(9, 3, 200000, 0)
    dev = pts.cuda()
    ws = torch.empty(kd.ops.build_workspace_bytes(*dev.shape), dtype=torch.uint8, device="cuda")
    perm = torch.empty(dev.shape[0], dtype=torch.int32, device="cuda")
    tp = torch.empty_like(dev)
    ref = kd.ops.build(pts)[0]
    for _ in range(3):
        perm.fill_(-1)
        kd.ops.build_into(dev, perm, tp, ws, verify=True)
        assert torch.equal(perm.cpu(), ref)
    assert kd.ops.last_launch_count() > 0


def test_large_build_rule_check(kd):
    # 2^24 x 3 (BASELINE.json config 2 size): compare with the multi-threaded CPU builder
    n = 1 << 24
    pts = kd.ops.generate_uniform(1, 3, n)
    perm, _ = kd.ops.build(pts.cuda())
    ref = kd.ops.build(pts)[0]
    assert torch.equal(perm.cpu(), ref)


def test_graphed_build_replays_give_the_same_trees(kd):
    # the whole launch sequence captured in a CUDA graph; new points are copied into the object's input tensor
    n, dim = 300000, 3
    gb = kd.ops.GraphedBuild(n, dim)
    for seed in (3, 4, 5):
        pts = kd.ops.generate_uniform(seed, dim, n)
        gb.points.copy_(pts)
        gb.run(verify=True)
        ref = kd.ops.build(pts)[0]
        assert torch.equal(gb.perm.cpu(), ref)
        assert torch.equal(gb.tree_points.cpu(), pts[ref.long()])
