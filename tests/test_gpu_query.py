"""sm_100a query kernels against the native CPU core (bit-for-bit)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed,dim,n,nq", [(1, 3, 1000, 64), (7, 3, 50000, 500), (3, 2, 200000, 2000),
                                           (2, 8, 50000, 300), (4, 1, 3000, 50), (5, 16, 20000, 100),
                                           (6, 2, 2, 5), (6, 2, 1, 5), (8, 128, 5000, 10)])
def test_nearest_bits_equal_cpu(kd, seed, dim, n, nq):
    pts, qs = kd.ops.generate_problem(seed, dim, n, nq)
    cpu = kd.KDTree(pts).nearest(qs)
    tree = kd.KDTree(pts.cuda())
    res = tree.nearest(qs.cuda())
    assert torch.equal(res.index.cpu(), cpu.index)
    assert torch.equal(res.dist2.cpu().view(torch.int32), cpu.dist2.view(torch.int32))
    assert torch.equal(res.dist.cpu().view(torch.int32), cpu.dist.view(torch.int32))
    assert torch.equal(res.visited.cpu(), cpu.visited)


@pytest.mark.parametrize("seed,dim,n,nq", [(1, 128, 20000, 37), (2, 16, 30000, 100), (3, 33, 4097, 16), (4, 3, 5000, 50),
                                           (5, 64, 1, 3), (6, 200, 300, 20)])
def test_dense_path_equals_dfs_path(kd, seed, dim, n, nq):
    # node-parallel path (+ its DFS fallback where pruning cannot be excluded, e.g. the 3-D case) == thread-per-query DFS
    pts, qs = kd.ops.generate_problem(seed, dim, n, nq)
    tree = kd.KDTree(pts.cuda())
    a, b = tree.nearest(qs.cuda(), mode="dfs"), tree.nearest(qs.cuda(), mode="dense")
    cpu = kd.KDTree(pts).nearest(qs)
    for x in (a, b):
        assert torch.equal(x.index.cpu(), cpu.index)
        assert torch.equal(x.dist2.cpu().view(torch.int32), cpu.dist2.view(torch.int32))
        assert torch.equal(x.dist.cpu().view(torch.int32), cpu.dist.view(torch.int32))
        assert torch.equal(x.visited.cpu(), cpu.visited)


def test_dense_path_orders_equal_distances_like_the_traversal(kd):
    # duplicated points: several nodes attain the smallest distance; the reference keeps the one its DFS meets first
    g = torch.Generator().manual_seed(5)
    base = torch.randn(500, 40, generator=g) * 30
    pts = base[torch.randint(0, 500, (6000,), generator=g)].contiguous()
    qs = base[:64] + 0.25
    cpu = kd.KDTree(pts).nearest(qs)
    res = kd.KDTree(pts.cuda()).nearest(qs.cuda(), mode="dense")
    assert torch.equal(res.index.cpu(), cpu.index) and torch.equal(res.visited.cpu(), cpu.visited)
    assert torch.equal(res.dist2.cpu().view(torch.int32), cpu.dist2.view(torch.int32))


def test_ops_reject_or_move_foreign_tensors(kd):
    pts, qs = kd.ops.generate_problem(1, 3, 1000, 5)
    tree = kd.KDTree(pts.cuda())
    res = kd.ops.nearest(tree.tree_points, tree.perm.cpu(), qs)          # CPU perm / queries are moved, not dereferenced
    assert torch.equal(res[0].cpu(), kd.KDTree(pts).nearest(qs).index)
    with pytest.raises(ValueError):
        kd.ops.build_into(pts.cuda(), torch.empty(999, dtype=torch.int32, device="cuda"), None)
    with pytest.raises(ValueError):
        kd.ops.build_into(pts.cuda(), torch.empty(1000, dtype=torch.int32), None)     # perm on the wrong device


def test_ordered_batches_give_the_same_answers(kd):
    # >= 4096 queries: grouped by the sub-tree they enter first (own counting sort); results are order independent
    pts, _ = kd.ops.generate_problem(3, 3, 300000, 0)
    qs = kd.ops.generate_uniform(9, 3, 20000)
    tc, tg = kd.KDTree(pts), kd.KDTree(pts.cuda())
    order = kd.ops.query_order(tg.tree_points, qs.cuda())
    assert torch.equal(order.long().sort().values.cpu(), torch.arange(20000))         # a permutation
    cpu, res = tc.nearest(qs), tg.nearest(qs.cuda(), mode="dfs")
    assert torch.equal(res.index.cpu(), cpu.index) and torch.equal(res.visited.cpu(), cpu.visited)
    assert torch.equal(res.dist.cpu().view(torch.int32), cpu.dist.view(torch.int32))
    ic, dc = kd.ops.knn(tc.tree_points, tc.perm, qs[:6000], 8)
    ig, dg = kd.ops.knn(tg.tree_points, tg.perm, qs[:6000].cuda(), 8)
    assert torch.equal(ig.cpu(), ic) and torch.equal(dg.cpu().view(torch.int32), dc.view(torch.int32))


def test_default_problem_golden_lines_on_gpu(kd):
    # the graded problem (PDF p.4): 500000 x 128, seed 1; golden = unmodified ./sequential
    golden = ["707.539", "653.765", "693.995", "715.386", "732.924", "702.161", "739.194", "693.53", "698.66",
              "736.767"]
    tree, qs = kd.KDTree.from_seed(1, 128, 500000, "cuda")
    res = tree.nearest(qs)
    assert [kd.utils.format_float(float(d)) for d in res.dist.cpu()] == golden
    assert res.visited.min().item() == 500000      # 128-D: no pruning ever fires


def test_materialize_and_bbox(kd):
    pts, _ = kd.ops.generate_problem(3, 5, 12345, 0)
    t_cpu, t_gpu = kd.KDTree(pts), kd.KDTree(pts.cuda())
    for a, b in zip(t_cpu.nodes(), t_gpu.nodes()):
        assert torch.equal(a, b.cpu())
    mn, mx = t_gpu.bbox()
    assert torch.equal(mn.cpu(), pts.min(0).values) and torch.equal(mx.cpu(), pts.max(0).values)
    assert torch.equal(kd.ops.gather_rows(pts.cuda(), t_gpu.perm).cpu(), pts[t_cpu.perm.long()])


def test_forest_on_gpu_matches_cpu_forest(kd):
    forest_c, qs = kd.models.Forest.from_seed(2, 3, 100003, 4)
    forest_g, _ = kd.models.Forest.from_seed(2, 3, 100003, 4, devices=["cuda"])
    dc, ic = forest_c.nearest(qs)
    dg, ig = forest_g.nearest(qs)
    assert torch.equal(dc, dg) and torch.equal(ic, ig)


@pytest.mark.parametrize("dim,n,nq,k", [(3, 50000, 500, 1), (3, 50000, 500, 8), (8, 100000, 300, 16), (2, 3, 10, 5)])
def test_knn_gpu_equals_cpu_definition(kd, dim, n, nq, k):
    pts, qs = kd.ops.generate_problem(4, dim, n, nq)
    tc, tg = kd.KDTree(pts), kd.KDTree(pts.cuda())
    ic, dc = kd.ops.knn(tc.tree_points, tc.perm, qs, k)            # squared distances: exact on both sides
    ig, dg = kd.ops.knn(tg.tree_points, tg.perm, qs.cuda(), k)
    assert torch.equal(ig.cpu(), ic)
    assert torch.equal(dg.cpu().view(torch.int32), dc.view(torch.int32))
