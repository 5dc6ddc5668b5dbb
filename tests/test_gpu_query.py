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
