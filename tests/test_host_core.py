"""CPU host core against the unmodified reference functions (oracle B) — SURVEY.md §4 / §7.2 steps 1-2."""
import numpy as np
import pytest
import torch

from conftest import write_points


@pytest.mark.parametrize("seed,dim,n", [(1, 3, 1000), (7, 1, 17), (5, 128, 40), (9, 2, 4097)])
def test_generator_bit_exact(kd, oracle, tmp_path, seed, dim, n):
    ref = oracle.points(seed, dim, n, 10, tmp_path)
    pts, qs = kd.ops.generate_problem(seed, dim, n, 10)
    assert np.array_equal(ref[:n].view(np.uint32), pts.numpy().view(np.uint32))
    assert np.array_equal(ref[n:].view(np.uint32), qs.numpy().view(np.uint32))


def test_generator_slice_mode_matches_sequential_stream(kd):
    # kdtree_mpi.cpp:19-41: a rank's slice is a window of the sequential stream
    full = kd.ops.generate_uniform(11, 5, 1000)
    for first, cnt in [(0, 10), (123, 456), (990, 10)]:
        part = kd.ops.generate_uniform(11, 5, cnt, first)
        assert torch.equal(full[first:first + cnt], part)


@pytest.mark.parametrize("seed,dim,n", [(1, 2, 1), (1, 2, 2), (1, 2, 3), (2, 3, 4), (3, 2, 7), (4, 3, 100),
                                        (5, 1, 257), (6, 8, 1000), (7, 3, 4096), (8, 2, 10007), (9, 128, 300),
                                        (10, 3, 50000)])
def test_cpu_build_equals_reference_tree(kd, oracle, tmp_path, seed, dim, n):
    ref_ids = oracle.inorder_ids(f"seed:{seed}", dim, n, tmp_path)
    pts, _ = kd.ops.generate_problem(seed, dim, n, 0)
    tree = kd.KDTree(pts)
    assert np.array_equal(ref_ids, tree.ids_inorder().numpy())
    assert tree.check() == 0
    assert torch.equal(kd.ops.build_cpu_sortref(pts), tree.perm)


def test_cpu_build_with_heavy_ties_matches_sort_formulation(kd):
    g = torch.Generator().manual_seed(0)
    for n, dim in [(5000, 2), (3001, 3), (70000, 3)]:
        pts = torch.randint(-3, 4, (n, dim), generator=g).float()   # 7 distinct values per axis
        pts[::17] = -0.0                                             # -0 and +0 must tie
        perm = kd.ops.build(pts)[0]
        assert torch.equal(perm, kd.ops.build_cpu_sortref(pts))
        assert kd.ops.check_tree(pts, perm) == 0


def test_print_tree_and_head_text_identical(kd, oracle):
    for seed, dim, n in [(3, 2, 7), (4, 6, 9), (5, 7, 64), (6, 1, 5), (2, 5, 33)]:
        pts, _ = kd.ops.generate_problem(seed, dim, n, 0)
        tree = kd.KDTree(pts)
        assert tree.format_tree() == oracle.tree_text(f"seed:{seed}", dim, n)
        assert tree.format_head_and_leaves() == oracle.head_text(f"seed:{seed}", dim, n)


@pytest.mark.parametrize("seed,dim,n,nq", [(1, 3, 1000, 50), (7, 3, 50000, 200), (3, 2, 20000, 200), (2, 8, 5000, 50),
                                           (4, 1, 3000, 50), (5, 16, 2000, 20), (6, 2, 2, 5), (6, 2, 1, 5)])
def test_cpu_nearest_equals_reference(kd, oracle, seed, dim, n, nq):
    pts, qs = kd.ops.generate_problem(seed, dim, n, nq)
    tree = kd.KDTree(pts)
    res = tree.nearest(qs)
    ref = oracle.queryx(f"seed:{seed}", dim, n, nq)
    assert [r[1] for r in ref] == res.ids.tolist()
    assert [r[2] for r in ref] == res.dist2.numpy().view(np.uint32).tolist()
    assert [r[3] for r in ref] == res.dist.numpy().view(np.uint32).tolist()
    lines = "".join(kd.utils.result_line(n + q, float(res.dist[q])) + "\n" for q in range(nq))
    assert lines == oracle.lines(f"seed:{seed}", dim, n, nq)


def test_clustered_points_against_reference(kd, oracle, tmp_path):
    n, dim, nq = 20000, 3, 100
    allp = kd.ops.generate_clustered(5, dim, n + nq, n_clusters=8, sigma=3.0)
    src = write_points(tmp_path / "c.bin", allp.numpy())
    tree = kd.KDTree(allp[:n])
    ref_ids = oracle.inorder_ids(src, dim, n, tmp_path)
    ours = tree.ids_inorder().numpy()
    if not np.array_equal(ref_ids, ours):
        # std::sort is unstable: the reference may order equal keys differently
        assert tree.check() == 0
    res = tree.nearest(allp[n:])
    ref = oracle.queryx(src, dim, n, nq)
    assert [r[3] for r in ref] == res.dist.numpy().view(np.uint32).tolist()


def test_materialize_nodes_cpu(kd):
    pts, _ = kd.ops.generate_problem(3, 2, 7, 0)
    tree = kd.KDTree(pts)
    node_point, left, right = tree.nodes()
    root = tree.root_slot
    assert root == 3 and left[root] == 1 and right[root] == 5
    assert left[1] == 0 and right[1] == 2 and left[0] == -1 and right[6] == -1
    assert torch.equal(node_point, tree.perm)
    # preorder walk over the arrays reproduces the printed tree order (IDs 2 4 3 1 5 6 7)
    order, stack = [], [root]
    while stack:
        s = stack.pop()
        if s < 0:
            continue
        order.append(int(node_point[s]) + 1)
        stack += [int(right[s]), int(left[s])]
    assert order == [2, 4, 3, 1, 5, 6, 7]


def test_forest_equals_reference_mpi_semantics(kd, oracle):
    for seed, dim, n, world in [(1, 3, 1000, 4), (2, 2, 1003, 8), (3, 3, 20000, 2), (4, 5, 777, 3)]:
        forest, qs = kd.models.Forest.from_seed(seed, dim, n, world)
        dist, _ = forest.nearest(qs)
        lines = "".join(kd.utils.result_line(n + q, float(dist[q])) + "\n" for q in range(10))
        assert lines == oracle.forest_lines(seed, dim, n, 10, world)


def test_forest_rejects_empty_rank(kd):
    with pytest.raises(ValueError):
        kd.models.slice_bounds(3, 8, 0)


def test_state_dict_roundtrip(kd):
    pts, qs = kd.ops.generate_problem(2, 3, 500, 10)
    tree = kd.KDTree(pts)
    again = kd.KDTree.from_state_dict(tree.state_dict())
    assert torch.equal(again.perm, tree.perm)
    import os, tempfile
    with tempfile.TemporaryDirectory() as tmp:
        tree.save(os.path.join(tmp, "tree.pt"))
        loaded = kd.KDTree.load(os.path.join(tmp, "tree.pt"))
    assert torch.equal(loaded.perm, tree.perm) and torch.equal(loaded.tree_points, tree.tree_points)
    assert torch.equal(again.nearest(qs).dist, tree.nearest(qs).dist)


def test_knn_extension_cpu(kd):
    # k = 1 is the reference's 1-NN; for k > 1 every list is ascending and, in 1-D (where the reference's
    # trees are valid kd-trees, SURVEY.md §0: 0 % wrong answers), equals brute force
    pts, qs = kd.ops.generate_problem(3, 3, 5000, 40)
    tree = kd.KDTree(pts)
    ids1, d1 = tree.knn(qs, 1)
    ref = tree.nearest(qs)
    assert torch.equal(ids1[:, 0], ref.ids) and torch.equal(d1[:, 0], ref.dist)
    ids, d = tree.knn(qs, 8)
    assert bool((d[:, 1:] >= d[:, :-1]).all())
    pts, qs = kd.ops.generate_problem(5, 1, 3000, 25)
    tree = kd.KDTree(pts)
    ids, d = tree.knn(qs, 5)
    brute = torch.cdist(qs.double(), pts.double()).topk(5, largest=False)
    assert torch.equal(ids.long() - 1, brute.indices)
    ids, d = kd.KDTree(pts[:3]).knn(qs, 5)
    assert (ids[:, 3:] == -1).all() and torch.isinf(d[:, 3:]).all()


def test_generator_arithmetic_matches_libstdcxx_objects(kd):
    for seed in (1, 7, 12345):
        assert kd.lib().kd_generator_selftest(seed, 3_000_000) == 0
