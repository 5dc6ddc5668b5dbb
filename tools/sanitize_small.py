"""Small end-to-end run for `compute-sanitizer --tool memcheck`: every kernel family once on small problems."""
import sys
sys.path.insert(0, ".")
import torch
import parallel_kd_tree_b200 as kd
from parallel_kd_tree_b200.parallel.dist_build import EmulatedDistBuild

ok = True
for n, dim in [(70001, 3), (40000, 8), (20000, 16), (3000, 2)]:
    pts, qs = kd.ops.generate_problem(3, dim, n, 64)
    ref = kd.KDTree(pts)
    t = kd.KDTree(pts.cuda())
    ok &= torch.equal(t.perm.cpu(), ref.perm)
    ok &= torch.equal(t.nearest(qs.cuda()).index.cpu(), ref.nearest(qs).index)
    ok &= torch.equal(t.knn(qs.cuda(), 4)[0].cpu(), ref.knn(qs, 4)[0])
    t.nodes(); t.bbox()
pts = kd.ops.generate_uniform(5, 3, 150000)
outs = EmulatedDistBuild(150000, 3, 4).build(pts)
ok &= torch.equal(outs[0][0].cpu(), kd.ops.build(pts)[0])
torch.cuda.synchronize()
print("sanitize_small:", "OK" if ok else "MISMATCH")
