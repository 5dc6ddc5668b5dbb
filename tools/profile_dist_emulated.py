"""Launch list of the distributed kernels on ONE GPU: W virtual ranks (EmulatedDistBuild) of a 2^log2n build, for
`ncu --metrics gpu__time_duration.sum --profile-from-start off`.  python tools/profile_dist_emulated.py [log2n] [world] [dim]"""
import sys

sys.path.insert(0, ".")
import torch

import parallel_kd_tree_b200 as kd
from parallel_kd_tree_b200.parallel.dist_build import EmulatedDistBuild

log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 26
world = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dim = int(sys.argv[3]) if len(sys.argv) > 3 else 3
n = 1 << log2n
pts = kd.ops.generate_uniform(1, dim, n).cuda()
eb = EmulatedDistBuild(n, dim, world)
eb.build(pts)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
eb.build(pts)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("ok")
