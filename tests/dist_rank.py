"""One rank of a process-per-worker team for tests/test_gpu_engine.py: python dist_rank.py <rank> <world> <shm> <device>.
Opens the persistent contexts (bb_*_dist_open), runs the searches every rank must make in the same order and prints one
JSON line with what rank 0 got back."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

bb = load_package()
INT_MAX = 2**31 - 1
rank, world, shm, device = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], int(sys.argv[4])
out = {}
d = bb.Dist(rank, world, shm + "_p", instance=bb.Instance(inst=30), lb=0, M=1 << 16, device=device)
for k in range(2):
    res, mine = d.search()
    out[f"ta030_lb1d_{k}"] = [res.tree, res.sol, res.best, int(mine.tree > 0)]
res, mine = d.search(ws=0)
out["ta030_lb1d_static"] = [res.tree, res.sol, res.best, int(mine.tree > 0)]
d.close()
d = bb.Dist(rank, world, shm + "_u", instance=bb.Instance(inst=14), lb=2, M=1 << 14, device=device)
res, _ = d.search(best0=INT_MAX)
out["ta014_lb2_u0_best"] = res.best
res, _ = d.search()
out["ta014_lb2"] = [res.tree, res.sol, res.best]
d.close()
d = bb.Dist(rank, world, shm + "_q", N=14, G=1, M=1 << 16, device=device)
res, mine = d.search()
out["nq14"] = [res.tree, res.sol, int(mine.tree > 0)]
d.close()
print(json.dumps(out))
