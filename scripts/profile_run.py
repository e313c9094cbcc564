"""One short search for ncu: python scripts/profile_run.py pfsp 29 2 [M]  |  nq 15 [M]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
bb = load_package()
kind = sys.argv[1]
if kind == "pfsp":
    inst, lb = int(sys.argv[2]), int(sys.argv[3])
    M = int(sys.argv[4]) if len(sys.argv) > 4 else 1 << 20
    res, per = bb.pfsp_search(bb.Instance(inst=inst), lb, M=M)
else:
    N = int(sys.argv[2])
    M = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 20
    res, per = bb.nq_search(N, 1, M=M)
print(res.tree, res.sol, res.best, "kernel_ms", res.kernel_ms_max, "iters", per[0].iterations)
