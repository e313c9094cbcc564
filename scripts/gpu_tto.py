"""time-to-optimum of searches that start WITHOUT an upper bound (-u 0, PFSP_lib.c:234-241): one B200, whole search
from ub = INT_MAX to proven optimum.  usage: python scripts/gpu_tto.py <inst> <lb> [M]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
bb = load_package()
inst, lb = int(sys.argv[1]), int(sys.argv[2])
M = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 18
ins = bb.Instance(inst=inst)
bb.pfsp_search(bb.Instance(inst=14), 2)  # context, module load
t = time.perf_counter()
res, per = bb.pfsp_search(ins, lb, best0=bb.INT_MAX, M=M)
dt = time.perf_counter() - t
print(f"ta{inst:03d} lb={lb} -u 0 M={M}: optimum {res.best} (best known {ins.best_known}) tree {res.tree} sol {res.sol} "
      f"search {res.search_s:.4f} s call {dt:.4f} s")
