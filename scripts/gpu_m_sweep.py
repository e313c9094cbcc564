import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
bb = load_package()
for M in (1 << 20, 1 << 22, 1 << 23, 1 << 24):
    res, per = bb.nq_search(17, 1, M=M)
    p = per[0]
    print("nq17 M=%d" % M, res.tree == 8017021931, "kernel %.1f ms iters %d maxpool %d wall %.3f" % (res.kernel_ms_max, p.iterations, p.max_pool, res.elapsed_s), flush=True)
for M in (1 << 21, 1 << 22):
    res, per = bb.pfsp_search(bb.Instance(inst=21), 2, M=M, capacity=100 << 20)
    p = per[0]
    print("ta021 M=%d" % M, res.tree == 3944527267, "kernel %.1f ms iters %d maxpool %d wall %.3f setup %.3f" % (res.kernel_ms_max, p.iterations, p.max_pool, res.elapsed_s, res.setup_s), flush=True)
