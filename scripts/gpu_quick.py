"""quick timing + parity of the headline searches: python scripts/gpu_quick.py [M]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
bb = load_package()
M = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
GOLD = {("pfsp", 14, 1): (2573652, 2648), ("pfsp", 14, 0): (2573652, 2648), ("pfsp", 30, 0): (111887480, 82), ("pfsp", 22, 0): (711291688, 4346), ("pfsp", 29, 2): (9499307, 0),
        ("pfsp", 21, 2): (3944527267, 0), ("pfsp", 24, 2): (2173092255, 0), ("pfsp", 22, 1): (711291688, 4346),
        ("nq", 15): (171129071, 2279184), ("nq", 16): (1141190302, 14772512), ("nq", 17): (8017021931, 95815104)}
for case in [("pfsp", 14, 1), ("pfsp", 30, 0), ("pfsp", 22, 0), ("pfsp", 22, 1), ("pfsp", 29, 2), ("pfsp", 24, 2), ("pfsp", 21, 2), ("nq", 15), ("nq", 16), ("nq", 17)]:
    t = time.time()
    if case[0] == "pfsp":
        res, per = bb.pfsp_search(bb.Instance(inst=case[1]), case[2], M=M)
    else:
        res, per = bb.nq_search(case[1], 1, M=M)
    dt = time.time() - t
    p = per[0]
    ok = "OK" if GOLD[case] == (res.tree, res.sol) else "MISMATCH %s" % (GOLD[case],)
    print(case, (res.tree, res.sol), ok, "wall %.3fs kernel %.2fms iters %d maxpool %d  %.2f Gnodes/s" % (
        dt, res.kernel_ms_max, p.iterations, p.max_pool, res.tree / max(res.kernel_ms_max, 1e-6) / 1e6), flush=True)
