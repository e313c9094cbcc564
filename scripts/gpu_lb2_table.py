import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
bb = load_package()
GOLD = {(1, 2): (0, 0), (2, 2): (7, 0), (3, 2): (80062, 0), (14, 2): (144639, 0), (29, 2): (9499307, 0), (30, 2): (13228600, 0), (22, 2): (14561974, 0),
        (17, 2): (17472239, 0), (21, 2): (3944527267, 0), (24, 2): (2173092255, 0)}
def run(inst, M, **kw):
    ins = bb.Instance(inst=inst)
    t = time.time(); res, per = bb.pfsp_search(ins, 2, M=M, **kw); dt = time.time() - t
    p = per[0]
    ok = "OK" if GOLD.get((inst, 2)) in (None, (res.tree, res.sol)) else "MISMATCH %s" % (GOLD[(inst, 2)],)
    print("ta%03d M=%d %s" % (inst, M, kw), (res.tree, res.sol, res.best), ok,
          "wall %.3fs kernel %.2fms iters %d lb2e %d  %.1f Mnodes/s" % (
              dt, res.kernel_ms_max, p.iterations, p.lb2_evals, res.tree / max(res.kernel_ms_max, 1e-6) / 1e3), flush=True)
for inst in (1, 2, 3, 14, 29, 30, 22, 17):
    run(inst, 1 << 20, lb2_table=1)
run(29, 1 << 20, lb2_table=1, lb2_prefilter=0)
run(29, 50000, lb2_table=1)
run(29, 1 << 20, lb2_table=0)
run(24, 1 << 20, lb2_table=1)
run(21, 1 << 20, lb2_table=1)
run(21, 1 << 22, lb2_table=1)
