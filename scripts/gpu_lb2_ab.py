"""A/B of the LB2 kernels on the GPU box: counts must agree with the golden values, timings are printed."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
bb = load_package()
GOLD = {(14, 2): (144639, 0), (29, 2): (9499307, 0), (30, 2): (13228600, 0), (22, 2): (14561974, 0),
        (17, 2): (17472239, 0), (21, 2): (3944527267, 0), (24, 2): (2173092255, 0)}
def run(inst, M, **kw):
    ins = bb.Instance(inst=inst)
    t = time.time(); res, per = bb.pfsp_search(ins, 2, M=M, **kw); dt = time.time() - t
    p = per[0]
    ok = "OK" if GOLD.get((inst, 2)) in (None, (res.tree, res.sol)) else "MISMATCH %s" % (GOLD[(inst, 2)],)
    print("ta%03d M=%d %s" % (inst, M, kw), (res.tree, res.sol, res.best), ok,
          "wall %.3fs kernel %.2fms iters %d lb2e %d pairs/ev %.1f  %.1f Mnodes/s" % (
              dt, res.kernel_ms_max, p.iterations, p.lb2_evals, p.lb2_pairs / max(p.lb2_evals, 1),
              res.tree / res.kernel_ms_max / 1e3), flush=True)
which = sys.argv[1] if len(sys.argv) > 1 else "small"
if which == "small":
    for inst in (14, 29, 30, 22, 17):
        run(inst, 1 << 20, lb2_staged=0)
        for ka in (4, 8, 16):
            run(inst, 1 << 20, lb2_staged=1, lb2_stage_a=ka)
    run(29, 50000, lb2_staged=1)
    run(29, 1 << 20, lb2_staged=1, lb2_prefilter=0)
else:
    for ka in (8,):
        run(24, 1 << 20, lb2_staged=1, lb2_stage_a=ka)
        run(21, 1 << 20, lb2_staged=1, lb2_stage_a=ka)
