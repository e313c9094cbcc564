import os
import sys, time, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),'tests'))
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
bb=load_package()
def run_pfsp(inst,lb,M,exp=None,**kw):
    ins=bb.Instance(inst=inst)
    t=time.time(); res,per=bb.pfsp_search(ins,lb,M=M,**kw); dt=time.time()-t
    p=per[0]
    print("ta%03d lb%d M=%d"%(inst,lb,M),(res.tree,res.sol,res.best),("OK" if exp is None or (res.tree,res.sol)==exp else "MISMATCH %s"%(exp,)),
          "wall %.3fs kernel %.2fms iters %d maxpool %d evals %d lb2e %d pairs/ev %.1f  %.2f Gnodes/s"%(dt,res.kernel_ms_max,p.iterations,p.max_pool,p.evals,p.lb2_evals,(p.lb2_pairs/max(p.lb2_evals,1)),res.tree/res.kernel_ms_max/1e6),flush=True)
def run_nq(N,M,exp=None):
    t=time.time(); res,per=bb.nq_search(N,1,M=M); dt=time.time()-t; p=per[0]
    print("nq %d M=%d"%(N,M),(res.tree,res.sol),("OK" if exp is None or (res.tree,res.sol)==exp else "MISMATCH"),
          "wall %.3fs kernel %.2fms iters %d maxpool %d %.2f Gnodes/s"%(dt,res.kernel_ms_max,p.iterations,p.max_pool,res.tree/res.kernel_ms_max/1e6),flush=True)
run_pfsp(14,0,50000)
for M in (50000,200000,1000000):
    run_pfsp(14,1,M,(2573652,2648))
    run_pfsp(14,2,M,(144639,0))
    run_pfsp(30,0,M,(111887480,82))
    run_pfsp(29,2,M,(9499307,0))
run_pfsp(29,2,200000,(9499307,0),lb2_prefilter=0)
run_pfsp(17,2,200000,(17472239,0))
run_pfsp(22,0,1000000,(711291688,4346))
for M in (50000,200000,1000000,4000000):
    run_nq(15,M,(171129071,2279184))
run_nq(16,1000000,(1141190302,14772512))
run_nq(17,1000000,(8017021931,95815104))
run_nq(17,4000000,(8017021931,95815104))
