import sys, time
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
bb=load_package()
def run_pfsp(inst,lb,M,**kw):
    ins=bb.Instance(inst=inst)
    t=time.time(); res,per=bb.pfsp_search(ins,lb,M=M,**kw); dt=time.time()-t
    p=per[0]
    print("ta%03d lb%d M=%d"%(inst,lb,M),(res.tree,res.sol,res.best),
          "wall %.3fs kernel %.2fms iters %d maxpool %d evals %d lb2e %d pairs/ev %.1f  %.3f Gnodes/s"%(dt,res.kernel_ms_max,p.iterations,p.max_pool,p.evals,p.lb2_evals,(p.lb2_pairs/max(p.lb2_evals,1)),res.tree/res.kernel_ms_max/1e6),flush=True)
run_pfsp(14,1,50000)
run_pfsp(24,2,1000000)
run_pfsp(21,2,1000000)
run_pfsp(24,0,1000000)
run_pfsp(21,0,1000000)
