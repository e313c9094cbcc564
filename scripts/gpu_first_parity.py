import os
import sys, time, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),'tests'))
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
bb=load_package()
import orc
for inst,lb,exp in [(14,0,(2573652,2648,1377)),(14,1,(2573652,2648,1377)),(14,2,(144639,0,1377)),(2,0,(30,0,1359)),(3,0,(2573133,5689,1081)),(1,2,(0,0,1278))]:
    ins=bb.Instance(inst=inst)
    t=time.time(); res,per=bb.pfsp_search(ins,lb); dt=time.time()-t
    print("ta%03d lb%d"%(inst,lb),(res.tree,res.sol,res.best),"exp",exp,"OK" if (res.tree,res.sol,res.best)==exp else "MISMATCH","%.3fs kernel %.2fms iters %d maxpool %d"%(dt,res.kernel_ms_max,per[0].iterations,per[0].max_pool),flush=True)
for N,exp in [(8,(2056,92)),(12,(856188,14200)),(14,(27358552,365596)),(16,(1141190302,14772512))]:
    t=time.time(); res,per=bb.nq_search(N,1,M=200000 if N>12 else 50000); dt=time.time()-t
    print("nq",N,(res.tree,res.sol),"exp",exp,"OK" if (res.tree,res.sol)==exp else "MISMATCH","%.3fs kernel %.2fms iters %d maxpool %d"%(dt,res.kernel_ms_max,per[0].iterations,per[0].max_pool),flush=True)
