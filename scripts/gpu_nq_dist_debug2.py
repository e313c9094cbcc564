import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
bb = load_package()
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
for i, (M, ipl) in enumerate([(1 << 21, 32), (1 << 21, 32)]):
    name = f"/bbnq2_{os.environ.get('MASTER_PORT','0')}_{i}"
    res, mine = bb.nq_search_dist(17, 1, rank, world, name, device=lr, M=M, iters_per_launch=ipl)
    print("step %d rank %d: kernel %.1f ms iters %d | search %.1f ms" % (i, rank, mine.kernel_ms, mine.iterations, res.search_s * 1e3), flush=True)
    time.sleep(0.5)
