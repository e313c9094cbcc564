import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
bb = load_package()
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
for i, (M, ipl) in enumerate([(1 << 21, 32), (1 << 21, 32), (1 << 21, 4), (1 << 21, 64)]):
    name = f"/bbnq_{os.getppid()}_{i}"
    t = time.time()
    res, mine = bb.nq_search_dist(17, 1, rank, world, name, device=lr, M=M, iters_per_launch=ipl)
    dt = time.time() - t
    print("step %d M=%d ipl=%d rank %d: own tree %d kernel %.1f ms iters %d steals %d/%d | search %.1f ms total %.1f ms" % (
        i, M, ipl, rank, mine.tree, mine.kernel_ms, mine.iterations, mine.steals_ok, mine.steals, res.search_s * 1e3, dt * 1e3), flush=True)
