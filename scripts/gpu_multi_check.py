"""Multi-GPU parity/timing check. In-process threads:  python scripts/gpu_multi_check.py threads D
   One process per GPU:  torchrun --nproc-per-node D scripts/gpu_multi_check.py dist"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
bb = load_package()
GOLD = {("pfsp", 14, 0): (2573652, 2648), ("pfsp", 29, 2): (9499307, 0), ("pfsp", 30, 0): (111887480, 82),
        ("pfsp", 22, 0): (711291688, 4346), ("pfsp", 21, 2): (3944527267, 0), ("pfsp", 24, 2): (2173092255, 0),
        ("nq", 15): (171129071, 2279184), ("nq", 16): (1141190302, 14772512), ("nq", 17): (8017021931, 95815104)}
mode = sys.argv[1]
cases = [("pfsp", 14, 0), ("pfsp", 29, 2), ("pfsp", 30, 0), ("nq", 15), ("nq", 16), ("pfsp", 22, 0)]
if len(sys.argv) > 3 and sys.argv[3] == "big":
    cases = [("pfsp", 24, 2), ("pfsp", 21, 2), ("nq", 17)]
def report(tag, case, res, dt):
    ok = "OK" if GOLD[case] == (res.tree, res.sol) else "MISMATCH %s" % (GOLD[case],)
    print(tag, case, (res.tree, res.sol, res.best), ok, "wall %.3fs search %.3fs kernel_max %.1fms steals %d/%d moved %d  %.1f Mnodes/s" % (
        dt, res.search_s, res.kernel_ms_max, res.steals_ok, res.steals, res.nodes_moved, res.tree / max(res.search_s, 1e-9) / 1e6), flush=True)
if mode == "threads":
    D = int(sys.argv[2])
    for case in cases:
        for d in sorted({1, D}):
            t = time.time()
            if case[0] == "pfsp":
                res, per = bb.pfsp_search(bb.Instance(inst=case[1]), case[2], D=d, M=1 << 20)
            else:
                res, per = bb.nq_search(case[1], 1, D=d, M=1 << 20)
            report("threads D=%d" % d, case, res, time.time() - t)
else:
    import torch, torch.distributed as dist
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl")
    for i, case in enumerate(cases):
        name = [f"/bbdist_{os.getpid()}_{i}"]
        dist.broadcast_object_list(name, src=0)
        dist.barrier()
        t = time.time()
        if case[0] == "pfsp":
            res, mine = bb.pfsp_search_dist(bb.Instance(inst=case[1]), case[2], rank, world, name[0], device=lr, M=1 << 20)
        else:
            res, mine = bb.nq_search_dist(case[1], 1, rank, world, name[0], device=lr, M=1 << 20)
        dt = time.time() - t
        print("   rank %d own tree %d kernel %.1f ms" % (rank, mine.tree, mine.kernel_ms), flush=True)
        if rank == 0:
            report("dist world=%d" % world, case, res, dt)
    dist.destroy_process_group()
