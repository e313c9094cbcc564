import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
bb = load_package()
import torch, torch.distributed as dist
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
ins = bb.Instance(inst=21)
for i, kw in enumerate([dict(), dict(), dict(capacity=32 << 20), dict(capacity=32 << 20)]):
    name = [f"/bbph_{os.getpid()}_{i}"]
    dist.broadcast_object_list(name, src=0)
    dist.barrier()
    if rank == 0: print("---- step", i, kw, flush=True)
    t = time.time()
    res, mine = bb.pfsp_search_dist(ins, 2, rank, world, name[0], device=lr, M=1 << 20, **kw)
    if rank == 0: print("total %.1f ms search %.1f setup %.1f warm %.1f tree %d" % ((time.time()-t)*1e3, res.search_s*1e3, res.setup_s*1e3, res.warm_s*1e3, res.tree), flush=True)
dist.destroy_process_group()
