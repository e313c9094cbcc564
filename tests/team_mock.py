"""Host stand-in for a GPU engine, used ONLY by tests/test_multi_protocol.py to drive the product's team protocol
(bb_team_worker in libbbgpu.so: work stealing, termination detection) with world_size-2 gloo processes on a CPU box.

The stand-in solves N-Queens exactly as nqueens_c.c:99-117 does (explored tree = every safe placement, solutions =
depth-N nodes) on a pool that lives in a multiprocessing SharedMemory block, so that a peer process can copy donated
nodes straight out of it - the CPU analogue of the NVLink peer copy the GPU engine performs.
"""
import ctypes as C
import os
import sys
from multiprocessing import shared_memory

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))

CAP = 1 << 16


class MockEngine:
    def __init__(self, bb, N, rank, world, tag, batch=48):
        self.bb, self.N, self.rank, self.world, self.tag, self.batch = bb, N, rank, world, tag, batch
        self.shm = shared_memory.SharedMemory(name=f"{tag}_pool{rank}", create=True, size=CAP * (N + 1))
        self.pool = np.ndarray((CAP, N + 1), dtype=np.uint8, buffer=self.shm.buf)
        self.bot = self.top = self.hold = 0
        self.tree = self.sol = 0
        self.pulled = 0
        self.peers = {}
        self.ops = bb.TeamOps()
        self.ops.self = None
        self._keep = (bb.TeamOps.RUN(self.run), bb.TeamOps.DONATE(self.donate), bb.TeamOps.DONE(self.donate_done),
                      bb.TeamOps.PULL(self.pull), bb.TeamOps.SETBEST(lambda s, b: 0))
        (self.ops.run, self.ops.donate, self.ops.donate_done, self.ops.pull, self.ops.set_best) = self._keep

    def push(self, node):
        self.pool[self.top] = node
        self.top += 1

    def expand(self, node):
        N, depth = self.N, int(node[0])
        if depth == N:
            self.sol += 1
            return
        board = node[1:]
        for j in range(depth, N):
            cand = int(board[j])
            if all(int(board[i]) != cand - (depth - i) and int(board[i]) != cand + (depth - i) for i in range(depth)):
                child = node.copy()
                child[1 + depth], child[1 + j] = board[j], board[depth]
                child[0] = depth + 1
                self.push(child)
                self.tree += 1

    # --- bb_team_ops ---
    def run(self, _self, max_iters, pool_size, best, busy_ms):
        for _ in range(max_iters):
            if self.top == self.bot:
                break
            n = min(self.batch, self.top - self.bot)
            batch = self.pool[self.top - n:self.top].copy()
            self.top -= n
            for node in batch:
                self.expand(node)
        if self.top == self.bot and self.hold == 0:
            self.top = self.bot = 0
        pool_size[0] = self.top - self.bot
        busy_ms[0] = 0.0
        return 0

    def donate(self, _self, min_keep, max_count, stack, start, count):
        avail = self.top - self.bot
        count[0] = 0
        if avail < 2 * min_keep:
            return 0
        give = min(avail // 2, max_count)
        stack[0], start[0], count[0] = 0, self.bot, give
        self.bot += give
        self.hold += 1
        return 0

    def donate_done(self, _self, stack):
        self.hold -= 1
        if self.top == self.bot and self.hold == 0:
            self.top = self.bot = 0
        return 0

    def pull(self, _self, victim, stack, start, count):
        if victim not in self.peers:
            shm = shared_memory.SharedMemory(name=f"{self.tag}_pool{victim}")
            self.peers[victim] = (shm, np.ndarray((CAP, self.N + 1), dtype=np.uint8, buffer=shm.buf))
        src = self.peers[victim][1]
        self.pool[self.top:self.top + count] = src[start:start + count]
        self.top += count
        self.pulled += count
        return 0

    def close(self):
        for shm, _ in self.peers.values():
            shm.close()
        self.peers = {}
        self.pool = None
        self.shm.close()
        self.shm.unlink()


def worker_main(rank, world, port, N, tag, lopsided, out_q):
    """one rank of the gloo job"""
    import torch
    import torch.distributed as dist
    from __graft_entry__ import load_package
    bb = load_package()
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    nbytes = bb.lib.bb_team_bytes(world)
    if rank == 0:
        team = shared_memory.SharedMemory(name=f"{tag}_team", create=True, size=nbytes)
        team.buf[:nbytes] = bytes(nbytes)
    dist.barrier()
    if rank != 0:
        team = shared_memory.SharedMemory(name=f"{tag}_team")
    mem = C.addressof(C.c_char.from_buffer(team.buf))
    if rank == 0:
        assert bb.lib.bb_team_init(mem, world) == 0
    else:
        assert bb.lib.bb_team_wait_ready(mem, world, 30.0) == 0
    eng = MockEngine(bb, N, rank, world, tag)
    # warm-up: every rank expands the root two levels; the share is by content hash, or everything to rank 0
    root = np.zeros(N + 1, np.uint8)
    root[1:] = np.arange(N)
    eng.push(root)
    for _ in range(2):
        level = eng.pool[eng.bot:eng.top].copy()
        eng.top = eng.bot
        for node in level:
            eng.expand(node)
    warm_tree = eng.tree
    nodes = eng.pool[:eng.top].copy()
    eng.top = eng.bot = 0
    eng.tree = warm_tree if rank == 0 else 0
    for nd in nodes:
        owner = 0 if lopsided else (int(nd[1]) * 31 + int(nd[2])) % world
        if owner == rank:
            eng.push(nd)
    assert bb.lib.bb_team_barrier(mem, 30.0) == 0
    res = bb.TeamResult()
    rc = bb.lib.bb_team_worker(mem, rank, C.byref(eng.ops), 1, 4, 1 << 20, bb.INT_MAX, 4, C.byref(res))
    assert rc == 0, bb.last_error()
    assert eng.top == eng.bot, "worker left nodes behind"
    t = torch.tensor([eng.tree, eng.sol, res.steals_ok, eng.pulled], dtype=torch.int64)
    mine = t.clone()
    dist.all_reduce(t)
    assert bb.lib.bb_team_barrier(mem, 30.0) == 0
    out_q.put((rank, [int(v) for v in mine], [int(v) for v in t]))
    dist.barrier()
    eng.close()
    del mem
    team.close()
    if rank == 0:
        team.unlink()
    dist.destroy_process_group()
