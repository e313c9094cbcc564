"""bb_pfsp_expand_static - the allocation-free expand step - against the engine entry points, through the C ABI.

Why the entry point exists: some callers of libbbgpu.so may not allocate device memory while they work - host
threads sharing one GPU (a cudaMalloc by one stalls all of them), hosts whose own pool already owns the HBM,
latency-critical loops. bb_pfsp_expand_static serves them from a fixed __device__ array inside the library, so it
has a size limit and keeps state between calls. This test therefore does what such a caller would do: all device
buffers are created up front, then the entry point is called twice in a row on different batches (the second one
smaller, so anything the first call left behind would show) and each result is compared with what
bb_engine_push_nodes + bb_engine_run(1) + bb_engine_export_nodes give for the same batch, and with what the CPU
oracle's evaluate + generate_children give for it: the same set of children and the same counters, bit for bit
(integer work, no tolerance).

The file name sorts last on purpose: with `pytest -x` every other test has already run if this one fails.
"""
import ctypes as C

import numpy as np
import pytest

import orc
from __graft_entry__ import load_package

pytestmark = pytest.mark.gpu
bb = load_package()


def node_set(a, width):
    return sorted(tuple(int(v) for v in r[:width]) for r in a)


def engine_step(ins, lb, parents, best):
    eng = bb.Engine.pfsp(ins, lb, M=100000)
    eng.reset(best)
    eng.push_nodes(parents)
    st = eng.run(1)
    kids = eng.export_nodes()
    out = (node_set(kids, 2 + ins.n), st.pool_size, st.tree, st.sol, st.best)
    eng.close()
    return out


@pytest.mark.parametrize("inst,lb", [(14, 0), (21, 1), (1, 0)])
def test_static_step_matches_engine_on_two_calls_in_a_row(inst, lb):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    torch.cuda.set_device(0)
    ins = bb.Instance(inst=inst)
    n = ins.n
    rng = np.random.default_rng(7 * inst + lb)
    limit = bb.STATIC_MAX_PARENTS
    batches = []
    # first batch: the full limit, inner nodes only; second: smaller, with parents whose children are leaves
    # (incumbent = optimum there, so no leaf can move it and pruning stays deterministic)
    for count, best, max_depth in ((limit, ins.best_known + 40, n - 2), (limit // 3 + 5, ins.best_known, n - 1)):
        parents = orc.random_pfsp_nodes(rng, n, count, 0, max_depth)
        batches.append((parents, best))
    # everything the caller owns is allocated before the first call; the calls themselves allocate nothing
    p_d = [torch.from_numpy(np.ascontiguousarray(p).reshape(-1)).cuda() for p, _ in batches]
    k_d = [torch.zeros(len(p) * n * bb.NODE_I16, dtype=torch.int16, device="cuda") for p, _ in batches]
    r_d = [torch.zeros(4, dtype=torch.int64, device="cuda") for _ in batches]
    torch.cuda.synchronize()
    for i, (parents, best) in enumerate(batches):  # two calls in a row, no synchronisation in between
        rc = bb.lib.bb_pfsp_expand_static(ins.ptr, lb, best, p_d[i].data_ptr(), len(parents), k_d[i].data_ptr(),
                                          len(parents) * n, r_d[i].data_ptr(), None)
        assert rc == 0, bb.last_error()
    torch.cuda.synchronize()
    L = orc.oracle()
    h = L.orc_pfsp_new(inst)
    for i, (parents, best) in enumerate(batches):
        produced, tree, sol, bst = (int(v) for v in r_d[i].cpu().numpy())
        kids = k_d[i].cpu().numpy().reshape(-1, bb.NODE_I16)[:produced]
        exp_kids, exp_count, exp_tree, exp_sol, exp_best = engine_step(ins, lb, parents, best)
        assert (produced, tree, sol, bst) == (exp_count, exp_tree, exp_sol, exp_best), (inst, lb, i)
        assert produced == tree  # every surviving child is one exploredTree increment (PFSP_lib.h:78-86)
        assert node_set(kids, 2 + n) == exp_kids, (inst, lb, i)
        # and directly against the CPU oracle (the reference's bounds + generate_children restated)
        o_kids, o_tree, o_sol, o_best, _ = orc.oracle_one_level(h, n, lb, parents, best)
        assert (produced, tree, sol, bst) == (len(o_kids), o_tree, o_sol, o_best), (inst, lb, i)
        assert node_set(kids, 2 + n) == node_set(o_kids, 2 + n), (inst, lb, i)
    L.orc_pfsp_free(h)
    ins.close()


def test_static_step_rejects_what_it_cannot_hold():
    import torch
    torch.cuda.set_device(0)
    ins = bb.Instance(inst=14)
    n = ins.n
    buf = torch.zeros(8 * bb.NODE_I16, dtype=torch.int16, device="cuda")
    res = torch.zeros(4, dtype=torch.int64, device="cuda")
    too_many = bb.STATIC_MAX_PARENTS + 1
    rc = bb.lib.bb_pfsp_expand_static(ins.ptr, 0, ins.best_known, buf.data_ptr(), too_many, buf.data_ptr(), 8,
                                      res.data_ptr(), None)
    assert rc == -1 and "limit" in bb.last_error(), bb.last_error()  # BB_ERR_ARG, nothing was launched
    rc = bb.lib.bb_pfsp_expand_static(ins.ptr, 2, ins.best_known, buf.data_ptr(), 1, buf.data_ptr(), 8,
                                      res.data_ptr(), None)
    assert rc == -1 and "not supported" in bb.last_error(), bb.last_error()
    # a parent that is not a node of this instance (depth out of range, job id out of range) can only be seen on
    # the device: the whole batch is refused - no children, all-ones in the two counters
    good = orc.random_pfsp_nodes(np.random.default_rng(3), n, 8, 0, n - 2)
    for field, value in ((0, n), (0, -1), (2 + 5, n), (2 + 5, -3)):
        bad = good.copy()
        bad[3, field] = value
        b_d = torch.from_numpy(np.ascontiguousarray(bad).reshape(-1)).cuda()
        k_d = torch.zeros(8 * n * bb.NODE_I16, dtype=torch.int16, device="cuda")
        rc = bb.lib.bb_pfsp_expand_static(ins.ptr, 0, ins.best_known + 100, b_d.data_ptr(), 8, k_d.data_ptr(), 8 * n,
                                          res.data_ptr(), None)
        assert rc == 0, bb.last_error()
        torch.cuda.synchronize()
        got = [int(v) for v in res.cpu().numpy()]
        assert got[0] == 0 and got[1] == -1 and got[2] == -1, (field, value, got)
    # ... and the next well-formed batch is served as usual (the error flag is cleared per call)
    b_d = torch.from_numpy(np.ascontiguousarray(good).reshape(-1)).cuda()
    k_d = torch.zeros(8 * n * bb.NODE_I16, dtype=torch.int16, device="cuda")
    rc = bb.lib.bb_pfsp_expand_static(ins.ptr, 0, ins.best_known + 100, b_d.data_ptr(), 8, k_d.data_ptr(), 8 * n,
                                      res.data_ptr(), None)
    assert rc == 0, bb.last_error()
    torch.cuda.synchronize()
    L = orc.oracle()
    h = L.orc_pfsp_new(14)
    o_kids, o_tree, o_sol, o_best, _ = orc.oracle_one_level(h, n, 0, good, ins.best_known + 100)
    L.orc_pfsp_free(h)
    got = [int(v) for v in res.cpu().numpy()]
    assert got == [len(o_kids), o_tree, o_sol, o_best]
    # an empty batch is fine and reports zeros (and the incumbent unchanged)
    rc = bb.lib.bb_pfsp_expand_static(ins.ptr, 0, 1234, None, 0, buf.data_ptr(), 8, res.data_ptr(), None)
    assert rc == 0, bb.last_error()
    torch.cuda.synchronize()
    assert [int(v) for v in res.cpu().numpy()] == [0, 0, 0, 1234]
    ins.close()
