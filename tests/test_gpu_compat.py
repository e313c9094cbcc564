"""GPU parity of the drop-in boundary (evaluate_gpu / lb*_alloc_gpu / nq_evaluate_gpu, the reference's own
signatures) against the CPU oracle, bit-exact on bounds[] / labels[]. torch only provides the device buffers."""
import ctypes as C

import numpy as np
import pytest

import orc
from __graft_entry__ import load_package

pytestmark = pytest.mark.gpu
bb = load_package()
INT_MAX = 2**31 - 1


@pytest.fixture(scope="module")
def torch():
    import torch as t
    assert t.cuda.is_available(), "GPU tests need a CUDA device"
    t.cuda.set_device(0)
    return t


def device_tables(ins):
    """lb1_alloc_gpu / lb2_alloc_gpu on the host instance tables, exactly as pfsp_multigpu_cuda.c:188-194 does"""
    c = ins.ptr.contents
    h1 = bb.Lb1BoundData(c.p_times, c.min_heads, c.min_tails, c.nb_jobs, c.nb_machines)
    h2 = bb.Lb2BoundData(c.johnson_schedules, c.lags, c.machine_pairs_1, c.machine_pairs_2, c.machine_pair_order,
                         c.nb_machine_pairs, c.nb_jobs, c.nb_machines)
    d1, d2 = bb.Lb1BoundData(), bb.Lb2BoundData()
    bb.lib.lb1_alloc_gpu(C.byref(d1), C.byref(h1), None, None, None, c.nb_jobs, c.nb_machines)
    bb.lib.lb2_alloc_gpu(C.byref(d2), C.byref(h2), None, None, None, None, None, c.nb_jobs, c.nb_machines)
    assert bb.lib.bb_last_cuda_error() == 0, bb.last_error()
    return d1, d2


def gpu_chunk(torch, ins, d1, d2, lb, parents, best):
    n = ins.n
    count = len(parents)
    so, ni = np.zeros(max(count, 1), np.int32), np.zeros(max(count * n, 1), np.int32)
    nb = orc.oracle().orc_batch_index(parents, count, n, so, ni) if count else 0
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    p_d, so_d, ni_d = dev(parents.reshape(-1) if count else np.zeros(1, np.int16)), dev(so), dev(ni)
    b_d = torch.full((max(nb, 1),), -7, dtype=torch.int32, device="cuda")
    bst = C.c_int(best)
    bb.lib.evaluate_gpu(n, lb, nb, (nb + 511) // 512, count, C.byref(bst), d1, d2, p_d.data_ptr(), b_d.data_ptr(),
                        so_d.data_ptr(), ni_d.data_ptr())
    torch.cuda.synchronize()
    assert bb.lib.bb_last_cuda_error() == 0, bb.last_error()
    return b_d.cpu().numpy()[:nb], nb


@pytest.mark.parametrize("inst", [1, 14, 21, 24])
def test_evaluate_gpu_matches_oracle(torch, inst):
    L = orc.oracle()
    ins = bb.Instance(inst=inst)
    d1, d2 = device_tables(ins)
    h = L.orc_pfsp_new(inst)
    rng = np.random.default_rng(inst)
    n = ins.n
    parents = orc.random_pfsp_nodes(rng, n, 3000)
    parents[0, :2] = (0, -1)
    parents[0, 2:2 + n] = np.arange(n)                 # the root
    parents[1:40, 0] = n - 1                           # parents of leaves
    parents[1:40, 1] = n - 2
    ub = ins.best_known
    for lb in (0, 1, 2):
        for best in ((ub,) if lb < 2 else (ub, INT_MAX, ub - 60, 0)):
            got, nb = gpu_chunk(torch, ins, d1, d2, lb, parents, best)
            exp = np.zeros(nb, np.int32)
            L.orc_evaluate_chunk(h, lb, parents, len(parents), best, exp)
            assert np.array_equal(got, exp), (inst, lb, best, int((got != exp).sum()))
    L.orc_pfsp_free(h)


def test_evaluate_gpu_matches_golden_bounds(torch, golden):
    for inst_s, g in golden("bounds").items():
        ins = bb.Instance(inst=int(inst_s))
        d1, d2 = device_tables(ins)
        parents = np.array(g["nodes"], dtype=np.int16)
        flat = lambda rows: np.array([v for r in rows for v in r], dtype=np.int32)
        for lb, best, key in [(0, g["ub"], "lb1_d"), (1, g["ub"], "lb1"), (2, g["ub"], "lb2_ub"),
                              (2, INT_MAX, "lb2_inf"), (2, g["ub"] - 40, "lb2_tight")]:
            got, _ = gpu_chunk(torch, ins, d1, d2, lb, parents, best)
            assert np.array_equal(got, flat(g[key])), (inst_s, key)


def test_evaluate_gpu_edge_batches(torch):
    L = orc.oracle()
    ins = bb.Instance(inst=14)
    d1, d2 = device_tables(ins)
    h = L.orc_pfsp_new(14)
    rng = np.random.default_rng(0)
    for count in (0, 1, 2, 511, 513, 50000):  # empty, ragged around the reference's BLOCK_SIZE, the default -M
        parents = orc.random_pfsp_nodes(rng, 20, count) if count else np.zeros((0, orc.NODE_I16), np.int16)
        for lb in (0, 1, 2):
            got, nb = gpu_chunk(torch, ins, d1, d2, lb, parents, 1377)
            exp = np.zeros(max(nb, 1), np.int32)
            if count:
                L.orc_evaluate_chunk(h, lb, parents, count, 1377, exp)
            assert np.array_equal(got, exp[:nb]), (count, lb)
    L.orc_pfsp_free(h)


@pytest.mark.parametrize("N,G", [(1, 1), (8, 1), (14, 1), (17, 3), (20, 2)])
def test_nq_evaluate_gpu_matches_oracle(torch, N, G):
    L = orc.oracle()
    rng = np.random.default_rng(N)
    nodes = orc.random_nq_nodes(rng, N, 4000, 0, N)
    exp = np.full(len(nodes) * N, 9, np.uint8)
    L.orc_nq_labels(N, G, nodes, len(nodes), exp)
    p_d = torch.from_numpy(nodes.reshape(-1)).cuda()
    l_d = torch.full((len(nodes) * N,), 9, dtype=torch.uint8, device="cuda")
    bb.lib.nq_evaluate_gpu(N, G, p_d.data_ptr(), l_d.data_ptr(), len(nodes) * N)
    torch.cuda.synchronize()
    assert bb.lib.bb_last_cuda_error() == 0, bb.last_error()
    assert np.array_equal(l_d.cpu().numpy(), exp)  # slots below depth stay untouched (9) in both


def test_nq_evaluate_gpu_matches_golden(torch, golden):
    for Ns, g in golden("nq_nodes").items():
        N = int(Ns)
        nodes = np.array(g["nodes"], dtype=np.uint8)
        p_d = torch.from_numpy(nodes.reshape(-1)).cuda()
        l_d = torch.zeros(len(nodes) * N, dtype=torch.uint8, device="cuda")
        bb.lib.nq_evaluate_gpu(N, 1, p_d.data_ptr(), l_d.data_ptr(), len(nodes) * N)
        torch.cuda.synchronize()
        lab = l_d.cpu().numpy()
        for i, (nd, e) in enumerate(zip(nodes, g["labels"])):
            assert lab[i * N + int(nd[0]):(i + 1) * N].tolist() == e
