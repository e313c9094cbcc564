"""GPU parity of the native device-resident engine (bb_engine_* / bb_pfsp_search / bb_nq_search through the C ABI)
against the CPU oracle and the golden counts produced by the compiled reference."""
import ctypes as C

import numpy as np
import pytest

import orc
from __graft_entry__ import load_package

pytestmark = pytest.mark.gpu
bb = load_package()
INT_MAX = 2**31 - 1


def node_set(a, width):
    return sorted(tuple(int(v) for v in r[:width]) for r in a)


oracle_one_level = orc.oracle_one_level


@pytest.mark.parametrize("inst", [1, 14, 21])
@pytest.mark.parametrize("lb", [0, 1, 2])
def test_one_iteration_children_match_oracle(inst, lb):
    """bound + prune + branch of one batch == evaluate_gpu + generate_children of the reference (as a set:
    the device appends survivors in warp order)"""
    L = orc.oracle()
    ins = bb.Instance(inst=inst)
    n = ins.n
    h = L.orc_pfsp_new(inst)
    rng = np.random.default_rng(100 * inst + lb)
    eng = bb.Engine.pfsp(ins, lb, M=100000)
    for best in (ins.best_known, ins.best_known + 150, ins.best_known - 30):
        parents = orc.random_pfsp_nodes(rng, n, 5000, 0, n - 2)  # inner nodes only: `best` stays constant
        parents[0, :2] = (0, -1)
        parents[0, 2:2 + n] = np.arange(n)
        eng.reset(best)
        eng.push_nodes(parents)
        assert node_set(eng.export_nodes(), 2 + n) == node_set(parents, 2 + n)  # seed/export round trip
        st = eng.run(1)
        kids, tree, sol, b, nb = oracle_one_level(h, n, lb, parents, best)
        assert (st.tree, st.sol, st.best, st.parents, st.evals) == (tree, sol, b, len(parents), nb)
        assert st.pool_size == len(kids)
        assert node_set(eng.export_nodes(), 2 + n) == node_set(kids, 2 + n)
    eng.close()
    L.orc_pfsp_free(h)


def test_leaves_update_incumbent():
    """parents one job short of complete: every child is a leaf, counts a solution, and the best makespan wins
    (PFSP_lib.h:64-71); with ub = inf this is how -u 0 finds the optimum"""
    L = orc.oracle()
    ins = bb.Instance(inst=14)
    h = L.orc_pfsp_new(14)
    rng = np.random.default_rng(8)
    parents = orc.random_pfsp_nodes(rng, 20, 3000, 19, 19)
    exp_best = min(L.orc_makespan(h, p[2:].astype(np.int32).copy()) for p in parents)
    for lb in (0, 1, 2):
        eng = bb.Engine.pfsp(ins, lb)
        eng.reset(INT_MAX)
        eng.push_nodes(parents)
        st = eng.run(0)
        assert (st.sol, st.tree, st.best, st.pool_size) == (3000, 0, exp_best, 0)
        eng.close()
    L.orc_pfsp_free(h)


def test_full_runs_match_reference_counts(golden):
    g = golden("runs")
    for r in g["pfsp"] + [s for s in g["survey"]["pfsp"] if s["tree"] < 200_000_000]:
        ins = bb.Instance(inst=r["inst"])
        res, per = bb.pfsp_search(ins, r["lb"], M=200000)
        assert (res.tree, res.sol, res.best) == (r["tree"], r["sol"], r["makespan"]), r
        assert per[0].parents == res.tree + 1  # every explored node is decomposed exactly once, plus the root
        assert per[0].pool_size == 0


@pytest.mark.parametrize("M,iters", [(25, 8), (1000, 32), (50000, 32), (1 << 20, 1)])
def test_counts_do_not_depend_on_batching(M, iters):
    ins = bb.Instance(inst=14)
    for lb, exp in ((1, (2573652, 2648, 1377)), (2, (144639, 0, 1377))):
        if M == 25 and lb == 1:
            continue  # 100k+ tiny iterations: covered by LB2 at this size
        res, _ = bb.pfsp_search(ins, lb, M=M, iters_per_launch=iters)
        assert (res.tree, res.sol, res.best) == exp


def test_lb2_prefilter_does_not_change_counts():
    ins = bb.Instance(inst=14)
    a, pa = bb.pfsp_search(ins, 2, lb2_prefilter=1)
    b, pb = bb.pfsp_search(ins, 2, lb2_prefilter=0)
    assert (a.tree, a.sol, a.best) == (b.tree, b.sol, b.best) == (144639, 0, 1377)
    assert pb[0].lb2_evals == pb[0].evals == 1298094 > pa[0].lb2_evals  # SURVEY appendix A.2 census


LB2_VARIANTS = {"table": dict(lb2_table=1), "staged": dict(lb2_table=0, lb2_staged=1, lb2_stage_a=4),
                "ordered": dict(lb2_table=0, lb2_staged=0)}


@pytest.mark.parametrize("variant", sorted(LB2_VARIANTS))
@pytest.mark.parametrize("inst", [14, 21])
def test_lb2_kernels_one_iteration_children_match_oracle(variant, inst):
    """the three LB2 evaluations - subset table (default), staged pair loop with learned order, pair loop in the
    reference's order - keep exactly the children generate_children keeps on the reference's bounds"""
    L = orc.oracle()
    ins = bb.Instance(inst=inst)
    n = ins.n
    h = L.orc_pfsp_new(inst)
    rng = np.random.default_rng(17 * inst)
    eng = bb.Engine.pfsp(ins, 2, M=100000, **LB2_VARIANTS[variant])
    for best in (ins.best_known, ins.best_known + 90):
        parents = orc.random_pfsp_nodes(rng, n, 6000, 0, n - 2)
        eng.reset(best)
        eng.push_nodes(parents)
        st = eng.run(1)
        kids, tree, sol, b, nb = oracle_one_level(h, n, 2, parents, best)
        assert (st.tree, st.sol, st.best, st.parents, st.evals) == (tree, sol, b, len(parents), nb)
        assert node_set(eng.export_nodes(), 2 + n) == node_set(kids, 2 + n)
    eng.close()
    L.orc_pfsp_free(h)


def test_lb2_kernels_agree_on_whole_searches():
    for inst, exp in ((14, (144639, 0, 1377)), (3, (80062, 0, 1081)), (29, (9499307, 0, 2237))):
        ins = bb.Instance(inst=inst)
        for name, kw in LB2_VARIANTS.items():
            res, _ = bb.pfsp_search(ins, 2, M=1 << 18, **kw)
            assert (res.tree, res.sol, res.best) == exp, (inst, name)


def _gpu_count():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("D", [2, 4, 8])
def test_multi_gpu_threads_same_counts(D):
    """D GPUs in one process (one host thread per GPU, peer-memory stealing): the counts are those of one GPU"""
    if _gpu_count() < D:
        pytest.skip(f"needs {D} GPUs")
    for ws in (1, 0):
        res, per = bb.pfsp_search(bb.Instance(inst=30), 0, D=D, ws=ws, M=1 << 18)
        assert (res.tree, res.sol, res.best) == (111887480, 82, 2178)
        assert sum(p.tree for p in per) + res.warm_tree == res.tree
        res, per = bb.pfsp_search(bb.Instance(inst=29), 2, D=D, ws=ws, M=1 << 18)
        assert (res.tree, res.sol, res.best) == (9499307, 0, 2237)
    res, per = bb.nq_search(15, 1, D=D, M=1 << 18)
    assert (res.tree, res.sol) == (171129071, 2279184)


@pytest.mark.parametrize("D", [2, 3])
def test_several_workers_on_one_gpu_same_counts(D, monkeypatch):
    """the multi-worker paths on a one-GPU box: with BB_WORKERS_SHARE_GPUS the D workers of a search run on GPU
    i mod visible. Everything else is the multi-GPU code: common warm-up, device-side split by prefix hash, host
    protocol, donations enqueued on the victim's stream, pulls out of the victim's transfer buffer, incumbent exchange,
    termination. Counts must be those of one worker; with work stealing off the static split alone must be complete."""
    monkeypatch.setenv("BB_WORKERS_SHARE_GPUS", "1")
    for ws in (1, 0):
        res, per = bb.pfsp_search(bb.Instance(inst=30), 0, D=D, ws=ws, M=1 << 16)
        assert (res.tree, res.sol, res.best) == (111887480, 82, 2178)
        assert sum(p.tree for p in per) + res.warm_tree == res.tree
        assert all(p.tree > 0 for p in per)  # every worker got a share
        res, per = bb.pfsp_search(bb.Instance(inst=29), 2, D=D, ws=ws, M=1 << 16)
        assert (res.tree, res.sol, res.best) == (9499307, 0, 2237)
    res, per = bb.nq_search(14, 1, D=D, M=1 << 16)
    assert (res.tree, res.sol) == (27358552, 365596)
    assert all(p.tree > 0 for p in per)
    # incumbents are exchanged between the workers (checkBest, pfsp_multigpu_cuda.c:30-50)
    for inst, lb, opt in ((14, 2, 1377), (3, 0, 1081)):
        res, _ = bb.pfsp_search(bb.Instance(inst=inst), lb, best0=INT_MAX, D=D, M=1 << 14)
        assert res.best == opt
    # a 100-job instance through the wide record (solved during the warm-up: the reference prints 11 nodes)
    res, _ = bb.pfsp_search(bb.Instance(inst=66), 0, D=D, M=1 << 14)
    assert (res.tree, res.sol, res.best) == (11, 0, 5135)


def test_process_per_worker_team_on_one_gpu_same_counts():
    """the process-per-GPU path of bench.py --gpus N (bb_*_dist_open / bb_dist_search: shared-memory team block, CUDA
    IPC mappings of the peers' transfer buffers, persistent contexts) with two processes that both use GPU 0: every
    search returns the one-worker counts on every rank, with and without work stealing, for PFSP, -u 0 and N-Queens"""
    import json
    import os
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    shm = f"/bbtest_{os.getpid()}_2p"
    procs = [subprocess.Popen([sys.executable, os.path.join(here, "dist_rank.py"), str(r), "2", shm, "0"],
                              stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for r in range(2)]
    outs = []
    for p in procs:
        try:
            o, e = p.communicate(timeout=300)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        assert p.returncode == 0, e[-2000:]
        outs.append(json.loads(o.strip().splitlines()[-1]))
    for o in outs:
        for k in ("ta030_lb1d_0", "ta030_lb1d_1", "ta030_lb1d_static"):
            assert o[k][:3] == [111887480, 82, 2178], (k, o[k])
            assert o[k][3] == 1  # this rank explored a share itself
        assert o["ta014_lb2_u0_best"] == 1377
        assert o["ta014_lb2"] == [144639, 0, 1377]
        assert o["nq14"][:2] == [27358552, 365596] and o["nq14"][2] == 1


@pytest.mark.parametrize("inst,lb,opt", [(14, 2, 1377), (3, 0, 1081), (14, 1, 1377)])
def test_search_from_infinite_upper_bound_finds_the_optimum(inst, lb, opt):
    """-u 0 (PFSP_lib.c:234-241): the incumbent starts at INT_MAX and is improved by leaves through the device
    atomicMin; only the optimum is pinned (node counts depend on the traversal order)"""
    res, _ = bb.pfsp_search(bb.Instance(inst=inst), lb, best0=INT_MAX, M=1 << 16)
    assert res.best == opt and res.sol >= 1
    if _gpu_count() >= 2:  # incumbents are exchanged between the workers (checkBest, pfsp_multigpu_cuda.c:30-50)
        res, _ = bb.pfsp_search(bb.Instance(inst=inst), lb, best0=INT_MAX, D=2, M=1 << 16)
        assert res.best == opt


def test_custom_instance_generic_kernel_and_unknown_optimum():
    """a 9x7 matrix exercises the runtime-m kernel; start from ub = inf to find the optimum, then both
    implementations must walk the same tree from ub = optimum"""
    L = orc.oracle()
    rng = np.random.default_rng(42)
    p = rng.integers(1, 100, size=(7, 9)).astype(np.int32)
    ins = bb.Instance(p_times=p)
    h = L.orc_pfsp_from_ptimes(9, 7, np.ascontiguousarray(p.reshape(-1)))
    st = orc.RunStats()
    L.orc_pfsp_search(h, 0, INT_MAX, 0, C.byref(st))
    opt = st.best
    for lb in (0, 1, 2):
        res, _ = bb.pfsp_search(ins, lb, best0=INT_MAX)
        assert res.best == opt                       # -u 0: only the optimum is pinned
        L.orc_pfsp_search(h, lb, opt, 0, C.byref(st))
        res, _ = bb.pfsp_search(ins, lb, best0=opt)
        assert (res.tree, res.sol, res.best) == (st.tree, st.sol, opt)
    L.orc_pfsp_free(h)


def test_small_capacity_is_safe():
    """near capacity the planner shrinks batches; if even that cannot fit the engine reports BB_ERR_OVERFLOW.
    It never returns wrong counts."""
    ins = bb.Instance(inst=14)
    try:
        res, _ = bb.pfsp_search(ins, 1, M=50000, capacity=120000)
        assert (res.tree, res.sol, res.best) == (2573652, 2648, 1377)
    except bb.BBError as ex:
        assert "overflow" in str(ex)
    try:  # a pool of 1024 nodes per stack: batches shrink to ~50 parents, still exact
        res, per = bb.pfsp_search(ins, 2, M=50000, capacity=1024)
        assert (res.tree, res.sol, res.best) == (144639, 0, 1377) and per[0].max_pool <= 2048
    except bb.BBError as ex:
        assert "overflow" in str(ex)
    with pytest.raises(bb.BBError, match="too small"):
        bb.pfsp_search(ins, 1, capacity=100)


@pytest.mark.parametrize("N", [1, 2, 3, 4, 8, 12, 14, 15, 16, 17, 20])
def test_nqueens_one_iteration_matches_oracle(N):
    L = orc.oracle()
    rng = np.random.default_rng(N)
    nodes = orc.random_nq_nodes(rng, N, 4000, 0, max(N - 1, 0))
    labels = np.zeros(len(nodes) * N, np.uint8)
    L.orc_nq_labels(N, 1, nodes, len(nodes), labels)
    tree, sol = C.c_ulonglong(0), C.c_ulonglong(0)
    kids = np.zeros((len(nodes) * N, orc.NQ_REC), np.uint8)
    k = L.orc_nq_generate_children(N, nodes, len(nodes), labels, C.byref(tree), C.byref(sol), kids)
    inner = kids[:k][kids[:k, 0] < N]
    eng = bb.Engine.nqueens(N, 1, packed=2)  # every node through the pool: one iteration = one level
    eng.reset()
    eng.push_nodes(nodes)
    assert node_set(eng.export_nodes(), 1 + N) == node_set(nodes, 1 + N)
    st = eng.run(1)
    # depth-N children are counted (tree and sol) but not stored
    assert st.tree == tree.value and st.sol == k - len(inner) and st.pool_size == len(inner)
    assert node_set(eng.export_nodes(), 1 + N) == node_set(inner, 1 + N)
    eng.close()


@pytest.mark.parametrize("N,G", [(4, 1), (8, 1), (11, 3), (12, 1), (14, 1), (17, 1), (20, 1)])
def test_nqueens_one_iteration_with_the_last_levels_in_registers(N, G):
    """the default engine (bb_config.packed = 3): parents with at most min(12, N - 7) columns left are enumerated to the end by
    the lane that read them, the others produce their children. After ONE iteration over random parents of every depth
    the pool must hold exactly the checker's children of the shallow parents, and the counters must be the checker's
    for one level of the shallow parents plus the whole subtrees of the deep ones (walked level by level here)."""
    L = orc.oracle()
    rng = np.random.default_rng(100 + N)
    tail = min(12, max(0, N - 7))
    # deep parents: consistent boards (random placements of d non-attacking queens are rare, so walk down from the root)
    level = np.zeros((1, orc.NQ_REC), np.uint8)
    level[0, 1:1 + N] = np.arange(N)
    picked = [level]
    for d in range(N - 1):
        labels = np.zeros(len(level) * N, np.uint8)
        L.orc_nq_labels(N, 1, level, len(level), labels)
        t, s_ = C.c_ulonglong(0), C.c_ulonglong(0)
        kids = np.zeros((len(level) * N, orc.NQ_REC), np.uint8)
        k = L.orc_nq_generate_children(N, level, len(level), labels, C.byref(t), C.byref(s_), kids)
        kids = kids[:k][kids[:k, 0] < N]
        if len(kids) == 0:
            break
        level = kids[rng.permutation(len(kids))[:60]]
        if d + 1 < N - tail or d + 1 >= N - 6:  # shallow parents, and deep ones whose subtrees the checker can walk
            picked.append(level[:25])
        elif d + 1 == N - tail:
            picked.append(level[:2])            # two whole subtrees of the full height
    rnd = orc.random_nq_nodes(rng, N, 300, 0, max(N - 1, 0))
    rnd = rnd[(rnd[:, 0] < N - tail) | (rnd[:, 0] >= N - 6)]
    nodes = np.concatenate(picked + [rnd])
    deep = nodes[:, 0] >= N - tail
    shallow = nodes[~deep]

    def one_level(par):
        labels = np.zeros(max(len(par), 1) * N, np.uint8)
        L.orc_nq_labels(N, G, par, len(par), labels)
        t, s_ = C.c_ulonglong(0), C.c_ulonglong(0)
        kids = np.zeros((max(len(par), 1) * N, orc.NQ_REC), np.uint8)
        k = L.orc_nq_generate_children(N, par, len(par), labels, C.byref(t), C.byref(s_), kids)
        kids = kids[:k]
        return kids[kids[:, 0] < N], t.value, int((kids[:, 0] == N).sum())

    exp_kids, tree, sol = one_level(shallow) if len(shallow) else (nodes[:0], 0, 0)
    parents = len(shallow)
    lvl = nodes[deep]
    while len(lvl):
        parents += len(lvl)
        lvl, t, s_ = one_level(lvl)
        tree, sol = tree + t, sol + s_
    eng = bb.Engine.nqueens(N, G)
    eng.reset()
    eng.push_nodes(nodes)
    st = eng.run(1)
    assert (st.tree, st.sol, st.parents) == (tree, sol, parents)
    assert st.pool_size == len(exp_kids)
    if len(exp_kids):
        assert node_set(eng.export_nodes(), 1 + N) == node_set(exp_kids, 1 + N)
    eng.close()


def test_nqueens_counts(golden):
    g = golden("runs")
    for r in g["nqueens"] + [s for s in g["survey"]["nqueens"] if s["N"] <= 16]:
        for G in ((1, 4) if r["N"] <= 12 else (1,)):
            res, per = bb.nq_search(r["N"], G, M=1 << 20)
            assert (res.tree, res.sol) == (r["tree"], r["sol"]), r
            assert per[0].parents == res.tree - res.sol + 1


def test_nqueens_17_full_size(golden):
    r = [s for s in golden("runs")["survey"]["nqueens"] if s["N"] == 17][0]
    res, _ = bb.nq_search(17, 1, M=1 << 22)
    assert (res.tree, res.sol) == (r["tree"], r["sol"]) == (8017021931, 95815104)


def test_pfsp_full_size_properties():
    """ta022 / LB1_d, 711 M nodes: counts from the reference run"""
    ins = bb.Instance(inst=22)
    res, per = bb.pfsp_search(ins, 0, M=1 << 20)
    assert (res.tree, res.sol, res.best) == (711291688, 4346, 2099)
    assert per[0].parents == res.tree + 1


def test_persistent_context_searches_repeat():
    """bb_*_dist_open / bb_dist_search / bb_dist_close with a team of one: several searches on one context give the
    reference's counts every time (nothing of a finished search leaks into the next), timed from the root"""
    import os
    d = bb.Dist(0, 1, f"/bbtest_{os.getpid()}_a", instance=bb.Instance(inst=14), lb=2, M=1 << 16)
    for _ in range(3):
        res, mine = d.search()
        assert (res.tree, res.sol, res.best) == (144639, 0, 1377)
        assert res.search_s > 0 and mine.parents + 0 >= 1
    res, _ = d.search(best0=INT_MAX)  # -u 0 on the same context
    assert res.best == 1377 and res.sol >= 1
    d.close()
    d = bb.Dist(0, 1, f"/bbtest_{os.getpid()}_b", N=13, G=1, M=1 << 18)
    for _ in range(2):
        res, _ = d.search()
        assert (res.tree, res.sol) == (4674889, 73712)
    d.close()
