"""The 50-job Taillard classes (ta031-ta060). The reference handles them by editing MAX_JOBS in pfsp/lib/macro.h:10
and rebuilding (pfsp/README.md:52); the checker is the same restatement built with 50-job node records
(oracle/liboracle50.so), pinned here against that rebuilt reference (oracle/_ref/libref_pfsp50.so, make -C oracle
ref50) where it is present, and against tests/golden/wide.json (generated from it by tests/golden/make_golden_wide.py)
everywhere. The GPU tests compare the engine's wide kernel with the checker through the C ABI: the child SET of one
iteration for all three bounds, and whole searches."""
import ctypes as C

import numpy as np
import pytest

import orc
from conftest import load_golden

W = 2 + 50
INT_MAX = 2**31 - 1


def chunk_bounds(L, h, n, lb, parents, best):
    nb = int((n - parents[:, 0]).sum())
    bounds = np.zeros(max(nb, 1), np.int32)
    L.orc_evaluate_chunk(h, lb, parents, len(parents), best, bounds)
    return bounds[:nb]


def golden_nodes(g):
    return np.array(g["parents"], dtype=np.int16).reshape(-1, W)


@pytest.mark.parametrize("inst", [31, 41, 51])
def test_oracle50_matches_reference_golden(inst):
    """tables and chunk-level bounds of the 50-job checker against what the rebuilt reference produced"""
    g = load_golden("wide")[str(inst)]
    L = orc.oracle(50)
    h = L.orc_pfsp_new(inst)
    n, m = L.orc_taillard_jobs(inst), L.orc_taillard_machines(inst)
    t = orc.Tables(n, m, m * (m - 1) // 2)
    L.orc_pfsp_tables(h, *t.arrays())
    assert {k: "%016x" % orc.fnv(a) for k, a in zip(("p", "heads", "tails", "lags", "johnson"), t.arrays())} == g["fnv"]
    assert L.orc_taillard_best_ub(inst) == g["best"]
    parents = golden_nodes(g)
    for lb in (0, 1, 2):
        got = chunk_bounds(L, h, n, lb, parents, g["best"])
        assert got.tolist() == g["bounds"][str(lb)], (inst, lb)
    L.orc_pfsp_free(h)


@pytest.mark.skipif(not orc.have_ref50(), reason="oracle/_ref/libref_pfsp50.so not built (make -C oracle ref50)")
@pytest.mark.parametrize("inst", [31, 45, 52])
def test_oracle50_matches_reference_directly(inst):
    L, R = orc.oracle(50), orc.ref_pfsp(50)
    h, r = L.orc_pfsp_new(inst), R.ref_pfsp_open(inst)
    n = L.orc_taillard_jobs(inst)
    best = L.orc_taillard_best_ub(inst)
    rng = np.random.default_rng(inst)
    for _ in range(60):
        d = int(rng.integers(0, n - 1))
        prmu = rng.permutation(n).astype(np.int32)
        lbb_o, lbb_r = np.zeros(n, np.int32), np.zeros(n, np.int32)
        L.orc_lb1_children(h, prmu, d - 1, n, lbb_o)
        R.ref_lb1_children_bounds(r, prmu, d - 1, n, lbb_r)
        assert lbb_o.tolist() == lbb_r.tolist()
        child = prmu.copy()
        k = int(rng.integers(d, n))
        child[d], child[k] = child[k], child[d]
        assert L.orc_lb1(h, child, d, n) == R.ref_lb1_bound(r, child, d, n)
        for b in (best, INT_MAX):
            assert L.orc_lb2(h, child, d, n, b) == R.ref_lb2_bound(r, child, d, n, b)
    L.orc_pfsp_free(h)
    R.ref_pfsp_close(r)


# ------------------------------------------------------------------ GPU

@pytest.mark.gpu
@pytest.mark.timeout(240, method="thread")
@pytest.mark.parametrize("inst", [31, 41, 51])
@pytest.mark.parametrize("lb", [0, 1, 2])
def test_wide_one_iteration_children_match_oracle(inst, lb):
    from __graft_entry__ import load_package
    bb = load_package()
    L = orc.oracle(50)
    ins = bb.Instance(inst=inst)
    n = ins.n
    h = L.orc_pfsp_new(inst)
    rng = np.random.default_rng(1000 * inst + lb)
    eng = bb.Engine.pfsp(ins, lb, M=100000)
    for best in (ins.best_known, ins.best_known + 400):
        parents = orc.random_pfsp_nodes(rng, n, 3000, 0, n - 2, width=W)
        parents[0, :2] = (0, -1)
        parents[0, 2:2 + n] = np.arange(n)
        eng.reset(best)
        eng.push_nodes(parents)
        as_set = lambda a: sorted(tuple(int(v) for v in r[:2 + n]) for r in a)
        assert as_set(eng.export_nodes()) == as_set(parents)  # seed / export round trip
        st = eng.run(1)
        kids, tree, sol, b, nb = orc.oracle_one_level(h, n, lb, parents, best, L=L)
        assert (st.tree, st.sol, st.best, st.parents, st.evals) == (tree, sol, b, len(parents), nb)
        assert as_set(eng.export_nodes()) == as_set(kids)
    eng.close()
    L.orc_pfsp_free(h)


@pytest.mark.gpu
@pytest.mark.timeout(240, method="thread")
def test_wide_whole_searches_match_oracle():
    """whole trees: the 50-job instances the reference solves at the root (ta032, ta034: 3 nodes, counts printed by
    the rebuilt pfsp_c, tests/golden/wide.json), and searches of ta031 / ta041 / ta051 and of random 26..40-job
    instances below an upper bound that no leaf can improve (best0 a little above the root's bound: the tree is then
    the set of nodes with lb < best0, a constant of (instance, bound, best0) like the reference's trees at the optimum)"""
    from __graft_entry__ import load_package
    bb = load_package()
    L = orc.oracle(50)
    for r in load_golden("wide")["runs"]:
        res, per = bb.pfsp_search(bb.Instance(inst=r["inst"]), r["lb"], M=1 << 16)
        assert (res.tree, res.sol, res.best) == (r["tree"], r["sol"], r["makespan"]), r
    rng = np.random.default_rng(5)
    cases = [(bb.Instance(inst=i), L.orc_pfsp_new(i)) for i in (31, 41, 51)]
    for n, m in ((26, 5), (33, 10), (40, 7)):
        p = rng.integers(1, 100, size=(m, n)).astype(np.int32)
        cases.append((bb.Instance(p_times=p), L.orc_pfsp_from_ptimes(n, m, np.ascontiguousarray(p.reshape(-1)))))
    compared = 0
    for ins, h in cases:
        n = ins.n
        root = np.arange(n, dtype=np.int32)
        lbb = np.zeros(n, np.int32)
        L.orc_lb1_children(h, root, -1, n, lbb)
        base = int(lbb.min())
        st = orc.RunStats()
        for lb in (0, 1, 2):
            for delta in (3, 12, 30):
                best0 = base + delta
                L.orc_pfsp_search(h, lb, best0, 400_000, C.byref(st))
                if st.parents >= 400_000 or st.best != best0:
                    break  # the checker's budget ran out, or a leaf improved the bound: not a constant tree
                res, per = bb.pfsp_search(ins, lb, best0=best0, M=1 << 14)
                assert (res.tree, res.sol, res.best) == (st.tree, st.sol, best0), (n, ins.m, lb, delta)
                compared += 1
        L.orc_pfsp_free(h)
    assert compared >= 12
