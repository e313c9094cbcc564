"""The 50-, 100- and 200-job Taillard classes (ta031-ta110). The reference handles them by editing MAX_JOBS in
pfsp/lib/macro.h:10 and rebuilding (pfsp/README.md:52); the checker is the same restatement built with 50- / 100-job
node records (oracle/liboracle50.so, liboracle100.so, liboracle200.so), pinned here against that rebuilt reference
(oracle/_ref/libref_pfsp<NJ>.so, make -C oracle ref50 ref100 ref200) where it is present, and against tests/golden/wide.json
(generated from it by tests/golden/make_golden_wide.py) everywhere. The GPU tests compare the engine's wide kernel with
the checker through the C ABI: the child SET of one iteration for all three bounds, and whole searches."""
import ctypes as C

import numpy as np
import pytest

import orc
from conftest import load_golden

INT_MAX = 2**31 - 1


def klass(inst):
    """job-capacity class of a Taillard instance: MAX_JOBS of the reference build that handles it"""
    return 50 if inst <= 60 else 100 if inst <= 90 else 200


def chunk_bounds(L, h, n, lb, parents, best):
    nb = int((n - parents[:, 0]).sum())
    bounds = np.zeros(max(nb, 1), np.int32)
    L.orc_evaluate_chunk(h, lb, parents, len(parents), best, bounds)
    return bounds[:nb]


def golden_nodes(g, inst):
    return np.array(g["parents"], dtype=np.int16).reshape(-1, 2 + klass(inst))


@pytest.mark.parametrize("inst", [31, 41, 51, 61, 71, 81, 91, 101])
def test_wide_oracle_matches_reference_golden(inst):
    """tables and chunk-level bounds of the 50- / 100-job checker against what the rebuilt reference produced"""
    g = load_golden("wide")[str(inst)]
    L = orc.oracle(klass(inst))
    h = L.orc_pfsp_new(inst)
    n, m = L.orc_taillard_jobs(inst), L.orc_taillard_machines(inst)
    t = orc.Tables(n, m, m * (m - 1) // 2)
    L.orc_pfsp_tables(h, *t.arrays())
    assert {k: "%016x" % orc.fnv(a) for k, a in zip(("p", "heads", "tails", "lags", "johnson"), t.arrays())} == g["fnv"]
    assert L.orc_taillard_best_ub(inst) == g["best"]
    parents = golden_nodes(g, inst)
    for lb in (0, 1, 2):
        got = chunk_bounds(L, h, n, lb, parents, g["best"])
        assert got.tolist() == g["bounds"][str(lb)], (inst, lb)
    L.orc_pfsp_free(h)


@pytest.mark.parametrize("inst", [31, 45, 52, 62, 77, 84, 93, 108])
def test_wide_oracle_matches_reference_directly(inst):
    if not orc.have_ref_wide(klass(inst)):
        pytest.skip("oracle/_ref/libref_pfsp%d.so not built (make -C oracle ref%d)" % (klass(inst), klass(inst)))
    L, R = orc.oracle(klass(inst)), orc.ref_pfsp(klass(inst))
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
@pytest.mark.parametrize("inst", [31, 41, 51, 61, 71, 81, 91, 101])
@pytest.mark.parametrize("lb", [0, 1, 2])
def test_wide_one_iteration_children_match_oracle(inst, lb):
    from __graft_entry__ import load_package
    bb = load_package()
    L = orc.oracle(klass(inst))
    W = 2 + klass(inst)
    ins = bb.Instance(inst=inst)
    n = ins.n
    h = L.orc_pfsp_new(inst)
    rng = np.random.default_rng(1000 * inst + lb)
    eng = bb.Engine.pfsp(ins, lb, M=100000)
    for best in (ins.best_known, ins.best_known + 400):
        parents = orc.random_pfsp_nodes(rng, n, 3000 if n <= 50 else 1200 if n <= 100 else 400, 0, n - 2, width=W)
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
@pytest.mark.parametrize("nj", [50, 100, 200])
def test_wide_whole_searches_match_oracle(nj):
    """whole trees: the instances the reference solves at the root (ta032, ta034, ta061, ta063, ta065, ta066, ta069:
    counts printed by the rebuilt pfsp_c, tests/golden/wide.json), and searches of ta031 / ta041 / ta051 (ta061 / ta071
    / ta081) and of random 26..40-job (60..90-job) instances below an upper bound that no leaf can improve (best0 a
    little above the root's bound: the tree is then the set of nodes with lb < best0, a constant of (instance, bound,
    best0) like the reference's trees at the optimum)"""
    from __graft_entry__ import load_package
    bb = load_package()
    L = orc.oracle(nj)
    for r in load_golden("wide")["runs"]:
        if klass(r["inst"]) != nj:
            continue
        res, per = bb.pfsp_search(bb.Instance(inst=r["inst"]), r["lb"], M=1 << 16)
        assert (res.tree, res.sol, res.best) == (r["tree"], r["sol"], r["makespan"]), r
    rng = np.random.default_rng(5 + nj)
    cases = [(bb.Instance(inst=i), L.orc_pfsp_new(i))
             for i in {50: (31, 41, 51), 100: (61, 71, 81), 200: (91, 101)}[nj]]
    for n, m in {50: ((26, 5), (33, 10), (40, 7)), 100: ((60, 5), (75, 10), (90, 7)), 200: ((130, 5), (180, 12))}[nj]:
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
                budget = {50: 400_000, 100: 150_000, 200: (150_000, 5_000, 1_000)[lb]}[nj]  # the checker's LB2 is O(n^2 m^2)
                L.orc_pfsp_search(h, lb, best0, budget, C.byref(st))
                if st.parents >= budget or st.best != best0:
                    break  # the checker's budget ran out, or a leaf improved the bound: not a constant tree
                res, per = bb.pfsp_search(ins, lb, best0=best0, M=1 << 14)
                assert (res.tree, res.sol, res.best) == (st.tree, st.sol, best0), (n, ins.m, lb, delta)
                compared += 1
        L.orc_pfsp_free(h)
    assert compared >= {50: 12, 100: 6, 200: 3}[nj]


@pytest.mark.gpu
@pytest.mark.timeout(240, method="thread")
@pytest.mark.parametrize("inst", [41, 51, 71, 81, 91, 101])
@pytest.mark.parametrize("lb", [0, 1, 2])
def test_wide_subtrees_below_deep_seeds_match_oracle(inst, lb):
    """searches to exhaustion that are not trivial at 50 / 100 jobs: 40 random nodes with six jobs left are seeded and
    everything below them is explored. (a) Without an upper bound and with a batch that takes a whole level (all
    seeds have the same depth, so both sides work level by level and the first leaves appear in the last level: 49 440
    nodes, 28 800 leaves, nothing order dependent). (b) With the upper bound AT the best leaf found in (a) and small
    batches: no leaf can improve it, the explored set is a constant of (seeds, bound). Counts and incumbent must be
    identical in both."""
    from __graft_entry__ import load_package
    bb = load_package()
    nj = klass(inst)
    L = orc.oracle(nj)
    ins = bb.Instance(inst=inst)
    n = ins.n
    h = L.orc_pfsp_new(inst)
    rng = np.random.default_rng(77 * inst + lb)
    seeds = orc.random_pfsp_nodes(rng, n, 40, n - 6, n - 6, width=2 + nj)

    def checker(best0):
        level, tree, sol, best = seeds, 0, 0, best0
        while len(level):
            level, t, s, best, _ = orc.oracle_one_level(h, n, lb, level, best, L=L)
            tree, sol = tree + t, sol + s
        return tree, sol, best

    full = checker(INT_MAX)
    assert full[0] == 40 * (6 + 30 + 120 + 360 + 720) and full[1] == 40 * 720
    mn = full[2]
    eng = bb.Engine.pfsp(ins, lb, M=1 << 16)
    eng.reset(INT_MAX)
    eng.push_nodes(seeds)
    st = eng.run()
    assert (st.tree, st.sol, st.best) == full
    eng.close()
    bounded = checker(mn)
    assert bounded[2] == mn
    eng = bb.Engine.pfsp(ins, lb, M=4096)
    for _ in range(2):
        eng.reset(mn)
        eng.push_nodes(seeds)
        st = eng.run()
        assert (st.tree, st.sol, st.best) == bounded
    # small batches without an upper bound: the tree is order dependent, the optimum below the seeds is not
    eng.reset(INT_MAX)
    eng.push_nodes(seeds)
    assert eng.run().best == mn
    eng.close()
    L.orc_pfsp_free(h)
