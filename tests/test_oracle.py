"""Pins the CPU checker (oracle/) against the reference: golden vectors generated from the compiled reference
(tests/golden/make_golden.py) and, when oracle/_ref is present, the reference's own functions directly."""
import ctypes as C

import numpy as np
import pytest

import orc

INT_MAX = 2**31 - 1


def fnv_hex(a):
    return "%016x" % orc.fnv(a)


def test_tables_match_reference_hashes(golden):
    g = golden("tables")
    for inst in range(1, 121):
        t, e = orc.oracle_tables(inst), g[str(inst)]
        assert (t.n, t.m, t.P) == (e["n"], e["m"], e["P"])
        assert orc.oracle().orc_taillard_best_ub(inst) == e["ub"]
        assert fnv_hex(t.p) == e["p"], inst
        assert fnv_hex(t.heads) == e["heads"] and fnv_hex(t.tails) == e["tails"], inst
        assert fnv_hex(t.lags) == e["lags"] and fnv_hex(t.johnson) == e["johnson"], inst
        assert fnv_hex(np.concatenate([t.pa, t.pb, t.order])) == e["pairs"], inst
        assert int(t.lags.max()) == e["max_lag"]


def test_survey_known_rows(golden):
    # SURVEY.md appendix A.3 (Taillard's published ta001 first row, heads/tails of ta014)
    t = orc.oracle_tables(1)
    assert t.p[:20].tolist() == [54, 83, 15, 71, 77, 36, 53, 38, 27, 87, 76, 91, 14, 29, 12, 77, 32, 87, 68, 94]
    t = orc.oracle_tables(14)
    assert t.heads.tolist() == [0, 5, 32, 39, 108, 139, 187, 203, 248, 289]
    assert t.tails.tolist() == [280, 205, 115, 110, 100, 78, 50, 31, 8, 0]
    assert t.johnson[:20].tolist() == [15, 2, 14, 8, 18, 4, 7, 5, 12, 13, 9, 3, 6, 19, 1, 10, 16, 17, 0, 11]
    L = orc.oracle()
    h = L.orc_pfsp_new(14)
    assert L.orc_makespan(h, np.arange(20, dtype=np.int32)) == 1726
    L.orc_pfsp_free(h)


@pytest.mark.parametrize("inst", [1, 5, 14, 17, 21, 24, 29])
def test_bounds_match_golden(golden, inst):
    g = golden("bounds")[str(inst)]
    L = orc.oracle()
    n, _ = orc.taillard_dims(inst)
    h = L.orc_pfsp_new(inst)
    ub = g["ub"]
    for nd, e_d, e_1, e_2, e_2i, e_2t in zip(g["nodes"], g["lb1_d"], g["lb1"], g["lb2_ub"], g["lb2_inf"],
                                             g["lb2_tight"]):
        d = nd[0]
        prmu = np.array(nd[2:], dtype=np.int32)
        lbb = np.zeros(n, np.int32)
        L.orc_lb1_children(h, prmu, d - 1, n, lbb)
        assert [int(lbb[prmu[k]]) for k in range(d, n)] == e_d
        for c, k in enumerate(range(d, n)):
            ch = prmu.copy()
            ch[d], ch[k] = ch[k], ch[d]
            assert L.orc_lb1(h, ch, d, n) == e_1[c]
            assert L.orc_lb2(h, ch, d, n, ub) == e_2[c]
            assert L.orc_lb2(h, ch, d, n, INT_MAX) == e_2i[c]
            assert L.orc_lb2(h, ch, d, n, ub - 40) == e_2t[c]
        # LB1 and LB1_d coincide under forward-only branching
        assert e_d == e_1
    L.orc_pfsp_free(h)


def test_chunk_protocol_matches_per_child_bounds():
    L = orc.oracle()
    rng = np.random.default_rng(3)
    for inst in (14, 21):
        n, _ = orc.taillard_dims(inst)
        ub = L.orc_taillard_best_ub(inst)
        h = L.orc_pfsp_new(inst)
        parents = orc.random_pfsp_nodes(rng, n, 50)
        so, ni = np.zeros(50, np.int32), np.zeros(50 * n, np.int32)
        nb = L.orc_batch_index(parents, 50, n, so, ni)
        assert nb == int((n - parents[:, 0]).sum()) and so[-1] == nb
        assert all(ni[c] == np.searchsorted(so, c, side="right") for c in range(nb))
        res = {}
        for lb in (0, 1, 2):
            b = np.zeros(nb, np.int32)
            L.orc_evaluate_chunk(h, lb, parents, 50, ub, b)
            res[lb] = b
        assert np.array_equal(res[0], res[1])
        assert np.all((res[2] >= res[0]) | (res[2] > ub))
        # prune+branch
        tree, sol, best = C.c_ulonglong(0), C.c_ulonglong(0), C.c_int(ub)
        kids = np.zeros((nb, orc.NODE_I16), np.int16)
        k = L.orc_generate_children(parents, 50, n, res[0], C.byref(tree), C.byref(sol), C.byref(best), kids)
        leaves = int((parents[:, 0] == n - 1).sum())
        assert sol.value == leaves
        inner = np.repeat(parents[:, 0] < n - 1, n - parents[:, 0])
        assert k == tree.value == int(((res[0] < ub) & inner).sum())
        for ch in kids[:k]:
            assert sorted(ch[2:2 + n].tolist()) == list(range(n)) and ch[1] == ch[0] - 1
        L.orc_pfsp_free(h)


def test_full_runs_match_golden(golden):
    L = orc.oracle()
    for r in golden("runs")["pfsp"]:
        if r["tree"] > 3_000_000:
            continue  # keep the CPU suite short; the big ones are GPU parity targets
        h = L.orc_pfsp_new(r["inst"])
        st = orc.RunStats()
        L.orc_pfsp_search(h, r["lb"], L.orc_taillard_best_ub(r["inst"]), 0, C.byref(st))
        assert (st.tree, st.sol, st.best) == (r["tree"], r["sol"], r["makespan"]), r
        L.orc_pfsp_free(h)


def test_batched_search_is_order_independent(golden):
    # the reference's 3-step batched structure must give the same counts as plain DFS (ub = optimum)
    L = orc.oracle()
    exp = {(r["inst"], r["lb"]): r for r in golden("runs")["pfsp"]}
    for inst, lb, m, M in [(14, 0, 25, 50000), (14, 2, 25, 50000), (3, 0, 25, 1000), (4, 2, 5, 64)]:
        h = L.orc_pfsp_new(inst)
        st, nb = orc.RunStats(), C.c_ulonglong(0)
        L.orc_pfsp_search_batched(h, lb, L.orc_taillard_best_ub(inst), m, M, C.byref(st), C.byref(nb))
        r = exp[(inst, lb)]
        assert (st.tree, st.sol, st.best) == (r["tree"], r["sol"], r["makespan"])
        L.orc_pfsp_free(h)


def test_ta014_search_shape(golden):
    # parents by depth and child-bound evaluations of the LB1_d run
    L = orc.oracle()
    h = L.orc_pfsp_new(14)
    st = orc.RunStats()
    L.orc_pfsp_search(h, 0, 1377, 0, C.byref(st))
    assert st.parents == 2573653 == golden("census")["14_0"]["parents"] and st.evals == 19121906
    assert list(st.depth_hist[:20]) == [1, 9, 39, 165, 639, 2252, 7003, 19626, 50445, 116575, 231523, 366493,
                                        464303, 466099, 366546, 240078, 144463, 72917, 21829, 2648]
    assert st.max_pool == 74
    L.orc_pfsp_free(h)


def test_nqueens_counts_match_golden(golden):
    L = orc.oracle()
    for r in golden("runs")["nqueens"]:
        if r["N"] > 12:
            continue
        for G in (1, 3):
            st = orc.RunStats()
            L.orc_nq_search(r["N"], G, 0, C.byref(st))
            assert (st.tree, st.sol) == (r["tree"], r["sol"])


def test_nqueens_labels_match_golden(golden):
    L = orc.oracle()
    for Ns, g in golden("nq_nodes").items():
        N = int(Ns)
        nodes = np.array(g["nodes"], dtype=np.uint8)
        labels = np.full(len(nodes) * N, 7, np.uint8)
        L.orc_nq_labels(N, 1, nodes, len(nodes), labels)
        for i, (nd, e) in enumerate(zip(nodes, g["labels"])):
            d = int(nd[0])
            assert labels[i * N + d:(i + 1) * N].tolist() == e
            assert np.all(labels[i * N:i * N + d] == 7)  # slots below depth are never written (K4)


@pytest.mark.skipif(not orc.have_ref(), reason="oracle/_ref not built (make -C oracle ref)")
def test_against_compiled_reference_random():
    L, R, Q = orc.oracle(), orc.ref_pfsp(), orc.ref_nq()
    rng = np.random.default_rng(99)
    for inst in (3, 14, 21, 27):
        n, _ = orc.taillard_dims(inst)
        ub = L.orc_taillard_best_ub(inst)
        ho, hr = L.orc_pfsp_new(inst), R.ref_pfsp_open(inst)
        for nd in orc.random_pfsp_nodes(rng, n, 150):
            d = int(nd[0])
            prmu = nd[2:].astype(np.int32).copy()
            a, b = np.zeros(n, np.int32), np.zeros(n, np.int32)
            L.orc_lb1_children(ho, prmu, d - 1, n, a)
            R.ref_lb1_children_bounds(hr, prmu, d - 1, n, b)
            assert np.array_equal(a, b)
            l2 = int(rng.integers(d, n + 1))
            for best in (ub, INT_MAX, ub - 25):
                assert L.orc_lb2(ho, prmu, d - 1, l2, best) == R.ref_lb2_bound(hr, prmu, d - 1, l2, best)
            assert L.orc_lb1(ho, prmu, d - 1, l2) == R.ref_lb1_bound(hr, prmu, d - 1, l2)
        # chunk level: reference generate_children on oracle bounds == oracle generate_children
        parents = orc.random_pfsp_nodes(rng, n, 40)
        nb = int((n - parents[:, 0]).sum())
        bounds = np.zeros(nb, np.int32)
        L.orc_evaluate_chunk(ho, 2, parents, 40, ub, bounds)
        outs = []
        for fn in (L.orc_generate_children, R.ref_generate_children):
            tree, sol, best = C.c_ulonglong(0), C.c_ulonglong(0), C.c_int(ub + 30)
            kids = np.zeros((nb, orc.NODE_I16), np.int16)
            k = fn(parents, 40, n, bounds, C.byref(tree), C.byref(sol), C.byref(best), kids)
            outs.append((k, tree.value, sol.value, best.value, kids[:k].tobytes()))
        assert outs[0] == outs[1]
        L.orc_pfsp_free(ho)
        R.ref_pfsp_close(hr)
    for N in (6, 13, 20):
        for nd in orc.random_nq_nodes(rng, N, 100):
            d = int(nd[0])
            board = np.ascontiguousarray(nd[1:])
            for k in range(d, N):
                assert L.orc_nq_is_safe(2, board, d, int(board[k])) == Q.ref_nq_is_safe(2, board, d, int(board[k]))
