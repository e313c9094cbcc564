#!/usr/bin/env python
"""Regenerate tests/golden/*.json from the UNMODIFIED reference compiled into oracle/_ref (make -C oracle ref).

Run in the build container (needs /root/reference): python tests/golden/make_golden.py
Every value in the fixtures is an output of the reference's own CPU code (libref_pfsp.so / libref_nqueens.so /
pfsp_c.out / nqueens_c.out); the oracle is NOT used to produce them.
"""
import ctypes as C
import json
import os
import re
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import orc  # noqa: E402

INT_MAX = 2**31 - 1


def fnv(a):
    h = 1469598103934665603
    for v in np.asarray(a, dtype=np.int64):
        h ^= int(v) & 0xFFFFFFFF
        h = (h * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return "%016x" % h


def tables():
    out = {}
    for inst in range(1, 121):
        t = orc.ref_tables(inst)
        out[str(inst)] = {"n": t.n, "m": t.m, "P": t.P, "ub": t.best_ub, "p": fnv(t.p), "heads": fnv(t.heads),
                          "tails": fnv(t.tails), "lags": fnv(t.lags), "johnson": fnv(t.johnson),
                          "pairs": fnv(np.concatenate([t.pa, t.pb, t.order])), "max_lag": int(t.lags.max())}
        if inst in (1, 14, 21, 24):
            out[str(inst)]["p_row0"] = t.p[:t.n].tolist()
            out[str(inst)]["heads_v"] = t.heads.tolist()
            out[str(inst)]["tails_v"] = t.tails.tolist()
            out[str(inst)]["johnson_pair0"] = t.johnson[:t.n].tolist()
    return out


def bounds():
    R = orc.ref_pfsp()
    rng = np.random.default_rng(20261106)
    out = {}
    for inst in (1, 5, 14, 17, 21, 24, 29):
        h = R.ref_pfsp_open(inst)
        n, m = orc.taillard_dims(inst)
        ub = orc.ref_tables(inst).best_ub
        nodes = orc.random_pfsp_nodes(rng, n, 96)
        nodes[0, 0], nodes[0, 1] = 0, -1
        nodes[0, 2:2 + n] = np.arange(n)  # the root
        nodes[1, 0], nodes[1, 1] = n - 1, n - 2  # a parent of one leaf
        rec = {"ub": ub, "nodes": nodes.tolist(), "lb1_d": [], "lb1": [], "lb2_ub": [], "lb2_inf": [],
               "lb2_tight": []}
        for nd in nodes:
            d = int(nd[0])
            prmu = nd[2:].astype(np.int32).copy()
            lbb = np.zeros(n, np.int32)
            R.ref_lb1_children_bounds(h, prmu, d - 1, n, lbb)
            row_d, row_1, row_2, row_2i, row_2t = [], [], [], [], []
            for k in range(d, n):
                row_d.append(int(lbb[prmu[k]]))
                ch = prmu.copy()
                ch[d], ch[k] = ch[k], ch[d]
                row_1.append(R.ref_lb1_bound(h, ch, d, n))
                row_2.append(R.ref_lb2_bound(h, ch, d, n, ub))
                row_2i.append(R.ref_lb2_bound(h, ch, d, n, INT_MAX))
                row_2t.append(R.ref_lb2_bound(h, ch, d, n, ub - 40))
            rec["lb1_d"].append(row_d)
            rec["lb1"].append(row_1)
            rec["lb2_ub"].append(row_2)
            rec["lb2_inf"].append(row_2i)
            rec["lb2_tight"].append(row_2t)
        R.ref_pfsp_close(h)
        out[str(inst)] = rec
    return out


def run_exe(name, args):
    txt = subprocess.run([os.path.join(orc.REF_DIR, name)] + args, capture_output=True, text=True, check=True,
                         cwd="/tmp").stdout
    g = lambda pat: re.search(pat, txt)
    r = {"tree": int(g(r"Size of the explored tree: (\d+)").group(1)),
         "sol": int(g(r"Number of explored solutions: (\d+)").group(1))}
    mk = g(r"Optimal makespan: (\d+)")
    if mk:
        r["makespan"] = int(mk.group(1))
    return r


def runs():
    out = {"pfsp": [], "nqueens": [], "survey": {}}
    for inst, lb in [(14, 0), (14, 1), (14, 2), (2, 0), (2, 2), (3, 0), (3, 2), (4, 0), (4, 2), (1, 2), (19, 0),
                     (19, 2), (11, 0), (12, 0), (13, 0), (15, 0), (16, 0)]:
        r = run_exe("pfsp_c.out", ["-i", str(inst), "-l", str(lb), "-u", "1"])
        r.update(inst=inst, lb=lb, ub=1)
        out["pfsp"].append(r)
        print(r, flush=True)
    for N in range(4, 15):
        r = run_exe("nqueens_c.out", ["-N", str(N), "-g", "1"])
        r.update(N=N)
        out["nqueens"].append(r)
        print(r, flush=True)
    # long runs measured once with the same binaries (hours of CPU, not regenerated here)
    out["survey"]["pfsp"] = [
        {"inst": 11, "lb": 2, "tree": 438563, "sol": 0, "makespan": 1582},
        {"inst": 12, "lb": 2, "tree": 666935, "sol": 0, "makespan": 1659},
        {"inst": 13, "lb": 2, "tree": 1300194, "sol": 0, "makespan": 1496},
        {"inst": 15, "lb": 2, "tree": 457191, "sol": 0, "makespan": 1419},
        {"inst": 16, "lb": 2, "tree": 2646205, "sol": 0, "makespan": 1397},
        {"inst": 17, "lb": 0, "tree": 577714583, "sol": 7805, "makespan": 1484},
        {"inst": 17, "lb": 2, "tree": 17472239, "sol": 0, "makespan": 1484},
        {"inst": 18, "lb": 0, "tree": 750624238, "sol": 14, "makespan": 1538},
        {"inst": 20, "lb": 0, "tree": 859257178, "sol": 3764, "makespan": 1591},
        {"inst": 20, "lb": 2, "tree": 4870386, "sol": 0, "makespan": 1591},
        {"inst": 29, "lb": 2, "tree": 9499307, "sol": 0, "makespan": 2237},
        {"inst": 29, "lb": 0, "tree": 3018629097, "sol": 85542, "makespan": 2237},
        {"inst": 30, "lb": 2, "tree": 13228600, "sol": 0, "makespan": 2178},
        {"inst": 30, "lb": 0, "tree": 111887480, "sol": 82, "makespan": 2178},
        {"inst": 22, "lb": 2, "tree": 14561974, "sol": 0, "makespan": 2099},
        {"inst": 22, "lb": 0, "tree": 711291688, "sol": 4346, "makespan": 2099},
    ]
    out["survey"]["nqueens"] = [
        {"N": 15, "tree": 171129071, "sol": 2279184},
        {"N": 16, "tree": 1141190302, "sol": 14772512},
        {"N": 17, "tree": 8017021931, "sol": 95815104},
        {"N": 18, "tree": 59365844490, "sol": 666090624},  # printed by nqueens_c.out -N 18 -g 1 (ref_gpu_runs.json, 9067 s)
    ]
    return out


def census():
    """search-shape statistics of the reference DFS on ta014 (parents by depth, evaluations)"""
    R = orc.ref_pfsp()
    out = {}
    for inst, lb in [(14, 0), (14, 2)]:
        h = R.ref_pfsp_open(inst)
        ub = orc.ref_tables(inst).best_ub
        n, _ = orc.taillard_dims(inst)
        # the reference has no counters of its own; count around its decompose() through the budgeted loop
        best, tree, sol = C.c_int(ub), C.c_ulonglong(0), C.c_ulonglong(0)
        parents = R.ref_pfsp_search(h, lb, C.byref(best), C.byref(tree), C.byref(sol), 0)
        out["%d_%d" % (inst, lb)] = {"parents": int(parents), "tree": tree.value, "sol": sol.value}
        R.ref_pfsp_close(h)
    return out


def nq_nodes():
    Q = orc.ref_nq()
    rng = np.random.default_rng(7)
    out = {}
    for N in (8, 14, 17, 20):
        nodes = orc.random_nq_nodes(rng, N, 64, 0, N)
        labels = []
        for nd in nodes:
            d = int(nd[0])
            board = np.ascontiguousarray(nd[1:])
            labels.append([int(Q.ref_nq_is_safe(1, board, d, int(board[k]))) for k in range(d, N)])
        out[str(N)] = {"nodes": nodes.tolist(), "labels": labels}
    return out


if __name__ == "__main__":
    assert orc.have_ref(), "build oracle/_ref first: make -C oracle ref"
    for name, fn in (("tables", tables), ("bounds", bounds), ("census", census), ("nq_nodes", nq_nodes),
                     ("runs", runs)):
        with open(os.path.join(HERE, name + ".json"), "w") as f:
            json.dump(fn(), f, separators=(",", ":"))
        print("wrote", name)
