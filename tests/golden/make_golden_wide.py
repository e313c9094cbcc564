#!/usr/bin/env python
"""tests/golden/wide.json: golden vectors for the 50-, 100- and 200-job Taillard classes from the REFERENCE rebuilt with
MAX_JOBS 50 / 100 / 200 (oracle/_ref/libref_pfsp<NJ>.so, oracle/_ref/pfsp_c<NJ>.out, `make -C oracle ref50 ref100 ref200`:
pfsp/README.md:52 asks for exactly that edit). Table hashes, the bounds[] a chunk of random parents gets from
lb1_children_bounds / lb1_bound / lb2_bound (c_bound_simple.c:143-244, c_bound_johnson.c:239-254) in evaluate_gpu's
layout, and the printed counts of the instances the reference solves at the root. Run in the build container:
python tests/golden/make_golden_wide.py"""
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

def klass(inst):
    return 50 if inst <= 60 else 100 if inst <= 90 else 200


def main():
    out = {"what": "reference rebuilt with MAX_JOBS 50 / 100 / 200 (make -C oracle ref50 ref100 ref200); see make_golden_wide.py"}
    for inst in (31, 41, 51, 61, 71, 81, 91, 101):
        R = orc.ref_pfsp(klass(inst))
        W = 2 + klass(inst)
        t = orc.ref_tables_of(R, inst)
        r = R.ref_pfsp_open(inst)
        n, best = t.n, t.best_ub
        rng = np.random.default_rng(7000 + inst)
        parents = orc.random_pfsp_nodes(rng, n, 24 if inst <= 60 else 10 if inst <= 90 else 5, 0, n - 2, width=W)
        bounds = {"0": [], "1": [], "2": []}
        for nd in parents:
            d = int(nd[0])
            prmu = nd[2:2 + n].astype(np.int32)
            lbb = np.zeros(n, np.int32)
            R.ref_lb1_children_bounds(r, prmu, d - 1, n, lbb)
            for k in range(d, n):
                child = prmu.copy()
                child[d], child[k] = child[k], child[d]
                bounds["0"].append(int(lbb[prmu[k]]))
                bounds["1"].append(int(R.ref_lb1_bound(r, child, d, n)))
                bounds["2"].append(int(R.ref_lb2_bound(r, child, d, n, best)))
        R.ref_pfsp_close(r)
        out[str(inst)] = {"fnv": {k: "%016x" % orc.fnv(a) for k, a in
                                  zip(("p", "heads", "tails", "lags", "johnson"), t.arrays())},
                          "best": best, "parents": parents.reshape(-1).tolist(), "bounds": bounds}
    runs = []
    # ta034 under LB1_d and most 100-job instances do not finish in minutes on a CPU; these are solved at the root
    for inst, lb in ((32, 0), (32, 1), (32, 2), (34, 2), (66, 0), (66, 1), (66, 2), (69, 0), (69, 2), (63, 2), (65, 2),
                     (61, 2), (94, 0), (94, 2), (97, 0), (97, 1), (97, 2), (100, 0), (100, 2)):
        exe = os.path.join(orc.REF_DIR, "pfsp_c%d.out" % klass(inst))
        txt = subprocess.run([exe, "-i", str(inst), "-l", str(lb), "-u", "1"], capture_output=True, text=True,
                             cwd="/tmp", timeout=120).stdout
        runs.append({"inst": inst, "lb": lb,
                     "tree": int(re.findall(r"Size of the explored tree: (\d+)", txt)[-1]),
                     "sol": int(re.findall(r"Number of explored solutions: (\d+)", txt)[-1]),
                     "makespan": int(re.findall(r"Optimal makespan: (\d+)", txt)[-1])})
    out["runs"] = runs
    json.dump(out, open(os.path.join(HERE, "wide.json"), "w"))
    print({k: (v if k == "runs" else "...") for k, v in out.items()})


if __name__ == "__main__":
    main()
