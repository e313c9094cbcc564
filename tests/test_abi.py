"""CPU-only checks of the product's C ABI: the library loads, exports every symbol include/bb_gpu.h declares,
its host-side table precompute (C, no GPU) is bit-identical to the reference's, and compute entry points fail
loudly without a CUDA device (no CPU fallback)."""
import os
import re

import numpy as np
import pytest

import orc
from __graft_entry__ import ROOT, load_package

bb = load_package()


def fnv_hex(a):
    return "%016x" % orc.fnv(a)


def declared_functions():
    src = open(os.path.join(ROOT, "include", "bb_gpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"\{[^{}]*\}", "{}", src)  # struct / enum bodies (function-pointer members are not exports)
    names = re.findall(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\([^;{}]*\)\s*;", src)
    return sorted(set(n for n in names if not n.startswith("defined")))


def test_library_exports_every_declared_symbol():
    names = declared_functions()
    assert len(names) >= 30
    for n in names:
        assert hasattr(bb.lib, n), f"libbbgpu.so does not export {n}"
    assert set(names) == set(bb.EXPORTS), set(names) ^ set(bb.EXPORTS)


def test_product_never_references_the_oracle():
    # the oracle is a checker: nothing under the package or include/ may name it
    pkg = os.path.join(ROOT, "dist-gpu-accelerated-tree-search_b200")
    for base, _, files in os.walk(pkg):
        if "build" in base.split(os.sep):
            continue
        for f in files:
            if f.endswith((".c", ".cu", ".cuh", ".h", ".py", "Makefile")):
                txt = open(os.path.join(base, f), errors="ignore").read()
                assert "liboracle" not in txt and "oracle/" not in txt and "import orc" not in txt, (base, f)


def test_host_tables_match_reference_hashes(golden):
    g = golden("tables")
    for inst in range(1, 121):
        ins = bb.Instance(inst=inst)
        t, e = ins.tables(), g[str(inst)]
        assert (ins.n, ins.m, ins.P, ins.best_known) == (e["n"], e["m"], e["P"], e["ub"])
        assert fnv_hex(t["p"]) == e["p"], inst
        assert fnv_hex(t["heads"]) == e["heads"] and fnv_hex(t["tails"]) == e["tails"], inst
        assert fnv_hex(t["lags"]) == e["lags"] and fnv_hex(t["johnson"]) == e["johnson"], inst
        assert fnv_hex(np.concatenate([t["pa"], t["pb"], t["order"]])) == e["pairs"], inst
        assert bb.lib.bb_taillard_nb_jobs(inst) == e["n"] and bb.lib.bb_taillard_nb_machines(inst) == e["m"]
        ins.close()


def test_host_tables_match_oracle_on_custom_matrices():
    rng = np.random.default_rng(5)
    L = orc.oracle()
    for n, m in [(3, 2), (7, 3), (12, 7), (20, 20), (9, 1)]:
        p = rng.integers(1, 100, size=(m, n)).astype(np.int32)
        p[:, : n // 2] = p[:, n // 2: 2 * (n // 2)]  # force ties in Johnson's rule
        ins = bb.Instance(p_times=p)
        h = L.orc_pfsp_from_ptimes(n, m, np.ascontiguousarray(p.reshape(-1)))
        t = orc.Tables(n, m, m * (m - 1) // 2)
        L.orc_pfsp_tables(h, *t.arrays())
        got = ins.tables()
        for name, arr in zip("p heads tails lags johnson pa pb order".split(), t.arrays()):
            assert np.array_equal(got[name], arr[:got[name].size]), (n, m, name)
        L.orc_pfsp_free(h)


def test_op_model_matches_reference_formulas():
    # PFSP_gpu_lib.cu:213-267 at ta021, lim = 12 (7308 ops)
    assert bb.lib.P_of(20) == 190
    assert bb.lib.flop_lb2(20, 20, 12) == 12 * 39 + 8 * 20 + 190 * 36 == 7468
    assert bb.lib.flop_lb1(20, 10, 5) == 5 * 19 + 15 * 10 + 2 + 36
    assert bb.lib.bytes_per_inv_lb1(20, 10) == 4 * 201
    assert bb.lib.bytes_per_inv_lb2(20, 20, 12) == 4 * (400 + 3 * 190 * 8 + 1)


def test_unsupported_sizes_are_refused_loudly():
    ins = bb.Instance(inst=111)  # 500 jobs: beyond the 50-, 100- and 200-job classes this build covers (pfsp/README.md:52)
    with pytest.raises(bb.BBError, match="500 jobs"):
        bb.Engine.pfsp(ins, 0)
    with pytest.raises(bb.BBError):
        bb.Engine.nqueens(21)
    with pytest.raises(bb.BBError):
        bb.Instance(inst=121)


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    ins = bb.Instance(inst=14)
    with pytest.raises(bb.BBError, match="CUDA"):
        bb.Engine.pfsp(ins, 1)
    with pytest.raises(bb.BBError, match="CUDA"):
        bb.nq_search(8)
