"""Whole-tree counts of the headline configurations against the numbers the REFERENCE ITSELF printed
(tests/golden/ref_gpu_runs.json: the reference's GPU drivers rebuilt for sm_100a and run on the B200 box, and its CPU
programs run in the build container - scripts/gpu_pin_reference.sh, scripts/pins_to_golden.py). With the upper bound
at the optimum the explored tree is a constant of (instance, bound), so these are exact parity targets
(pfsp_multigpu_cuda.c:503-510, nqueens_multigpu_cuda.cu:364-369)."""
import pytest

from conftest import load_golden
from __graft_entry__ import load_package

pytestmark = pytest.mark.gpu
bb = load_package()
RUNS = load_golden("ref_gpu_runs")["runs"]


@pytest.mark.parametrize("key", sorted(RUNS))
def test_whole_tree_equals_what_the_reference_printed(key):
    r = RUNS[key]
    parts = key.split(":")
    if parts[0] == "pfsp":
        inst, lb = int(parts[1]), int(parts[2])
        res, per = bb.pfsp_search(bb.Instance(inst=inst), lb, M=1 << 22)
        assert (res.tree, res.sol, res.best) == (r["tree"], r["sol"], r["makespan"]), (key, r["command"])
        assert per[0].parents == res.tree + 1
    else:
        N = int(parts[1])
        res, per = bb.nq_search(N, 1, M=1 << 23)
        assert (res.tree, res.sol) == (r["tree"], r["sol"]), (key, r["command"])
        assert per[0].parents == res.tree - res.sol + 1


@pytest.mark.parametrize("packed", [0, 1])
def test_earlier_kernel_generations_agree(packed):
    """bb_config.packed selects the kernel generation; all of them walk the same tree"""
    ins = bb.Instance(inst=14)
    for lb, exp in ((0, (2573652, 2648, 1377)), (1, (2573652, 2648, 1377)), (2, (144639, 0, 1377))):
        res, _ = bb.pfsp_search(ins, lb, M=1 << 18, packed=packed)
        assert (res.tree, res.sol, res.best) == exp, (packed, lb)
    res, _ = bb.nq_search(13, 1, M=1 << 18, packed=packed)
    assert (res.tree, res.sol) == (4674889, 73712)
    res, _ = bb.nq_search(17 if packed else 16, 1, M=1 << 20, packed=packed)
    assert (res.tree, res.sol) == ((8017021931, 95815104) if packed else (1141190302, 14772512))


def test_every_nqueens_node_through_the_pool_agrees():
    """bb_config.packed = 2: the kernel of the default engine without the in-register enumeration of the last levels"""
    for N, exp in ((13, (4674889, 73712)), (16, (1141190302, 14772512))):
        res, per = bb.nq_search(N, 1, M=1 << 20, packed=2)
        assert (res.tree, res.sol) == exp and per[0].parents == res.tree - res.sol + 1
