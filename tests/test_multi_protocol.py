"""The N > 1 path on a CPU box: two gloo ranks run the product's team protocol (bb_team_worker: requests, donations,
peer copies, acknowledgements, termination detection) around a host stand-in engine (tests/team_mock.py). The sum of
the explored trees must be the reference's N-Queens count however the work moved between the ranks."""
import multiprocessing as mp
import os
import random

import pytest

import team_mock

NQ = {8: (2056, 92), 9: (8393, 352)}  # nqueens_c.out -N 8 / -N 9 (tests/golden/runs.json holds 8; 9 = OEIS A000170)


def run_job(N, lopsided, world=2):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = random.randint(20000, 40000)
    tag = f"bbt{os.getpid()}_{random.randint(0, 1 << 30)}"
    procs = [ctx.Process(target=team_mock.worker_main, args=(r, world, port, N, tag, lopsided, q)) for r in range(world)]
    for p in procs:
        p.start()
    outs = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return sorted(outs)


@pytest.mark.timeout(300)
def test_two_ranks_share_by_hash_and_terminate():
    outs = run_job(8, lopsided=False)
    total = outs[0][2]
    assert (total[0], total[1]) == NQ[8]
    assert outs[0][2] == outs[1][2]


@pytest.mark.timeout(300)
def test_idle_rank_steals_from_the_loaded_one():
    """rank 1 starts with an empty pool: everything it explores it must have stolen through the protocol"""
    outs = run_job(9, lopsided=True)
    (r0, mine0, total), (r1, mine1, _) = outs
    assert (total[0], total[1]) == NQ[9]
    assert mine1[0] > 0 and mine1[2] > 0 and mine1[3] > 0, mine1  # explored, successful steals, nodes pulled


@pytest.mark.timeout(300)
def test_four_ranks_one_loaded():
    """three idle ranks hunt one loaded rank (and later each other): requests are serialised per victim, donations
    go one at a time, every rank leaves through the epoch-checked termination, and no node is lost or duplicated"""
    outs = run_job(9, lopsided=True, world=4)
    total = outs[0][2]
    assert (total[0], total[1]) == NQ[9]
    assert all(o[2] == total for o in outs)
    assert sum(1 for o in outs[1:] if o[1][0] > 0) >= 2  # at least two of the idle ranks got work
