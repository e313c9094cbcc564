"""bb_engine_keep_share: the device-side warm-up split of the multi-GPU searches (k_partition), for every record layout
the engines use - 20-job PFSP (m = 10, 20), the 50- / 100- / 200-job classes, and the three N-Queens records. Engines
that hold the same nodes and keep share `me` of `parts` must end up with disjoint shares whose union is the set, and
the share of a node may depend on its placed prefix only (that is what lets every GPU split without communication).
Runs on one GPU."""
import numpy as np
import pytest

import orc
from __graft_entry__ import load_package

pytestmark = pytest.mark.gpu
bb = load_package()


def _key(kind, n, row):
    d = int(row[0])
    return (d,) + tuple(int(v) for v in (row[2:2 + n] if kind == "pfsp" else row[1:1 + n]))


def _prefix(kind, n, row):
    d = int(row[0])
    return (d,) + tuple(int(v) for v in (row[2:2 + d] if kind == "pfsp" else row[1:1 + d]))


CASES = [("pfsp", dict(inst=14), {}), ("pfsp", dict(inst=21), {}), ("pfsp", dict(inst=31), {}), ("pfsp", dict(inst=61), {}),
         ("pfsp", dict(inst=91), {}), ("nq", dict(N=12), {}), ("nq", dict(N=17), {}), ("nq", dict(N=12), dict(packed=1)),
         ("nq", dict(N=18), dict(packed=1)), ("nq", dict(N=17), dict(packed=0))]


@pytest.mark.parametrize("kind,what,cfg", CASES)
def test_shares_are_disjoint_cover_the_pool_and_depend_on_the_prefix_only(kind, what, cfg):
    rng = np.random.default_rng(11)
    if kind == "pfsp":
        ins = bb.Instance(**what)
        n = ins.n
        eng = bb.Engine.pfsp(ins, 0, M=1 << 16, **cfg)
        nodes = orc.random_pfsp_nodes(rng, n, 3000, 0, n - 1, width=eng.node_i16)
        # the same prefixes again with the unscheduled jobs in another order: must land in the same share
        twins = nodes[:500].copy()
        for r in twins:
            d = int(r[0])
            r[2 + d:2 + n] = rng.permutation(r[2 + d:2 + n])
    else:
        n = what["N"]
        eng = bb.Engine.nqueens(n, 1, M=1 << 16, **cfg)
        nodes = orc.random_nq_nodes(rng, n, 3000, 0, n - 1)
        twins = nodes[:500].copy()
        for r in twins:
            d = int(r[0])
            r[1 + d:1 + n] = rng.permutation(r[1 + d:1 + n])
    allnodes = np.concatenate([nodes, twins])
    everything = sorted(_key(kind, n, r) for r in allnodes)
    for parts in (2, 5):
        shares, where = [], {}
        for me in range(parts):
            eng.reset()
            eng.push_nodes(allnodes)
            eng.keep_share(parts, me)
            got = eng.export_nodes() if eng.stats().pool_size else allnodes[:0]
            shares.append(sorted(_key(kind, n, r) for r in got))
            for r in got:
                assert where.setdefault(_prefix(kind, n, r), me) == me  # one prefix, one share
        assert sorted(k for s in shares for k in s) == everything       # disjoint (as multisets) and complete
        assert min(len(s) for s in shares) > len(everything) // (4 * parts)  # and roughly balanced
        # an engine keeps working on its share: the counters of the shares add up to those of the whole
        if kind == "nq":
            eng.reset()
            eng.push_nodes(allnodes)
            whole = eng.run(1)
            tot = 0
            for me in range(parts):
                eng.reset()
                eng.push_nodes(allnodes)
                eng.keep_share(parts, me)
                tot += eng.run(1).tree
            assert tot == whole.tree
    eng.close()
