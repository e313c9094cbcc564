"""ctypes bindings for the CPU checker (oracle/liboracle.so) and, when present, the compiled reference
(oracle/_ref/libref_*.so). TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs. Never by the product package."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
REFERENCE_SRC = "/root/reference"

MAX_JOBS = 20
NODE_I16 = 2 + MAX_JOBS
MAX_QUEENS = 20
NQ_REC = 1 + MAX_QUEENS

i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
i16p = np.ctypeslib.ndpointer(dtype=np.int16, flags="C_CONTIGUOUS")
u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
ullp = C.POINTER(C.c_ulonglong)


class RunStats(C.Structure):
    _fields_ = [("tree", C.c_ulonglong), ("sol", C.c_ulonglong), ("parents", C.c_ulonglong),
                ("evals", C.c_ulonglong), ("depth_hist", C.c_ulonglong * 64), ("best", C.c_int),
                ("max_pool", C.c_int)]


def build_oracle(ref=False):
    """(re)build liboracle.so; with ref=True also oracle/_ref when the reference sources are present."""
    subprocess.run(["make", "-C", ORACLE_DIR, "-s"], check=True)
    if ref and os.path.isdir(REFERENCE_SRC):
        subprocess.run(["make", "-C", ORACLE_DIR, "-s", "ref"], check=True, stderr=subprocess.DEVNULL)


_oracle = {}


def oracle(max_jobs=20):
    """the CPU checker; max_jobs = 50 / 100 loads the same code built with 50- / 100-job node records (liboracle<NJ>.so)"""
    if max_jobs in _oracle:
        return _oracle[max_jobs]
    path = os.path.join(ORACLE_DIR, "liboracle.so" if max_jobs == 20 else f"liboracle{max_jobs}.so")
    if not os.path.exists(path):
        build_oracle()
    L = C.CDLL(path)
    L.orc_pfsp_new.restype = C.c_void_p
    L.orc_pfsp_new.argtypes = [C.c_int]
    L.orc_pfsp_from_ptimes.restype = C.c_void_p
    L.orc_pfsp_from_ptimes.argtypes = [C.c_int, C.c_int, i32p]
    L.orc_pfsp_free.argtypes = [C.c_void_p]
    L.orc_pfsp_tables.argtypes = [C.c_void_p] + [i32p] * 8
    L.orc_taillard_ptimes.argtypes = [C.c_int, i32p]
    L.orc_taillard_seed.restype = C.c_long
    L.orc_fnv1a_i32.restype = C.c_uint64
    L.orc_fnv1a_i32.argtypes = [i32p, C.c_int]
    L.orc_makespan.argtypes = [C.c_void_p, i32p]
    L.orc_lb1.argtypes = [C.c_void_p, i32p, C.c_int, C.c_int]
    L.orc_lb1_children.argtypes = [C.c_void_p, i32p, C.c_int, C.c_int, i32p]
    L.orc_lb2.argtypes = [C.c_void_p, i32p, C.c_int, C.c_int, C.c_int]
    L.orc_lb2_census.argtypes = [C.c_void_p, i32p, C.c_int, C.c_int, C.c_int, ullp, ullp]
    L.orc_batch_index.argtypes = [i16p, C.c_int, C.c_int, i32p, i32p]
    L.orc_evaluate_chunk.argtypes = [C.c_void_p, C.c_int, i16p, C.c_int, C.c_int, i32p]
    L.orc_generate_children.argtypes = [i16p, C.c_int, C.c_int, i32p, ullp, ullp, C.POINTER(C.c_int), i16p]
    L.orc_pfsp_search.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_longlong, C.POINTER(RunStats)]
    L.orc_pfsp_search_batched.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(RunStats), ullp]
    L.orc_nq_is_safe.argtypes = [C.c_int, u8p, C.c_int, C.c_int]
    L.orc_nq_labels.argtypes = [C.c_int, C.c_int, u8p, C.c_int, u8p]
    L.orc_nq_generate_children.argtypes = [C.c_int, u8p, C.c_int, u8p, ullp, ullp, u8p]
    L.orc_nq_search.argtypes = [C.c_int, C.c_int, C.c_longlong, C.POINTER(RunStats)]
    L.node_i16 = 2 + max_jobs
    _oracle[max_jobs] = L
    return L


def have_ref():
    return os.path.exists(os.path.join(REF_DIR, "libref_pfsp.so"))


_ref_pfsp = {}
_ref_nq = None


def have_ref_wide(max_jobs=50):
    return os.path.exists(os.path.join(REF_DIR, f"libref_pfsp{max_jobs}.so"))


def have_ref50():
    return have_ref_wide(50)


def ref_pfsp(max_jobs=20):
    """the compiled reference; max_jobs = 50 / 100 is the build with MAX_JOBS edited (make -C oracle ref50 ref100)"""
    if max_jobs not in _ref_pfsp:
        L = C.CDLL(os.path.join(REF_DIR, "libref_pfsp.so" if max_jobs == 20 else f"libref_pfsp{max_jobs}.so"))
        L.ref_pfsp_open.restype = C.c_void_p
        L.ref_pfsp_open.argtypes = [C.c_int]
        L.ref_pfsp_close.argtypes = [C.c_void_p]
        L.ref_pfsp_dims.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 4
        L.ref_pfsp_tables.argtypes = [C.c_void_p] + [i32p] * 8
        L.ref_lb1_bound.argtypes = [C.c_void_p, i32p, C.c_int, C.c_int]
        L.ref_lb1_children_bounds.argtypes = [C.c_void_p, i32p, C.c_int, C.c_int, i32p]
        L.ref_lb2_bound.argtypes = [C.c_void_p, i32p, C.c_int, C.c_int, C.c_int]
        L.ref_eval_solution.argtypes = [C.c_void_p, i32p]
        L.ref_pfsp_search.restype = C.c_ulonglong
        L.ref_pfsp_search.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), ullp, ullp, C.c_longlong]
        L.ref_pfsp_decompose.argtypes = [C.c_void_p, C.c_int, i16p, C.c_int, C.POINTER(C.c_int), ullp, ullp, i16p]
        L.ref_generate_children.argtypes = [i16p, C.c_int, C.c_int, i32p, ullp, ullp, C.POINTER(C.c_int), i16p]
        L.node_i16 = 2 + max_jobs
        _ref_pfsp[max_jobs] = L
    return _ref_pfsp[max_jobs]


def ref_nq():
    global _ref_nq
    if _ref_nq is None:
        L = C.CDLL(os.path.join(REF_DIR, "libref_nqueens.so"))
        L.ref_nq_is_safe.argtypes = [C.c_int, u8p, C.c_int, C.c_int]
        L.ref_nq_decompose.argtypes = [C.c_int, C.c_int, u8p, C.c_int, ullp, ullp, u8p]
        L.ref_nq_search.restype = C.c_ulonglong
        L.ref_nq_search.argtypes = [C.c_int, C.c_int, ullp, ullp, C.c_longlong]
        _ref_nq = L
    return _ref_nq


class Tables:
    """numpy view of one instance's precomputed tables from either implementation."""

    def __init__(self, n, m, P):
        self.n, self.m, self.P = n, m, P
        z = lambda k: np.zeros(max(k, 1), dtype=np.int32)
        self.p, self.heads, self.tails = z(n * m), z(m), z(m)
        self.lags, self.johnson = z(P * n), z(P * n)
        self.pa, self.pb, self.order = z(P), z(P), z(P)

    def arrays(self):
        return [self.p, self.heads, self.tails, self.lags, self.johnson, self.pa, self.pb, self.order]


def taillard_dims(inst):
    L = oracle()
    return L.orc_taillard_jobs(inst), L.orc_taillard_machines(inst)


def oracle_tables(inst):
    L = oracle()
    n, m = taillard_dims(inst)
    h = L.orc_pfsp_new(inst)
    t = Tables(n, m, m * (m - 1) // 2)
    L.orc_pfsp_tables(h, *t.arrays())
    L.orc_pfsp_free(h)
    return t


def ref_tables(inst):
    return ref_tables_of(ref_pfsp(), inst)


def ref_tables_of(R, inst):
    h = R.ref_pfsp_open(inst)
    n, m, P, ub = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    R.ref_pfsp_dims(h, n, m, P, ub)
    t = Tables(n.value, m.value, P.value)
    R.ref_pfsp_tables(h, *t.arrays())
    R.ref_pfsp_close(h)
    t.best_ub = ub.value
    return t


def fnv(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return int(oracle().orc_fnv1a_i32(a, a.size))


def random_pfsp_nodes(rng, n, count, min_depth=0, max_depth=None, width=NODE_I16):
    """random partial permutations as flat reference Node records [depth, limit1, prmu[MAX_JOBS]]"""
    max_depth = n - 1 if max_depth is None else max_depth
    out = np.zeros((count, width), dtype=np.int16)
    for i in range(count):
        d = int(rng.integers(min_depth, max_depth + 1))
        out[i, 0] = d
        out[i, 1] = d - 1
        out[i, 2:2 + n] = rng.permutation(n)
    return out


def random_nq_nodes(rng, N, count, min_depth=0, max_depth=None):
    max_depth = N if max_depth is None else max_depth
    out = np.zeros((count, NQ_REC), dtype=np.uint8)
    for i in range(count):
        out[i, 0] = int(rng.integers(min_depth, max_depth + 1))
        out[i, 1:1 + N] = rng.permutation(N)
    return out


def oracle_one_level(inst_handle, n, lb, parents, best, L=None):
    """one batch through the oracle: evaluate_gpu's bounds[] (orc_evaluate_chunk) then generate_children
    (PFSP_lib.h:51-95). Returns (children, tree increment, sol increment, incumbent, child bounds evaluated)."""
    L = L or oracle()
    parents = np.ascontiguousarray(parents, dtype=np.int16)
    nb = int((n - parents[:, 0]).sum())
    bounds = np.zeros(max(nb, 1), np.int32)
    L.orc_evaluate_chunk(inst_handle, lb, parents, len(parents), best, bounds)
    tree, sol, b = C.c_ulonglong(0), C.c_ulonglong(0), C.c_int(best)
    kids = np.zeros((max(nb, 1), L.node_i16), np.int16)
    k = L.orc_generate_children(parents, len(parents), n, bounds, C.byref(tree), C.byref(sol), C.byref(b), kids)
    return kids[:k], tree.value, sol.value, b.value, nb
