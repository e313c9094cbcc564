"""ctypes binding of libbbgpu.so (C ABI in include/bb_gpu.h). Mirrors the C names one to one."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbbgpu.so")

MAX_JOBS = 20
MAX_MACHINES = 20
MAX_QUEENS = 20
MAX_JOBS_WIDE = 200
NODE_I16_WIDE = 2 + 50  # the reference's Node when built with MAX_JOBS 50 (engines of 21..50-job instances)
NODE_I16_WIDE100 = 2 + 100  # ... with MAX_JOBS 100 (engines of 51..100-job instances)
NODE_I16_WIDE200 = 2 + 200  # ... with MAX_JOBS 200 (engines of 101..200-job instances)
NODE_I16 = 2 + MAX_JOBS     # reference PFSP Node: int16 depth, limit1, prmu[20]  (pfsp/lib/PFSP_node.h:15-20)
NQ_REC = 1 + MAX_QUEENS     # reference N-Queens Node: uint8 depth, board[20]     (nqueens/lib/NQueens_node.h:13-17)
INT_MAX = 2**31 - 1


class BBError(RuntimeError):
    pass


if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `make -C {_HERE}/csrc` (or __graft_entry__.build()). "
        "There is no CPU fallback for this package.")

lib = C.CDLL(LIB_PATH)


class Lb1BoundData(C.Structure):  # pfsp/lib/c_bound_simple.h:14-21
    _fields_ = [("p_times", C.POINTER(C.c_int)), ("min_heads", C.POINTER(C.c_int)),
                ("min_tails", C.POINTER(C.c_int)), ("nb_jobs", C.c_int), ("nb_machines", C.c_int)]


class Lb2BoundData(C.Structure):  # pfsp/lib/c_bound_johnson.h:16-27
    _fields_ = [("johnson_schedules", C.POINTER(C.c_int)), ("lags", C.POINTER(C.c_int)),
                ("machine_pairs_1", C.POINTER(C.c_int)), ("machine_pairs_2", C.POINTER(C.c_int)),
                ("machine_pair_order", C.POINTER(C.c_int)), ("nb_machine_pairs", C.c_int), ("nb_jobs", C.c_int),
                ("nb_machines", C.c_int)]


class PfspInstance(C.Structure):
    _fields_ = [("nb_jobs", C.c_int), ("nb_machines", C.c_int), ("nb_machine_pairs", C.c_int),
                ("best_known", C.c_int), ("inst", C.c_int), ("p_times", C.POINTER(C.c_int)),
                ("min_heads", C.POINTER(C.c_int)), ("min_tails", C.POINTER(C.c_int)),
                ("machine_pairs_1", C.POINTER(C.c_int)), ("machine_pairs_2", C.POINTER(C.c_int)),
                ("machine_pair_order", C.POINTER(C.c_int)), ("lags", C.POINTER(C.c_int)),
                ("johnson_schedules", C.POINTER(C.c_int))]


class Config(C.Structure):
    _fields_ = [("device", C.c_int), ("M", C.c_int), ("m", C.c_int), ("capacity", C.c_longlong),
                ("iters_per_launch", C.c_int), ("lb2_prefilter", C.c_int), ("lb2_staged", C.c_int),
                ("lb2_stage_a", C.c_int), ("lb2_table", C.c_int), ("packed", C.c_int)]


class Stats(C.Structure):
    _fields_ = [("tree", C.c_ulonglong), ("sol", C.c_ulonglong), ("parents", C.c_ulonglong),
                ("evals", C.c_ulonglong), ("lb2_evals", C.c_ulonglong), ("lb2_pairs", C.c_ulonglong),
                ("iterations", C.c_ulonglong), ("launches", C.c_ulonglong), ("pool_size", C.c_longlong),
                ("max_pool", C.c_longlong), ("best", C.c_int), ("kernel_ms", C.c_double), ("span_ms", C.c_double),
                ("h2d_bytes", C.c_ulonglong), ("d2h_bytes", C.c_ulonglong), ("table_bytes", C.c_ulonglong),
                ("grid", C.c_int), ("smem_bytes", C.c_int), ("record_bytes", C.c_int),
                ("steals", C.c_ulonglong), ("steals_ok", C.c_ulonglong), ("donated", C.c_ulonglong)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class SearchResult(C.Structure):
    _fields_ = [("tree", C.c_ulonglong), ("sol", C.c_ulonglong), ("best", C.c_int), ("elapsed_s", C.c_double),
                ("kernel_ms_max", C.c_double), ("steals", C.c_ulonglong), ("steals_ok", C.c_ulonglong),
                ("search_s", C.c_double), ("nodes_moved", C.c_ulonglong), ("warm_tree", C.c_ulonglong),
                ("warm_sol", C.c_ulonglong), ("warm_s", C.c_double), ("setup_s", C.c_double)]


class TeamResult(C.Structure):
    _fields_ = [("best", C.c_int), ("steals", C.c_ulonglong), ("steals_ok", C.c_ulonglong),
                ("donated", C.c_ulonglong), ("busy_ms", C.c_double)]


class TeamOps(C.Structure):  # bb_team_ops: how the team protocol reaches a worker's engine
    RUN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_longlong, C.POINTER(C.c_longlong), C.POINTER(C.c_int),
                      C.POINTER(C.c_double))
    DONATE = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_longlong, C.c_longlong, C.POINTER(C.c_int),
                         C.POINTER(C.c_longlong), C.POINTER(C.c_longlong))
    DONE = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int)
    PULL = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_longlong, C.c_longlong)
    SETBEST = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int)
    _fields_ = [("self", C.c_void_p), ("run", RUN), ("donate", DONATE), ("donate_done", DONE), ("pull", PULL),
                ("set_best", SETBEST)]


_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_vp = C.c_void_p

# every symbol include/bb_gpu.h declares; tests/test_abi.py checks the library exports all of them
EXPORTS = [
    "bb_last_error", "bb_last_cuda_error", "evaluate_gpu", "lb1_alloc_gpu", "lb2_alloc_gpu", "P_of", "flop_lb1",
    "bytes_per_inv_lb1", "flop_lb2", "bytes_per_inv_lb2", "nq_evaluate_gpu", "bb_taillard_nb_jobs",
    "bb_taillard_nb_machines", "bb_taillard_best_ub", "bb_taillard_seed", "bb_taillard_processing_times",
    "bb_pfsp_instance_taillard", "bb_pfsp_instance_from_ptimes", "bb_pfsp_instance_free", "bb_config_default",
    "bb_pfsp_engine_create", "bb_nq_engine_create", "bb_engine_destroy", "bb_release_cached_memory", "bb_engine_reset", "bb_engine_push_root",
    "bb_engine_push_nodes", "bb_engine_export_nodes", "bb_engine_set_best", "bb_engine_get_stats", "bb_engine_run",
    "bb_engine_keep_share",
    "bb_pfsp_search", "bb_nq_search", "bb_pfsp_expand_static", "bb_pfsp_search_dist", "bb_nq_search_dist", "bb_pfsp_dist_open", "bb_nq_dist_open", "bb_dist_search",
    "bb_dist_close",
    "bb_team_bytes", "bb_team_init", "bb_team_reset", "bb_team_wait_ready", "bb_team_barrier", "bb_team_fail", "bb_team_worker",
    "bb_team_best", "bb_measure_slot_step_peak",
]

lib.bb_last_error.restype = C.c_char_p
lib.bb_taillard_seed.restype = C.c_long
lib.bb_taillard_processing_times.argtypes = [C.c_int, _i32p]
lib.bb_pfsp_instance_taillard.restype = C.POINTER(PfspInstance)
lib.bb_pfsp_instance_taillard.argtypes = [C.c_int]
lib.bb_pfsp_instance_from_ptimes.restype = C.POINTER(PfspInstance)
lib.bb_pfsp_instance_from_ptimes.argtypes = [C.c_int, C.c_int, _i32p]
lib.bb_pfsp_instance_free.argtypes = [C.POINTER(PfspInstance)]
lib.bb_config_default.argtypes = [C.POINTER(Config)]
lib.bb_pfsp_engine_create.restype = _vp
lib.bb_pfsp_engine_create.argtypes = [C.POINTER(PfspInstance), C.c_int, C.POINTER(Config)]
lib.bb_nq_engine_create.restype = _vp
lib.bb_nq_engine_create.argtypes = [C.c_int, C.c_int, C.POINTER(Config)]
lib.bb_engine_destroy.argtypes = [_vp]
lib.bb_release_cached_memory.argtypes = [C.c_int]
lib.bb_engine_reset.argtypes = [_vp, C.c_int]
lib.bb_engine_push_root.argtypes = [_vp]
lib.bb_engine_push_nodes.argtypes = [_vp, _vp, C.c_int]
lib.bb_engine_keep_share.argtypes = [_vp, C.c_int, C.c_int]
lib.bb_engine_export_nodes.restype = C.c_longlong
lib.bb_engine_export_nodes.argtypes = [_vp, _vp, C.c_longlong]
lib.bb_engine_set_best.argtypes = [_vp, C.c_int]
lib.bb_engine_get_stats.argtypes = [_vp, C.POINTER(Stats)]
lib.bb_engine_run.argtypes = [_vp, C.c_longlong, C.POINTER(Stats)]
lib.bb_pfsp_search.argtypes = [C.POINTER(PfspInstance), C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(Config),
                               C.POINTER(SearchResult), C.POINTER(Stats)]
lib.bb_nq_search.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(Config), C.POINTER(SearchResult),
                             C.POINTER(Stats)]
lib.bb_pfsp_search_dist.argtypes = [C.POINTER(PfspInstance), C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int,
                                    C.POINTER(Config), C.POINTER(SearchResult), C.POINTER(Stats)]
lib.bb_nq_search_dist.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int, C.POINTER(Config),
                                  C.POINTER(SearchResult), C.POINTER(Stats)]
lib.bb_pfsp_dist_open.restype = _vp
lib.bb_pfsp_dist_open.argtypes = [C.POINTER(PfspInstance), C.c_int, C.c_int, C.c_int, C.c_char_p, C.POINTER(Config)]
lib.bb_nq_dist_open.restype = _vp
lib.bb_nq_dist_open.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.POINTER(Config)]
lib.bb_dist_search.argtypes = [_vp, C.c_int, C.c_int, C.POINTER(SearchResult), C.POINTER(Stats)]
lib.bb_dist_close.restype = None
lib.bb_dist_close.argtypes = [_vp]
lib.bb_team_bytes.restype = C.c_size_t
lib.bb_team_bytes.argtypes = [C.c_int]
lib.bb_team_init.argtypes = [_vp, C.c_int]
lib.bb_team_reset.argtypes = [_vp]
lib.bb_team_wait_ready.argtypes = [_vp, C.c_int, C.c_double]
lib.bb_team_barrier.argtypes = [_vp, C.c_double]
lib.bb_team_fail.restype = None
lib.bb_team_fail.argtypes = [_vp]
lib.bb_team_worker.argtypes = [_vp, C.c_int, C.POINTER(TeamOps), C.c_int, C.c_longlong, C.c_longlong, C.c_int, C.c_int,
                               C.POINTER(TeamResult)]
lib.bb_team_best.argtypes = [_vp]
lib.bb_measure_slot_step_peak.argtypes = [C.c_int, C.POINTER(C.c_double)]
lib.bb_pfsp_expand_static.argtypes = [C.POINTER(PfspInstance), C.c_int, C.c_int, _vp, C.c_int, _vp, C.c_int, _vp, _vp]
STATIC_MAX_PARENTS = 2048  # BB_STATIC_MAX_PARENTS
lib.evaluate_gpu.restype = None
lib.evaluate_gpu.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), Lb1BoundData,
                             Lb2BoundData, _vp, _vp, _vp, _vp]
lib.lb1_alloc_gpu.restype = None
lib.lb1_alloc_gpu.argtypes = [C.POINTER(Lb1BoundData), C.POINTER(Lb1BoundData), _vp, _vp, _vp, C.c_int, C.c_int]
lib.lb2_alloc_gpu.restype = None
lib.lb2_alloc_gpu.argtypes = [C.POINTER(Lb2BoundData), C.POINTER(Lb2BoundData), _vp, _vp, _vp, _vp, _vp, C.c_int,
                              C.c_int]
lib.nq_evaluate_gpu.restype = None
lib.nq_evaluate_gpu.argtypes = [C.c_int, C.c_int, _vp, _vp, C.c_int]


def last_error():
    return (lib.bb_last_error() or b"").decode()


def check(rc, what=""):
    if rc is None or (isinstance(rc, int) and rc < 0):
        raise BBError(f"{what}: {last_error()} (code {rc})")
    return rc


def default_config(**kw):
    cfg = Config()
    lib.bb_config_default(C.byref(cfg))
    for k, v in kw.items():
        setattr(cfg, k, v)
    return cfg


class Instance:
    """host-side PFSP instance tables (bb_pfsp_instance)"""

    def __init__(self, inst=None, p_times=None):
        if inst is not None:
            self.ptr = lib.bb_pfsp_instance_taillard(int(inst))
        else:
            p = np.ascontiguousarray(p_times, dtype=np.int32)
            m, n = p.shape
            self.ptr = lib.bb_pfsp_instance_from_ptimes(n, m, p.reshape(-1))
        if not self.ptr:
            raise BBError("could not build the PFSP instance")
        c = self.ptr.contents
        self.n, self.m, self.P, self.best_known = c.nb_jobs, c.nb_machines, c.nb_machine_pairs, c.best_known

    def array(self, name, count):
        return np.ctypeslib.as_array(getattr(self.ptr.contents, name), shape=(max(count, 1),))[:count].copy()

    def tables(self):
        n, m, P = self.n, self.m, self.P
        return {"p": self.array("p_times", n * m), "heads": self.array("min_heads", m),
                "tails": self.array("min_tails", m), "lags": self.array("lags", P * n),
                "johnson": self.array("johnson_schedules", P * n), "pa": self.array("machine_pairs_1", P),
                "pb": self.array("machine_pairs_2", P), "order": self.array("machine_pair_order", P)}

    def close(self):
        if getattr(self, "ptr", None):
            lib.bb_pfsp_instance_free(self.ptr)
            self.ptr = None

    def __del__(self):
        self.close()


class Engine:
    """one device-resident pool engine (bb_engine)"""

    def __init__(self, handle, kind):
        if not handle:
            raise BBError("engine creation failed: " + last_error())
        self.h, self.kind = handle, kind

    @classmethod
    def pfsp(cls, instance, lb, **cfg):
        e = cls(lib.bb_pfsp_engine_create(instance.ptr, lb, C.byref(default_config(**cfg))), "pfsp")
        e.instance = instance  # keep the host tables alive
        e.node_i16 = (NODE_I16 if instance.n <= MAX_JOBS else NODE_I16_WIDE if instance.n <= 50
                      else NODE_I16_WIDE100 if instance.n <= 100 else NODE_I16_WIDE200)
        return e

    @classmethod
    def nqueens(cls, N, G=1, **cfg):
        return cls(lib.bb_nq_engine_create(N, G, C.byref(default_config(**cfg))), "nq")

    def reset(self, best=INT_MAX):
        check(lib.bb_engine_reset(self.h, best), "bb_engine_reset")

    def push_root(self):
        check(lib.bb_engine_push_root(self.h), "bb_engine_push_root")

    def push_nodes(self, nodes):
        dt, w = (np.int16, getattr(self, "node_i16", NODE_I16)) if self.kind == "pfsp" else (np.uint8, NQ_REC)
        a = np.ascontiguousarray(nodes, dtype=dt).reshape(-1, w)
        check(lib.bb_engine_push_nodes(self.h, a.ctypes.data, a.shape[0]), "bb_engine_push_nodes")

    def export_nodes(self, max_count=1 << 22):
        dt, w = (np.int16, getattr(self, "node_i16", NODE_I16)) if self.kind == "pfsp" else (np.uint8, NQ_REC)
        st = self.stats()
        cnt = max(int(st.pool_size), 1)
        if cnt > max_count:
            raise BBError(f"pool has {cnt} nodes > max_count {max_count}")
        out = np.zeros((cnt, w), dtype=dt)
        got = check(lib.bb_engine_export_nodes(self.h, out.ctypes.data, cnt), "bb_engine_export_nodes")
        return out[:got]

    def set_best(self, best):
        check(lib.bb_engine_set_best(self.h, best), "bb_engine_set_best")

    def keep_share(self, parts, me):
        check(lib.bb_engine_keep_share(self.h, parts, me), "bb_engine_keep_share")

    def stats(self):
        st = Stats()
        check(lib.bb_engine_get_stats(self.h, C.byref(st)), "bb_engine_get_stats")
        return st

    def run(self, max_iters=0):
        st = Stats()
        check(lib.bb_engine_run(self.h, max_iters, C.byref(st)), "bb_engine_run")
        return st

    def close(self):
        if getattr(self, "h", None):
            lib.bb_engine_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


def pfsp_search(instance, lb, best0=None, D=1, ws=1, **cfg):
    res, per = SearchResult(), (Stats * D)()
    best0 = instance.best_known if best0 is None else best0
    check(lib.bb_pfsp_search(instance.ptr, lb, best0, D, ws, C.byref(default_config(**cfg)), C.byref(res), per),
          "bb_pfsp_search")
    return res, list(per)


def pfsp_search_dist(instance, lb, rank, world, shm_name, best0=None, ws=1, **cfg):
    res, mine = SearchResult(), Stats()
    best0 = instance.best_known if best0 is None else best0
    check(lib.bb_pfsp_search_dist(instance.ptr, lb, best0, rank, world, shm_name.encode(), ws,
                                  C.byref(default_config(**cfg)), C.byref(res), C.byref(mine)), "bb_pfsp_search_dist")
    return res, mine


def nq_search_dist(N, G, rank, world, shm_name, ws=1, **cfg):
    res, mine = SearchResult(), Stats()
    check(lib.bb_nq_search_dist(N, G, rank, world, shm_name.encode(), ws, C.byref(default_config(**cfg)),
                                C.byref(res), C.byref(mine)), "bb_nq_search_dist")
    return res, mine


class Dist:
    """persistent multi-process team (bb_dist): open once, search many times; every rank makes the same calls"""

    def __init__(self, rank, world, shm_name, instance=None, lb=None, N=None, G=1, **cfg):
        self.instance = instance
        c = default_config(**cfg)
        if instance is not None:
            self.h = lib.bb_pfsp_dist_open(instance.ptr, lb, rank, world, shm_name.encode(), C.byref(c))
        else:
            self.h = lib.bb_nq_dist_open(N, G, rank, world, shm_name.encode(), C.byref(c))
        if not self.h:
            raise BBError("bb_*_dist_open failed: " + last_error())

    def search(self, best0=None, ws=1):
        res, mine = SearchResult(), Stats()
        if best0 is None:
            best0 = self.instance.best_known if self.instance is not None else INT_MAX
        check(lib.bb_dist_search(self.h, best0, ws, C.byref(res), C.byref(mine)), "bb_dist_search")
        return res, mine

    def close(self):
        if getattr(self, "h", None):
            lib.bb_dist_close(self.h)
            self.h = None


def nq_search(N, G=1, D=1, ws=1, **cfg):
    res, per = SearchResult(), (Stats * D)()
    check(lib.bb_nq_search(N, G, D, ws, C.byref(default_config(**cfg)), C.byref(res), per), "bb_nq_search")
    return res, list(per)
