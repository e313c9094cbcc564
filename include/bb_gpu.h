/*
 * bb_gpu.h - C ABI of libbbgpu.so: the B200-native (sm_100a) bound -> prune -> branch engine.
 *
 * Two levels, both plain C (pointers + sizes, no C++/torch types):
 *
 *  1. COMPAT level  - the exact extern "C" symbols the reference's C drivers bind
 *                     (pfsp/lib/PFSP_gpu_lib.cuh:13-29), so pfsp_multigpu_cuda.c links against this library
 *                     unchanged, plus the analogous N-Queens entry the reference keeps inside its .cu driver.
 *  2. NATIVE level  - a device-resident structure-of-arrays node pool per GPU with the whole
 *                     bound+prune+branch+compaction loop on the device; host code only seeds, polls, steals.
 *
 * File:line citations are relative to the reference tree (ivantag13/dist-GPU-accelerated-tree-search).
 * Every int-returning function returns 0 on success and a negative code on failure; bb_last_error() gives the
 * message for the calling thread. There is NO CPU fallback anywhere in this library: without a CUDA device every
 * compute entry point fails with BB_ERR_CUDA.
 */
#ifndef BB_GPU_H
#define BB_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BB_MAX_JOBS 20     /* pfsp/lib/macro.h:10 */
#define BB_MAX_JOBS_WIDE 200 /* the 50-, 100- and 200-job Taillard classes ta031-ta110: the reference rebuilt with
                                MAX_JOBS 50 / 100 / 200 (pfsp/README.md:52). Engines of such instances exchange nodes
                                with the host as that build's Node: int16 [depth, limit1, prmu[NJ]] with NJ = 50 for
                                21..50 jobs, 100 for 51..100, 200 for 101..200; the compat level stays at 20 */
#define BB_NODE_JOBS(n) ((n) <= BB_MAX_JOBS ? BB_MAX_JOBS : (n) <= 50 ? 50 : (n) <= 100 ? 100 : 200) /* prmu[] entries of a host node */
#define BB_MAX_MACHINES 20 /* pfsp/lib/macro.h:11 */
#define BB_MAX_QUEENS 20   /* nqueens/lib/NQueens_node.h:11 */

enum
{
  BB_OK = 0,
  BB_ERR_ARG = -1,      /* bad argument (size, null pointer, unsupported instance class) */
  BB_ERR_CUDA = -2,     /* a CUDA call failed; bb_last_cuda_error() holds the cudaError_t */
  BB_ERR_OVERFLOW = -3, /* device pool capacity exceeded */
  BB_ERR_STATE = -4     /* call not valid in the engine's current state */
};

const char *bb_last_error(void);
int bb_last_cuda_error(void);

/* =====================================================================================================
 * COMPAT LEVEL (drop-in for pfsp/lib/PFSP_gpu_lib.cu)
 * The struct layouts are the reference's; when the reference's own headers are included first their
 * definitions are used.
 * ===================================================================================================== */

#ifndef C_BOUND_SIMPLE_H_ /* pfsp/lib/c_bound_simple.h:14-21 */
typedef struct lb1_bound_data
{
  int *p_times;
  int *min_heads;
  int *min_tails;
  int nb_jobs;
  int nb_machines;
} lb1_bound_data;
#endif

#ifndef C_BOUND_JOHNSON_H /* pfsp/lib/c_bound_johnson.h:16-27 */
typedef struct lb2_bound_data
{
  int *johnson_schedules;
  int *lags;
  int *machine_pairs_1;
  int *machine_pairs_2;
  int *machine_pair_order;
  int nb_machine_pairs;
  int nb_jobs;
  int nb_machines;
} lb2_bound_data;
#endif

#if !defined(PFSP_NODE_H) && !defined(NQUEENS_NODE_H) /* pfsp/lib/PFSP_node.h:15-20 : 44-byte AoS node */
typedef struct
{
  int16_t depth;
  int16_t limit1;
  int16_t prmu[BB_MAX_JOBS];
} Node;
#endif

/* nqueens/lib/NQueens_node.h:13-17 : 21-byte node. (Named apart: the reference reuses the name `Node`.) */
typedef struct
{
  uint8_t depth;
  uint8_t board[BB_MAX_QUEENS];
} bb_nq_node;

/* Replaces PFSP_gpu_lib.cu:129-152 (called at pfsp_multigpu_cuda.c:295). Launches, does not synchronise.
 * lb: 0 = LB1_d over parentsSize parents, 1 = LB1 / 2 = LB2 over `size` children. All pointers except `best`
 * are device pointers owned by the caller; `best` is a host pointer read once. bounds[] layout and values are
 * bit-identical to the reference kernels, including the LB2 early-exit value. */
void evaluate_gpu(const int jobs, const int lb, const int size, const int nbBlocks, const int parentsSize,
                  int *best, const lb1_bound_data lbound1, const lb2_bound_data lbound2, Node *parents,
                  int *bounds, int *sumOffSets_d, int *nodeIndex_d);

/* Replace PFSP_gpu_lib.cu:154-200 (called at pfsp_multigpu_cuda.c:190,194): device deep copy of the host
 * tables. As in the reference the trailing pointer arguments are passed by value and ignored. */
void lb1_alloc_gpu(lb1_bound_data *lbound1_d, lb1_bound_data *lbound1, int *p_times_d, int *min_heads_d,
                   int *min_tails_d, int jobs, int machines);
void lb2_alloc_gpu(lb2_bound_data *lbound2_d, lb2_bound_data *lbound2, int *johnson_schedule_d, int *lags_d,
                   int *machine_pairs_1_d, int *machine_pairs_2_d, int *machine_pair_order_d, int jobs,
                   int machines);

/* The authors' op/byte model, PFSP_gpu_lib.cu:213-267. */
int P_of(int M);
int flop_lb1(int N, int M, int lim);
int bytes_per_inv_lb1(int N, int M);
int flop_lb2(int N, int M, int lim);
int bytes_per_inv_lb2(int N, int M, int lim);

/* Replaces the kernel + launch the reference keeps inside its driver (nqueens_multigpu_cuda.cu:158-186,
 * launched at :301): labels_d[k + i*N] for slot k >= depth of parent i; slots below depth are not written. */
void nq_evaluate_gpu(const int N, const int G, const bb_nq_node *parents_d, uint8_t *labels_d, const int size);

/* =====================================================================================================
 * HOST PRECOMPUTE (C, no GPU): replaces c_taillard.c:46-105, c_bound_simple.c:277-322,
 * c_bound_johnson.c:48-178. Tables are bit-identical to the reference's (tests compare FNV hashes).
 * ===================================================================================================== */

typedef struct bb_pfsp_instance
{
  int nb_jobs;
  int nb_machines;
  int nb_machine_pairs;
  int best_known;          /* Taillard best-known makespan, or INT_MAX for custom matrices */
  int inst;                /* Taillard id 1..120, 0 for custom */
  int *p_times;            /* [machines*jobs] machine-major */
  int *min_heads;          /* [machines] */
  int *min_tails;          /* [machines] */
  int *machine_pairs_1;    /* [pairs] */
  int *machine_pairs_2;    /* [pairs] */
  int *machine_pair_order; /* [pairs] */
  int *lags;               /* [pairs*jobs] */
  int *johnson_schedules;  /* [pairs*jobs] */
} bb_pfsp_instance;

int bb_taillard_nb_jobs(int inst);
int bb_taillard_nb_machines(int inst);
int bb_taillard_best_ub(int inst);
long bb_taillard_seed(int inst);
void bb_taillard_processing_times(int inst, int *ptm);
bb_pfsp_instance *bb_pfsp_instance_taillard(int inst);
bb_pfsp_instance *bb_pfsp_instance_from_ptimes(int nb_jobs, int nb_machines, const int *p_times);
void bb_pfsp_instance_free(bb_pfsp_instance *ins);

/* =====================================================================================================
 * NATIVE LEVEL: device-resident pool engine (one per GPU)
 * Replaces the host pool (Pool_atom.c), the per-batch PCIe round trip and index prep
 * (pfsp_multigpu_cuda.c:227-230, 275-301), host prune+branch (PFSP_lib.h:51-95,
 * nqueens_multigpu_cuda.cu:189-210) and the CPU warm-up / drain steps (pfsp_multigpu_cuda.c:111-118, 487-495).
 * ===================================================================================================== */

typedef struct bb_engine bb_engine;

typedef struct bb_config
{
  int device;             /* CUDA device ordinal */
  int M;                  /* max parents per device iteration (-M, PFSP_lib.c:252-259); <=0: 50000 */
  int m;                  /* min pool size a worker keeps when asked to share (-m); <=0: 25 */
  long long capacity;     /* nodes per stack (two stacks are allocated); <=0: chosen from M */
  int iters_per_launch;   /* device iterations chained in one CUDA-graph launch; <=0: 32 */
  int lb2_prefilter;      /* LB2 only: 1 = skip the pair loop for children LB1_d already prunes (default 1) */
  int lb2_staged;         /* LB2 only: 1 = staged kernel with learned pair order and candidate compaction (default),
                             0 = first-generation kernel walking the pairs in the reference's order */
  int lb2_stage_a;        /* LB2 staged kernel: pairs tried before the survivors are compacted; <=0: 8 */
  int lb2_table;          /* LB2 only: 1 = evaluate LB2 through the per-instance subset table (default; 2^n rows,
                             42 MB at 20x20, built on the device at engine creation), 0 = walk the machine pairs */
  int packed;             /* kernel generation. 3 (default): PFSP one lane per parent, two children per step in packed
                             16-bit halves; N-Queens attack masks in the record, the last levels of the tree enumerated
                             in registers by the lane that read the parent. 2: the same kernels, every N-Queens node
                             through the pool (one iteration = one level of the popped parents). 1: second generation
                             (two sibling children per lane on the warp tile; N-Queens safe-row masks rebuilt from the
                             board). 0: first generation, one child per lane. All walk the same tree. */
} bb_config;

typedef struct bb_stats
{
  unsigned long long tree;        /* exploredTree increments made on this device */
  unsigned long long sol;         /* exploredSol increments */
  unsigned long long parents;     /* nodes popped and decomposed */
  unsigned long long evals;       /* child bounds / safety evaluations */
  unsigned long long lb2_evals;   /* children that entered the LB2 pair loop */
  unsigned long long lb2_pairs;   /* machine pairs evaluated in total */
  unsigned long long iterations;  /* non-empty device iterations */
  unsigned long long launches;    /* kernels launched by this engine */
  long long pool_size;            /* nodes left in the pool */
  long long max_pool;             /* high-water mark */
  int best;                       /* incumbent */
  double kernel_ms;               /* CUDA-event time of all expand launches (events on the launching stream) */
  double span_ms;                 /* CUDA-event time from the first to the last launch of bb_engine_run calls */
  unsigned long long h2d_bytes;   /* host->device bytes copied since the last reset (seeded nodes) */
  unsigned long long d2h_bytes;   /* device->host bytes copied since the last reset (control-block polls) */
  unsigned long long table_bytes; /* instance tables uploaded at creation */
  int grid;                       /* CTAs per launch (a multiple of the SM count) */
  int smem_bytes;                 /* dynamic shared memory per CTA */
  int record_bytes;               /* bytes one node occupies in the device pool */
  unsigned long long steals, steals_ok, donated; /* whole searches on D > 1 GPUs: this worker's steal attempts,
                                                    successes, and nodes it gave away */
} bb_stats;

void bb_config_default(bb_config *cfg);

/* lb: 0 LB1_d, 1 LB1, 2 LB2 (PFSP_lib.c:225-232). Uploads the tables (a6) and allocates the pool. Instances with up
 * to 20 jobs run the tuned kernels; 21..200 jobs (ta031-ta110) run a wider record and kernel in three job-capacity
 * classes, 50, 100 and 200 (LB2 as the machine-pair loop: a 2^n table does not exist there); more than 200 jobs (the
 * 500-job class ta111-ta120: job ids no longer fit the record's bytes, and the reference's GPU path narrows the depth to
 * uint8_t, PFSP_lib.h:60) or 20 machines are refused. */
bb_engine *bb_pfsp_engine_create(const bb_pfsp_instance *ins, int lb, const bb_config *cfg);
bb_engine *bb_nq_engine_create(int N, int G, const bb_config *cfg);
void bb_engine_destroy(bb_engine *e);

/* Pools and LB2 tables of destroyed engines stay cached in a memory pool owned by this library (so that the next
 * engine on the same GPU does not pay the allocation again); this returns the cached blocks of `device` to the driver.
 * The device's default memory pool, which other code in the process may use, is never touched. */
int bb_release_cached_memory(int device);

int bb_engine_reset(bb_engine *e, int best);            /* empty pool, zero counters, set incumbent */
int bb_engine_push_root(bb_engine *e);                  /* initRoot: PFSP_node.c:7-14 / NQueens_node.c:7-13 */
/* Seed from HOST nodes in the reference's AoS layouts (Node / bb_nq_node); the engine derives its cached
 * per-node state on the device. */
int bb_engine_push_nodes(bb_engine *e, const void *host_nodes, int count);
/* Copy the pool back to HOST in the reference's AoS layout (tests, checkpointing). Returns the count. */
long long bb_engine_export_nodes(bb_engine *e, void *host_nodes, long long max_count);
int bb_engine_set_best(bb_engine *e, int best);
int bb_engine_get_stats(bb_engine *e, bb_stats *out);
/* Keep this engine's share of the pool, on the device and in place: the nodes whose placed prefix hashes to `me`
 * of `parts` (FNV-1a over the scheduled jobs / placed rows), dropping the others. D engines that hold the same set of
 * nodes and each call this with their own `me` end up with disjoint shares whose union is the set - the warm-up split
 * of the multi-GPU searches (replaces roundRobin_distribution, Pool_atom.c:14-36) without any communication. */
int bb_engine_keep_share(bb_engine *e, int parts, int me);

/* Run device iterations until the pool is empty (max_iters <= 0) or max_iters iterations were done. */
int bb_engine_run(bb_engine *e, long long max_iters, bb_stats *out);

/* Whole search on D GPUs of this process (one host thread per GPU, peer-memory work stealing when ws != 0):
 * the equivalent of pfsp_search() / nqueens_search() in the reference drivers. per_gpu may be NULL or [D]. */
typedef struct bb_search_result
{
  unsigned long long tree;
  unsigned long long sol;
  int best;
  double elapsed_s;     /* wall time of the call: engine creation, table upload, search, teardown; the one-time
                           CUDA context creation (and peer enabling for D > 1) of the process is not included */
  double kernel_ms_max; /* max over GPUs of device time in expand kernels */
  unsigned long long steals, steals_ok;
  double search_s;      /* max over GPUs of the wall time from the common start barrier (pools seeded, tables
                           resident) to the detection of global termination */
  unsigned long long nodes_moved; /* nodes that changed GPU through work stealing */
  /* phase split, as the reference drivers print it (pfsp_multigpu_cuda.c:120-125, 476-481, 503-508) */
  unsigned long long warm_tree, warm_sol; /* counted during the common warm-up (step 1; on the GPU here) */
  double warm_s;                          /* engine creation excluded: warm-up expansion + partition */
  double setup_s;                         /* engine creation: pool cudaMalloc, table upload, LB2 table build */
} bb_search_result;

int bb_pfsp_search(const bb_pfsp_instance *ins, int lb, int best0, int D, int ws, const bb_config *cfg,
                   bb_search_result *res, bb_stats *per_gpu);
int bb_nq_search(int N, int G, int D, int ws, const bb_config *cfg, bb_search_result *res, bb_stats *per_gpu);

/* The same searches with one PROCESS per GPU (rank = 0..world-1 on one node, e.g. under torchrun): every rank calls
 * with the same arguments and its own cfg->device; ranks meet in the POSIX shared-memory object `shm_name`
 * (created by whoever arrives first, unlinked by rank 0 at the end; use a fresh name per call), exchange CUDA IPC
 * handles of their transfer buffers through it and steal from each other's HBM over NVLink. Every rank returns the global
 * totals in `res`; `mine` (may be NULL) gets the caller's own engine statistics. */
int bb_pfsp_search_dist(const bb_pfsp_instance *ins, int lb, int best0, int rank, int world, const char *shm_name,
                        int ws, const bb_config *cfg, bb_search_result *res, bb_stats *mine);
int bb_nq_search_dist(int N, int G, int rank, int world, const char *shm_name, int ws, const bb_config *cfg,
                      bb_search_result *res, bb_stats *mine);

/* The same with a PERSISTENT team: open once per (instance, bound) - engine, pool, tables, transfer buffer and the
 * CUDA IPC mappings of every peer's buffer are created here - then run any number of searches; nothing is allocated,
 * mapped or freed per search. Every rank calls each function with the same arguments (collective calls). A search is
 * timed from the root: res->search_s is the max over ranks of the span from just before the root is seeded (reset,
 * root, common warm-up, device-side partition) to the detection of global termination; res->elapsed_s is the wall
 * time of the bb_dist_search call on the calling rank. world = 1 is allowed (no stealing). */
typedef struct bb_dist bb_dist;
bb_dist *bb_pfsp_dist_open(const bb_pfsp_instance *ins, int lb, int rank, int world, const char *shm_name,
                           const bb_config *cfg);
bb_dist *bb_nq_dist_open(int N, int G, int rank, int world, const char *shm_name, const bb_config *cfg);
int bb_dist_search(bb_dist *d, int best0, int ws, bb_search_result *res, bb_stats *mine);
void bb_dist_close(bb_dist *d);

/* ----- the team protocol underneath (host code only; replaces the steal loop pfsp_multigpu_cuda.c:343-404, the
 * allIdle termination common/util.c:19-38 and checkBest pfsp_multigpu_cuda.c:30-50). `mem` is a zero-filled block
 * of bb_team_bytes(world) bytes visible to every worker (heap for threads, a shared mapping for processes).
 * A worker reaches its engine only through bb_team_ops. */
typedef struct bb_team_ops
{
  void *self;
  /* run at most max_iters pool iterations; report nodes left, the incumbent and device time spent */
  int (*run)(void *self, long long max_iters, long long *pool_size, int *best, double *busy_ms);
  /* reserve the oldest nodes for a thief: count = 0 when fewer than 2*min_keep nodes are left (or no transfer slot is
   * free), count = -1 when the answer is not known yet - the protocol then calls again at its next poll (the GPU
   * engine only enqueues the donation on its stream and never waits for the device). `stack` names the stack / the
   * transfer-buffer slot (0 or 1) the nodes sit in until donate_done. */
  int (*donate)(void *self, long long min_keep, long long max_count, int *stack, long long *start, long long *count);
  int (*donate_done)(void *self, int stack); /* the thief has copied the range out */
  /* thief side: copy [start, start+count) of `stack` out of worker `victim`'s memory into the own pool */
  int (*pull)(void *self, int victim, int stack, long long start, long long count);
  int (*set_best)(void *self, int best);
} bb_team_ops;

typedef struct bb_team_result
{
  int best;
  unsigned long long steals, steals_ok, donated;
  double busy_ms;
} bb_team_result;

size_t bb_team_bytes(int world);
int bb_team_init(void *mem, int world);                          /* by one worker, before the others use mem */
int bb_team_reset(void *mem);                                    /* by one worker, between two searches of the team */
int bb_team_wait_ready(void *mem, int world, double timeout_s);  /* by the others */
int bb_team_barrier(void *mem, double timeout_s);
void bb_team_fail(void *mem);
int bb_team_worker(void *mem, int me, const bb_team_ops *ops, int ws, long long m_keep, long long max_steal,
                   int best0, int iters_per_poll, bb_team_result *out);
int bb_team_best(void *mem);

/* Measured ceiling for the LB2 roofline (SURVEY section 8(d) asks for a measured integer-pipe peak; the reference's own
 * op model is PFSP_gpu_lib.cu:244-254): sustained rate of the LB2 inner step - one job-bit test plus one predicated
 * fused add-max per lane - on `device`, in slot steps per second. */
int bb_measure_slot_step_peak(int device, double *slot_steps_per_s);

/* =====================================================================================================
 * ALLOCATION-FREE STEP. This entry point has no counterpart in the reference.
 *
 * One bound -> prune -> branch step over a batch of parents (what bb_engine_push_nodes + bb_engine_run(1) +
 * bb_engine_export_nodes do) for callers that must not allocate device memory while they work: host threads
 * that share a GPU with other threads (cudaMalloc/cudaFree synchronise the device and would stall them all),
 * hosts that have already given every free byte of HBM to their own pool, and latency-critical sections that must
 * not enter the driver's allocator. It calls no cudaMalloc, cudaFree or
 * other allocating routine: its whole working set - control block, instance tables, both node stacks - is one
 * file-scope __device__ array of about 4.6 MiB inside libbbgpu.so (present once on every GPU that loads the
 * library, until the process exits).
 *
 *   ins         host instance tables (bb_pfsp_instance_*); lb: 0 = LB1_d, 1 = LB1 (LB2 is not offered here)
 *   best        incumbent to prune against
 *   parents_d   DEVICE pointer, `count` nodes in the reference layout; count <= BB_STATIC_MAX_PARENTS, larger
 *               batches are refused with BB_ERR_ARG (split them). Every parent must be an inner node of this
 *               instance: 0 <= depth < nb_jobs and prmu[0..nb_jobs) a permutation of 0..nb_jobs-1. The nodes are in
 *               device memory, so this is checked on the device: a batch holding a parent with a depth or a job id
 *               out of range is refused as a whole - result_d[0] = 0 children and result_d[1] = result_d[2] =
 *               all-ones
 *   children_d  DEVICE pointer, room for max_children nodes; survivors beyond max_children are dropped, so
 *               compare result_d[0] with max_children (count * nb_jobs always suffices)
 *   result_d    DEVICE pointer to 4 x unsigned long long: children produced, exploredTree increment,
 *               exploredSol increment, incumbent after the step
 *   stream      a cudaStream_t (NULL = default stream); work is only enqueued, the caller synchronises
 *
 * The array is shared by every call on that GPU and keeps what the last call left in it; each call clears the
 * parts that must start at zero, but two calls whose GPU work can overlap race on it - keep them ordered.
 * The first call on each device also resolves the array's device address and the launch geometry: make that
 * call before any stream capture begins.
 * ===================================================================================================== */
#define BB_STATIC_MAX_PARENTS 2048
int bb_pfsp_expand_static(const bb_pfsp_instance *ins, int lb, int best, const Node *parents_d, int count,
                          Node *children_d, int max_children, unsigned long long *result_d, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* BB_GPU_H */
