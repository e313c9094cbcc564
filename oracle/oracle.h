/*
 * oracle.h - CPU restatement of the reference's bound -> prune -> branch path.
 *
 * TEST INFRASTRUCTURE ONLY. This is the checker the CUDA path is compared against; it is never linked into,
 * imported by or executed from the product (libbbgpu.so, the host executables, the python bindings).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
 *
 * Parity status: PINNED. Every function here is checked (tests/test_oracle.py) against the reference's own
 * CPU code compiled in place into oracle/_ref/ and against golden vectors generated from it (tests/golden/).
 *
 * All citations are relative to the reference tree (ivantag13/dist-GPU-accelerated-tree-search).
 */
#ifndef ORACLE_H
#define ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- PFSP instance tables (pfsp/lib/c_bound_simple.h:14-21, c_bound_johnson.h:16-27) ---- */
typedef struct
{
  int n;           /* jobs */
  int m;           /* machines */
  int P;           /* machine pairs = m(m-1)/2 (LB2_FULL, c_bound_johnson.c:15) */
  int *p;          /* [m*n] machine-major processing times */
  int *heads;      /* [m] min_heads */
  int *tails;      /* [m] min_tails */
  int *pair_a;     /* [P] first machine of pair */
  int *pair_b;     /* [P] second machine of pair */
  int *pair_order; /* [P] evaluation order (identity) */
  int *lags;       /* [P*n] */
  int *johnson;    /* [P*n] Johnson order of job ids per pair */
} orc_pfsp;

int orc_taillard_jobs(int inst);
int orc_taillard_machines(int inst);
int orc_taillard_best_ub(int inst);
long orc_taillard_seed(int inst);
void orc_taillard_ptimes(int inst, int *out /* [m*n] */);

orc_pfsp *orc_pfsp_new(int inst);
orc_pfsp *orc_pfsp_from_ptimes(int n, int m, const int *p);
void orc_pfsp_free(orc_pfsp *t);
/* copy tables out (ctypes convenience) */
void orc_pfsp_tables(const orc_pfsp *t, int *p, int *heads, int *tails, int *lags, int *johnson, int *pa, int *pb,
                     int *order);
uint64_t orc_fnv1a_i32(const int *v, int count);

/* ---- bounds ---- */
int orc_makespan(const orc_pfsp *t, const int *prmu);
int orc_lb1(const orc_pfsp *t, const int *prmu, int limit1, int limit2);
void orc_lb1_children(const orc_pfsp *t, const int *prmu, int limit1, int limit2, int *lb_begin /* [n] */);
int orc_lb2(const orc_pfsp *t, const int *prmu, int limit1, int limit2, int best);
/* LB2 with census: pairs visited and inner iterations on unscheduled jobs (SURVEY A.2 counters) */
int orc_lb2_census(const orc_pfsp *t, const int *prmu, int limit1, int limit2, int best,
                   unsigned long long *pairs_visited, unsigned long long *inner_unsched);

/* ---- batch level: the reference's chunk protocol ---- */
/* Node records are flat int16: [depth, limit1, prmu[ORC_MAX_JOBS]] = the reference's 44-byte Node. */
#ifndef ORC_MAX_JOBS /* liboracle50.so is this same code built with -DORC_MAX_JOBS=50 (the 50-job Taillard classes) */
#define ORC_MAX_JOBS 20
#endif
#define ORC_NODE_I16 (2 + ORC_MAX_JOBS)
/* nodeIndex/sumOffSets construction, pfsp_multigpu_cuda.c:275-283. returns numBounds */
int orc_batch_index(const int16_t *parents, int nparents, int jobs, int *sum_offsets, int *node_index);
/* what evaluate_gpu writes into bounds[] for lb = 0,1,2 (PFSP_gpu_lib.cu:43-127) */
void orc_evaluate_chunk(const orc_pfsp *t, int lb, const int16_t *parents, int nparents, int best, int *bounds);
/* host prune+branch, PFSP_lib.h:51-95. returns number of children written */
int orc_generate_children(const int16_t *parents, int nparents, int jobs, const int *bounds,
                          unsigned long long *tree, unsigned long long *sol, int *best, int16_t *children);

/* ---- run level ---- */
typedef struct
{
  unsigned long long tree;      /* exploredTree */
  unsigned long long sol;       /* exploredSol */
  unsigned long long parents;   /* nodes decomposed */
  unsigned long long evals;     /* child bounds evaluated */
  unsigned long long depth_hist[64]; /* parents by depth */
  int best;
  int max_pool;
} orc_run_stats;
/* sequential DFS, pfsp_c.c:55-63 + decompose (PFSP_lib.c:7-129). budget<=0: to completion */
void orc_pfsp_search(const orc_pfsp *t, int lb, int best0, long long budget, orc_run_stats *out);
/* the 3-step batched structure of pfsp_multigpu_cuda.c (one worker): warm-up BFS to >= m nodes, pop-back batches
 * of <= M parents through orc_evaluate_chunk + orc_generate_children, drain. */
void orc_pfsp_search_batched(const orc_pfsp *t, int lb, int best0, int m, int M, orc_run_stats *out,
                             unsigned long long *batches);

/* ---- N-Queens (nqueens/nqueens_c.c:80-140, nqueens_multigpu_cuda.cu:158-210) ---- */
#define ORC_MAX_QUEENS 20
int orc_nq_is_safe(int G, const uint8_t *board, int queen_num, int row_pos);
/* labels[k + i*N] for k >= depth (K4); slots k < depth are left untouched */
void orc_nq_labels(int N, int G, const uint8_t *parents /* [size][1+ORC_MAX_QUEENS] */, int size, uint8_t *labels);
int orc_nq_generate_children(int N, const uint8_t *parents, int size, const uint8_t *labels,
                             unsigned long long *tree, unsigned long long *sol, uint8_t *children);
void orc_nq_search(int N, int G, long long budget, orc_run_stats *out);

#ifdef __cplusplus
}
#endif
#endif
