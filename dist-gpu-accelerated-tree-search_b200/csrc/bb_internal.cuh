/*
 * bb_internal.cuh - shared device/host definitions of the native engine (not part of the C ABI).
 *
 * HBM layout of the node pool
 * ---------------------------
 * A pool is TWO LIFO stacks (ping-pong: an iteration pops a batch of parents from the top of one and appends the
 * surviving children to the top of the other, so a batch never leaves a hole and nothing is copied twice).
 * A stack stores fixed-size records as PLANES of 16-byte vectors: plane q of node i lives at
 *     base + (q * capacity + i) * 16
 * i.e. structure-of-arrays over 16-byte chunks. Thread i of a warp touching plane q of node i is a fully
 * coalesced 512-byte uint4 access.
 *
 * PFSP record (32-bit words; mp = machines rounded up to even):
 *     w0..w4   prmu[20] as uint8         (first `depth` entries = scheduled prefix)
 *     w5       depth (byte 0)
 *     w6..     front[mp]  uint16         completion time of the prefix on each machine (cached, not recomputed)
 *     ..       remain[mp] uint16         total unscheduled work on each machine (cached)
 *   -> 6 + mp words: 3 planes (48 B) at m=5, 4 planes (64 B) at m=10, 7 planes (112 B) at m=20.
 *   limit1 is not stored (== depth-1 under the reference's forward-only branching, PFSP_node.c:9-10).
 *
 * N-Queens record: plane 0 = board[0..15]; N >= 16 adds a second array of `tail words` 32-bit words per node
 *   (board[16..] then depth in the last byte); N <= 15 keeps depth in byte 15 of plane 0.
 */
#ifndef BB_INTERNAL_CUH
#define BB_INTERNAL_CUH

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "bb_gpu.h"

#define BB_WARPS_PER_CTA 8
#define BB_CTA_THREADS (BB_WARPS_PER_CTA * 32)
#define BB_TILE_MAX 32 /* parents per warp tile */

/* Device-resident control block: everything an iteration needs to know is here, so iterations can be chained in
 * a CUDA graph with no host round trip (replaces popBackBulk/pushBackBulk + the index prep of the reference). */
struct __align__(256) bb_ctrl
{
  /* ---- first 128-byte line: read-mostly during a launch ----
   * iteration plan, written by the last CTA of the previous iteration (or by the host when seeding) */
  int size[2];         /* top of each stack: live nodes are [bot, size) */
  int bot[2];          /* bottom of each stack: nodes below were donated to another GPU (oldest nodes go first) */
  int src;             /* stack this iteration pops from */
  int n;               /* parents in this iteration's batch (0 = nothing to do) */
  int base;            /* index of the first parent of the batch in stack src */
  int tile;            /* parents per warp tile for this iteration */
  /* search state */
  int best;            /* incumbent upper bound (atomicMin) */
  int M;               /* max parents per iteration */
  int capacity;        /* nodes per stack */
  int maxkids;         /* most children one parent can have (jobs / queens): bounds the growth of one iteration */
  int max_pool;        /* high-water mark of size[0]+size[1] */
  /* the last donation (k_donate_plan): which stack, from which index, how many nodes */
  int don_stack, don_from, don_count;
  unsigned int next_round; /* dynamic hand-out of tiles to warps (nq_expand5); lives here, away from the cursor's
                            * line: the plan is only read at the start of a launch */
  unsigned int next_deep;  /* nq_expand5<1>, pass A: next chunk of the batch to hand to a warp */
  int pad0[14];
  /* ---- second line: what every warp of a launch hammers with atomics. It is kept apart from `best` and the plan:
   * loads of the first line used to queue behind the cursor atomics in the L2 (ncu: a third of the stall samples
   * of the LB1_d kernel sat on the load of `best`). ---- */
  unsigned int tail;   /* append cursor of stack dst = 1 - src */
  unsigned int ticket; /* CTAs finished in this iteration */
  unsigned int next_tile; /* dynamic tile hand-out of this iteration (staged LB2 kernel) */
  int error;           /* sticky: 1 = pool overflow */
  /* counters */
  unsigned long long tree, sol, parents, evals, lb2_evals, lb2_pairs, iterations;
};
static_assert(sizeof(bb_ctrl) == 256, "bb_ctrl is two 128-byte lines");
static_assert(offsetof(bb_ctrl, tail) == 128, "the cursor starts the second line");

struct bb_pool
{
  uint4 *planes[2];   /* per stack: [nplanes][capacity] uint4 */
  uint32_t *tailw[2]; /* per stack: N-Queens tail words [capacity][tail_words] (may be null) */
  int nplanes;
  int tail_words;
  int capacity;
};

/* instance tables in device global memory, repacked for the kernels (staged into shared memory per CTA) */
struct bb_pfsp_dev
{
  int n, m, mp, P;
  const uint8_t *p8;     /* [m][BB_MAX_JOBS] processing times (1..99), zero padded */
  const int *tails;      /* [m] min_tails */
  const int *heads;      /* [m] min_heads */
  /* LB2: per (pair, slot) in Johnson order, always BB_MAX_JOBS slots (null slots have jbit = 0) */
  const uint32_t *jbit;  /* 1 << job */
  const int *ja;         /* p[m2][job] - p[m1][job] */
  const int *jb;         /* lag + p[m2][job] */
  const int2 *pairs;     /* (m1, m2) in the reference's evaluation order */
  /* staged LB2 kernel: the same four tables in the LEARNED pair order (rewritten on the device between launches),
   * the order itself, per-pair scores and the kill histogram by position (+1 slot: samples since the last sort) */
  uint32_t *jbit_act;
  int *ja_act;
  int *jb_act;
  int2 *pairs_act;
  int *order;
  unsigned *score;
  unsigned *hist;
  /* subset-table LB2 (see k_lb2_build_table in bb_engine.cu): g16[U * grow + a] = max over pairs (a, b) of the
   * head-independent part of the pair's Johnson makespan on the unscheduled set U, plus min_tails[b] */
  const uint16_t *g16;
  int grow; /* u16 per row (m - 1 rounded up to a multiple of 8) */
  /* packed kernel (bb_pfsp_packed.cuh): min_tails[i] in both 16-bit halves; lives in the kernel parameter block, so
   * the unrolled machine loop reads it as a constant-bank operand */
  uint32_t tails2[BB_MAX_MACHINES];
  int best_known;
  const uint32_t *prow; /* [ceil(mu/4)][BB_MAX_JOBS]: word q of job j = its times on machines 4q..4q+3 (bb_pfsp_tpp.cuh) */
  uint32_t one; /* 1 at run time: multiplier of the IMAD that does an add on the FMA pipe (bb_pfsp_tpp.cuh) */
};

#define BB_CUDA_TRY(call)                                         \
  do                                                              \
  {                                                               \
    cudaError_t err__ = (call);                                   \
    if (err__ != cudaSuccess)                                     \
      return bb_fail_cuda(err__, #call, __FILE__, __LINE__);      \
  } while (0)

int bb_fail_cuda(cudaError_t err, const char *what, const char *file, int line);
int bb_fail(int code, const char *fmt, ...);
void bb_note_cuda(cudaError_t err, const char *what);

#endif
