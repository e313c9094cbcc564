/*
 * bb_engine.cu - host side of the native engine: device pool, table upload, CUDA-graph iteration chains,
 * seeding / export in the reference's node layouts, statistics. C ABI declared in include/bb_gpu.h.
 *
 * What of the reference this replaces (all host code there):
 *   SinglePool_atom push/pop/realloc              pfsp/lib/Pool_atom.c:4-258        -> two device stacks + bb_ctrl
 *   lb1_alloc_gpu / lb2_alloc_gpu table upload    pfsp/lib/PFSP_gpu_lib.cu:154-200  -> upload_pfsp_tables (repacked)
 *   the while(1) batch loop                       pfsp/pfsp_multigpu_cuda.c:221-320 -> bb_engine_run (graph chains)
 *   CPU warm-up and drain (steps 1 and 3)         pfsp_multigpu_cuda.c:111-118, 487-495 -> same device loop
 */
#include "bb_internal.cuh"
#include "bb_nq_kernels.cuh"
#include "bb_pfsp_kernels.cuh"
#include "bb_pfsp_lb2.cuh"
#include "bb_pfsp_packed.cuh"
#include "bb_pfsp_tpp.cuh"
#include "bb_pfsp_wide.cuh"

#include <climits>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

/* ------------------------------------------------------------------ errors */

static thread_local char g_err[512] = "";
static thread_local int g_cuda_err = 0;

extern "C" const char *bb_last_error(void) { return g_err; }
extern "C" int bb_last_cuda_error(void) { return g_cuda_err; }

int bb_fail(int code, const char *fmt, ...)
{
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int bb_fail_cuda(cudaError_t err, const char *what, const char *file, int line)
{
  g_cuda_err = (int)err;
  snprintf(g_err, sizeof(g_err), "CUDA error %d (%s) at %s:%d: %s", (int)err, cudaGetErrorString(err), file, line,
           what);
  return BB_ERR_CUDA;
}

void bb_note_cuda(cudaError_t err, const char *what)
{
  if (err != cudaSuccess)
  {
    g_cuda_err = (int)err;
    snprintf(g_err, sizeof(g_err), "CUDA error %d (%s): %s", (int)err, cudaGetErrorString(err), what);
  }
}

/* ------------------------------------------------------------------ engine object */

enum
{
  KIND_PFSP = 1,
  KIND_NQ = 2
};

#define BB_STAGE_SLOTS 2

struct bb_engine
{
  int kind;
  int device;
  cudaStream_t stream;
  bb_ctrl *ctrl;   /* device */
  bb_ctrl *ctrl_h; /* pinned host mirror */
  bb_pool pool;
  void *pool_mem[2];
  void *tail_mem[2];
  void *stage_mem;        /* transfer buffer for donations (multi-GPU searches only): BB_STAGE_SLOTS slots */
  long long stage_nodes;  /* nodes per slot */
  int slot_busy[BB_STAGE_SLOTS]; /* slot holds a donation its thief has not acknowledged yet */
  int don_slot;           /* slot of the donation enqueued on the stream whose outcome is not fetched yet, or -1 */
  volatile int *mailbox_h; /* pinned + mapped: (stack, count) of the last donation, written by k_donate_plan */
  int *mailbox_d;
  cudaEvent_t don_ev;
  /* PFSP */
  bb_pfsp_dev T;
  void *table_mem;
  int lb, n, m, prefilter;
  int table;           /* LB2: subset-table kernel (pfsp_expand<3,MM>) */
  int packed;          /* LB1_d / LB1 / LB2-table: 2 = lane-per-parent kernel (bb_pfsp_tpp.cuh, default), 1 = two
                          children per lane on the warp tile (bb_pfsp_packed.cuh), 0 = first generation */
  void *g16_mem;
  int staged, stage_a; /* LB2: staged kernel with learned pair order (bb_pfsp_lb2.cuh), pairs tried in stage A */
  unsigned long long table_bytes;
  /* PFSP, 21..200 jobs (bb_pfsp_wide.cuh): wider record, its own tables; 0, or the job-capacity class 50 / 100 / 200 */
  int wide;
  bb_pfsp_wide_dev TW;
  void *wide_mem;
  /* N-Queens */
  int N, G;
  int nq_gen; /* 3 = masks travel with the node (nq_expand3, 32-byte record), 2 / 1 = earlier kernels (board + depth) */
  int nq_tail; /* nq_expand5: parents with at most this many columns left are enumerated in registers (0 = never) */
  int nq_arg;  /* the kernel's last argument: G, with nq_tail << 16 for nq_expand5 */
  /* launch */
  const void *kernel;
  bool pdl; /* the kernel waits for its predecessor itself (pdl_wait_then_release): chained with programmatic launch */
  int grid, smem;
  int threads; /* per CTA */
  int iters_per_launch;
  cudaGraphExec_t graph;
  cudaGraphExec_t small_graph[5]; /* chains of 1, 2, 4, 8, 16 iterations (built on first use): polling workers */
  /* pipelined polling (bb_engine_run_pipelined): one batch stays in flight while the host looks at the previous one */
  bb_ctrl *snap_h;                /* pinned, 2 snapshots of the control block */
  cudaEvent_t pipe_k0[2], pipe_k1[2], pipe_done[2];
  int pipe_slot, pipe_inflight;
  cudaEvent_t ev0, ev1, ev_begin, ev_end;
  double kernel_ms, span_ms;
  unsigned long long h2d_bytes, d2h_bytes;
  unsigned long long launches;
  int M, m_keep;
};

#define ENGINE_SET_DEVICE(e) BB_CUDA_TRY(cudaSetDevice((e)->device))

extern "C" void bb_config_default(bb_config *cfg)
{
  memset(cfg, 0, sizeof(*cfg));
  cfg->device = 0;
  cfg->M = 50000; /* PFSP_lib.c:179 */
  cfg->m = 25;    /* PFSP_lib.c:178 */
  cfg->capacity = 0;
  cfg->iters_per_launch = 32;
  cfg->lb2_prefilter = 1;
  cfg->lb2_staged = 1;
  cfg->lb2_stage_a = 8;
  cfg->lb2_table = 1;
  cfg->packed = 3;
}

/* ------------------------------------------------------------------ small kernels: plan / seed / export */

__global__ void k_plan(bb_ctrl *c, int add0, int total_warps)
{
  if (threadIdx.x == 0 && blockIdx.x == 0)
  {
    c->size[0] += add0;
    bbk::plan_next(c, 0, total_warps);
  }
}

__global__ void k_reset(bb_ctrl *c, int best, int M, int capacity, int maxkids)
{
  if (threadIdx.x == 0 && blockIdx.x == 0)
  {
    int *raw = reinterpret_cast<int *>(c);
    for (unsigned i = 0; i < sizeof(bb_ctrl) / sizeof(int); i++)
      raw[i] = 0;
    c->best = best;
    c->M = M;
    c->capacity = capacity;
    c->maxkids = maxkids;
    c->tile = 1;
  }
}

__global__ void k_set_best(bb_ctrl *c, int best) { atomicMin(&c->best, best); }

/* Subset table of LB2 (bb_pfsp_kernels.cuh, LB = 3): one thread per unscheduled set U walks every machine pair's
 * Johnson order once with the head-independent recurrence (D starts at -inf, so only the max_k(b_k + suffix a)
 * term survives) and keeps, per first machine, the largest value plus the second machine's tail. */
__global__ void __launch_bounds__(256) k_lb2_build_table(bb_pfsp_dev T, uint16_t *__restrict__ g16, int grow)
{
  extern __shared__ __align__(16) unsigned char sm[];
  const int P = T.P, n = T.n, m = T.m;
  uint32_t *jbit = reinterpret_cast<uint32_t *>(sm);
  int *ja = reinterpret_cast<int *>(jbit + P * BB_MAX_JOBS);
  int *jb = ja + P * BB_MAX_JOBS;
  int2 *pairs = reinterpret_cast<int2 *>(jb + P * BB_MAX_JOBS);
  int *tails = reinterpret_cast<int *>(pairs + P);
  int *G = tails + BB_MAX_MACHINES; /* [BB_MAX_MACHINES][256] */
  for (int i = threadIdx.x; i < P * BB_MAX_JOBS; i += blockDim.x)
  {
    jbit[i] = T.jbit[i];
    ja[i] = T.ja[i];
    jb[i] = T.jb[i];
  }
  for (int i = threadIdx.x; i < P; i += blockDim.x)
    pairs[i] = T.pairs[i];
  for (int i = threadIdx.x; i < m; i += blockDim.x)
    tails[i] = T.tails[i];
  for (int a = 0; a < BB_MAX_MACHINES; a++)
    G[a * 256 + threadIdx.x] = 0;
  __syncthreads();
  const uint32_t U = blockIdx.x * blockDim.x + threadIdx.x;
  if (U >= (1u << n))
    return;
  const int NEG = -(1 << 28);
  for (int s = 0; s < P; s++)
  {
    int D = NEG;
    const uint4 *vbit = reinterpret_cast<const uint4 *>(jbit + s * BB_MAX_JOBS);
    const int4 *va = reinterpret_cast<const int4 *>(ja + s * BB_MAX_JOBS);
    const int4 *vb = reinterpret_cast<const int4 *>(jb + s * BB_MAX_JOBS);
#pragma unroll
    for (int q = 0; q < BB_MAX_JOBS / 4; q++)
    {
      const uint4 bits = vbit[q];
      const int4 a4 = va[q];
      const int4 b4 = vb[q];
      if (U & bits.x) D = bbk::add_max(D, a4.x, b4.x);
      if (U & bits.y) D = bbk::add_max(D, a4.y, b4.y);
      if (U & bits.z) D = bbk::add_max(D, a4.z, b4.z);
      if (U & bits.w) D = bbk::add_max(D, a4.w, b4.w);
    }
    const int2 ab = pairs[s];
    const int v = D + tails[ab.y];
    if (v > G[ab.x * 256 + threadIdx.x])
      G[ab.x * 256 + threadIdx.x] = v;
  }
  for (int a = 0; a < grow; a++)
  {
    const int v = a < m - 1 ? G[a * 256 + threadIdx.x] : 0;
    g16[(size_t)U * grow + a] = (uint16_t)(v > 65535 ? 65535 : v);
  }
}

/* LB2 staged kernel: forget the learned pair order (identity order, zero scores and histogram) */
__global__ void k_lb2_reset_order(bb_pfsp_dev T)
{
  const int P = T.P;
  for (int l = threadIdx.x; l < P; l += blockDim.x)
  {
    T.order[l] = l;
    T.score[l] = 0u;
    T.pairs_act[l] = T.pairs[l];
  }
  for (int l = threadIdx.x; l <= P; l += blockDim.x)
    T.hist[l] = 0u;
  for (int i = threadIdx.x; i < P * BB_MAX_JOBS; i += blockDim.x)
  {
    T.jbit_act[i] = T.jbit[i];
    T.ja_act[i] = T.ja[i];
    T.jb_act[i] = T.jb[i];
  }
}

/* reference Node (int16 depth, limit1, prmu[20]) -> device record, deriving front/remain/mask
 * (schedule_front + sum_unscheduled of c_bound_simple.c:51-68, 108-124, done once per seeded node only) */
__global__ void k_seed_pfsp(const int16_t *__restrict__ nodes, int count, bb_pool pool, bb_pfsp_dev T, int at,
                            int mp, int nplanes)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count)
    return;
  const int16_t *nd = nodes + (size_t)i * (2 + BB_MAX_JOBS);
  const int depth = nd[0];
  uint32_t w[28];
  for (int q = 0; q < 28; q++)
    w[q] = 0;
  uint32_t mask = 0;
  for (int j = 0; j < T.n; j++)
  {
    const uint32_t job = (uint32_t)nd[2 + j] & 0xffu;
    w[j >> 2] |= job << (8 * (j & 3));
    if (j < depth)
      mask |= 1u << job;
  }
  w[5] = (uint32_t)depth | (mask << 8);
  int front[BB_MAX_MACHINES], remain[BB_MAX_MACHINES];
  for (int k = 0; k < T.m; k++)
  {
    front[k] = 0;
    remain[k] = 0;
  }
  for (int j = 0; j < T.n; j++)
  {
    const int job = nd[2 + j];
    if (j < depth)
    {
      front[0] += T.p8[job];
      for (int k = 1; k < T.m; k++)
        front[k] = max(front[k - 1], front[k]) + T.p8[k * BB_MAX_JOBS + job];
    }
    else
      for (int k = 0; k < T.m; k++)
        remain[k] += T.p8[k * BB_MAX_JOBS + job];
  }
  uint16_t *h = reinterpret_cast<uint16_t *>(w);
  for (int k = 0; k < T.m; k++)
  {
    h[12 + k] = (uint16_t)front[k];
    h[12 + mp + k] = (uint16_t)remain[k];
  }
  for (int q = 0; q < nplanes; q++)
    pool.planes[0][(size_t)q * pool.capacity + at + i] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
}

__global__ void k_export_pfsp(int16_t *__restrict__ nodes, bb_pool pool, int stack, int from, int count, int out_at)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count)
    return;
  const uint4 q0 = pool.planes[stack][from + i];
  const uint4 q1 = pool.planes[stack][(size_t)pool.capacity + from + i];
  const uint32_t w[6] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y};
  int16_t *nd = nodes + (size_t)(out_at + i) * (2 + BB_MAX_JOBS);
  const int depth = w[5] & 0xff;
  nd[0] = (int16_t)depth;
  nd[1] = (int16_t)(depth - 1);
  for (int j = 0; j < BB_MAX_JOBS; j++)
    nd[2 + j] = (int16_t)((w[j >> 2] >> (8 * (j & 3))) & 0xff);
}

/* 21..200 jobs: the reference's Node built with MAX_JOBS NJ (int16 depth, limit1, prmu[NJ]) <-> the wide record */
template <int NJ>
__global__ void k_seed_pfsp_wide(const int16_t *__restrict__ nodes, int count, bb_pool pool, bb_pfsp_wide_dev T, int at)
{
  using G = wide_geom<NJ>;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count)
    return;
  const int16_t *nd = nodes + (size_t)i * (2 + NJ);
  const int depth = nd[0];
  uint32_t w[G::NPL * 4];
  for (int q = 0; q < G::NPL * 4; q++)
    w[q] = 0;
  int front[BB_MAX_MACHINES], remain[BB_MAX_MACHINES];
  for (int k = 0; k < T.m; k++)
    front[k] = remain[k] = 0;
  for (int j = 0; j < T.n; j++)
  {
    const int job = nd[2 + j] & 0xff;
    w[j >> 2] |= (uint32_t)job << (8 * (j & 3));
    if (j < depth)
    {
      w[G::W_MASK + (job >> 5)] |= 1u << (job & 31);
      front[0] += T.p8[job];
      for (int k = 1; k < T.m; k++)
        front[k] = max(front[k - 1], front[k]) + T.p8[k * G::STRIDE + job];
    }
    else
      for (int k = 0; k < T.m; k++)
        remain[k] += T.p8[k * G::STRIDE + job];
  }
  w[G::W_DEPTH] = (uint32_t)depth;
  uint16_t *h = reinterpret_cast<uint16_t *>(w);
  for (int k = 0; k < T.m; k++)
  {
    h[2 * G::W_FRONT + k] = (uint16_t)front[k];
    h[2 * G::W_REMAIN + k] = (uint16_t)remain[k];
  }
  for (int q = 0; q < G::NPL; q++)
    pool.planes[0][(size_t)q * pool.capacity + at + i] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
}

template <int NJ>
__global__ void k_export_pfsp_wide(int16_t *__restrict__ nodes, bb_pool pool, int stack, int from, int count, int out_at)
{
  using G = wide_geom<NJ>;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count)
    return;
  constexpr int HP = (G::W_DEPTH + 4) / 4; /* planes holding prmu[] and the depth word */
  uint32_t w[HP * 4];
  for (int q = 0; q < HP; q++)
  {
    const uint4 v = pool.planes[stack][(size_t)q * pool.capacity + from + i];
    w[4 * q] = v.x;
    w[4 * q + 1] = v.y;
    w[4 * q + 2] = v.z;
    w[4 * q + 3] = v.w;
  }
  int16_t *nd = nodes + (size_t)(out_at + i) * (2 + NJ);
  const int depth = w[G::W_DEPTH] & 0xff;
  nd[0] = (int16_t)depth;
  nd[1] = (int16_t)(depth - 1);
  for (int j = 0; j < NJ; j++)
    nd[2 + j] = (int16_t)((w[j >> 2] >> (8 * (j & 3))) & 0xff);
}

/* reference N-Queens Node (uint8 depth, board[20]) <-> device record */
__global__ void k_seed_nq(const uint8_t *__restrict__ nodes, int count, bb_pool pool, int N, int at)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count)
    return;
  const uint8_t *nd = nodes + (size_t)i * (1 + BB_MAX_QUEENS);
  unsigned char rec[24];
  for (int b = 0; b < 24; b++)
    rec[b] = 0;
  for (int j = 0; j < N; j++)
    rec[j] = nd[1 + j];
  rec[N <= 15 ? 15 : N] = nd[0];
  const uint32_t *w = reinterpret_cast<const uint32_t *>(rec);
  uint32_t ww[6];
  for (int q = 0; q < 6; q++)
    ww[q] = (uint32_t)rec[4 * q] | ((uint32_t)rec[4 * q + 1] << 8) | ((uint32_t)rec[4 * q + 2] << 16) |
            ((uint32_t)rec[4 * q + 3] << 24);
  (void)w;
  pool.planes[0][at + i] = make_uint4(ww[0], ww[1], ww[2], ww[3]);
  for (int t = 0; t < pool.tail_words; t++)
    pool.tailw[0][(size_t)(at + i) * pool.tail_words + t] = ww[4 + t];
}

__global__ void k_export_nq(uint8_t *__restrict__ nodes, bb_pool pool, int stack, int from, int count, int out_at,
                            int N)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count)
    return;
  const uint4 q0 = pool.planes[stack][from + i];
  uint32_t ww[6] = {q0.x, q0.y, q0.z, q0.w, 0u, 0u};
  for (int t = 0; t < pool.tail_words; t++)
    ww[4 + t] = pool.tailw[stack][(size_t)(from + i) * pool.tail_words + t];
  uint8_t *nd = nodes + (size_t)(out_at + i) * (1 + BB_MAX_QUEENS);
  const int db = N <= 15 ? 15 : N;
  nd[0] = (uint8_t)((ww[db >> 2] >> (8 * (db & 3))) & 0xff);
  for (int j = 0; j < BB_MAX_QUEENS; j++)
    nd[1 + j] = j < N ? (uint8_t)((ww[j >> 2] >> (8 * (j & 3))) & 0xff) : 0;
}

/* third-generation N-Queens record (bb_nq_kernels.cuh, nq_expand3): board[20] + used rows | depth << 24 + the rows
 * of the next column attacked along the two diagonal directions, derived once per seeded node */
__global__ void k_seed_nq3(const uint8_t *__restrict__ nodes, int count, bb_pool pool, int N, int at)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count)
    return;
  const uint8_t *nd = nodes + (size_t)i * (1 + BB_MAX_QUEENS);
  const int depth = nd[0];
  uint32_t w[5] = {0u, 0u, 0u, 0u, 0u};
  for (int j = 0; j < N; j++)
    w[j >> 2] |= (uint32_t)nd[1 + j] << (8 * (j & 3));
  const uint32_t full = (1u << N) - 1u;
  uint32_t used = 0u, up = 0u, down = 0u;
  for (int c = 0; c < depth; c++)
  {
    const int b = nd[1 + c], d = depth - c;
    used |= 1u << b;
    if (b + d < 32)
      up |= 1u << (b + d);
    if (b >= d)
      down |= 1u << (b - d);
  }
  pool.planes[0][at + i] = make_uint4(w[0], w[1], w[2], w[3]);
  pool.planes[0][(size_t)pool.capacity + at + i] = make_uint4(w[4], (used & full) | ((uint32_t)depth << 24), up & full, down & full);
}

__global__ void k_export_nq3(uint8_t *__restrict__ nodes, bb_pool pool, int stack, int from, int count, int out_at,
                             int N)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count)
    return;
  const uint4 q0 = pool.planes[stack][from + i];
  const uint4 q1 = pool.planes[stack][(size_t)pool.capacity + from + i];
  const uint32_t ww[5] = {q0.x, q0.y, q0.z, q0.w, q1.x};
  uint8_t *nd = nodes + (size_t)(out_at + i) * (1 + BB_MAX_QUEENS);
  nd[0] = (uint8_t)(q1.y >> 24);
  for (int j = 0; j < BB_MAX_QUEENS; j++)
    nd[1 + j] = j < N ? (uint8_t)((ww[j >> 2] >> (8 * (j & 3))) & 0xff) : 0;
}

/* ------------------------------------------------------------------ creation */

/* The node stacks and the LB2 table come from a stream-ordered memory pool that BELONGS TO THIS LIBRARY (one per
 * device, created on first use, unlimited release threshold): the first engine of a process pays cudaMalloc-class
 * cost for its multi-GB pool, later engines on the same GPU (one search after another, which is how the library is
 * used) get the cached blocks back in microseconds. The device's default pool - which the embedding application and
 * other libraries share - is left alone, and bb_release_cached_memory() hands the cached blocks back to the driver. */
static cudaMemPool_t g_mempool[64];
static std::mutex g_mempool_mutex;

static int library_pool(int device, cudaMemPool_t *out)
{
  if (device < 0 || device >= 64)
    return bb_fail(BB_ERR_ARG, "device ordinal %d not supported (max 63)", device);
  std::lock_guard<std::mutex> lock(g_mempool_mutex);
  if (!g_mempool[device])
  {
    cudaMemPoolProps props;
    memset(&props, 0, sizeof(props));
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = device;
    cudaMemPool_t mp;
    BB_CUDA_TRY(cudaMemPoolCreate(&mp, &props));
    unsigned long long keep = ~0ull;
    BB_CUDA_TRY(cudaMemPoolSetAttribute(mp, cudaMemPoolAttrReleaseThreshold, &keep));
    g_mempool[device] = mp;
  }
  *out = g_mempool[device];
  return BB_OK;
}

static int pool_alloc(bb_engine *e, void **ptr, size_t bytes)
{
  cudaMemPool_t mp;
  const int rc = library_pool(e->device, &mp);
  if (rc)
    return rc;
  BB_CUDA_TRY(cudaMallocFromPoolAsync(ptr, bytes, mp, e->stream));
  return BB_OK;
}

extern "C" int bb_release_cached_memory(int device)
{
  if (device < 0 || device >= 64)
    return bb_fail(BB_ERR_ARG, "device ordinal %d not supported (max 63)", device);
  std::lock_guard<std::mutex> lock(g_mempool_mutex);
  if (g_mempool[device])
  {
    BB_CUDA_TRY(cudaSetDevice(device));
    BB_CUDA_TRY(cudaDeviceSynchronize());
    BB_CUDA_TRY(cudaMemPoolTrimTo(g_mempool[device], 0));
  }
  return BB_OK;
}

static void pool_free(bb_engine *e, void *ptr)
{
  if (ptr)
    cudaFreeAsync(ptr, e->stream);
}

static int alloc_common(bb_engine *e, const bb_config *cfg, int nplanes, int tail_words)
{
  bb_config c;
  if (cfg)
    c = *cfg;
  else
    bb_config_default(&c);
  e->device = c.device;
  e->M = c.M > 0 ? c.M : 50000;
  e->m_keep = c.m > 0 ? c.m : 25;
  e->iters_per_launch = c.iters_per_launch > 0 ? c.iters_per_launch : 32;
  ENGINE_SET_DEVICE(e);
  long long capacity = c.capacity;
  if (capacity <= 0)
  {
    /* depth-first batches of M leave up to M*(branching-1) nodes behind per level; measured high-water marks are
     * 20-75 x M nodes over both stacks, and the planner keeps room for maxkids children per popped parent. Size each
     * stack to 48 x M nodes (at least 4 M nodes), within a third of the free HBM: pool size drives allocation /
     * release / mapping time, which is most of the set-up cost of a short search, and the planner shrinks batches
     * instead of overflowing near capacity. */
    size_t free_b = 0, total_b = 0;
    BB_CUDA_TRY(cudaMemGetInfo(&free_b, &total_b));
    const long long rec = (long long)nplanes * 16 + (long long)tail_words * 4;
    const long long by_mem = (long long)(free_b / 3) / (2 * rec);
    capacity = (long long)e->M * 48;
    if (capacity < (1 << 22))
      capacity = 1 << 22;
    if (capacity > by_mem)
      capacity = by_mem;
  }
  if (capacity > (1ll << 30))
    capacity = 1ll << 30;
  if (capacity < 1024)
    return bb_fail(BB_ERR_ARG, "pool capacity %lld is too small", capacity);
  ENGINE_SET_DEVICE(e);
  BB_CUDA_TRY(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
  BB_CUDA_TRY(cudaEventCreate(&e->ev0));
  BB_CUDA_TRY(cudaEventCreate(&e->ev1));
  BB_CUDA_TRY(cudaEventCreate(&e->ev_begin));
  BB_CUDA_TRY(cudaEventCreate(&e->ev_end));
  BB_CUDA_TRY(cudaMalloc(&e->ctrl, sizeof(bb_ctrl)));
  BB_CUDA_TRY(cudaMallocHost(&e->ctrl_h, sizeof(bb_ctrl)));
  e->pool.capacity = (int)capacity;
  e->pool.nplanes = nplanes;
  e->pool.tail_words = tail_words;
  for (int s = 0; s < 2; s++)
  {
    int rc = pool_alloc(e, &e->pool_mem[s], (size_t)capacity * nplanes * sizeof(uint4));
    if (rc)
      return rc;
    e->pool.planes[s] = (uint4 *)e->pool_mem[s];
    e->pool.tailw[s] = nullptr;
    if (tail_words)
    {
      rc = pool_alloc(e, &e->tail_mem[s], (size_t)capacity * tail_words * sizeof(uint32_t));
      if (rc)
        return rc;
      e->pool.tailw[s] = (uint32_t *)e->tail_mem[s];
    }
  }
  return BB_OK;
}

static int setup_launch(bb_engine *e, const void *kernel, int smem, int threads = BB_CTA_THREADS)
{
  e->kernel = kernel;
  e->pdl = false;
  e->smem = smem;
  e->threads = threads;
  if (smem > 48 * 1024)
    BB_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  int per_sm = 0, sms = 0;
  BB_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem));
  BB_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e->device));
  if (per_sm < 1)
    return bb_fail(BB_ERR_CUDA, "kernel does not fit on an SM (smem %d B)", smem);
  e->grid = per_sm * sms; /* persistent-style: every CTA resident, a multiple of the SM count */
  return BB_OK;
}

/* Kernels that call pdl_wait_then_release() before they touch the plan or the pool (pfsp_expand3, nq_expand5) are
 * chained with programmatic dependent launch: iteration i+1 is scheduled - and runs its prologue - while the last
 * CTAs of iteration i are still finishing, instead of after the grid has drained and the launch latency on top. */
static int launch_expand(bb_engine *e, cudaStream_t s)
{
  cudaLaunchConfig_t lc = {};
  lc.gridDim = dim3(e->grid);
  lc.blockDim = dim3(e->threads);
  lc.dynamicSmemBytes = (size_t)e->smem;
  lc.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  lc.attrs = attr;
  lc.numAttrs = (e->pdl && getenv("BB_NO_PDL") == nullptr) ? 1 : 0;
  if (e->kind == KIND_PFSP && e->wide)
  {
    void *args[] = {&e->ctrl, &e->pool, &e->TW, &e->prefilter};
    BB_CUDA_TRY(cudaLaunchKernelExC(&lc, e->kernel, args));
  }
  else if (e->kind == KIND_PFSP)
  {
    void *args[] = {&e->ctrl, &e->pool, &e->T, &e->prefilter, &e->stage_a};
    BB_CUDA_TRY(cudaLaunchKernelExC(&lc, e->kernel, args));
  }
  else
  {
    e->nq_arg = e->G | (e->nq_tail << 16);
    void *args[] = {&e->ctrl, &e->pool, &e->N, &e->nq_arg};
    BB_CUDA_TRY(cudaLaunchKernelExC(&lc, e->kernel, args));
  }
  return BB_OK;
}

static int build_graph_n(bb_engine *e, int iters, cudaGraphExec_t *out)
{
  cudaGraph_t g;
  BB_CUDA_TRY(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
  for (int i = 0; i < iters; i++)
  {
    int rc = launch_expand(e, e->stream);
    if (rc)
    {
      cudaStreamEndCapture(e->stream, &g);
      return rc;
    }
  }
  BB_CUDA_TRY(cudaStreamEndCapture(e->stream, &g));
  BB_CUDA_TRY(cudaGraphInstantiate(out, g, 0));
  BB_CUDA_TRY(cudaGraphDestroy(g));
  return BB_OK;
}

static int build_graph(bb_engine *e) { return build_graph_n(e, e->iters_per_launch, &e->graph); }

/* enqueue `iters` pool iterations as graph chains: the full chain, then powers of two */
static int enqueue_iterations(bb_engine *e, long long iters)
{
  while (iters > 0)
  {
    if (iters >= e->iters_per_launch)
    {
      BB_CUDA_TRY(cudaGraphLaunch(e->graph, e->stream));
      iters -= e->iters_per_launch;
      e->launches += e->iters_per_launch;
      continue;
    }
    int b = 4;
    while (b > 0 && ((1 << b) > iters || (1 << b) >= e->iters_per_launch))
      b--;
    if (!e->small_graph[b])
    {
      const int rc = build_graph_n(e, 1 << b, &e->small_graph[b]);
      if (rc)
        return rc;
    }
    BB_CUDA_TRY(cudaGraphLaunch(e->small_graph[b], e->stream));
    iters -= 1 << b;
    e->launches += 1 << b;
  }
  return BB_OK;
}

template <int LB>
static const void *pick_pfsp_kernel(int mm)
{
  switch (mm)
  {
  case 5:
    return (const void *)bbk::pfsp_expand<LB, 5>;
  case 10:
    return (const void *)bbk::pfsp_expand<LB, 10>;
  case 20:
    return (const void *)bbk::pfsp_expand<LB, 20>;
  default:
    return (const void *)bbk::pfsp_expand<LB, 0>;
  }
}

template <int LB>
static const void *pick_pfsp_kernel2(int mm)
{
  switch (mm)
  {
  case 5:
    return (const void *)bbk::pfsp_expand2<LB, 5>;
  case 10:
    return (const void *)bbk::pfsp_expand2<LB, 10>;
  case 20:
    return (const void *)bbk::pfsp_expand2<LB, 20>;
  default:
    return (const void *)bbk::pfsp_expand2<LB, 0>;
  }
}

template <int LB>
static const void *pick_pfsp_kernel3(int mm)
{
  switch (mm)
  {
  case 5:
    return (const void *)bbk::pfsp_expand3<LB, 5>;
  case 10:
    return (const void *)bbk::pfsp_expand3<LB, 10>;
  case 20:
    return (const void *)bbk::pfsp_expand3<LB, 20>;
  default:
    return (const void *)bbk::pfsp_expand3<LB, 0>;
  }
}

static int upload_pfsp_tables(bb_engine *e, const bb_pfsp_instance *ins)
{
  const int n = ins->nb_jobs, m = ins->nb_machines, P = ins->nb_machine_pairs;
  const size_t sz_p8 = ((size_t)m * BB_MAX_JOBS + 15) & ~(size_t)15;
  const size_t sz_vec = ((size_t)m * 4 + 15) & ~(size_t)15;
  const size_t sz_slot = ((size_t)(P > 0 ? P : 1) * BB_MAX_JOBS * 4 + 15) & ~(size_t)15;
  const size_t sz_pairs = ((size_t)(P > 0 ? P : 1) * 8 + 15) & ~(size_t)15;
  const size_t sz_ord = ((size_t)((P > 0 ? P : 1) + 1) * 4 + 15) & ~(size_t)15;
  const size_t canon = sz_p8 + 2 * sz_vec + 3 * sz_slot + sz_pairs;
  /* + the learned-order copies of the four pair tables, order[], score[], hist[] (device-written) */
  /* + the lane-per-parent kernel's processing-time rows, [word][job] with four machines per word, ready to be copied
   * into shared memory as they are (one TMA bulk copy per CTA) */
  const size_t off_prow = canon + 3 * sz_slot + sz_pairs + 3 * sz_ord;
  const size_t sz_prow = (size_t)((BB_MAX_MACHINES + 3) / 4) * BB_MAX_JOBS * 4;
  const size_t total = off_prow + sz_prow;
  std::vector<unsigned char> h(total, 0);
  uint8_t *p8 = h.data();
  int *tails = (int *)(h.data() + sz_p8);
  int *heads = (int *)(h.data() + sz_p8 + sz_vec);
  uint32_t *jbit = (uint32_t *)(h.data() + sz_p8 + 2 * sz_vec);
  int *ja = (int *)(h.data() + sz_p8 + 2 * sz_vec + sz_slot);
  int *jb = (int *)(h.data() + sz_p8 + 2 * sz_vec + 2 * sz_slot);
  int2 *pairs = (int2 *)(h.data() + sz_p8 + 2 * sz_vec + 3 * sz_slot);
  for (int k = 0; k < m; k++)
  {
    for (int j = 0; j < n; j++)
    {
      const int v = ins->p_times[k * n + j];
      if (v < 0 || v > 255)
        return bb_fail(BB_ERR_ARG, "processing time %d out of the supported 0..255 range", v);
      p8[k * BB_MAX_JOBS + j] = (uint8_t)v;
    }
    tails[k] = ins->min_tails[k];
    heads[k] = ins->min_heads[k];
  }
  for (int l = 0; l < P; l++)
  {
    const int s = ins->machine_pair_order[l];
    const int m1 = ins->machine_pairs_1[s], m2 = ins->machine_pairs_2[s];
    pairs[l] = make_int2(m1, m2);
    for (int q = 0; q < n; q++)
    {
      const int job = ins->johnson_schedules[s * n + q];
      jbit[l * BB_MAX_JOBS + q] = 1u << job;
      ja[l * BB_MAX_JOBS + q] = ins->p_times[m2 * n + job] - ins->p_times[m1 * n + job];
      jb[l * BB_MAX_JOBS + q] = ins->lags[s * n + job] + ins->p_times[m2 * n + job];
    }
  }
  {
    const int mu = bbk::pfsp_mm(m) ? bbk::pfsp_mm(m) : BB_MAX_MACHINES;
    uint32_t *prow = (uint32_t *)(h.data() + off_prow);
    for (int q = 0; q < (mu + 3) / 4; q++)
      for (int j = 0; j < BB_MAX_JOBS; j++)
        for (int b = 0; b < 4; b++)
          if (4 * q + b < m)
            prow[q * BB_MAX_JOBS + j] |= (uint32_t)p8[(4 * q + b) * BB_MAX_JOBS + j] << (8 * b);
  }
  BB_CUDA_TRY(cudaMalloc(&e->table_mem, total));
  BB_CUDA_TRY(cudaMemcpy(e->table_mem, h.data(), total, cudaMemcpyHostToDevice));
  e->table_bytes = total;
  unsigned char *d = (unsigned char *)e->table_mem;
  e->T.n = n;
  e->T.m = m;
  e->T.mp = bbk::pfsp_mp(m);
  e->T.P = P;
  e->T.p8 = d;
  e->T.tails = (const int *)(d + sz_p8);
  e->T.heads = (const int *)(d + sz_p8 + sz_vec);
  e->T.jbit = (const uint32_t *)(d + sz_p8 + 2 * sz_vec);
  e->T.ja = (const int *)(d + sz_p8 + 2 * sz_vec + sz_slot);
  e->T.jb = (const int *)(d + sz_p8 + 2 * sz_vec + 2 * sz_slot);
  e->T.pairs = (const int2 *)(d + sz_p8 + 2 * sz_vec + 3 * sz_slot);
  e->T.jbit_act = (uint32_t *)(d + canon);
  e->T.ja_act = (int *)(d + canon + sz_slot);
  e->T.jb_act = (int *)(d + canon + 2 * sz_slot);
  e->T.pairs_act = (int2 *)(d + canon + 3 * sz_slot);
  e->T.order = (int *)(d + canon + 3 * sz_slot + sz_pairs);
  e->T.score = (unsigned *)(d + canon + 3 * sz_slot + sz_pairs + sz_ord);
  e->T.hist = (unsigned *)(d + canon + 3 * sz_slot + sz_pairs + 2 * sz_ord);
  e->T.prow = (const uint32_t *)(d + off_prow);
  e->T.best_known = ins->best_known;
  e->T.one = 1u;
  for (int k = 0; k < BB_MAX_MACHINES; k++)
    e->T.tails2[k] = k < m ? (uint32_t)ins->min_tails[k] * 0x10001u : 0u;
  return BB_OK;
}

/* 21..200 jobs: tables of bb_pfsp_wide_dev in one allocation, the wide kernel of the requested bound and class */
static int create_pfsp_wide(bb_engine *e, const bb_pfsp_instance *ins, int lb, const bb_config *cfg)
{
  e->kind = KIND_PFSP;
  e->wide = wide_class(ins->nb_jobs); /* job-capacity class of the record and of the host's Node: 50, 100 or 200 */
  e->lb = lb;
  e->n = ins->nb_jobs;
  e->m = ins->nb_machines;
  e->prefilter = cfg ? (cfg->lb2_prefilter != 0) : 1;
  const int NJ = e->wide, WS = wide_stride(NJ);
  int rc = alloc_common(e, cfg, wide_planes(NJ), 0);
  if (rc)
    return rc;
  const int n = e->n, m = e->m, P = ins->nb_machine_pairs > 0 ? ins->nb_machine_pairs : 0, P1 = P > 0 ? P : 1;
  const size_t sz_p8 = (size_t)m * WS, sz_tails = ((size_t)m * 4 + 15) & ~(size_t)15;
  const size_t sz_jjob = (size_t)P1 * WS, sz_int = (size_t)P1 * WS * 4, sz_pairs = (size_t)P1 * 8;
  const size_t o_tails = (sz_p8 + 15) & ~(size_t)15, o_jjob = o_tails + sz_tails, o_ja = (o_jjob + sz_jjob + 15) & ~(size_t)15;
  const size_t o_jb = o_ja + sz_int, o_pairs = o_jb + sz_int, total = o_pairs + sz_pairs;
  std::vector<unsigned char> h(total, 0);
  for (int k = 0; k < m; k++)
  {
    for (int j = 0; j < n; j++)
    {
      const int v = ins->p_times[k * n + j];
      if (v < 0 || v > 255)
        return bb_fail(BB_ERR_ARG, "processing time %d out of the supported 0..255 range", v);
      h[(size_t)k * WS + j] = (unsigned char)v;
    }
    ((int *)(h.data() + o_tails))[k] = ins->min_tails[k];
  }
  memset(h.data() + o_jjob, 0xff, sz_jjob);
  for (int l = 0; l < P; l++)
  {
    const int sidx = ins->machine_pair_order[l];
    const int m1 = ins->machine_pairs_1[sidx], m2 = ins->machine_pairs_2[sidx];
    ((int2 *)(h.data() + o_pairs))[l] = make_int2(m1, m2);
    for (int q = 0; q < n; q++)
    {
      const int job = ins->johnson_schedules[sidx * n + q];
      h[o_jjob + (size_t)l * WS + q] = (unsigned char)job;
      ((int *)(h.data() + o_ja))[(size_t)l * WS + q] = ins->p_times[m2 * n + job] - ins->p_times[m1 * n + job];
      ((int *)(h.data() + o_jb))[(size_t)l * WS + q] = ins->lags[sidx * n + job] + ins->p_times[m2 * n + job];
    }
  }
  BB_CUDA_TRY(cudaMalloc(&e->wide_mem, total));
  BB_CUDA_TRY(cudaMemcpy(e->wide_mem, h.data(), total, cudaMemcpyHostToDevice));
  e->table_bytes = total;
  unsigned char *d = (unsigned char *)e->wide_mem;
  e->TW.n = n;
  e->TW.m = m;
  e->TW.P = P;
  e->TW.p8 = d;
  e->TW.tails = (const int *)(d + o_tails);
  e->TW.jjob = d + o_jjob;
  e->TW.ja = (const int *)(d + o_ja);
  e->TW.jb = (const int *)(d + o_jb);
  e->TW.pairs = (const int2 *)(d + o_pairs);
  e->T.n = n; /* the generic code paths read n / m from here */
  e->T.m = m;
  e->T.P = P;
  const void *k = NJ == 50    ? (lb == 0   ? (const void *)bbk::pfsp_expand_wide<0, 50>
                                 : lb == 1 ? (const void *)bbk::pfsp_expand_wide<1, 50>
                                           : (const void *)bbk::pfsp_expand_wide<2, 50>)
                  : NJ == 100 ? (lb == 0   ? (const void *)bbk::pfsp_expand_wide<0, 100>
                                 : lb == 1 ? (const void *)bbk::pfsp_expand_wide<1, 100>
                                           : (const void *)bbk::pfsp_expand_wide<2, 100>)
                              : (lb == 0   ? (const void *)bbk::pfsp_expand_wide<0, 200>
                                 : lb == 1 ? (const void *)bbk::pfsp_expand_wide<1, 200>
                                           : (const void *)bbk::pfsp_expand_wide<2, 200>);
  rc = setup_launch(e, k, bbk::wide_layout(m, NJ).total);
  if (rc)
    return rc;
  return build_graph(e);
}

static int create_pfsp(bb_engine *e, const bb_pfsp_instance *ins, int lb, const bb_config *cfg)
{
  if (ins->nb_jobs > BB_MAX_JOBS)
    return create_pfsp_wide(e, ins, lb, cfg);
  e->kind = KIND_PFSP;
  e->lb = lb;
  e->n = ins->nb_jobs;
  e->m = ins->nb_machines;
  e->prefilter = cfg ? (cfg->lb2_prefilter != 0) : 1;
  e->packed = cfg ? cfg->packed : 2;
  if (e->packed < 0 || e->packed > 2)
    e->packed = 2;
  /* the lane-per-parent kernel packs two processing times per PRMT with the sign-replicate trick: times <= 127 */
  for (int i = 0; e->packed == 2 && i < ins->nb_jobs * ins->nb_machines; i++)
    if (ins->p_times[i] > 127)
      e->packed = 1;
  e->table = (lb == 2) && (cfg ? (cfg->lb2_table != 0) : 1) && ins->nb_machine_pairs > 0;
  e->staged = (lb == 2) && !e->table && (cfg ? (cfg->lb2_staged != 0) : 1) && ins->nb_machine_pairs > 0;
  e->stage_a = cfg && cfg->lb2_stage_a > 0 ? cfg->lb2_stage_a : 8;
  int rc = alloc_common(e, cfg, bbk::pfsp_planes(e->m), 0);
  if (rc)
    return rc;
  rc = upload_pfsp_tables(e, ins);
  if (rc)
    return rc;
  const int mm = bbk::pfsp_mm(e->m);
  if (e->table)
  {
    /* build the subset table once per engine: 2^n rows of grow uint16 (part of the instance tables, like the
     * Johnson schedules the reference precomputes on the host) */
    /* row length: what the kernel class loads ((MU - 1) entries rounded up to whole 16-byte pieces, MU = 20 for
     * machine counts without a specialised kernel), so that a row read never leaves the table */
    const int mu = bbk::pfsp_mm(e->m) ? bbk::pfsp_mm(e->m) : BB_MAX_MACHINES;
    const int grow = ((mu - 1) + 7) & ~7;
    const size_t rows = (size_t)1 << e->n;
    rc = pool_alloc(e, &e->g16_mem, rows * grow * sizeof(uint16_t));
    if (rc)
      return rc;
    /* built on the device from the uploaded tables: HBM footprint, not a host-to-device copy */
    const int P = e->T.P;
    const size_t sm = (size_t)P * BB_MAX_JOBS * 12 + (size_t)P * 8 + BB_MAX_MACHINES * 4 + BB_MAX_MACHINES * 256 * 4;
    BB_CUDA_TRY(cudaFuncSetAttribute(k_lb2_build_table, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    k_lb2_build_table<<<(unsigned)((rows + 255) / 256), 256, sm, e->stream>>>(e->T, (uint16_t *)e->g16_mem, grow);
    BB_CUDA_TRY(cudaGetLastError());
    BB_CUDA_TRY(cudaStreamSynchronize(e->stream));
    e->T.g16 = (const uint16_t *)e->g16_mem;
    e->T.grow = grow;
    if (e->packed == 2)
    {
      rc = setup_launch(e, pick_pfsp_kernel3<3>(mm), bbk::tpp_layout(e->m).total);
      /* not chained programmatically: measured 0.3 % slower for this variant (launches of 200+ us, 2 CTAs/SM) */
    }
    else
      rc = setup_launch(e, e->packed ? pick_pfsp_kernel2<3>(mm) : pick_pfsp_kernel<3>(mm),
                        bbk::pfsp_layout(e->m, e->T.P, 3, e->packed ? 2 : 1).total);
    if (rc)
      return rc;
    return build_graph(e);
  }
  if (e->staged)
  {
    const void *k = mm == 5    ? (const void *)bbk::pfsp_expand_lb2<5>
                    : mm == 10 ? (const void *)bbk::pfsp_expand_lb2<10>
                    : mm == 20 ? (const void *)bbk::pfsp_expand_lb2<20>
                               : (const void *)bbk::pfsp_expand_lb2<0>;
    rc = setup_launch(e, k, bbk::lb2_layout(e->m, e->T.P).total, BB_LB2_THREADS);
    if (rc)
      return rc;
    return build_graph(e);
  }
  if (e->packed == 2 && lb != 2)
  {
    rc = setup_launch(e, lb == 0 ? pick_pfsp_kernel3<0>(mm) : pick_pfsp_kernel3<1>(mm), bbk::tpp_layout(e->m).total);
    e->pdl = true;
    if (rc)
      return rc;
    return build_graph(e);
  }
  const void *k = lb == 0   ? (e->packed ? pick_pfsp_kernel2<0>(mm) : pick_pfsp_kernel<0>(mm))
                  : lb == 1 ? (e->packed ? pick_pfsp_kernel2<1>(mm) : pick_pfsp_kernel<1>(mm))
                            : pick_pfsp_kernel<2>(mm);
  const int smem = bbk::pfsp_layout(e->m, e->T.P, lb, (e->packed && lb != 2) ? 2 : 1).total;
  rc = setup_launch(e, k, smem);
  if (rc)
    return rc;
  return build_graph(e);
}

static int create_nq(bb_engine *e, int N, int G, const bb_config *cfg)
{
  e->kind = KIND_NQ;
  e->N = N;
  e->G = G;
  e->nq_gen = cfg ? (cfg->packed >= 2 || cfg->packed < 0 ? 3 : cfg->packed + 1) : 3; /* `packed` selects the generation */
  if (e->nq_gen == 3)
  {
    const bool c_nq4 = getenv("BB_NQ_PER_TILE_CURSOR") == nullptr; /* debugging aid: nq_expand3 reserves per tile */
    int rc = alloc_common(e, cfg, 2, 0);
    if (rc)
      return rc;
    const bool c_nq5 = getenv("BB_NQ_CTA_ROUNDS") == nullptr; /* debugging aid: nq_expand4 reserves per CTA round */
    /* packed = 2: every node goes through the pool (one launch = one level of every popped parent); packed = 3 (the
     * default): the last levels are enumerated in registers. BB_NQ_TAIL overrides the number of levels (tuning aid). */
    e->nq_tail = 0;
    if (c_nq4 && c_nq5 && (!cfg || cfg->packed >= 3 || cfg->packed < 0))
    {
      /* measured best on B200 (scripts/gpu_nq_tail_n.sh, gpu_nq_tail.sh): 7 / 8 / 9 / 10 / 11 levels at N = 14 / 15 /
       * 16 / 17 / 18; deeper subtrees leave too few of them to keep every lane busy, shallower ones send more nodes
       * through the pool */
      e->nq_tail = N - 7;
      if (cfg && cfg->packed >= 4)
        e->nq_tail -= cfg->packed - 3; /* workers of a team of >= 4: each holds a share of the parents only */
      if (const char *t = getenv("BB_NQ_TAIL"))
        e->nq_tail = atoi(t);
      if (e->nq_tail > NQ5_MAX_TAIL)
        e->nq_tail = NQ5_MAX_TAIL;
      if (e->nq_tail < 0)
        e->nq_tail = 0;
    }
    if (G > 0xffff)
      e->nq_tail = 0; /* the kernel argument has 16 bits for G next to the tail */
    rc = setup_launch(e,
                      !c_nq4  ? (const void *)bbk::nq_expand3
                      : c_nq5 ? (e->nq_tail ? (const void *)bbk::nq_expand5<1> : (const void *)bbk::nq_expand5<0>)
                              : (const void *)bbk::nq_expand4,
                      0);
    e->pdl = c_nq4 && c_nq5;
    if (rc)
      return rc;
    return build_graph(e);
  }
  const int tw = bbk::nq_tail_words(N);
  int rc = alloc_common(e, cfg, 1, tw);
  if (rc)
    return rc;
  const bool gen2 = e->nq_gen == 2;
  const void *k = gen2 ? (tw == 0   ? (const void *)bbk::nq_expand2<0>
                          : tw == 1 ? (const void *)bbk::nq_expand2<1>
                                    : (const void *)bbk::nq_expand2<2>)
                       : (tw == 0   ? (const void *)bbk::nq_expand<0>
                          : tw == 1 ? (const void *)bbk::nq_expand<1>
                                    : (const void *)bbk::nq_expand<2>);
  rc = setup_launch(e, k, 0);
  if (rc)
    return rc;
  return build_graph(e);
}

extern "C" bb_engine *bb_pfsp_engine_create(const bb_pfsp_instance *ins, int lb, const bb_config *cfg)
{
  if (!ins || lb < 0 || lb > 2)
  {
    bb_fail(BB_ERR_ARG, "bb_pfsp_engine_create: bad instance or lower bound %d", lb);
    return nullptr;
  }
  if (ins->nb_jobs > BB_MAX_JOBS_WIDE || ins->nb_machines > BB_MAX_MACHINES)
  {
    /* the reference needs a rebuild with a larger MAX_JOBS for these (pfsp/README.md:52); refuse loudly */
    bb_fail(BB_ERR_ARG, "instance has %d jobs x %d machines; this build supports at most %d x %d", ins->nb_jobs,
            ins->nb_machines, BB_MAX_JOBS_WIDE, BB_MAX_MACHINES);
    return nullptr;
  }
  bb_engine *e = (bb_engine *)calloc(1, sizeof(bb_engine));
  if (create_pfsp(e, ins, lb, cfg) != BB_OK)
  {
    bb_engine_destroy(e);
    return nullptr;
  }
  if (bb_engine_reset(e, INT_MAX) != BB_OK)
  {
    bb_engine_destroy(e);
    return nullptr;
  }
  return e;
}

extern "C" bb_engine *bb_nq_engine_create(int N, int G, const bb_config *cfg)
{
  if (N < 1 || N > BB_MAX_QUEENS || G < 1)
  {
    bb_fail(BB_ERR_ARG, "bb_nq_engine_create: need 1 <= N <= %d and G >= 1 (got N=%d G=%d)", BB_MAX_QUEENS, N, G);
    return nullptr;
  }
  bb_engine *e = (bb_engine *)calloc(1, sizeof(bb_engine));
  if (create_nq(e, N, G, cfg) != BB_OK)
  {
    bb_engine_destroy(e);
    return nullptr;
  }
  if (bb_engine_reset(e, INT_MAX) != BB_OK)
  {
    bb_engine_destroy(e);
    return nullptr;
  }
  return e;
}

extern "C" void bb_engine_destroy(bb_engine *e)
{
  if (!e)
    return;
  cudaSetDevice(e->device);
  if (e->stream)
    cudaStreamSynchronize(e->stream);
  if (e->graph)
    cudaGraphExecDestroy(e->graph);
  for (int b = 0; b < 5; b++)
    if (e->small_graph[b])
      cudaGraphExecDestroy(e->small_graph[b]);
  if (e->snap_h)
  {
    cudaFreeHost(e->snap_h);
    for (int s = 0; s < 2; s++)
    {
      cudaEventDestroy(e->pipe_k0[s]);
      cudaEventDestroy(e->pipe_k1[s]);
      cudaEventDestroy(e->pipe_done[s]);
    }
  }
  for (int s = 0; s < 2; s++)
  {
    pool_free(e, e->pool_mem[s]);
    pool_free(e, e->tail_mem[s]);
    if (s == 0 && e->stage_mem)
    {
      cudaFree(e->stage_mem);
      cudaFreeHost((void *)e->mailbox_h);
      cudaEventDestroy(e->don_ev);
    }
  }
  if (e->table_mem)
    cudaFree(e->table_mem);
  if (e->wide_mem)
    cudaFree(e->wide_mem);
  pool_free(e, e->g16_mem);
  if (e->stream)
    cudaStreamSynchronize(e->stream);
  if (e->ctrl)
    cudaFree(e->ctrl);
  if (e->ctrl_h)
    cudaFreeHost(e->ctrl_h);
  if (e->ev0)
    cudaEventDestroy(e->ev0);
  if (e->ev1)
    cudaEventDestroy(e->ev1);
  if (e->ev_begin)
    cudaEventDestroy(e->ev_begin);
  if (e->ev_end)
    cudaEventDestroy(e->ev_end);
  if (e->stream)
    cudaStreamDestroy(e->stream);
  free(e);
}

/* ------------------------------------------------------------------ state */

static int fetch_ctrl(bb_engine *e)
{
  BB_CUDA_TRY(cudaMemcpyAsync(e->ctrl_h, e->ctrl, sizeof(bb_ctrl), cudaMemcpyDeviceToHost, e->stream));
  BB_CUDA_TRY(cudaStreamSynchronize(e->stream));
  e->d2h_bytes += sizeof(bb_ctrl);
  return BB_OK;
}

extern "C" int bb_engine_reset(bb_engine *e, int best)
{
  if (!e)
    return bb_fail(BB_ERR_ARG, "null engine");
  ENGINE_SET_DEVICE(e);
  k_reset<<<1, 1, 0, e->stream>>>(e->ctrl, best, e->M, e->pool.capacity, e->kind == KIND_PFSP ? e->n : e->N);
  if (e->kind == KIND_PFSP && e->staged)
    k_lb2_reset_order<<<1, 256, 0, e->stream>>>(e->T); /* a new search learns its pair order from scratch */
  BB_CUDA_TRY(cudaGetLastError());
  if (e->stage_mem)
  { /* a new search: no donation outstanding */
    if (e->don_slot >= 0)
      cudaEventSynchronize(e->don_ev);
    e->don_slot = -1;
    for (int s = 0; s < BB_STAGE_SLOTS; s++)
      e->slot_busy[s] = 0;
  }
  e->kernel_ms = 0;
  e->span_ms = 0;
  e->launches = 0;
  e->h2d_bytes = 0;
  e->d2h_bytes = 0;
  return fetch_ctrl(e);
}

extern "C" int bb_engine_set_best(bb_engine *e, int best)
{
  if (!e)
    return bb_fail(BB_ERR_ARG, "null engine");
  ENGINE_SET_DEVICE(e);
  k_set_best<<<1, 1, 0, e->stream>>>(e->ctrl, best);
  BB_CUDA_TRY(cudaGetLastError());
  return BB_OK;
}

extern "C" int bb_engine_push_nodes(bb_engine *e, const void *host_nodes, int count)
{
  if (!e || (!host_nodes && count > 0) || count < 0)
    return bb_fail(BB_ERR_ARG, "bb_engine_push_nodes: bad arguments");
  if (count == 0)
    return BB_OK;
  ENGINE_SET_DEVICE(e);
  int rc = fetch_ctrl(e);
  if (rc)
    return rc;
  const int at = e->ctrl_h->size[0];
  if ((long long)at + count > e->pool.capacity)
    return bb_fail(BB_ERR_OVERFLOW, "pool capacity %d exceeded while seeding %d nodes", e->pool.capacity, count);
  const size_t rec = e->kind == KIND_PFSP ? sizeof(int16_t) * (2 + (e->wide ? e->wide : BB_MAX_JOBS)) : (1 + BB_MAX_QUEENS);
  if (e->kind == KIND_PFSP)
  {
    /* nodes the engine cannot represent are rejected here rather than silently mis-handled */
    const int16_t *nd = (const int16_t *)host_nodes;
    const size_t st16 = 2 + (e->wide ? e->wide : BB_MAX_JOBS);
    for (int i = 0; i < count; i++)
      if (nd[(size_t)i * st16] < 0 || nd[(size_t)i * st16] >= e->n)
        return bb_fail(BB_ERR_ARG, "node %d has depth %d outside 0..%d", i, nd[(size_t)i * st16], e->n - 1);
  }
  else
  {
    const uint8_t *nd = (const uint8_t *)host_nodes;
    for (int i = 0; i < count; i++)
      if (nd[(size_t)i * (1 + BB_MAX_QUEENS)] >= e->N)
        return bb_fail(BB_ERR_ARG, "node %d has depth %d outside 0..%d", i, nd[(size_t)i * (1 + BB_MAX_QUEENS)], e->N - 1);
  }
  void *stage = nullptr; /* from the library's stream-ordered pool: no device-wide synchronisation per call */
  rc = pool_alloc(e, &stage, rec * count);
  if (rc)
    return rc;
  cudaError_t err = cudaMemcpyAsync(stage, host_nodes, rec * count, cudaMemcpyHostToDevice, e->stream);
  e->h2d_bytes += rec * count;
  if (err == cudaSuccess)
  {
    const int blocks = (count + 127) / 128;
    if (e->kind == KIND_PFSP && e->wide == 50)
      k_seed_pfsp_wide<50><<<blocks, 128, 0, e->stream>>>((const int16_t *)stage, count, e->pool, e->TW, at);
    else if (e->kind == KIND_PFSP && e->wide == 100)
      k_seed_pfsp_wide<100><<<blocks, 128, 0, e->stream>>>((const int16_t *)stage, count, e->pool, e->TW, at);
    else if (e->kind == KIND_PFSP && e->wide)
      k_seed_pfsp_wide<200><<<blocks, 128, 0, e->stream>>>((const int16_t *)stage, count, e->pool, e->TW, at);
    else if (e->kind == KIND_PFSP)
      k_seed_pfsp<<<blocks, 128, 0, e->stream>>>((const int16_t *)stage, count, e->pool, e->T, at, e->T.mp,
                                                 e->pool.nplanes);
    else if (e->nq_gen == 3)
      k_seed_nq3<<<blocks, 128, 0, e->stream>>>((const uint8_t *)stage, count, e->pool, e->N, at);
    else
      k_seed_nq<<<blocks, 128, 0, e->stream>>>((const uint8_t *)stage, count, e->pool, e->N, at);
    k_plan<<<1, 1, 0, e->stream>>>(e->ctrl, count, e->grid * (e->threads / 32));
    err = cudaGetLastError();
  }
  pool_free(e, stage);
  if (err == cudaSuccess)
    err = cudaStreamSynchronize(e->stream);
  if (err != cudaSuccess)
    return bb_fail_cuda(err, "seeding", __FILE__, __LINE__);
  return BB_OK;
}

extern "C" int bb_engine_push_root(bb_engine *e)
{
  if (!e)
    return bb_fail(BB_ERR_ARG, "null engine");
  if (e->kind == KIND_PFSP)
  {
    int16_t nd[2 + BB_MAX_JOBS_WIDE]; /* an engine reads the first 2 + 20 / 50 / 100 / 200 entries */
    memset(nd, 0, sizeof(nd));
    nd[0] = 0;
    nd[1] = -1;
    for (int j = 0; j < e->n; j++)
      nd[2 + j] = (int16_t)j;
    return bb_engine_push_nodes(e, nd, 1);
  }
  uint8_t nd[1 + BB_MAX_QUEENS];
  memset(nd, 0, sizeof(nd));
  for (int j = 0; j < e->N; j++)
    nd[1 + j] = (uint8_t)j;
  return bb_engine_push_nodes(e, nd, 1);
}

extern "C" long long bb_engine_export_nodes(bb_engine *e, void *host_nodes, long long max_count)
{
  if (!e || !host_nodes)
    return bb_fail(BB_ERR_ARG, "bb_engine_export_nodes: bad arguments");
  if (cudaSetDevice(e->device) != cudaSuccess)
    return bb_fail(BB_ERR_CUDA, "cudaSetDevice failed");
  int rc = fetch_ctrl(e);
  if (rc)
    return rc;
  const long long total = (long long)(e->ctrl_h->size[0] - e->ctrl_h->bot[0]) + (e->ctrl_h->size[1] - e->ctrl_h->bot[1]);
  if (total > max_count)
    return bb_fail(BB_ERR_ARG, "export buffer holds %lld nodes, pool has %lld", max_count, total);
  if (total == 0)
    return 0;
  const size_t rec = e->kind == KIND_PFSP ? sizeof(int16_t) * (2 + (e->wide ? e->wide : BB_MAX_JOBS)) : (1 + BB_MAX_QUEENS);
  void *stage = nullptr;
  rc = pool_alloc(e, &stage, rec * total);
  if (rc)
    return rc;
  int at = 0;
  for (int s = 0; s < 2; s++)
  {
    const int from = e->ctrl_h->bot[s], cnt = e->ctrl_h->size[s] - from;
    if (!cnt)
      continue;
    const int blocks = (cnt + 127) / 128;
    if (e->kind == KIND_PFSP && e->wide == 50)
      k_export_pfsp_wide<50><<<blocks, 128, 0, e->stream>>>((int16_t *)stage, e->pool, s, from, cnt, at);
    else if (e->kind == KIND_PFSP && e->wide == 100)
      k_export_pfsp_wide<100><<<blocks, 128, 0, e->stream>>>((int16_t *)stage, e->pool, s, from, cnt, at);
    else if (e->kind == KIND_PFSP && e->wide)
      k_export_pfsp_wide<200><<<blocks, 128, 0, e->stream>>>((int16_t *)stage, e->pool, s, from, cnt, at);
    else if (e->kind == KIND_PFSP)
      k_export_pfsp<<<blocks, 128, 0, e->stream>>>((int16_t *)stage, e->pool, s, from, cnt, at);
    else if (e->nq_gen == 3)
      k_export_nq3<<<blocks, 128, 0, e->stream>>>((uint8_t *)stage, e->pool, s, from, cnt, at, e->N);
    else
      k_export_nq<<<blocks, 128, 0, e->stream>>>((uint8_t *)stage, e->pool, s, from, cnt, at, e->N);
    at += cnt;
  }
  cudaError_t err = cudaGetLastError();
  if (err == cudaSuccess)
    err = cudaMemcpyAsync(host_nodes, stage, rec * total, cudaMemcpyDeviceToHost, e->stream);
  pool_free(e, stage);
  if (err == cudaSuccess)
    err = cudaStreamSynchronize(e->stream);
  if (err != cudaSuccess)
    return bb_fail_cuda(err, "export", __FILE__, __LINE__);
  return total;
}

static void fill_stats(bb_engine *e, bb_stats *out)
{
  const bb_ctrl *c = e->ctrl_h;
  out->tree = c->tree;
  out->sol = c->sol;
  out->parents = c->parents;
  out->evals = c->evals;
  out->lb2_evals = c->lb2_evals;
  out->lb2_pairs = c->lb2_pairs;
  out->iterations = c->iterations;
  out->launches = e->launches;
  out->pool_size = (long long)(c->size[0] - c->bot[0]) + (c->size[1] - c->bot[1]);
  out->max_pool = c->max_pool;
  out->best = c->best;
  out->kernel_ms = e->kernel_ms;
  out->span_ms = e->span_ms;
  out->h2d_bytes = e->h2d_bytes;
  out->d2h_bytes = e->d2h_bytes;
  out->table_bytes = e->table_bytes;
  out->grid = e->grid;
  out->smem_bytes = e->smem;
  out->record_bytes = e->pool.nplanes * 16 + e->pool.tail_words * 4;
  out->steals = out->steals_ok = out->donated = 0;
}

extern "C" int bb_engine_get_stats(bb_engine *e, bb_stats *out)
{
  if (!e || !out)
    return bb_fail(BB_ERR_ARG, "bb_engine_get_stats: bad arguments");
  ENGINE_SET_DEVICE(e);
  int rc = fetch_ctrl(e);
  if (rc)
    return rc;
  fill_stats(e, out);
  return BB_OK;
}

/* ------------------------------------------------------------------ sharing work between GPUs (bb_multi.cu)
 * Replaces the host-memory steal of the reference (popBackBulkFree(victim, m, 5M, stolen, 2) + pushBackBulk,
 * pfsp_multigpu_cuda.c:355-404): the victim only RESERVES the oldest nodes of one of its stacks (they sit at the
 * bottom: shallow nodes, large sub-trees); the thief copies the planes straight out of the victim's HBM over
 * NVLink (peer access in one process, CUDA IPC between processes) and appends them to its own stack. */

/* Donation, device side. The victim's host thread never waits for its GPU: it ENQUEUES two kernels between two
 * batches of iterations and goes on launching; the result arrives in a pinned mailbox.
 *   k_donate_plan  decides on the state the pool really has at that point of the stream (not on the host's view,
 *                  which is one batch old): the oldest half of the fuller stack, capped by the transfer-buffer slot;
 *                  moves that stack's bottom up, re-plans the next iteration, publishes (stack, count);
 *   k_donate_copy  copies the reserved nodes' planes into the slot (compact, plane-major). */
__global__ void k_donate_plan(bb_ctrl *c, long long min_keep, long long max_count, int total_warps, volatile int *mailbox)
{
  if (threadIdx.x != 0 || blockIdx.x != 0)
    return;
  const long long a0 = c->size[0] - c->bot[0], a1 = c->size[1] - c->bot[1];
  long long give = 0;
  const int s = a0 >= a1 ? 0 : 1;
  if (!c->error && a0 + a1 >= 2 * min_keep && a0 + a1 >= 2)
  {
    give = (a0 + a1) / 2;
    const long long as = s == 0 ? a0 : a1;
    if (give > as)
      give = as;
    if (give > max_count)
      give = max_count;
  }
  c->don_stack = s;
  c->don_from = c->bot[s];
  c->don_count = (int)give;
  if (give > 0)
  {
    c->bot[s] += (int)give;
    bbk::plan_next(c, c->src, total_warps); /* the planned batch may have reached into the donated range */
  }
  mailbox[0] = s;
  mailbox[1] = (int)give;
}

__global__ void __launch_bounds__(256) k_donate_copy(const bb_ctrl *__restrict__ c, const bb_pool pool,
                                                    uint4 *__restrict__ slot)
{
  const int count = c->don_count, from = c->don_from, s = c->don_stack;
  if (count <= 0)
    return;
  const uint4 *__restrict__ src = pool.planes[s];
  const long long total = (long long)count * pool.nplanes;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x)
  {
    const int q = (int)(i / count), k = (int)(i - (long long)q * count);
    slot[i] = src[(size_t)q * pool.capacity + from + k];
  }
  if (pool.tail_words)
  {
    uint32_t *__restrict__ tw = reinterpret_cast<uint32_t *>(slot + total);
    const uint32_t *__restrict__ st = pool.tailw[s] + (size_t)from * pool.tail_words;
    const long long words = (long long)count * pool.tail_words;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (long long)gridDim.x * blockDim.x)
      tw[i] = st[i];
  }
}

/* thief side: append `count` records of a peer's transfer-buffer slot (peer or IPC-mapped pointer: plain loads over
 * NVLink) to the top of our stack 0, wherever that is on the device at this point of the stream */
__global__ void __launch_bounds__(256) k_pull_copy(bb_ctrl *__restrict__ c, const bb_pool pool,
                                                  const uint4 *__restrict__ slot, int count)
{
  const int at = c->size[0];
  if ((long long)at + count > pool.capacity)
  {
    if (threadIdx.x == 0 && blockIdx.x == 0)
      c->error = 1;
    return;
  }
  uint4 *__restrict__ dst = pool.planes[0];
  const long long total = (long long)count * pool.nplanes;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x)
  {
    const int q = (int)(i / count), k = (int)(i - (long long)q * count);
    dst[(size_t)q * pool.capacity + at + k] = slot[i];
  }
  if (pool.tail_words)
  {
    const uint32_t *__restrict__ tw = reinterpret_cast<const uint32_t *>(slot + total);
    uint32_t *__restrict__ dt = pool.tailw[0] + (size_t)at * pool.tail_words;
    const long long words = (long long)count * pool.tail_words;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (long long)gridDim.x * blockDim.x)
      dt[i] = tw[i];
  }
}

__global__ void k_pull_commit(bb_ctrl *c, int count, int total_warps)
{
  if (threadIdx.x == 0 && blockIdx.x == 0 && !c->error)
  {
    c->size[0] += count;
    bbk::plan_next(c, 0, total_warps);
  }
}

/* Transfer buffer of an engine that takes part in a multi-GPU search: BB_STAGE_SLOTS slots, so that two donations
 * can be outstanding. A donation is copied out of the pool into a slot (plane-major, compact) and the thief reads
 * the SLOT, not the pool. Only this allocation has to be visible to the peers: mapping a whole multi-GB pool with
 * cudaIpcOpenMemHandle costs ~50 ms per peer (measured), mapping the buffer a few. */
int bb_engine_stage_alloc(bb_engine *e, size_t bytes)
{
  ENGINE_SET_DEVICE(e);
  if (e->stage_mem)
    return BB_OK;
  const long long rec = (long long)e->pool.nplanes * 16 + (long long)e->pool.tail_words * 4;
  long long nodes = (long long)(bytes / BB_STAGE_SLOTS / (size_t)rec);
  if (nodes > e->pool.capacity)
    nodes = e->pool.capacity;
  nodes &= ~15ll; /* slots stay 256-byte aligned */
  if (nodes < 1024)
    return bb_fail(BB_ERR_ARG, "transfer buffer of %zu bytes is too small", bytes);
  BB_CUDA_TRY(cudaMalloc(&e->stage_mem, (size_t)nodes * rec * BB_STAGE_SLOTS));
  BB_CUDA_TRY(cudaHostAlloc((void **)&e->mailbox_h, 64, cudaHostAllocMapped));
  BB_CUDA_TRY(cudaHostGetDevicePointer((void **)&e->mailbox_d, (void *)e->mailbox_h, 0));
  BB_CUDA_TRY(cudaEventCreateWithFlags(&e->don_ev, cudaEventDisableTiming));
  e->stage_nodes = nodes;
  e->don_slot = -1;
  for (int s = 0; s < BB_STAGE_SLOTS; s++)
    e->slot_busy[s] = 0;
  return BB_OK;
}

void *bb_engine_stage_ptr(bb_engine *e) { return e->stage_mem; }
size_t bb_engine_stage_slot_bytes(bb_engine *e)
{
  return (size_t)e->stage_nodes * ((size_t)e->pool.nplanes * 16 + (size_t)e->pool.tail_words * 4);
}

/* victim side. First call: enqueue the donation behind the batch in flight and return *count = -1 ("pending; ask
 * again"). Later calls: -1 until the GPU has got there, then the slot (in *stack) and the number of nodes it holds
 * (0 = too little to share, or no free slot). The engine's stream is never synchronised. */
int bb_engine_donate(bb_engine *e, long long min_keep, long long max_count, int *stack, long long *start,
                     long long *count)
{
  *count = 0;
  *start = 0;
  if (!e->stage_mem)
    return bb_fail(BB_ERR_STATE, "bb_engine_donate: no transfer buffer");
  ENGINE_SET_DEVICE(e);
  if (e->don_slot < 0)
  {
    int slot = -1;
    for (int s = 0; s < BB_STAGE_SLOTS; s++)
      if (!e->slot_busy[s])
        slot = s;
    if (slot < 0)
      return BB_OK; /* every slot still waits for its thief's acknowledgement */
    if (max_count > e->stage_nodes)
      max_count = e->stage_nodes;
    uint4 *dst = (uint4 *)((char *)e->stage_mem + (size_t)slot * bb_engine_stage_slot_bytes(e));
    k_donate_plan<<<1, 1, 0, e->stream>>>(e->ctrl, min_keep, max_count, e->grid * (e->threads / 32), e->mailbox_d);
    k_donate_copy<<<128, 256, 0, e->stream>>>(e->ctrl, e->pool, dst);
    BB_CUDA_TRY(cudaGetLastError());
    BB_CUDA_TRY(cudaEventRecord(e->don_ev, e->stream));
    e->don_slot = slot;
    *count = -1;
    return BB_OK;
  }
  const cudaError_t q = cudaEventQuery(e->don_ev);
  if (q == cudaErrorNotReady)
  {
    *count = -1;
    return BB_OK;
  }
  if (q != cudaSuccess)
    return bb_fail_cuda(q, "donation", __FILE__, __LINE__);
  const int give = e->mailbox_h[1];
  *stack = e->don_slot;
  *count = give;
  if (give > 0)
    e->slot_busy[e->don_slot] = 1;
  e->don_slot = -1;
  return BB_OK;
}

int bb_engine_donate_done(bb_engine *e, int slot)
{
  if (slot >= 0 && slot < BB_STAGE_SLOTS)
    e->slot_busy[slot] = 0; /* the thief has copied the slot out */
  return BB_OK;
}

/* thief side: append `count` records from a peer's transfer-buffer slot (device pointer valid on this device) to
 * the top of our stack 0; returns when the copy is complete, so that the caller may acknowledge */
int bb_engine_pull(bb_engine *e, const void *peer_slot, long long count)
{
  ENGINE_SET_DEVICE(e);
  if (count <= 0)
    return BB_OK;
  k_pull_copy<<<128, 256, 0, e->stream>>>(e->ctrl, e->pool, (const uint4 *)peer_slot, (int)count);
  k_pull_commit<<<1, 1, 0, e->stream>>>(e->ctrl, (int)count, e->grid * (e->threads / 32));
  BB_CUDA_TRY(cudaGetLastError());
  int rc = fetch_ctrl(e);
  if (rc)
    return rc;
  if (e->ctrl_h->error)
    return bb_fail(BB_ERR_OVERFLOW, "pool capacity %d exceeded while taking over %lld nodes", e->pool.capacity, count);
  return BB_OK;
}

/* zero the search counters but keep the pool (workers other than 0 after the common warm-up) */
__global__ void k_zero_counters(bb_ctrl *c)
{
  c->tree = c->sol = c->parents = c->evals = c->lb2_evals = c->lb2_pairs = c->iterations = 0ull;
}

int bb_engine_zero_counters(bb_engine *e)
{
  ENGINE_SET_DEVICE(e);
  k_zero_counters<<<1, 1, 0, e->stream>>>(e->ctrl);
  BB_CUDA_TRY(cudaGetLastError());
  return fetch_ctrl(e);
}

/* Warm-up partition on the device (replaces roundRobin_distribution, Pool_atom.c:14-36): every worker has expanded
 * the root to the same SET of nodes; it keeps the nodes whose placed prefix hashes to its id and drops the others, in
 * place, stack by stack (one CTA: the warm-up pool is a few thousand nodes). A content hash needs no communication
 * and does not depend on the order in which the device appended the nodes.
 * mode 0: PFSP record, 1: N-Queens record with masks (nq_expand3/4/5), 2: N-Queens plane + tail words, 3 / 4 / 5: PFSP
 * record of the 50- / 100- / 200-job classes. */
__global__ void __launch_bounds__(1024) k_partition(bb_ctrl *c, bb_pool pool, int mode, int nq_dbyte, int D, int me,
                                                   int total_warps)
{
  __shared__ int warp_tot[32];
  __shared__ int chunk_base;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int s = 0; s < 2; s++)
  {
    const int from = c->bot[s], to = c->size[s];
    int out = from;
    for (int c0 = from; c0 < to; c0 += blockDim.x)
    {
      const int i = c0 + threadIdx.x;
      uint4 pl[wide_geom<200>::NPL]; /* the widest record */
      uint32_t tw[2] = {0u, 0u};
      bool keep = false;
      if (i < to)
      {
        for (int q = 0; q < pool.nplanes; q++)
          pl[q] = pool.planes[s][(size_t)q * pool.capacity + i];
        for (int t = 0; t < pool.tail_words; t++)
          tw[t] = pool.tailw[s][(size_t)i * pool.tail_words + t];
        /* the permutation starts at byte 0 of the record in every layout; only mode 2 continues in the tail words */
        const uint32_t *pw = reinterpret_cast<const uint32_t *>(pl);
        auto byte_at = [&](int b) -> unsigned {
          const uint32_t word = (mode == 2 && b >= 16) ? tw[(b - 16) >> 2] : pw[b >> 2];
          return (word >> (8 * (b & 3))) & 0xffu;
        };
        int depth;
        if (mode == 3)
          depth = (int)(pw[wide_geom<50>::W_DEPTH] & 0xffu);
        else if (mode == 4)
          depth = (int)(pw[wide_geom<100>::W_DEPTH] & 0xffu);
        else if (mode == 5)
          depth = (int)(pw[wide_geom<200>::W_DEPTH] & 0xffu);
        else if (mode == 0)
          depth = (int)(pl[1].y & 0xffu);
        else if (mode == 1)
          depth = (int)(pl[1].y >> 24);
        else
          depth = (int)byte_at(nq_dbyte);
        unsigned h = 2166136261u; /* FNV-1a over the placed prefix */
        for (int b = 0; b < depth; b++)
        {
          h ^= byte_at(b);
          h *= 16777619u;
        }
        h ^= h >> 15;
        keep = (int)(h % (unsigned)D) == me;
      }
      /* block-wide exclusive scan of the keep flags */
      const unsigned bal = __ballot_sync(0xffffffffu, keep);
      if (lane == 0)
        warp_tot[warp] = __popc(bal);
      __syncthreads();
      if (warp == 0)
      {
        const int v = lane < (int)(blockDim.x >> 5) ? warp_tot[lane] : 0;
        const int incl = bbk::warp_incl_scan(v, lane);
        warp_tot[lane] = incl - v;
        if (lane == 31)
          chunk_base = incl;
      }
      __syncthreads();
      /* every record of this chunk is in registers: writing at out + rank (<= i) cannot clobber unread nodes */
      if (keep)
      {
        const int at = out + warp_tot[warp] + __popc(bal & ((1u << lane) - 1u));
        for (int q = 0; q < pool.nplanes; q++)
          pool.planes[s][(size_t)q * pool.capacity + at] = pl[q];
        for (int t = 0; t < pool.tail_words; t++)
          pool.tailw[s][(size_t)at * pool.tail_words + t] = tw[t];
      }
      out += chunk_base;
      __syncthreads();
    }
    if (threadIdx.x == 0)
      c->size[s] = out;
    __syncthreads();
  }
  if (threadIdx.x == 0)
    bbk::plan_next(c, c->src, total_warps);
}

/* keep this worker's share of the (common) warm-up pool; zero_counters: the warm-up counts are kept by the caller */
int bb_engine_partition(bb_engine *e, int D, int me, int zero_counters)
{
  ENGINE_SET_DEVICE(e);
  int mode = e->wide == 50 ? 3 : e->wide == 100 ? 4 : e->wide ? 5 : 0, dbyte = 0;
  if (e->kind == KIND_NQ)
  {
    mode = e->nq_gen == 3 ? 1 : 2;
    dbyte = e->pool.tail_words ? 16 + (e->N - 16) : 15;
  }
  k_partition<<<1, 1024, 0, e->stream>>>(e->ctrl, e->pool, mode, dbyte, D, me, e->grid * (e->threads / 32));
  if (zero_counters)
    k_zero_counters<<<1, 1, 0, e->stream>>>(e->ctrl);
  BB_CUDA_TRY(cudaGetLastError());
  return BB_OK;
}

extern "C" int bb_engine_keep_share(bb_engine *e, int parts, int me)
{
  if (!e || parts < 1 || me < 0 || me >= parts)
    return bb_fail(BB_ERR_ARG, "bb_engine_keep_share: bad engine or share %d of %d", me, parts);
  int rc = bb_engine_partition(e, parts, me, 0);
  if (rc)
    return rc;
  return fetch_ctrl(e);
}

int bb_engine_kind_is_pfsp(const bb_engine *e) { return e->kind == KIND_PFSP; }

/* ------------------------------------------------------------------ the device loop */

extern "C" int bb_engine_run(bb_engine *e, long long max_iters, bb_stats *out)
{
  if (!e)
    return bb_fail(BB_ERR_ARG, "null engine");
  ENGINE_SET_DEVICE(e);
  int rc = fetch_ctrl(e);
  if (rc)
    return rc;
  const unsigned long long it0 = e->ctrl_h->iterations;
  BB_CUDA_TRY(cudaEventRecord(e->ev_begin, e->stream));
  while (e->ctrl_h->n > 0 && !e->ctrl_h->error)
  {
    long long left = max_iters > 0 ? max_iters - (long long)(e->ctrl_h->iterations - it0) : LLONG_MAX;
    if (left <= 0)
      break;
    BB_CUDA_TRY(cudaEventRecord(e->ev0, e->stream));
    rc = enqueue_iterations(e, left >= e->iters_per_launch ? e->iters_per_launch : left);
    if (rc)
      return rc;
    BB_CUDA_TRY(cudaEventRecord(e->ev1, e->stream));
    rc = fetch_ctrl(e);
    if (rc)
      return rc;
    float ms = 0;
    BB_CUDA_TRY(cudaEventElapsedTime(&ms, e->ev0, e->ev1));
    e->kernel_ms += ms;
  }
  BB_CUDA_TRY(cudaEventRecord(e->ev_end, e->stream));
  BB_CUDA_TRY(cudaEventSynchronize(e->ev_end));
  {
    float ms = 0;
    BB_CUDA_TRY(cudaEventElapsedTime(&ms, e->ev_begin, e->ev_end));
    e->span_ms += ms;
  }
  if (out)
    fill_stats(e, out);
  if (e->ctrl_h->error)
    return bb_fail(BB_ERR_OVERFLOW, "device pool overflow: capacity %d nodes per stack is too small (raise "
                                    "bb_config.capacity or lower M)", e->pool.capacity);
  return BB_OK;
}

/* Polling variant for the multi-GPU workers: every call enqueues one more batch of `iters` iterations and returns
 * the state after the PREVIOUS batch, so the GPU never waits for the host between batches (the device plans its
 * own iterations; an enqueued batch on an empty pool is a chain of kernels that return at once). The reported pool
 * size is one batch old; a size of 0 is final. Any other engine call synchronises the stream first. */
int bb_engine_run_pipelined(bb_engine *e, long long iters, bb_stats *out)
{
  ENGINE_SET_DEVICE(e);
  if (!e->snap_h)
  {
    BB_CUDA_TRY(cudaMallocHost(&e->snap_h, 2 * sizeof(bb_ctrl)));
    for (int s = 0; s < 2; s++)
    {
      BB_CUDA_TRY(cudaEventCreate(&e->pipe_k0[s]));
      BB_CUDA_TRY(cudaEventCreate(&e->pipe_k1[s]));
      BB_CUDA_TRY(cudaEventCreateWithFlags(&e->pipe_done[s], cudaEventDisableTiming));
    }
  }
  auto enqueue = [&](int slot) -> int {
    BB_CUDA_TRY(cudaEventRecord(e->pipe_k0[slot], e->stream));
    const int rc = enqueue_iterations(e, iters);
    if (rc)
      return rc;
    BB_CUDA_TRY(cudaEventRecord(e->pipe_k1[slot], e->stream));
    BB_CUDA_TRY(cudaMemcpyAsync(&e->snap_h[slot], e->ctrl, sizeof(bb_ctrl), cudaMemcpyDeviceToHost, e->stream));
    BB_CUDA_TRY(cudaEventRecord(e->pipe_done[slot], e->stream));
    e->d2h_bytes += sizeof(bb_ctrl);
    return BB_OK;
  };
  int rc;
  if (!e->pipe_inflight)
  {
    if ((rc = enqueue(e->pipe_slot)))
      return rc;
    e->pipe_inflight = 1;
  }
  const int cur = e->pipe_slot;
  if ((rc = enqueue(cur ^ 1)))
    return rc;
  BB_CUDA_TRY(cudaEventSynchronize(e->pipe_done[cur]));
  float ms = 0;
  BB_CUDA_TRY(cudaEventElapsedTime(&ms, e->pipe_k0[cur], e->pipe_k1[cur]));
  e->kernel_ms += ms;
  *e->ctrl_h = e->snap_h[cur];
  e->pipe_slot = cur ^ 1;
  if (e->ctrl_h->n == 0 || e->ctrl_h->error)
  { /* the pool was already empty (or failed) after that batch: the batch in flight does nothing */
    BB_CUDA_TRY(cudaEventSynchronize(e->pipe_done[cur ^ 1]));
    BB_CUDA_TRY(cudaEventElapsedTime(&ms, e->pipe_k0[cur ^ 1], e->pipe_k1[cur ^ 1]));
    e->kernel_ms += ms;
    *e->ctrl_h = e->snap_h[cur ^ 1];
    e->pipe_inflight = 0;
  }
  if (out)
    fill_stats(e, out);
  if (e->ctrl_h->error)
    return bb_fail(BB_ERR_OVERFLOW, "device pool overflow: capacity %d nodes per stack is too small (raise "
                                    "bb_config.capacity or lower M)", e->pool.capacity);
  return BB_OK;
}

/* ------------------------------------------------------------------ whole searches (single GPU here; D > 1 in bb_multi.cu) */

int bb_multi_search(int kind, const bb_pfsp_instance *ins, int lb, int best0, int N, int G, int D, int ws,
                    const bb_config *cfg, bb_search_result *res, bb_stats *per_gpu);

extern "C" int bb_pfsp_search(const bb_pfsp_instance *ins, int lb, int best0, int D, int ws, const bb_config *cfg,
                              bb_search_result *res, bb_stats *per_gpu)
{
  if (!ins || !res || D < 1)
    return bb_fail(BB_ERR_ARG, "bb_pfsp_search: bad arguments");
  return bb_multi_search(KIND_PFSP, ins, lb, best0, 0, 0, D, ws, cfg, res, per_gpu);
}

extern "C" int bb_nq_search(int N, int G, int D, int ws, const bb_config *cfg, bb_search_result *res,
                            bb_stats *per_gpu)
{
  if (!res || D < 1)
    return bb_fail(BB_ERR_ARG, "bb_nq_search: bad arguments");
  return bb_multi_search(KIND_NQ, nullptr, 0, INT_MAX, N, G, D, ws, cfg, res, per_gpu);
}
