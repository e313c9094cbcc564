/*
 * bb_static.cu - bb_pfsp_expand_static(): one bound -> prune -> branch step for callers that may not allocate.
 *
 * Who needs this. The engine entry points (bb_*_engine_create, bb_engine_push_nodes, bb_engine_export_nodes) call
 * cudaMalloc/cudaFree for the pool, the control block and staging buffers. Some users of this library cannot
 * afford that inside their loop:
 *   - a worker thread of a multi-threaded driver in the style of pfsp_multigpu_cuda.c, where several host threads
 *     share one GPU: cudaMalloc and cudaFree synchronise the whole device, so an allocation made by one thread
 *     stalls the kernels and copies of every other thread on that GPU;
 *   - a host that has already handed all free device memory to its own pool (a solver embedding this library
 *     next to other GPU consumers), where a late cudaMalloc simply fails;
 *   - a latency-critical section that must not take the driver's allocation lock.
 * For them this file offers the same expand step as the engine (the kernel is bbk::pfsp_expand, unchanged) with
 * every byte of working memory taken from ONE file-scope __device__ array that is part of the library image:
 * the control block (counters, append cursor), the repacked instance tables and the two structure-of-arrays
 * stacks the kernel works on. The caller passes only its inputs (parent nodes) and outputs (child nodes, result).
 *
 * Limits and rules (also in include/bb_gpu.h):
 *   - at most BB_STATIC_MAX_PARENTS parents per call, lb 0 (LB1_d) or 1 (LB1), <= 20 jobs x 20 machines;
 *   - the array (about 4.6 MiB) exists once per GPU that loads the library and lives until the process exits;
 *   - all calls on a GPU share it and it keeps the previous call's contents: the parts that must start from zero
 *     are cleared at the start of every call, and two calls whose GPU work overlaps (different streams, or
 *     different host threads without ordering) race on it - serialise them;
 *   - the FIRST call on each device resolves the array's device address (cudaGetSymbolAddress) and queries the
 *     launch geometry; make that first call outside any stream capture. Later calls only launch kernels.
 * This entry point has no counterpart in the reference.
 */
#include "bb_internal.cuh"
#include "bb_pfsp_kernels.cuh"

#include <cstring>
#include <mutex>

static constexpr int SP = BB_STATIC_MAX_PARENTS; /* parents per call */
static constexpr int SC = SP * BB_MAX_JOBS;      /* most children a call can produce */
static constexpr int SCAP = SP + SC;             /* plane stride shared by both stacks */
static constexpr int NPL_MAX = 7;                /* planes of the largest record (m = 20) */
static constexpr size_t OFF_CTRL = 0;            /* bb_ctrl, 256 B */
static constexpr size_t OFF_TABLES = 256;        /* p8[400] tails[20] heads[20], padded to 1 KiB */
static constexpr size_t OFF_POOL = 1280;         /* [NPL_MAX][SCAP] uint4 */
static constexpr size_t SCRATCH_BYTES = OFF_POOL + (size_t)NPL_MAX * SCAP * sizeof(uint4);

/* The only working memory of this entry point (file scope, one copy per GPU that loads the library). In HOST code
 * this name is just a handle for the runtime - it is not a device address - so the host side below never passes
 * it to a kernel; it passes the pointer obtained from cudaGetSymbolAddress. */
static __device__ __align__(256) unsigned char g_static_scratch[SCRATCH_BYTES];

namespace
{

struct static_tables
{
  uint8_t p8[BB_MAX_MACHINES * BB_MAX_JOBS];
  int tails[BB_MAX_MACHINES];
  int heads[BB_MAX_MACHINES];
};

/* clears what must start at zero (the whole control block: counters, cursor, error flag), installs the tables and
 * plans the single iteration over `count` parents */
__global__ void k_static_begin(unsigned char *scratch, const static_tables tb, int best, int count, int maxkids,
                               int total_warps)
{
  uint8_t *p8 = scratch + OFF_TABLES;
  int *tails = reinterpret_cast<int *>(scratch + OFF_TABLES + sizeof(tb.p8));
  int *heads = tails + BB_MAX_MACHINES;
  for (int i = threadIdx.x; i < (int)sizeof(tb.p8); i += blockDim.x)
    p8[i] = tb.p8[i];
  for (int i = threadIdx.x; i < BB_MAX_MACHINES; i += blockDim.x)
  {
    tails[i] = tb.tails[i];
    heads[i] = tb.heads[i];
  }
  if (threadIdx.x == 0)
  {
    bb_ctrl *c = reinterpret_cast<bb_ctrl *>(scratch + OFF_CTRL);
    int *raw = reinterpret_cast<int *>(c);
    for (unsigned i = 0; i < sizeof(bb_ctrl) / sizeof(int); i++)
      raw[i] = 0;
    c->best = best;
    c->M = SP;
    c->capacity = SCAP;
    c->maxkids = maxkids;
    c->size[0] = count;
    bbk::plan_next(c, 0, total_warps);
  }
}

/* caller's parents (reference Node layout, device memory) -> records of stack 0, deriving front/remain/mask */
__global__ void k_static_seed(unsigned char *scratch, const Node *__restrict__ parents, int count,
                              uint4 *__restrict__ planes0, int n, int m, int mp, int nplanes,
                              const uint8_t *__restrict__ p8)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count)
    return;
  const int16_t *nd = reinterpret_cast<const int16_t *>(parents + i);
  const int depth = nd[0];
  /* the parents are in device memory, so they can only be checked here: a depth outside 0..n-1 or a job id outside
   * 0..n-1 would make the expand kernel index its tables out of range. Such a batch is refused as a whole: the
   * control block's error flag turns the expand launch into a no-op and the result words report it. */
  bool bad = depth < 0 || depth >= n;
  for (int j = 0; j < n; j++)
    bad = bad || nd[2 + j] < 0 || nd[2 + j] >= n;
  if (bad)
  {
    bb_ctrl *c = reinterpret_cast<bb_ctrl *>(scratch + OFF_CTRL);
    c->n = 0; /* pfsp_expand returns at once when the planned batch is empty */
    atomicExch(&c->error, 1);
    return;
  }
  uint32_t w[NPL_MAX * 4];
  for (int q = 0; q < NPL_MAX * 4; q++)
    w[q] = 0;
  uint32_t mask = 0;
  int front[BB_MAX_MACHINES], remain[BB_MAX_MACHINES];
  for (int k = 0; k < m; k++)
    front[k] = remain[k] = 0;
  for (int j = 0; j < n; j++)
  {
    const int job = nd[2 + j] & 0xff;
    w[j >> 2] |= (uint32_t)job << (8 * (j & 3));
    if (j < depth)
    {
      mask |= 1u << job;
      front[0] += p8[job];
      for (int k = 1; k < m; k++)
        front[k] = max(front[k - 1], front[k]) + p8[k * BB_MAX_JOBS + job];
    }
    else
      for (int k = 0; k < m; k++)
        remain[k] += p8[k * BB_MAX_JOBS + job];
  }
  w[5] = (uint32_t)depth | (mask << 8);
  uint16_t *h = reinterpret_cast<uint16_t *>(w);
  for (int k = 0; k < m; k++)
  {
    h[12 + k] = (uint16_t)front[k];
    h[12 + mp + k] = (uint16_t)remain[k];
  }
  for (int q = 0; q < nplanes; q++)
    planes0[(size_t)q * SCAP + i] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
}

/* records of stack 1 -> caller's children (reference Node layout) and the four result words */
__global__ void k_static_export(const unsigned char *__restrict__ scratch, const uint4 *__restrict__ planes1,
                                Node *__restrict__ children, int max_children, unsigned long long *__restrict__ result)
{
  const bb_ctrl *c = reinterpret_cast<const bb_ctrl *>(scratch + OFF_CTRL);
  const int produced = c->error ? 0 : c->size[1];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0)
  {
    /* a refused batch (malformed parent, see k_static_seed) reports no children and all-ones in the counters */
    result[0] = (unsigned long long)produced;
    result[1] = c->error ? ~0ull : c->tree;
    result[2] = c->error ? ~0ull : c->sol;
    result[3] = (unsigned long long)(unsigned int)c->best;
  }
  if (i >= produced || i >= max_children)
    return;
  const uint4 q0 = planes1[i];
  const uint4 q1 = planes1[(size_t)SCAP + i];
  const uint32_t w[6] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y};
  int16_t *nd = reinterpret_cast<int16_t *>(children + i);
  const int depth = w[5] & 0xff;
  nd[0] = (int16_t)depth;
  nd[1] = (int16_t)(depth - 1);
  for (int j = 0; j < BB_MAX_JOBS; j++)
    nd[2 + j] = (int16_t)((w[j >> 2] >> (8 * (j & 3))) & 0xff);
}

struct static_dev
{
  unsigned char *scratch; /* device address of g_static_scratch on this device */
  int sms;
  int per_sm[2][4]; /* resident CTAs per SM for (lb, machine class) */
  bool ready;
};

constexpr int MAX_DEV = 64;
static_dev g_dev[MAX_DEV];
std::mutex g_dev_mutex;

template <int LB>
const void *static_kernel(int mm)
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

int mm_class(int mm) { return mm == 5 ? 0 : mm == 10 ? 1 : mm == 20 ? 2 : 3; }

/* once per device, with that device current: resolve the array's device address and the launch geometry */
int static_setup(int dev, static_dev **out)
{
  if (dev < 0 || dev >= MAX_DEV)
    return bb_fail(BB_ERR_ARG, "bb_pfsp_expand_static: device ordinal %d not supported (max %d)", dev, MAX_DEV - 1);
  std::lock_guard<std::mutex> lock(g_dev_mutex);
  static_dev *d = &g_dev[dev];
  if (!d->ready)
  {
    void *ptr = nullptr;
    BB_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_static_scratch));
    BB_CUDA_TRY(cudaDeviceGetAttribute(&d->sms, cudaDevAttrMultiProcessorCount, dev));
    const int classes[4] = {5, 10, 20, 0};
    for (int lb = 0; lb < 2; lb++)
      for (int c = 0; c < 4; c++)
      {
        const void *k = lb == 0 ? static_kernel<0>(classes[c]) : static_kernel<1>(classes[c]);
        const int m = classes[c] ? classes[c] : BB_MAX_MACHINES;
        BB_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d->per_sm[lb][c], k, BB_CTA_THREADS,
                                                                  bbk::pfsp_layout(m, 0, lb).total));
        if (d->per_sm[lb][c] < 1)
          return bb_fail(BB_ERR_CUDA, "bb_pfsp_expand_static: expand kernel does not fit on an SM");
      }
    d->scratch = static_cast<unsigned char *>(ptr);
    d->ready = true;
  }
  *out = d;
  return BB_OK;
}

} // namespace

extern "C" int bb_pfsp_expand_static(const bb_pfsp_instance *ins, int lb, int best, const Node *parents_d, int count,
                                     Node *children_d, int max_children, unsigned long long *result_d, void *stream)
{
  if (!ins || !result_d || count < 0 || max_children < 0 || (count > 0 && !parents_d) ||
      (max_children > 0 && !children_d))
    return bb_fail(BB_ERR_ARG, "bb_pfsp_expand_static: bad arguments");
  if (lb != 0 && lb != 1)
    return bb_fail(BB_ERR_ARG, "bb_pfsp_expand_static: lower bound %d not supported here (0 = LB1_d, 1 = LB1)", lb);
  if (count > SP)
    return bb_fail(BB_ERR_ARG, "bb_pfsp_expand_static: %d parents exceed the static scratch limit of %d per call",
                   count, SP);
  const int n = ins->nb_jobs, m = ins->nb_machines;
  if (n < 1 || n > BB_MAX_JOBS || m < 1 || m > BB_MAX_MACHINES)
    return bb_fail(BB_ERR_ARG, "bb_pfsp_expand_static: %d jobs x %d machines exceeds %d x %d", n, m, BB_MAX_JOBS,
                   BB_MAX_MACHINES);
  int dev = 0;
  BB_CUDA_TRY(cudaGetDevice(&dev));
  static_dev *d = nullptr;
  int rc = static_setup(dev, &d); /* first call on this device only: must not happen inside a stream capture */
  if (rc)
    return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);

  static_tables tb;
  memset(&tb, 0, sizeof(tb));
  for (int k = 0; k < m; k++)
  {
    for (int j = 0; j < n; j++)
    {
      const int v = ins->p_times[k * n + j];
      if (v < 0 || v > 255)
        return bb_fail(BB_ERR_ARG, "processing time %d out of the supported 0..255 range", v);
      tb.p8[k * BB_MAX_JOBS + j] = (uint8_t)v;
    }
    tb.tails[k] = ins->min_tails[k];
    tb.heads[k] = ins->min_heads[k];
  }

  const int mm = bbk::pfsp_mm(m), mp = bbk::pfsp_mp(m), nplanes = bbk::pfsp_planes(m);
  unsigned char *scratch = d->scratch;
  bb_ctrl *ctrl = reinterpret_cast<bb_ctrl *>(scratch + OFF_CTRL);
  uint4 *planes0 = reinterpret_cast<uint4 *>(scratch + OFF_POOL);
  bb_pool pool;
  pool.planes[0] = planes0;
  pool.planes[1] = planes0 + SP; /* same plane stride, shifted past the parents' columns */
  pool.tailw[0] = pool.tailw[1] = nullptr;
  pool.nplanes = nplanes;
  pool.tail_words = 0;
  pool.capacity = SCAP;
  bb_pfsp_dev T;
  memset(&T, 0, sizeof(T));
  T.n = n;
  T.m = m;
  T.mp = mp;
  T.P = 0;
  T.p8 = scratch + OFF_TABLES;
  T.tails = reinterpret_cast<const int *>(scratch + OFF_TABLES + sizeof(tb.p8));
  T.heads = T.tails + BB_MAX_MACHINES;
  T.best_known = ins->best_known;

  const int grid = d->per_sm[lb][mm_class(mm)] * d->sms;
  k_static_begin<<<1, 256, 0, s>>>(scratch, tb, best, count, n, grid * BB_WARPS_PER_CTA);
  if (count > 0)
  {
    k_static_seed<<<(count + 127) / 128, 128, 0, s>>>(scratch, parents_d, count, planes0, n, m, mp, nplanes, T.p8);
    const void *kernel = lb == 0 ? static_kernel<0>(mm) : static_kernel<1>(mm);
    int prefilter = 0;
    void *args[] = {&ctrl, &pool, &T, &prefilter};
    BB_CUDA_TRY(cudaLaunchKernel(kernel, dim3(grid), dim3(BB_CTA_THREADS), args, bbk::pfsp_layout(m, 0, lb).total, s));
  }
  k_static_export<<<(SC + 127) / 128, 128, 0, s>>>(scratch, pool.planes[1], children_d, max_children, result_d);
  BB_CUDA_TRY(cudaGetLastError());
  return BB_OK;
}
