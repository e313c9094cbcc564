/*
 * bb_multi.cu - whole searches on D GPUs of one node: one worker per GPU around the per-GPU engines
 * (include/bb_gpu.h: bb_pfsp_search / bb_nq_search with threads, bb_*_search_dist with one process per GPU).
 * Replaces pfsp_search() (pfsp/pfsp_multigpu_cuda.c:55-511) and nqueens_search()
 * (nqueens/nqueens_multigpu_cuda.cu:213-372).
 *
 * Phases (the reference's three steps, all on the GPU here):
 *   1. warm-up: EVERY worker expands the root on its own GPU until the pool holds >= D*m nodes (the pools are
 *      identical as sets; pfsp_multigpu_cuda.c:111-118 does this once on the CPU, pfsp_dist_multigpu_cuda.c:198-205
 *      redundantly on every rank like here) and keeps the nodes whose scheduled prefix hashes to its own id -
 *      a content-based stand-in for roundRobin_distribution (Pool_atom.c:14-36) that needs no communication and does
 *      not depend on the order in which the device appended the nodes. Worker 0 alone keeps the warm-up counters.
 *   2. parallel phase: bb_team_worker (bb_team.cpp): device iterations, polls, peer-memory work stealing, incumbent
 *      exchange, termination detection. A steal copies node planes from the victim's HBM straight into the thief's
 *      through a small per-GPU transfer buffer (cudaMemcpy2DAsync on peer-enabled / IPC-mapped pointers: NVLink, no host
 *      staging; only the 128 MB buffers are mapped between processes, not the multi-GB pools).
 *   3. there is no CPU drain step: every pool runs to empty on its GPU.
 */
#include "bb_internal.cuh"

#include <chrono>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

/* engine internals shared with bb_engine.cu (not part of the C ABI) */
int bb_engine_donate(bb_engine *e, long long min_keep, long long max_count, int *stack, long long *start,
                     long long *count);
int bb_engine_donate_done(bb_engine *e, int stack);
int bb_engine_pull(bb_engine *e, const void *peer_stage, long long count);
int bb_engine_stage_alloc(bb_engine *e, size_t bytes);
void *bb_engine_stage_ptr(bb_engine *e);
int bb_engine_run_pipelined(bb_engine *e, long long iters, bb_stats *out);

enum
{
  KIND_PFSP = 1,
  KIND_NQ = 2
};

namespace
{

double now_s()
{
  using namespace std::chrono;
  return duration<double>(steady_clock::now().time_since_epoch()).count();
}

/* span of the parallel phase on the device timeline: two events on the (idle) default stream of the current
 * device bracket bb_team_worker; the engine's own stream is non-blocking, so they do not serialise with it */
struct device_span
{
  cudaEvent_t a = nullptr, b = nullptr;
  void begin()
  {
    if (cudaEventCreate(&a) == cudaSuccess && cudaEventCreate(&b) == cudaSuccess)
      cudaEventRecord(a, 0);
  }
  double end_s(double fallback)
  {
    float ms = -1.f;
    if (a && b && cudaEventRecord(b, 0) == cudaSuccess && cudaEventSynchronize(b) == cudaSuccess)
      cudaEventElapsedTime(&ms, a, b);
    if (a)
      cudaEventDestroy(a);
    if (b)
      cudaEventDestroy(b);
    return ms >= 0.f ? ms * 1e-3 : fallback;
  }
};

#define BB_STAGE_BYTES ((size_t)128 << 20) /* transfer buffer per GPU: the most one steal can move */

struct ipc_slot
{
  cudaIpcMemHandle_t stage; /* the rank's transfer buffer */
  /* results */
  unsigned long long tree, sol, steals, steals_ok, donated, warm_tree, warm_sol;
  double kernel_ms, search_s, warm_s, setup_s;
  int best, rc;
};

struct gpu_worker
{
  bb_engine *e;
  int me;
  double kernel_ms_seen;
  void *peer_stage[64]; /* peers' transfer buffers as seen from this device */
  /* one process per GPU: a peer's buffer is mapped (cudaIpcOpenMemHandle) the first time we steal from it */
  const ipc_slot *ipc;
  bool opened[64];
};

int ensure_peer_mapped(gpu_worker *w, int j)
{
  if (!w->ipc || w->peer_stage[j])
    return BB_OK;
  const cudaError_t err = cudaIpcOpenMemHandle(&w->peer_stage[j], w->ipc[j].stage, cudaIpcMemLazyEnablePeerAccess);
  if (err != cudaSuccess)
    return bb_fail_cuda(err, "cudaIpcOpenMemHandle", __FILE__, __LINE__);
  w->opened[j] = true;
  return BB_OK;
}

int op_run(void *self, long long max_iters, long long *pool_size, int *best, double *busy_ms)
{
  gpu_worker *w = static_cast<gpu_worker *>(self);
  bb_stats st;
  static const bool no_pipeline = getenv("BB_NO_PIPELINE") != nullptr; /* debugging aid */
  const int rc = no_pipeline ? bb_engine_run(w->e, max_iters, &st) : bb_engine_run_pipelined(w->e, max_iters, &st);
  if (rc)
    return rc;
  *pool_size = st.pool_size;
  *best = st.best;
  *busy_ms = st.kernel_ms - w->kernel_ms_seen;
  w->kernel_ms_seen = st.kernel_ms;
  return BB_OK;
}

int op_donate(void *self, long long min_keep, long long max_count, int *stack, long long *start, long long *count)
{
  return bb_engine_donate(static_cast<gpu_worker *>(self)->e, min_keep, max_count, stack, start, count);
}

int op_donate_done(void *self, int stack) { return bb_engine_donate_done(static_cast<gpu_worker *>(self)->e, stack); }

int op_pull(void *self, int victim, int stack, long long start, long long count)
{
  gpu_worker *w = static_cast<gpu_worker *>(self);
  (void)stack;
  (void)start; /* the donation sits at the beginning of the victim's transfer buffer */
  const int rc = ensure_peer_mapped(w, victim);
  if (rc)
    return rc;
  return bb_engine_pull(w->e, w->peer_stage[victim], count);
}

int op_set_best(void *self, int best) { return bb_engine_set_best(static_cast<gpu_worker *>(self)->e, best); }

bb_team_ops gpu_ops(gpu_worker *w)
{
  bb_team_ops o;
  o.self = w;
  o.run = op_run;
  o.donate = op_donate;
  o.donate_done = op_donate_done;
  o.pull = op_pull;
  o.set_best = op_set_best;
  return o;
}

/* worker of a node = FNV-1a of its placed prefix, mod D */
inline unsigned prefix_hash(const unsigned char *bytes, int len)
{
  unsigned h = 2166136261u;
  for (int i = 0; i < len; i++)
  {
    h ^= bytes[i];
    h *= 16777619u;
  }
  return h ^ (h >> 15);
}

/* phase 1: common warm-up on this GPU, then keep this worker's share. Counters of the warm-up are returned in
 * base_* (the caller adds them for worker 0 only); the engine restarts with zero counters and the kept nodes. */
int seed_partition(bb_engine *e, int kind, int D, int me, int m_keep, int best0, unsigned long long *base_tree,
                   unsigned long long *base_sol, int *best)
{
  bb_stats st;
  memset(&st, 0, sizeof(st));
  int rc = bb_engine_reset(e, best0);
  if (!rc)
    rc = bb_engine_push_root(e);
  if (rc)
    return rc;
  for (int it = 0; it < 64; it++)
  {
    rc = bb_engine_run(e, 1, &st);
    if (rc)
      return rc;
    /* the reference stops at D*m nodes (pfsp_multigpu_cuda.c:111); a finer split costs nothing here (the
     * warm-up runs on the GPU) and evens out the sub-tree sizes the workers start from */
    if (st.pool_size == 0 || st.pool_size >= (long long)D * (m_keep > 2048 ? m_keep : 2048))
      break;
  }
  *base_tree = st.tree;
  *base_sol = st.sol;
  *best = st.best;
  if (st.pool_size == 0)
    return bb_engine_reset(e, st.best);
  const size_t rec = kind == KIND_PFSP ? sizeof(int16_t) * (2 + BB_MAX_JOBS) : (1 + BB_MAX_QUEENS);
  std::vector<unsigned char> all((size_t)st.pool_size * rec), keep;
  const long long got = bb_engine_export_nodes(e, all.data(), st.pool_size);
  if (got < 0)
    return (int)got;
  keep.reserve(all.size() / D + rec);
  for (long long i = 0; i < got; i++)
  {
    const unsigned char *nd = all.data() + (size_t)i * rec;
    unsigned h;
    if (kind == KIND_PFSP)
    {
      const int16_t *p = reinterpret_cast<const int16_t *>(nd);
      h = prefix_hash(reinterpret_cast<const unsigned char *>(p + 2), 2 * p[0]);
    }
    else
      h = prefix_hash(nd + 1, nd[0]);
    if ((int)(h % (unsigned)D) == me)
      keep.insert(keep.end(), nd, nd + rec);
  }
  rc = bb_engine_reset(e, st.best);
  if (!rc && !keep.empty())
    rc = bb_engine_push_nodes(e, keep.data(), (int)(keep.size() / rec));
  return rc;
}

bb_engine *create_engine(int kind, const bb_pfsp_instance *ins, int lb, int N, int G, const bb_config *cfg, int device)
{
  bb_config c;
  if (cfg)
    c = *cfg;
  else
    bb_config_default(&c);
  c.device = device;
  if (c.capacity <= 0)
  {
    /* a worker of a multi-GPU search holds only its share of the tree: 26 x M nodes per stack keep the planner's
     * worst-case room (maxkids children per popped parent) and halve the cudaMalloc / cudaFree time, which D
     * processes pay one after the other (device memory mapping is serialised system-wide) */
    const long long M = c.M > 0 ? c.M : 50000;
    c.capacity = 26 * M > (1ll << 22) ? 26 * M : (1ll << 22);
  }
  return kind == KIND_PFSP ? bb_pfsp_engine_create(ins, lb, &c) : bb_nq_engine_create(N, G, &c);
}

int search_single(int kind, const bb_pfsp_instance *ins, int lb, int best0, int N, int G, const bb_config *cfg,
                  bb_search_result *res, bb_stats *per_gpu)
{
  /* one-time per process and not part of a search: the CUDA context of the device */
  BB_CUDA_TRY(cudaSetDevice(cfg ? cfg->device : 0));
  BB_CUDA_TRY(cudaFree(nullptr));
  const double t0 = now_s();
  bb_engine *e = kind == KIND_PFSP ? bb_pfsp_engine_create(ins, lb, cfg) : bb_nq_engine_create(N, G, cfg);
  if (!e)
    return BB_ERR_CUDA;
  bb_stats st;
  memset(&st, 0, sizeof(st));
  int rc = bb_engine_reset(e, best0);
  if (!rc)
    rc = bb_engine_push_root(e);
  const double t1 = now_s();
  if (!rc)
    rc = bb_engine_run(e, 0, &st);
  const double t2 = now_s();
  bb_engine_destroy(e);
  if (rc)
    return rc;
  memset(res, 0, sizeof(*res));
  res->tree = st.tree;
  res->sol = st.sol;
  res->best = st.best;
  res->elapsed_s = now_s() - t0;
  res->search_s = t2 - t1;
  res->setup_s = t1 - t0;
  res->kernel_ms_max = st.kernel_ms;
  if (per_gpu)
    per_gpu[0] = st;
  return BB_OK;
}

struct thread_out
{
  double setup_s, warm_s;
  int rc;
  char err[256];
  bb_stats st;
  bb_team_result tr;
  unsigned long long base_tree, base_sol;
  double search_s;
};

} // namespace

/* ------------------------------------------------------------------ D GPUs, one process, one thread per GPU */

int bb_multi_search(int kind, const bb_pfsp_instance *ins, int lb, int best0, int N, int G, int D, int ws,
                    const bb_config *cfg, bb_search_result *res, bb_stats *per_gpu)
{
  if (D == 1)
    return search_single(kind, ins, lb, best0, N, G, cfg, res, per_gpu);
  int ndev = 0;
  BB_CUDA_TRY(cudaGetDeviceCount(&ndev));
  if (D > ndev || D > 64)
    return bb_fail(BB_ERR_ARG, "%d GPUs requested, %d visible", D, ndev); /* pfsp_multigpu_cuda.c:539-543 */
  bb_config c;
  if (cfg)
    c = *cfg;
  else
    bb_config_default(&c);
  const int m_keep = c.m > 0 ? c.m : 25;
  const long long M = c.M > 0 ? c.M : 50000;
  /* one-time per process, outside the timed search: create the D CUDA contexts (in parallel - a context takes a
   * few hundred ms and they are created serially otherwise) and enable peer access between them while no pool is
   * allocated yet */
  {
    std::vector<std::thread> init;
    std::vector<int> irc(D, 0);
    for (int i = 0; i < D; i++)
      init.emplace_back([&, i]() {
        if (cudaSetDevice(i) != cudaSuccess || cudaFree(nullptr) != cudaSuccess)
          irc[i] = 1;
      });
    for (auto &t : init)
      t.join();
    for (int i = 0; i < D; i++)
      if (irc[i])
        return bb_fail(BB_ERR_CUDA, "could not initialise GPU %d", i);
  }
  for (int i = 0; i < D; i++)
    for (int j = 0; j < D; j++)
      if (i != j)
      {
        BB_CUDA_TRY(cudaSetDevice(i));
        int can = 0;
        BB_CUDA_TRY(cudaDeviceCanAccessPeer(&can, i, j));
        if (!can)
          return bb_fail(BB_ERR_CUDA, "GPU %d cannot access GPU %d's memory (no peer path)", i, j);
        const cudaError_t err = cudaDeviceEnablePeerAccess(j, 0);
        if (err != cudaSuccess && err != cudaErrorPeerAccessAlreadyEnabled)
          return bb_fail_cuda(err, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
        cudaGetLastError();
      }
  const double t0 = now_s(); /* elapsed_s: engine creation, warm-up, search, teardown */
  void *team = nullptr;
  if (posix_memalign(&team, 128, bb_team_bytes(D)) != 0)
    return bb_fail(BB_ERR_ARG, "out of host memory");
  bb_team_init(team, D);
  std::vector<gpu_worker> W(D);
  std::vector<thread_out> out(D);
  std::vector<std::thread> th;
  for (int i = 0; i < D; i++)
    th.emplace_back([&, i]() {
      thread_out &o = out[i];
      memset(&o, 0, sizeof(o));
      gpu_worker &w = W[i];
      memset(&w, 0, sizeof(w));
      w.me = i;
      auto fail = [&](int rc) {
        o.rc = rc;
        strncpy(o.err, bb_last_error(), sizeof(o.err) - 1);
        bb_team_fail(team);
      };
      const double c0 = now_s();
      w.e = create_engine(kind, ins, lb, N, G, &c, i);
      if (!w.e)
        fail(BB_ERR_CUDA);
      else if (bb_engine_stage_alloc(w.e, BB_STAGE_BYTES) != BB_OK)
        fail(BB_ERR_CUDA);
      o.setup_s = now_s() - c0;
      if (bb_team_barrier(team, 120) != BB_OK || o.rc)
      {
        if (w.e)
          bb_engine_destroy(w.e);
        if (!o.rc)
          fail(BB_ERR_STATE);
        return;
      }
      for (int j = 0; j < D; j++)
        w.peer_stage[j] = bb_engine_stage_ptr(W[j].e); /* peer access is enabled: direct pointers */
      int best = best0;
      const double w0 = now_s();
      int rc = seed_partition(w.e, kind, D, i, m_keep, best0, &o.base_tree, &o.base_sol, &best);
      o.warm_s = now_s() - w0;
      if (rc)
        fail(rc);
      if (bb_team_barrier(team, 120) == BB_OK && !o.rc)
      {
        const double s0 = now_s();
        cudaSetDevice(i);
        device_span span;
        span.begin();
        bb_team_ops ops = gpu_ops(&w);
        rc = bb_team_worker(team, i, &ops, ws, m_keep, 5 * M, best, c.iters_per_launch, &o.tr);
        cudaSetDevice(i);
        o.search_s = span.end_s(now_s() - s0);
        if (rc)
          fail(rc);
        else
        {
          bb_engine_get_stats(w.e, &o.st);
          o.st.steals = o.tr.steals;
          o.st.steals_ok = o.tr.steals_ok;
          o.st.donated = o.tr.donated;
        }
      }
      bb_team_barrier(team, 120); /* nobody frees a pool a peer may still be reading */
      bb_engine_destroy(w.e);
    });
  for (auto &t : th)
    t.join();
  int rc = BB_OK;
  memset(res, 0, sizeof(*res));
  res->best = bb_team_best(team);
  for (int i = 0; i < D; i++)
  {
    if (out[i].rc && !rc)
      rc = bb_fail(out[i].rc, "GPU %d: %s", i, out[i].err);
    res->tree += out[i].st.tree;
    res->sol += out[i].st.sol;
    res->steals += out[i].tr.steals;
    res->steals_ok += out[i].tr.steals_ok;
    res->nodes_moved += out[i].tr.donated;
    if (out[i].st.kernel_ms > res->kernel_ms_max)
      res->kernel_ms_max = out[i].st.kernel_ms;
    if (out[i].search_s > res->search_s)
      res->search_s = out[i].search_s;
    if (out[i].st.best < res->best && out[i].st.best > 0)
      res->best = out[i].st.best;
    if (per_gpu)
      per_gpu[i] = out[i].st;
  }
  res->tree += out[0].base_tree;
  res->sol += out[0].base_sol;
  res->warm_tree = out[0].base_tree;
  res->warm_sol = out[0].base_sol;
  for (int i = 0; i < D; i++)
  {
    if (out[i].warm_s > res->warm_s)
      res->warm_s = out[i].warm_s;
    if (out[i].setup_s > res->setup_s)
      res->setup_s = out[i].setup_s;
  }
  res->elapsed_s = now_s() - t0;
  free(team);
  return rc;
}

/* ------------------------------------------------------------------ one process per GPU (torchrun): shm + CUDA IPC */

namespace
{


int search_dist(int kind, const bb_pfsp_instance *ins, int lb, int best0, int N, int G, int rank, int world,
                const char *shm_name, int ws, const bb_config *cfg, bb_search_result *res, bb_stats *mine)
{
  if (!res || !shm_name || world < 1 || world > 64 || rank < 0 || rank >= world)
    return bb_fail(BB_ERR_ARG, "search_dist: bad arguments");
  const double t0 = now_s();
  bb_config c;
  if (cfg)
    c = *cfg;
  else
    bb_config_default(&c);
  const int m_keep = c.m > 0 ? c.m : 25;
  const long long M = c.M > 0 ? c.M : 50000;
  const size_t team_sz = (bb_team_bytes(world) + 255) & ~(size_t)255;
  const size_t total = team_sz + sizeof(ipc_slot) * (size_t)world;
  const int fd = shm_open(shm_name, O_CREAT | O_RDWR, 0600);
  if (fd < 0)
    return bb_fail(BB_ERR_STATE, "shm_open(%s) failed", shm_name);
  if (ftruncate(fd, (off_t)total) != 0)
  {
    close(fd);
    return bb_fail(BB_ERR_STATE, "ftruncate(%s) failed", shm_name);
  }
  void *mem = mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (mem == MAP_FAILED)
    return bb_fail(BB_ERR_STATE, "mmap(%s) failed", shm_name);
  ipc_slot *I = reinterpret_cast<ipc_slot *>(static_cast<char *>(mem) + team_sz);
  int rc = rank == 0 ? bb_team_init(mem, world) : bb_team_wait_ready(mem, world, 120);
  const bool trace = getenv("BB_TRACE_PHASES") != nullptr; /* per-rank phase times on stderr */
  double tp[8];
  tp[0] = now_s();
  gpu_worker w;
  memset(&w, 0, sizeof(w));
  w.me = rank;
  bb_stats st;
  memset(&st, 0, sizeof(st));
  bb_team_result tr;
  memset(&tr, 0, sizeof(tr));
  unsigned long long base_tree = 0, base_sol = 0;
  double search_s = 0;
  auto fail = [&](int code) {
    rc = code;
    bb_team_fail(mem);
  };
  double setup_s = 0, warm_s = 0;
  if (!rc)
  {
    const double c0 = now_s();
    w.e = create_engine(kind, ins, lb, N, G, &c, c.device);
    setup_s = now_s() - c0;
    if (!w.e)
      fail(BB_ERR_CUDA);
  }
  if (!rc && bb_engine_stage_alloc(w.e, BB_STAGE_BYTES) != BB_OK)
    fail(BB_ERR_CUDA);
  if (!rc)
  {
    const cudaError_t err = cudaIpcGetMemHandle(&I[rank].stage, bb_engine_stage_ptr(w.e));
    if (err != cudaSuccess)
      fail(bb_fail_cuda(err, "cudaIpcGetMemHandle", __FILE__, __LINE__));
    w.peer_stage[rank] = bb_engine_stage_ptr(w.e);
  }
  tp[1] = now_s();
  if (bb_team_barrier(mem, 180) != BB_OK && !rc)
    rc = BB_ERR_STATE;
  tp[2] = now_s();
  w.ipc = I; /* peers' transfer buffers are mapped lazily, at the first steal (ensure_peer_mapped) */
  int best = best0;
  tp[3] = now_s();
  if (!rc)
  {
    const double w0 = now_s();
    const int r2 = seed_partition(w.e, kind, world, rank, m_keep, best0, &base_tree, &base_sol, &best);
    warm_s = now_s() - w0;
    if (r2)
      fail(r2);
  }
  if (bb_team_barrier(mem, 180) != BB_OK && !rc)
    rc = BB_ERR_STATE;
  if (!rc)
  {
    const double s0 = now_s();
    cudaSetDevice(c.device);
    device_span span;
    span.begin();
    bb_team_ops ops = gpu_ops(&w);
    const int r2 = bb_team_worker(mem, rank, &ops, ws, m_keep, 5 * M, best, c.iters_per_launch, &tr);
    cudaSetDevice(c.device);
    search_s = span.end_s(now_s() - s0);
    if (r2)
      fail(r2);
    else
      bb_engine_get_stats(w.e, &st);
  }
  tp[4] = now_s();
  I[rank].tree = st.tree + (rank == 0 ? base_tree : 0);
  I[rank].sol = st.sol + (rank == 0 ? base_sol : 0);
  I[rank].steals = tr.steals;
  I[rank].steals_ok = tr.steals_ok;
  I[rank].donated = tr.donated;
  I[rank].kernel_ms = st.kernel_ms;
  I[rank].search_s = search_s;
  I[rank].warm_s = warm_s;
  I[rank].setup_s = setup_s;
  I[rank].warm_tree = rank == 0 ? base_tree : 0;
  I[rank].warm_sol = rank == 0 ? base_sol : 0;
  I[rank].best = tr.best;
  st.steals = tr.steals;
  st.steals_ok = tr.steals_ok;
  st.donated = tr.donated;
  I[rank].rc = rc;
  const int b1 = bb_team_barrier(mem, 180);
  memset(res, 0, sizeof(*res));
  res->best = INT_MAX;
  if (b1 == BB_OK && !rc)
    for (int j = 0; j < world; j++)
    {
      res->tree += I[j].tree;
      res->sol += I[j].sol;
      res->steals += I[j].steals;
      res->steals_ok += I[j].steals_ok;
      res->nodes_moved += I[j].donated;
      if (I[j].kernel_ms > res->kernel_ms_max)
        res->kernel_ms_max = I[j].kernel_ms;
      if (I[j].search_s > res->search_s)
        res->search_s = I[j].search_s;
      if (I[j].warm_s > res->warm_s)
        res->warm_s = I[j].warm_s;
      if (I[j].setup_s > res->setup_s)
        res->setup_s = I[j].setup_s;
      res->warm_tree += I[j].warm_tree;
      res->warm_sol += I[j].warm_sol;
      if (I[j].best < res->best)
        res->best = I[j].best;
      if (I[j].rc && !rc)
        rc = bb_fail(I[j].rc, "rank %d failed", j);
    }
  else if (!rc)
    rc = BB_ERR_STATE;
  for (int j = 0; j < world; j++)
    if (w.opened[j])
      cudaIpcCloseMemHandle(w.peer_stage[j]);
  tp[5] = now_s();
  bb_team_barrier(mem, 60); /* every mapping of our pool is closed before we free it */
  if (w.e)
    bb_engine_destroy(w.e);
  tp[6] = now_s();
  if (trace)
    fprintf(stderr, "[bb rank %d] create+handles %.1f ms, barrier %.1f, ipc open %.1f, warm-up..search end %.1f "
                    "(search %.1f), reduce+close %.1f, destroy %.1f\n", rank, (tp[1] - tp[0]) * 1e3,
            (tp[2] - tp[1]) * 1e3, (tp[3] - tp[2]) * 1e3, (tp[4] - tp[3]) * 1e3, search_s * 1e3, (tp[5] - tp[4]) * 1e3,
            (tp[6] - tp[5]) * 1e3);
  if (mine)
    *mine = st;
  munmap(mem, total);
  if (rank == 0)
    shm_unlink(shm_name);
  res->elapsed_s = now_s() - t0;
  return rc;
}

} // namespace

extern "C" int bb_pfsp_search_dist(const bb_pfsp_instance *ins, int lb, int best0, int rank, int world,
                                   const char *shm_name, int ws, const bb_config *cfg, bb_search_result *res,
                                   bb_stats *mine)
{
  if (!ins || lb < 0 || lb > 2)
    return bb_fail(BB_ERR_ARG, "bb_pfsp_search_dist: bad instance or lower bound");
  return search_dist(KIND_PFSP, ins, lb, best0, 0, 0, rank, world, shm_name, ws, cfg, res, mine);
}

extern "C" int bb_nq_search_dist(int N, int G, int rank, int world, const char *shm_name, int ws, const bb_config *cfg,
                                 bb_search_result *res, bb_stats *mine)
{
  return search_dist(KIND_NQ, nullptr, 0, INT_MAX, N, G, rank, world, shm_name, ws, cfg, res, mine);
}
