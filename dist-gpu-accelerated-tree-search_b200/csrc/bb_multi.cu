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
int bb_engine_pull(bb_engine *e, const void *peer_slot, long long count);
int bb_engine_stage_alloc(bb_engine *e, size_t bytes);
void *bb_engine_stage_ptr(bb_engine *e);
size_t bb_engine_stage_slot_bytes(bb_engine *e);
int bb_engine_partition(bb_engine *e, int D, int me, int zero_counters);
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

#define BB_STAGE_BYTES ((size_t)256 << 20) /* transfer buffer per GPU: two slots, each the most one steal can move */

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
  (void)start; /* the donation sits at the beginning of slot `stack` of the victim's transfer buffer */
  const int rc = ensure_peer_mapped(w, victim);
  if (rc)
    return rc;
  const char *slot = static_cast<const char *>(w->peer_stage[victim]) + (size_t)stack * bb_engine_stage_slot_bytes(w->e);
  return bb_engine_pull(w->e, slot, count);
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

/* phase 1: common warm-up on this GPU, then keep this worker's share (k_partition: hash of the placed prefix mod D,
 * on the device, in place). Counters of the warm-up are returned in base_* (the caller adds them for worker 0 only);
 * the engine continues with zero counters and the kept nodes. */
int seed_partition(bb_engine *e, int kind, int D, int me, int m_keep, int best0, unsigned long long *base_tree,
                   unsigned long long *base_sol, int *best)
{
  (void)kind;
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
  return bb_engine_partition(e, D, me, 1);
}

bb_engine *create_engine(int kind, const bb_pfsp_instance *ins, int lb, int N, int G, const bb_config *cfg, int device,
                         int team = 1)
{
  bb_config c;
  if (cfg)
    c = *cfg;
  else
    bb_config_default(&c);
  c.device = device;
  /* N-Queens, last levels in registers: a worker of a team of four or more holds so few deep parents that whole
   * subtrees of the single-GPU height leave lanes idle at the end of a launch (8 GPUs, N = 17: 5 subtrees per lane);
   * one level less gives eight times as many, smaller ones */
  if (kind != KIND_PFSP && team >= 4 && c.packed == 3)
    c.packed = 4;
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
  /* BB_WORKERS_SHARE_GPUS=1: more workers than GPUs, worker i on GPU i mod visible. The team protocol, the warm-up
   * split, donations and pulls run exactly as across GPUs (the "peer" buffer is then a buffer of the same device);
   * it exists so that the multi-worker paths can be exercised on a one-GPU box, not for speed. */
  const bool share = getenv("BB_WORKERS_SHARE_GPUS") != nullptr && ndev > 0;
  if ((D > ndev && !share) || D > 64)
    return bb_fail(BB_ERR_ARG, "%d GPUs requested, %d visible", D, ndev); /* pfsp_multigpu_cuda.c:539-543 */
  auto dev_of = [ndev](int worker) { return worker % ndev; };
  const int used_dev = D < ndev ? D : ndev;
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
        if (cudaSetDevice(dev_of(i)) != cudaSuccess || cudaFree(nullptr) != cudaSuccess)
          irc[i] = 1;
      });
    for (auto &t : init)
      t.join();
    for (int i = 0; i < D; i++)
      if (irc[i])
        return bb_fail(BB_ERR_CUDA, "could not initialise GPU %d", i);
  }
  for (int i = 0; i < used_dev; i++)
    for (int j = 0; j < used_dev; j++)
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
      w.e = create_engine(kind, ins, lb, N, G, &c, dev_of(i), D);
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
      /* the search is timed from the root: warm-up expansion and partition are inside the span */
      const double s0 = now_s();
      cudaSetDevice(dev_of(i));
      device_span span;
      span.begin();
      int rc = seed_partition(w.e, kind, D, i, m_keep, best0, &o.base_tree, &o.base_sol, &best);
      o.warm_s = now_s() - s0;
      if (rc)
        fail(rc);
      if (!o.rc)
      {
        bb_team_ops ops = gpu_ops(&w);
        rc = bb_team_worker(team, i, &ops, ws, m_keep, 5 * M, best, c.iters_per_launch, &o.tr);
        cudaSetDevice(dev_of(i));
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

/* ------------------------------------------------------------------ one process per GPU (torchrun): shm + CUDA IPC
 * A persistent team: bb_*_dist_open creates the rank's engine, its transfer buffer and the mappings of every peer's
 * buffer ONCE; bb_dist_search then runs whole searches on it (reset, root, common warm-up, device-side partition,
 * stealing, termination) with nothing allocated, mapped or freed per search. bb_*_search_dist is open + search +
 * close in one call. */

struct bb_dist
{
  int kind, rank, world;
  bb_config c;
  int m_keep;
  long long M;
  bb_engine *e;
  gpu_worker w;
  void *mem;
  size_t total, team_sz;
  ipc_slot *I;
  char shm_name[200];
  double setup_s; /* engine creation (pool allocation, table upload, LB2 table build) */
  double open_s;  /* the whole open call */
  long long searches;
};

namespace
{

void dist_free(bb_dist *d, bool team_alive)
{
  if (!d)
    return;
  if (d->e)
    cudaSetDevice(d->c.device);
  for (int j = 0; j < d->world && j < 64; j++)
    if (d->w.opened[j])
      cudaIpcCloseMemHandle(d->w.peer_stage[j]);
  if (team_alive && d->mem)
    bb_team_barrier(d->mem, 60); /* every mapping of our buffer is closed before we free it */
  if (d->e)
    bb_engine_destroy(d->e);
  if (d->mem)
    munmap(d->mem, d->total);
  if (d->rank == 0 && d->shm_name[0])
    shm_unlink(d->shm_name);
  free(d);
}

bb_dist *dist_open(int kind, const bb_pfsp_instance *ins, int lb, int N, int G, int rank, int world,
                   const char *shm_name, const bb_config *cfg)
{
  if (!shm_name || world < 1 || world > 64 || rank < 0 || rank >= world || strlen(shm_name) >= 200)
  {
    bb_fail(BB_ERR_ARG, "dist_open: bad arguments");
    return nullptr;
  }
  const double t0 = now_s();
  bb_dist *d = static_cast<bb_dist *>(calloc(1, sizeof(bb_dist)));
  d->kind = kind;
  d->rank = rank;
  d->world = world;
  if (cfg)
    d->c = *cfg;
  else
    bb_config_default(&d->c);
  d->m_keep = d->c.m > 0 ? d->c.m : 25;
  d->M = d->c.M > 0 ? d->c.M : 50000;
  strcpy(d->shm_name, shm_name);
  d->team_sz = (bb_team_bytes(world) + 255) & ~(size_t)255;
  d->total = d->team_sz + sizeof(ipc_slot) * (size_t)world;
  const int fd = shm_open(shm_name, O_CREAT | O_RDWR, 0600);
  if (fd < 0 || ftruncate(fd, (off_t)d->total) != 0)
  {
    if (fd >= 0)
      close(fd);
    bb_fail(BB_ERR_STATE, "shm_open / ftruncate(%s) failed", shm_name);
    free(d);
    return nullptr;
  }
  d->mem = mmap(nullptr, d->total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (d->mem == MAP_FAILED)
  {
    bb_fail(BB_ERR_STATE, "mmap(%s) failed", shm_name);
    free(d);
    return nullptr;
  }
  d->I = reinterpret_cast<ipc_slot *>(static_cast<char *>(d->mem) + d->team_sz);
  int rc = rank == 0 ? bb_team_init(d->mem, world) : bb_team_wait_ready(d->mem, world, 120);
  d->w.me = rank;
  if (!rc)
  {
    const double c0 = now_s();
    d->e = create_engine(kind, ins, lb, N, G, &d->c, d->c.device, world);
    d->setup_s = now_s() - c0;
    if (!d->e)
      rc = BB_ERR_CUDA;
  }
  if (!rc && world > 1)
  {
    rc = bb_engine_stage_alloc(d->e, BB_STAGE_BYTES);
    if (!rc)
    {
      const cudaError_t err = cudaIpcGetMemHandle(&d->I[rank].stage, bb_engine_stage_ptr(d->e));
      if (err != cudaSuccess)
        rc = bb_fail_cuda(err, "cudaIpcGetMemHandle", __FILE__, __LINE__);
      d->w.peer_stage[rank] = bb_engine_stage_ptr(d->e);
    }
  }
  if (rc)
    bb_team_fail(d->mem);
  d->w.e = d->e;
  d->w.ipc = d->I;
  if (bb_team_barrier(d->mem, 180) != BB_OK && !rc)
    rc = BB_ERR_STATE;
  /* map every peer's transfer buffer now (a few ms each, once per team instead of inside the first steals) */
  for (int j = 0; !rc && world > 1 && j < world; j++)
    if (j != rank)
      rc = ensure_peer_mapped(&d->w, j);
  if (rc)
    bb_team_fail(d->mem);
  if (bb_team_barrier(d->mem, 180) != BB_OK && !rc)
    rc = BB_ERR_STATE;
  if (rc)
  {
    char msg[512];
    strncpy(msg, bb_last_error(), sizeof(msg) - 1);
    msg[sizeof(msg) - 1] = 0;
    dist_free(d, false);
    bb_fail(rc, "dist_open (rank %d): %s", rank, msg);
    return nullptr;
  }
  d->open_s = now_s() - t0;
  return d;
}

} // namespace

extern "C" int bb_dist_search(bb_dist *d, int best0, int ws, bb_search_result *res, bb_stats *mine)
{
  if (!d || !res)
    return bb_fail(BB_ERR_ARG, "bb_dist_search: bad arguments");
  const double t0 = now_s();
  const int rank = d->rank, world = d->world;
  ipc_slot *I = d->I;
  int rc = BB_OK;
  /* every rank has left the previous search; rank 0 clears the team block; then everybody starts together */
  if (bb_team_barrier(d->mem, 180) != BB_OK)
    return BB_ERR_STATE;
  if (rank == 0 && d->searches > 0)
    bb_team_reset(d->mem);
  if (bb_team_barrier(d->mem, 180) != BB_OK)
    return BB_ERR_STATE;
  d->searches++;
  bb_stats st;
  memset(&st, 0, sizeof(st));
  bb_team_result tr;
  memset(&tr, 0, sizeof(tr));
  unsigned long long base_tree = 0, base_sol = 0;
  int best = best0;
  /* the search is timed from the root: reset, root seeding, common warm-up and partition are inside the span */
  const double s0 = now_s();
  cudaSetDevice(d->c.device);
  device_span span;
  span.begin();
  rc = seed_partition(d->e, d->kind, world, rank, d->m_keep, best0, &base_tree, &base_sol, &best);
  const double warm_s = now_s() - s0;
  double search_s = 0;
  if (rc)
    bb_team_fail(d->mem);
  else
  {
    d->w.kernel_ms_seen = 0;
    bb_team_ops ops = gpu_ops(&d->w);
    rc = bb_team_worker(d->mem, rank, &ops, ws, d->m_keep, 5 * d->M, best, d->c.iters_per_launch, &tr);
    cudaSetDevice(d->c.device);
    search_s = span.end_s(now_s() - s0);
    if (rc)
      bb_team_fail(d->mem);
    else
      bb_engine_get_stats(d->e, &st);
  }
  I[rank].tree = st.tree + (rank == 0 ? base_tree : 0);
  I[rank].sol = st.sol + (rank == 0 ? base_sol : 0);
  I[rank].steals = tr.steals;
  I[rank].steals_ok = tr.steals_ok;
  I[rank].donated = tr.donated;
  I[rank].kernel_ms = st.kernel_ms;
  I[rank].search_s = search_s;
  I[rank].warm_s = warm_s;
  I[rank].setup_s = d->searches == 1 ? d->setup_s : 0.0;
  I[rank].warm_tree = rank == 0 ? base_tree : 0;
  I[rank].warm_sol = rank == 0 ? base_sol : 0;
  I[rank].best = tr.best;
  I[rank].rc = rc;
  st.steals = tr.steals;
  st.steals_ok = tr.steals_ok;
  st.donated = tr.donated;
  const int b1 = bb_team_barrier(d->mem, 180);
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
  if (mine)
    *mine = st;
  res->elapsed_s = now_s() - t0;
  if (getenv("BB_TRACE_PHASES"))
    fprintf(stderr, "[bb rank %d] call %.3f ms: entry barriers %.3f, warm-up %.3f, span %.3f, stats+final barrier %.3f\n",
            rank, res->elapsed_s * 1e3, (s0 - t0) * 1e3, warm_s * 1e3, search_s * 1e3,
            (now_s() - s0 - search_s) * 1e3);
  return rc;
}

extern "C" void bb_dist_close(bb_dist *d)
{
  if (d)
    dist_free(d, true);
}

extern "C" bb_dist *bb_pfsp_dist_open(const bb_pfsp_instance *ins, int lb, int rank, int world, const char *shm_name,
                                      const bb_config *cfg)
{
  if (!ins || lb < 0 || lb > 2)
  {
    bb_fail(BB_ERR_ARG, "bb_pfsp_dist_open: bad instance or lower bound");
    return nullptr;
  }
  return dist_open(KIND_PFSP, ins, lb, 0, 0, rank, world, shm_name, cfg);
}

extern "C" bb_dist *bb_nq_dist_open(int N, int G, int rank, int world, const char *shm_name, const bb_config *cfg)
{
  return dist_open(KIND_NQ, nullptr, 0, N, G, rank, world, shm_name, cfg);
}

static int search_dist_once(bb_dist *d, int best0, int ws, bb_search_result *res, bb_stats *mine, double t0)
{
  if (!d)
    return BB_ERR_STATE;
  const int rc = bb_dist_search(d, best0, ws, res, mine);
  bb_dist_close(d);
  if (res)
    res->elapsed_s = now_s() - t0;
  return rc;
}

extern "C" int bb_pfsp_search_dist(const bb_pfsp_instance *ins, int lb, int best0, int rank, int world,
                                   const char *shm_name, int ws, const bb_config *cfg, bb_search_result *res,
                                   bb_stats *mine)
{
  if (!res)
    return bb_fail(BB_ERR_ARG, "bb_pfsp_search_dist: bad arguments");
  const double t0 = now_s();
  return search_dist_once(bb_pfsp_dist_open(ins, lb, rank, world, shm_name, cfg), best0, ws, res, mine, t0);
}

extern "C" int bb_nq_search_dist(int N, int G, int rank, int world, const char *shm_name, int ws, const bb_config *cfg,
                                 bb_search_result *res, bb_stats *mine)
{
  if (!res)
    return bb_fail(BB_ERR_ARG, "bb_nq_search_dist: bad arguments");
  const double t0 = now_s();
  return search_dist_once(bb_nq_dist_open(N, G, rank, world, shm_name, cfg), INT_MAX, ws, res, mine, t0);
}
