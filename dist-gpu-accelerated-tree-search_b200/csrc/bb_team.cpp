/*
 * bb_team.cpp - work sharing, incumbent exchange and termination detection between the workers of one node
 * (one worker per GPU). Pure host code over a block of shared memory: ordinary heap memory when the workers are
 * threads of one process (bb_pfsp_search / bb_nq_search with D > 1), a POSIX shared-memory mapping when they are
 * processes (one rank per GPU under torchrun: bb_*_search_dist). The engine is reached only through bb_team_ops,
 * so the same protocol drives the GPU engines and the host stand-in the CPU tests use.
 *
 * What of the reference this replaces:
 *   steal loop (random victim, <= 10 lock tries, steal half, pfsp/pfsp_multigpu_cuda.c:343-404)
 *       -> the idle worker posts a REQUEST in the victim's slot; the victim answers at its next poll by reserving
 *          the oldest half of its pool (bb_team_ops.donate) and publishing where it is; the thief copies it out
 *          of the victim's memory itself (bb_team_ops.pull: NVLink peer copy for the GPU engine) and acknowledges.
 *          No lock is ever held across a transfer and the victim's GPU keeps running.
 *   eachTaskState / allIdle termination (pfsp_multigpu_cuda.c:408-426, common/util.c:19-38)
 *       -> three states (BUSY / HUNTING / IDLE) and an epoch counter bumped on every IDLE -> HUNTING transition;
 *          a worker declares the search finished when it reads every state IDLE between two equal epoch reads.
 *          A donor stays BUSY until its donation is acknowledged, so work in flight always has a BUSY owner.
 *   checkBest (pfsp_multigpu_cuda.c:30-50) -> every slot publishes its incumbent, workers adopt the minimum.
 */
#include "bb_gpu.h"

#include <limits.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

int bb_fail(int code, const char *fmt, ...);

namespace
{

enum
{
  ST_BUSY = 0,
  ST_HUNTING = 1,
  ST_IDLE = 2
};
enum
{
  RESP_NONE = 0,
  RESP_GRANT = 1,
  RESP_DENY = 2
};

constexpr unsigned TEAM_MAGIC = 0xbb7ea301u;

struct alignas(128) team_slot
{
  int state;
  int request; /* thief id + 1, 0 = none */
  int resp;    /* written by the victim into the THIEF's slot */
  int resp_victim, resp_stack;
  long long resp_start, resp_count;
  int acks[2]; /* written by thieves into the VICTIM's slot, one counter per stack */
  int best;
  long long pool_size;
  /* results */
  unsigned long long tree, sol, steals, steals_ok, donated;
  double busy_ms;
  int rc;
};

struct alignas(128) team_hdr
{
  unsigned magic;
  int world;
  int epoch;
  int finished;
  int failed;
  int bar_count, bar_sense;
};

inline team_slot *slots_of(void *mem) { return reinterpret_cast<team_slot *>(static_cast<char *>(mem) + sizeof(team_hdr)); }

inline int ld(const int *p) { return __atomic_load_n(p, __ATOMIC_SEQ_CST); }
inline void st(int *p, int v) { __atomic_store_n(p, v, __ATOMIC_SEQ_CST); }
inline long long ldll(const long long *p) { return __atomic_load_n(p, __ATOMIC_SEQ_CST); }
inline void stll(long long *p, long long v) { __atomic_store_n(p, v, __ATOMIC_SEQ_CST); }

inline double now_ms();

/* short waits are spun (nanosleep rounds a 20-50 us request up to 70-120 us with the default timer slack, and a
 * search on 8 GPUs lasts only tens of milliseconds); one worker per GPU leaves the host cores for it */
inline void nap(unsigned us)
{
  if (us >= 200)
  {
    struct timespec ts = {0, (long)us * 1000L};
    nanosleep(&ts, nullptr);
    return;
  }
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  do
  {
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#endif
    clock_gettime(CLOCK_MONOTONIC, &t1);
  } while ((t1.tv_sec - t0.tv_sec) * 1000000L + (t1.tv_nsec - t0.tv_nsec) / 1000L < (long)us);
}

inline double now_ms()
{
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

struct worker
{
  team_hdr *H;
  team_slot *S;
  int me, D;
  const bb_team_ops *ops;
  long long m_keep, max_steal;
  int outstanding[2], acks_seen[2];
  bool donation_pending; /* ops->donate answered "pending" for the request in our slot */
  int best;
  unsigned rng;

  team_slot &mine() { return S[me]; }

  int service_acks()
  {
    for (int s = 0; s < 2; s++)
      while (acks_seen[s] < ld(&mine().acks[s]))
      {
        int rc = ops->donate_done(ops->self, s);
        if (rc)
          return rc;
        acks_seen[s]++;
        outstanding[s]--;
      }
    return 0;
  }

  /* answer the request waiting in our slot, if any: grant the oldest half of the pool or deny. ops->donate may
   * answer "pending" (count = -1: the GPU engine has only enqueued the donation behind the batch in flight and never
   * waits for its device); the request then stays in our slot and is looked at again at the next poll. */
  int service_request(bool can_give)
  {
    const int req = ld(&mine().request);
    if (!req)
      return 0;
    team_slot &thief = S[req - 1];
    int stack = 0;
    long long start = 0, count = 0;
    /* a donation goes through one of the engine's two transfer-buffer slots, which stays busy until the thief
     * acknowledges; the engine answers 0 when both are taken */
    if (donation_pending || (can_give && !ld(&H->finished)))
    {
      int rc = ops->donate(ops->self, m_keep, max_steal, &stack, &start, &count);
      if (rc)
        return rc;
      donation_pending = count < 0;
      if (donation_pending)
        return 0;
    }
    if (count > 0)
    {
      outstanding[stack & 1]++;
      mine().donated += (unsigned long long)count;
      thief.resp_victim = me;
      thief.resp_stack = stack;
      thief.resp_start = start;
      thief.resp_count = count;
      st(&thief.resp, RESP_GRANT);
    }
    else
      st(&thief.resp, RESP_DENY);
    st(&mine().request, 0);
    return 0;
  }

  void publish_best()
  {
    st(&mine().best, best);
  }

  int adopt_best()
  {
    int g = best;
    for (int i = 0; i < D; i++)
    {
      const int b = ld(&S[i].best);
      if (b < g)
        g = b;
    }
    if (g < best)
    {
      best = g;
      return ops->set_best(ops->self, g);
    }
    return 0;
  }

  bool all_idle()
  {
    const int e1 = ld(&H->epoch);
    for (int i = 0; i < D; i++)
      if (ld(&S[i].state) != ST_IDLE)
        return false;
    return ld(&H->epoch) == e1;
  }

  /* one round over the other workers in random order; returns 1 when work was taken over */
  int hunt(int *got)
  {
    *got = 0;
    __atomic_add_fetch(&H->epoch, 1, __ATOMIC_SEQ_CST);
    st(&mine().state, ST_HUNTING);
    __atomic_add_fetch(&H->epoch, 1, __ATOMIC_SEQ_CST);
    int order[64];
    int nv = 0;
    for (int i = 0; i < D; i++)
      if (i != me)
        order[nv++] = i;
    for (int i = nv - 1; i > 0; i--) /* Fisher-Yates, like permute() in common/util.c:40-58 */
    {
      rng = rng * 1664525u + 1013904223u;
      const int j = (int)((rng >> 8) % (unsigned)(i + 1));
      const int t = order[i];
      order[i] = order[j];
      order[j] = t;
    }
    for (int vi = 0; vi < nv && !ld(&H->finished); vi++)
    {
      const int v = order[vi];
      if (ld(&S[v].state) != ST_BUSY || ldll(&S[v].pool_size) < 2 * m_keep)
        continue;
      mine().steals++;
      st(&mine().resp, RESP_NONE);
      int expect = 0;
      if (!__atomic_compare_exchange_n(&S[v].request, &expect, me + 1, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST))
        continue; /* another thief is already waiting there */
      int r;
      while ((r = ld(&mine().resp)) == RESP_NONE)
      {
        if (ld(&H->finished) || ld(&H->failed))
          return 0;
        int rc = service_request(false); /* we have nothing: deny whoever asks us, so two hunters never wait on
                                            each other */
        if (rc)
          return rc;
        rc = service_acks();
        if (rc)
          return rc;
        nap(20);
      }
      if (r == RESP_GRANT)
      {
        st(&mine().state, ST_BUSY); /* before the copy: the chunk has a BUSY owner at every instant */
        const int vs = mine().resp_stack;
        int rc = ops->pull(ops->self, mine().resp_victim, vs, mine().resp_start, mine().resp_count);
        __atomic_add_fetch(&S[v].acks[vs], 1, __ATOMIC_SEQ_CST); /* the victim may reuse the range (even on error) */
        if (rc)
          return rc;
        mine().steals_ok++;
        *got = 1;
        return 0;
      }
    }
    return 0;
  }
};

} // namespace

extern "C" size_t bb_team_bytes(int world)
{
  if (world < 1 || world > 64)
    return 0;
  return sizeof(team_hdr) + sizeof(team_slot) * (size_t)world;
}

extern "C" int bb_team_init(void *mem, int world)
{
  if (!mem || world < 1 || world > 64)
    return bb_fail(BB_ERR_ARG, "bb_team_init: need a memory block and 1 <= world <= 64");
  memset(mem, 0, bb_team_bytes(world));
  team_hdr *H = static_cast<team_hdr *>(mem);
  team_slot *S = slots_of(mem);
  for (int i = 0; i < world; i++)
  {
    S[i].state = ST_BUSY;
    S[i].best = INT_MAX;
  }
  H->world = world;
  __atomic_store_n(&H->magic, TEAM_MAGIC, __ATOMIC_SEQ_CST);
  return BB_OK;
}

/* between two searches of a persistent team (every worker has left bb_team_worker and met in a barrier; one worker
 * calls this, then a second barrier): forget the finished search, keep the barrier state */
extern "C" int bb_team_reset(void *mem)
{
  team_hdr *H = static_cast<team_hdr *>(mem);
  if (!H || __atomic_load_n(&H->magic, __ATOMIC_SEQ_CST) != TEAM_MAGIC)
    return bb_fail(BB_ERR_ARG, "bb_team_reset: not a team block");
  team_slot *S = slots_of(mem);
  for (int i = 0; i < H->world; i++)
  {
    memset(&S[i], 0, sizeof(team_slot));
    S[i].state = ST_BUSY;
    S[i].best = INT_MAX;
  }
  st(&H->epoch, 0);
  st(&H->finished, 0);
  st(&H->failed, 0);
  return BB_OK;
}

extern "C" int bb_team_wait_ready(void *mem, int world, double timeout_s)
{
  team_hdr *H = static_cast<team_hdr *>(mem);
  const double t0 = now_ms();
  while (__atomic_load_n(&H->magic, __ATOMIC_SEQ_CST) != TEAM_MAGIC || ld(&H->world) != world)
  {
    if (now_ms() - t0 > timeout_s * 1e3)
      return bb_fail(BB_ERR_STATE, "bb_team_wait_ready: the team block was not initialised within %.0f s", timeout_s);
    nap(100);
  }
  return BB_OK;
}

/* sense-reversing barrier over the team block (threads or processes) */
extern "C" int bb_team_barrier(void *mem, double timeout_s)
{
  team_hdr *H = static_cast<team_hdr *>(mem);
  const int sense = ld(&H->bar_sense);
  if (__atomic_add_fetch(&H->bar_count, 1, __ATOMIC_SEQ_CST) == H->world)
  {
    st(&H->bar_count, 0);
    st(&H->bar_sense, sense ^ 1);
    return BB_OK;
  }
  const double t0 = now_ms();
  while (ld(&H->bar_sense) == sense)
  {
    if (ld(&H->failed))
      return bb_fail(BB_ERR_STATE, "bb_team_barrier: another worker failed");
    if (now_ms() - t0 > timeout_s * 1e3)
    {
      st(&H->failed, 1);
      return bb_fail(BB_ERR_STATE, "bb_team_barrier: timed out after %.0f s", timeout_s);
    }
    nap(50);
  }
  return BB_OK;
}

extern "C" void bb_team_fail(void *mem) { st(&static_cast<team_hdr *>(mem)->failed, 1); }

extern "C" int bb_team_worker(void *mem, int me, const bb_team_ops *ops, int ws, long long m_keep, long long max_steal,
                              int best0, int iters_per_poll, bb_team_result *out)
{
  team_hdr *H = static_cast<team_hdr *>(mem);
  if (!H || !ops || !out || __atomic_load_n(&H->magic, __ATOMIC_SEQ_CST) != TEAM_MAGIC || me < 0 || me >= H->world)
    return bb_fail(BB_ERR_ARG, "bb_team_worker: bad team block or worker id");
  worker w;
  memset(&w, 0, sizeof(w));
  w.H = H;
  w.S = slots_of(mem);
  w.me = me;
  w.D = H->world;
  w.ops = ops;
  w.m_keep = m_keep > 0 ? m_keep : 1;
  w.max_steal = max_steal > 0 ? max_steal : 1;
  w.best = best0;
  w.rng = 2654435761u * (unsigned)(me + 1) + (unsigned)getpid();
  team_slot &me_s = w.mine();
  const int D = w.D;
  int rc = 0;
  long long k = iters_per_poll > 0 ? iters_per_poll : 32;
  const long long kmax = k;
  long long pool = -1;
  const bool trace = getenv("BB_TRACE_TEAM") != nullptr; /* per-worker time split on stderr */
  double t_run = 0, t_serve = 0, t_hunt = 0, t_idle = 0;
  long long n_run = 0, n_hunt = 0, k_sum = 0;
  const double t_begin = now_ms();

  while (!ld(&H->finished) && !ld(&H->failed))
  {
    if (pool != 0)
    {
      st(&me_s.state, ST_BUSY);
      int b = w.best;
      double ms = 0;
      const double t0 = now_ms();
      rc = ops->run(ops->self, k, &pool, &b, &ms);
      if (rc)
        break;
      const double dt = now_ms() - t0;
      t_run += dt;
      n_run++;
      k_sum += k;
      me_s.busy_ms += ms;
      if (D > 1)
      { /* poll often enough to answer thieves within about a millisecond */
        if (dt > 1.5 && k > 1)
          k /= 2;
        else if (dt < 0.4 && k < kmax)
          k *= 2;
      }
      if (b < w.best)
        w.best = b;
      stll(&me_s.pool_size, pool);
      w.publish_best();
      const double s0 = now_ms();
      if ((rc = w.adopt_best()) || (rc = w.service_acks()) || (rc = w.service_request(ws != 0)))
        break;
      t_serve += now_ms() - s0;
      continue;
    }
    /* our pool is empty */
    stll(&me_s.pool_size, 0);
    if ((rc = w.service_acks()) || (rc = w.service_request(false)))
      break;
    if (w.donation_pending)
    { /* the engine still owes the answer to a thief: do not hunt or go idle before it is given */
      nap(10);
      continue;
    }
    if (w.outstanding[0] + w.outstanding[1] > 0)
    { /* a thief is still copying out of our memory: stay BUSY until it acknowledges */
      nap(20);
      continue;
    }
    int got = 0;
    if (ws && D > 1)
    {
      const double h0 = now_ms();
      rc = w.hunt(&got);
      t_hunt += now_ms() - h0;
      n_hunt++;
      if (rc)
        break;
      if (got)
      {
        pool = -1; /* unknown but non-zero: run */
        continue;
      }
    }
    st(&me_s.state, ST_IDLE);
    if (w.all_idle())
    {
      st(&H->finished, 1);
      break;
    }
    const double i0 = now_ms();
    nap(D > 1 ? 50 : 1);
    t_idle += now_ms() - i0;
    if ((rc = w.adopt_best()))
      break;
  }
  if (trace)
    fprintf(stderr, "[bb team %d] total %.1f ms: run %.1f ms in %lld calls (avg k %.1f, device busy %.1f ms), serve %.1f, "
                    "hunt %.1f in %lld rounds, idle naps %.1f; steals %llu/%llu donated %llu\n", me, now_ms() - t_begin,
            t_run, n_run, n_run ? (double)k_sum / n_run : 0.0, me_s.busy_ms, t_serve, t_hunt, n_hunt, t_idle,
            me_s.steals_ok, me_s.steals, me_s.donated);
  if (rc)
    st(&H->failed, 1);
  /* leave nobody waiting on us, and keep our memory until every copy out of it is acknowledged */
  const double t_exit = now_ms();
  for (;;)
  {
    w.service_request(false);
    w.service_acks();
    if ((w.outstanding[0] + w.outstanding[1] == 0 && !w.donation_pending) || ld(&H->failed) ||
        now_ms() - t_exit > 30e3)
      break;
    nap(20);
  }
  st(&me_s.state, ST_IDLE);
  me_s.rc = rc;
  out->best = w.best;
  out->steals = me_s.steals;
  out->steals_ok = me_s.steals_ok;
  out->donated = me_s.donated;
  out->busy_ms = me_s.busy_ms;
  if (!rc && ld(&H->failed))
    return bb_fail(BB_ERR_STATE, "bb_team_worker: another worker of the team failed");
  return rc;
}

extern "C" int bb_team_best(void *mem)
{
  team_hdr *H = static_cast<team_hdr *>(mem);
  team_slot *S = slots_of(mem);
  int g = INT_MAX;
  for (int i = 0; i < H->world; i++)
  {
    const int b = ld(&S[i].best);
    if (b < g)
      g = b;
  }
  return g;
}
