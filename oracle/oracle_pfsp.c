/*
 * oracle_pfsp.c - CPU restatement of the reference PFSP path (tables, LB1 / LB1_d / LB2, chunk protocol, search).
 * TEST INFRASTRUCTURE ONLY (see oracle.h). Parity: PINNED against oracle/_ref and tests/golden.
 * Each function cites the reference file:line whose behaviour it restates.
 */
#include "oracle.h"
#include "taillard_table.inc"

#include <limits.h>
#include <stdlib.h>
#include <string.h>

static inline int imax(int a, int b) { return a > b ? a : b; }
static inline int imin(int a, int b) { return a < b ? a : b; }

/* ------------------------------------------------------------------ Taillard instances */

/* class table, pfsp/lib/c_taillard.c:46-69 */
int orc_taillard_jobs(int inst)
{
  static const int upto[] = {30, 60, 90, 110, 120};
  static const int jobs[] = {20, 50, 100, 200, 500};
  for (int c = 0; c < 5; c++)
    if (inst <= upto[c])
      return jobs[c];
  return 500;
}

int orc_taillard_machines(int inst)
{
  /* ten instances per (jobs x machines) class */
  static const int mach[12] = {5, 10, 20, 5, 10, 20, 5, 10, 20, 10, 20, 20};
  int cls = (inst - 1) / 10;
  if (cls < 0) cls = 0;
  if (cls > 11) cls = 11;
  return mach[cls];
}

int orc_taillard_best_ub(int inst) { return orc_taillard[inst - 1][1]; } /* c_taillard.c:71-74 */
long orc_taillard_seed(int inst) { return orc_taillard[inst - 1][0]; }

/* Lehmer generator with Schrage's decomposition and a FLOAT32 division, c_taillard.c:76-88 */
static long lcg_unif(long *seed, long lo, long hi)
{
  const long mod = 2147483647L, mul = 16807L, q = 127773L, r = 2836L;
  long k = *seed / q;
  *seed = mul * (*seed % q) - k * r;
  if (*seed < 0)
    *seed += mod;
  double u = (float)*seed / (float)mod; /* single-precision quotient, widened afterwards */
  return lo + (long)(u * (double)(hi - lo + 1));
}

/* machine-major fill, c_taillard.c:90-105 */
void orc_taillard_ptimes(int inst, int *out)
{
  int n = orc_taillard_jobs(inst), m = orc_taillard_machines(inst);
  long seed = orc_taillard_seed(inst);
  for (int k = 0; k < m; k++)
    for (int j = 0; j < n; j++)
      out[k * n + j] = (int)lcg_unif(&seed, 1, 99);
}

/* ------------------------------------------------------------------ precomputed tables */

/* c_bound_simple.c:277-322 */
static void fill_heads_tails(orc_pfsp *t)
{
  const int n = t->n, m = t->m;
  for (int k = 0; k < m; k++)
  {
    t->heads[k] = INT_MAX;
    t->tails[k] = INT_MAX;
  }
  t->heads[0] = 0;
  t->tails[m - 1] = 0;
  for (int j = 0; j < n; j++)
  {
    int acc = 0; /* time job j needs to reach machine k when started alone at 0 */
    for (int k = 1; k < m; k++)
    {
      acc += t->p[(k - 1) * n + j];
      t->heads[k] = imin(t->heads[k], acc);
    }
    acc = 0; /* time job j still needs after leaving machine k */
    for (int k = m - 2; k >= 0; k--)
    {
      acc += t->p[(k + 1) * n + j];
      t->tails[k] = imin(t->tails[k], acc);
    }
  }
}

/* c_bound_johnson.c:48-69 (LB2_FULL: all i<j pairs, lexicographic, identity order) */
static void fill_pairs(orc_pfsp *t)
{
  int c = 0;
  for (int a = 0; a < t->m - 1; a++)
    for (int b = a + 1; b < t->m; b++)
    {
      t->pair_a[c] = a;
      t->pair_b[c] = b;
      t->pair_order[c] = c;
      c++;
    }
}

/* c_bound_johnson.c:92-107: lag = work on the machines strictly between the pair */
static void fill_lags(orc_pfsp *t)
{
  for (int s = 0; s < t->P; s++)
    for (int j = 0; j < t->n; j++)
    {
      int q = 0;
      for (int k = t->pair_a[s] + 1; k < t->pair_b[s]; k++)
        q += t->p[k * t->n + j];
      t->lags[s * t->n + j] = q;
    }
}

/* Johnson's rule, c_bound_johnson.c:109-178. Keys a=p[m1]+lag, b=p[m2]+lag; set {a<b} first by ascending a,
 * then set {a>=b} by descending b. The reference sorts with libc qsort(); glibc's qsort is a merge sort for
 * arrays this small, i.e. stable, which is what the insertion sort below reproduces (tie order is checked
 * against oracle/_ref by hash in tests/test_oracle.py). */
static int johnson_before(int a1, int b1, int a2, int b2)
{
  int s1 = !(a1 < b1), s2 = !(a2 < b2);
  if (s1 != s2)
    return s1 < s2;
  return s1 == 0 ? a1 < a2 : b1 > b2;
}

static void fill_johnson(orc_pfsp *t)
{
  const int n = t->n;
  int *ka = malloc(sizeof(int) * n), *kb = malloc(sizeof(int) * n);
  for (int s = 0; s < t->P; s++)
  {
    int *ord = t->johnson + s * n;
    for (int j = 0; j < n; j++)
    {
      ka[j] = t->p[t->pair_a[s] * n + j] + t->lags[s * n + j];
      kb[j] = t->p[t->pair_b[s] * n + j] + t->lags[s * n + j];
      ord[j] = j;
    }
    for (int i = 1; i < n; i++)
    {
      int v = ord[i], k = i;
      while (k > 0 && johnson_before(ka[v], kb[v], ka[ord[k - 1]], kb[ord[k - 1]]))
      {
        ord[k] = ord[k - 1];
        k--;
      }
      ord[k] = v;
    }
  }
  free(ka);
  free(kb);
}

orc_pfsp *orc_pfsp_from_ptimes(int n, int m, const int *p)
{
  orc_pfsp *t = calloc(1, sizeof(*t));
  t->n = n;
  t->m = m;
  t->P = m * (m - 1) / 2;
  int P = t->P > 0 ? t->P : 1;
  t->p = malloc(sizeof(int) * n * m);
  memcpy(t->p, p, sizeof(int) * n * m);
  t->heads = malloc(sizeof(int) * m);
  t->tails = malloc(sizeof(int) * m);
  t->pair_a = malloc(sizeof(int) * P);
  t->pair_b = malloc(sizeof(int) * P);
  t->pair_order = malloc(sizeof(int) * P);
  t->lags = malloc(sizeof(int) * P * n);
  t->johnson = malloc(sizeof(int) * P * n);
  fill_heads_tails(t);
  fill_pairs(t);
  fill_lags(t);
  fill_johnson(t);
  return t;
}

orc_pfsp *orc_pfsp_new(int inst)
{
  int n = orc_taillard_jobs(inst), m = orc_taillard_machines(inst);
  int *p = malloc(sizeof(int) * n * m);
  orc_taillard_ptimes(inst, p);
  orc_pfsp *t = orc_pfsp_from_ptimes(n, m, p);
  free(p);
  return t;
}

void orc_pfsp_free(orc_pfsp *t)
{
  if (!t)
    return;
  free(t->p); free(t->heads); free(t->tails); free(t->pair_a); free(t->pair_b);
  free(t->pair_order); free(t->lags); free(t->johnson); free(t);
}

void orc_pfsp_tables(const orc_pfsp *t, int *p, int *heads, int *tails, int *lags, int *johnson, int *pa, int *pb,
                     int *order)
{
  memcpy(p, t->p, sizeof(int) * t->n * t->m);
  memcpy(heads, t->heads, sizeof(int) * t->m);
  memcpy(tails, t->tails, sizeof(int) * t->m);
  memcpy(lags, t->lags, sizeof(int) * t->P * t->n);
  memcpy(johnson, t->johnson, sizeof(int) * t->P * t->n);
  memcpy(pa, t->pair_a, sizeof(int) * t->P);
  memcpy(pb, t->pair_b, sizeof(int) * t->P);
  memcpy(order, t->pair_order, sizeof(int) * t->P);
}

/* FNV-1a over int32 elements as in SURVEY.md appendix A.3 */
uint64_t orc_fnv1a_i32(const int *v, int count)
{
  uint64_t h = 1469598103934665603ULL;
  for (int i = 0; i < count; i++)
  {
    h ^= (uint32_t)v[i];
    h *= 1099511628211ULL;
  }
  return h;
}

/* ------------------------------------------------------------------ one-machine bound pieces */

/* append one job to the front schedule: completion times per machine, c_bound_simple.c:31-38 */
static void push_front(const orc_pfsp *t, int job, int *front)
{
  front[0] += t->p[job];
  for (int k = 1; k < t->m; k++)
    front[k] = imax(front[k - 1], front[k]) + t->p[k * t->n + job];
}

/* prepend one job to the back schedule, c_bound_simple.c:40-49 */
static void push_back(const orc_pfsp *t, int job, int *back)
{
  int k = t->m - 1;
  back[k] += t->p[k * t->n + job];
  for (k = t->m - 2; k >= 0; k--)
    back[k] = imax(back[k], back[k + 1]) + t->p[k * t->n + job];
}

/* c_bound_simple.c:51-68: empty prefix -> min_heads */
static void head_times(const orc_pfsp *t, const int *prmu, int limit1, int *front)
{
  if (limit1 == -1)
  {
    memcpy(front, t->heads, sizeof(int) * t->m);
    return;
  }
  memset(front, 0, sizeof(int) * t->m);
  for (int i = 0; i <= limit1; i++)
    push_front(t, prmu[i], front);
}

/* c_bound_simple.c:70-89: empty suffix -> min_tails */
static void tail_times(const orc_pfsp *t, const int *prmu, int limit2, int *back)
{
  if (limit2 == t->n)
  {
    memcpy(back, t->tails, sizeof(int) * t->m);
    return;
  }
  memset(back, 0, sizeof(int) * t->m);
  for (int i = t->n - 1; i >= limit2; i--)
    push_back(t, prmu[i], back);
}

/* c_bound_simple.c:108-124 */
static void middle_work(const orc_pfsp *t, const int *prmu, int limit1, int limit2, int *remain)
{
  memset(remain, 0, sizeof(int) * t->m);
  for (int i = limit1 + 1; i < limit2; i++)
    for (int k = 0; k < t->m; k++)
      remain[k] += t->p[k * t->n + prmu[i]];
}

/* c_bound_simple.c:91-106 (the reference sizes its scratch by n but only uses m entries) */
int orc_makespan(const orc_pfsp *t, const int *prmu)
{
  int front[64] = {0};
  for (int i = 0; i < t->n; i++)
    push_front(t, prmu[i], front);
  return front[t->m - 1];
}

/* LB1, c_bound_simple.c:143-158 + :126-141: running max of (front+remain) carried down the machines */
int orc_lb1(const orc_pfsp *t, const int *prmu, int limit1, int limit2)
{
  int front[64], back[64], remain[64];
  head_times(t, prmu, limit1, front);
  tail_times(t, prmu, limit2, back);
  middle_work(t, prmu, limit1, limit2, remain);
  int carry = front[0] + remain[0];
  int lb = carry + back[0];
  for (int k = 1; k < t->m; k++)
  {
    carry = imax(carry, front[k] + remain[k]);
    lb = imax(lb, carry + back[k]);
  }
  return lb;
}

/* LB1_d for all children at once, c_bound_simple.c:160-216 (begin direction only) + :218-244 */
void orc_lb1_children(const orc_pfsp *t, const int *prmu, int limit1, int limit2, int *lb_begin)
{
  int front[64], back[64], remain[64];
  head_times(t, prmu, limit1, front);
  tail_times(t, prmu, limit2, back);
  middle_work(t, prmu, limit1, limit2, remain);
  memset(lb_begin, 0, sizeof(int) * t->n);
  for (int i = limit1 + 1; i < limit2; i++)
  {
    const int job = prmu[i];
    int lb = front[0] + remain[0] + back[0];
    int done = front[0] + t->p[job]; /* completion of `job` on the machine just handled */
    for (int k = 1; k < t->m; k++)
    {
      int start = imax(done, front[k]);
      lb = imax(lb, start + remain[k] + back[k]);
      done = start + t->p[k * t->n + job];
    }
    lb_begin[job] = lb;
  }
}

/* LB2, c_bound_johnson.c:239-254 + :180-237. Early exit once lb exceeds `best` (strictly). */
int orc_lb2_census(const orc_pfsp *t, const int *prmu, int limit1, int limit2, int best,
                   unsigned long long *pairs_visited, unsigned long long *inner_unsched)
{
  int front[64], back[64];
  unsigned char fixed[512];
  head_times(t, prmu, limit1, front);
  tail_times(t, prmu, limit2, back);
  memset(fixed, 0, (size_t)t->n);
  for (int i = 0; i <= limit1; i++)
    fixed[prmu[i]] = 1;
  for (int i = limit2; i < t->n; i++)
    fixed[prmu[i]] = 1;

  int lb = 0;
  for (int l = 0; l < t->P; l++)
  {
    const int s = t->pair_order[l];
    const int a = t->pair_a[s], b = t->pair_b[s];
    int ta = front[a], tb = front[b];
    for (int q = 0; q < t->n; q++)
    {
      const int job = t->johnson[s * t->n + q];
      if (fixed[job])
        continue;
      ta += t->p[a * t->n + job];
      tb = imax(tb, ta + t->lags[s * t->n + job]) + t->p[b * t->n + job];
      if (inner_unsched)
        (*inner_unsched)++;
    }
    lb = imax(lb, imax(tb + back[b], ta + back[a]));
    if (pairs_visited)
      (*pairs_visited)++;
    if (lb > best)
      break;
  }
  return lb;
}

int orc_lb2(const orc_pfsp *t, const int *prmu, int limit1, int limit2, int best)
{
  return orc_lb2_census(t, prmu, limit1, limit2, best, 0, 0);
}

/* ------------------------------------------------------------------ chunk protocol */

/* pfsp_multigpu_cuda.c:275-283 */
int orc_batch_index(const int16_t *parents, int nparents, int jobs, int *sum_offsets, int *node_index)
{
  int sum = 0;
  for (int i = 0; i < nparents; i++)
  {
    int kids = jobs - parents[(size_t)i * ORC_NODE_I16];
    for (int c = 0; c < kids; c++)
      node_index[sum + c] = i;
    sum += kids;
    sum_offsets[i] = sum;
  }
  return sum;
}

/* bounds[] exactly as the three reference kernels lay them out (PFSP_gpu_lib.cu:43-65, 73-102, 105-127):
 * child c of parent i (c = 0 .. jobs-depth-1, i.e. position k = depth + c swapped to the front) sits at
 * sumOffSets[i-1] + c. lb 1 / 2 swap prmu[depth]<->prmu[k] and bound with limit1+1; lb 0 scatters lb_begin[job]. */
void orc_evaluate_chunk(const orc_pfsp *t, int lb, const int16_t *parents, int nparents, int best, int *bounds)
{
  int base = 0;
  int prmu[ORC_MAX_JOBS], lbb[512];
  for (int i = 0; i < nparents; i++)
  {
    const int16_t *nd = parents + (size_t)i * ORC_NODE_I16;
    const int depth = nd[0], limit1 = nd[1];
    for (int q = 0; q < ORC_MAX_JOBS; q++)
      prmu[q] = nd[2 + q];
    if (lb == 0)
    {
      orc_lb1_children(t, prmu, limit1, t->n, lbb);
      for (int k = limit1 + 1; k < t->n; k++)
        bounds[base + (k - (limit1 + 1))] = lbb[prmu[k]];
    }
    else
    {
      for (int k = depth; k < t->n; k++)
      {
        int tmp = prmu[depth];
        prmu[depth] = prmu[k];
        prmu[k] = tmp;
        bounds[base + (k - depth)] = lb == 1 ? orc_lb1(t, prmu, limit1 + 1, t->n)
                                             : orc_lb2(t, prmu, limit1 + 1, t->n, best);
        tmp = prmu[depth];
        prmu[depth] = prmu[k];
        prmu[k] = tmp;
      }
    }
    base += t->n - depth;
  }
}

/* PFSP_lib.h:51-95. Leaves (depth+1 == jobs) count a solution and may improve best; inner children survive
 * iff bound < best. NB the reference narrows depth to uint8_t (:60). */
int orc_generate_children(const int16_t *parents, int nparents, int jobs, const int *bounds,
                          unsigned long long *tree, unsigned long long *sol, int *best, int16_t *children)
{
  int base = 0, out = 0;
  for (int i = 0; i < nparents; i++)
  {
    const int16_t *nd = parents + (size_t)i * ORC_NODE_I16;
    const int depth = (uint8_t)nd[0], limit1 = nd[1];
    for (int k = limit1 + 1; k < jobs; k++)
    {
      const int b = bounds[base + (k - (limit1 + 1))];
      if (depth + 1 == jobs)
      {
        (*sol)++;
        if (b < *best)
          *best = b;
      }
      else if (b < *best)
      {
        int16_t *ch = children + (size_t)out * ORC_NODE_I16;
        memcpy(ch, nd, sizeof(int16_t) * ORC_NODE_I16);
        ch[0] = (int16_t)(depth + 1);
        ch[1] = (int16_t)(limit1 + 1);
        int16_t tmp = ch[2 + depth];
        ch[2 + depth] = ch[2 + k];
        ch[2 + k] = tmp;
        out++;
        (*tree)++;
      }
    }
    base += jobs - depth;
  }
  return out;
}

/* ------------------------------------------------------------------ run level */

typedef struct
{
  int16_t *v;
  size_t cap, front, size;
} node_deque;

static void dq_push(node_deque *d, const int16_t *nd)
{
  if (d->front + d->size >= d->cap)
  {
    d->cap = d->cap ? d->cap * 2 : 1024;
    d->v = realloc(d->v, d->cap * ORC_NODE_I16 * sizeof(int16_t));
  }
  memcpy(d->v + (d->front + d->size) * ORC_NODE_I16, nd, sizeof(int16_t) * ORC_NODE_I16);
  d->size++;
}

static void root_node(int jobs, int16_t *nd) /* PFSP_node.c:7-14 */
{
  memset(nd, 0, sizeof(int16_t) * ORC_NODE_I16);
  nd[0] = 0;
  nd[1] = -1;
  for (int i = 0; i < jobs; i++)
    nd[2 + i] = (int16_t)i;
}

/* decompose one parent (PFSP_lib.c:7-129), pushing survivors; fused bound+prune+branch */
static void decompose_one(const orc_pfsp *t, int lb, const int16_t *parent, node_deque *pool, orc_run_stats *st)
{
  const int n = t->n, depth = parent[0], limit1 = parent[1];
  int prmu[ORC_MAX_JOBS], lbb[512];
  for (int q = 0; q < ORC_MAX_JOBS; q++)
    prmu[q] = parent[2 + q];
  if (lb == 0)
    orc_lb1_children(t, prmu, limit1, n, lbb);
  st->parents++;
  if (depth < 64)
    st->depth_hist[depth]++;
  for (int k = limit1 + 1; k < n; k++)
  {
    int bound;
    if (lb == 0)
      bound = lbb[prmu[k]];
    else
    {
      int tmp = prmu[depth];
      prmu[depth] = prmu[k];
      prmu[k] = tmp;
      bound = lb == 1 ? orc_lb1(t, prmu, limit1 + 1, n) : orc_lb2(t, prmu, limit1 + 1, n, st->best);
      tmp = prmu[depth];
      prmu[depth] = prmu[k];
      prmu[k] = tmp;
    }
    st->evals++;
    if (depth + 1 == n)
    {
      st->sol++;
      if (bound < st->best)
        st->best = bound;
    }
    else if (bound < st->best)
    {
      int16_t ch[ORC_NODE_I16];
      memcpy(ch, parent, sizeof(ch));
      ch[0] = (int16_t)(depth + 1);
      ch[1] = (int16_t)(limit1 + 1);
      int16_t tmp = ch[2 + depth];
      ch[2 + depth] = ch[2 + k];
      ch[2 + k] = tmp;
      dq_push(pool, ch);
      st->tree++;
    }
  }
}

void orc_pfsp_search(const orc_pfsp *t, int lb, int best0, long long budget, orc_run_stats *out)
{
  memset(out, 0, sizeof(*out));
  out->best = best0;
  node_deque pool = {0};
  int16_t nd[ORC_NODE_I16];
  root_node(t->n, nd);
  dq_push(&pool, nd);
  while (pool.size > 0)
  {
    pool.size--;
    memcpy(nd, pool.v + (pool.front + pool.size) * ORC_NODE_I16, sizeof(nd));
    decompose_one(t, lb, nd, &pool, out);
    if ((int)pool.size > out->max_pool)
      out->max_pool = (int)pool.size;
    if (budget > 0 && (long long)out->parents >= budget)
      break;
  }
  free(pool.v);
}

/* pfsp_multigpu_cuda.c:107-118 (step 1), :221-320 (step 2, GPU branch with the CPU standing in for the kernel),
 * :483-495 (step 3); popBackBulk semantics Pool_atom.c:154-178 with ratio 1. */
void orc_pfsp_search_batched(const orc_pfsp *t, int lb, int best0, int m, int M, orc_run_stats *out,
                             unsigned long long *batches)
{
  memset(out, 0, sizeof(*out));
  out->best = best0;
  *batches = 0;
  node_deque pool = {0};
  int16_t nd[ORC_NODE_I16];
  root_node(t->n, nd);
  dq_push(&pool, nd);
  while ((int)pool.size < m && pool.size > 0)
  {
    memcpy(nd, pool.v + pool.front * ORC_NODE_I16, sizeof(nd));
    pool.front++;
    pool.size--;
    decompose_one(t, lb, nd, &pool, out);
  }
  int16_t *parents = malloc(sizeof(int16_t) * ORC_NODE_I16 * (size_t)M);
  int16_t *children = malloc(sizeof(int16_t) * ORC_NODE_I16 * (size_t)M * t->n);
  int *bounds = malloc(sizeof(int) * (size_t)M * t->n);
  while ((int)pool.size >= m)
  {
    int take = imin((int)pool.size, M);
    pool.size -= take;
    memcpy(parents, pool.v + (pool.front + pool.size) * ORC_NODE_I16, sizeof(int16_t) * ORC_NODE_I16 * take);
    orc_evaluate_chunk(t, lb, parents, take, out->best, bounds);
    for (int i = 0; i < take; i++)
    {
      out->parents++;
      out->evals += t->n - parents[(size_t)i * ORC_NODE_I16];
    }
    int kids = orc_generate_children(parents, take, t->n, bounds, &out->tree, &out->sol, &out->best, children);
    for (int c = 0; c < kids; c++)
      dq_push(&pool, children + (size_t)c * ORC_NODE_I16);
    if ((int)pool.size > out->max_pool)
      out->max_pool = (int)pool.size;
    (*batches)++;
  }
  while (pool.size > 0)
  {
    pool.size--;
    memcpy(nd, pool.v + (pool.front + pool.size) * ORC_NODE_I16, sizeof(nd));
    decompose_one(t, lb, nd, &pool, out);
  }
  free(parents);
  free(children);
  free(bounds);
  free(pool.v);
}
