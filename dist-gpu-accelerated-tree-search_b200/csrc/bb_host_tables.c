/*
 * bb_host_tables.c - host-side (C, no GPU) instance generation and bound precompute for PFSP.
 *
 * Product code: what the reference computes once per run on the host before the search starts
 * (pfsp/pfsp_multigpu_cuda.c:96-105). Restated from the algorithms, bit-identical tables:
 *   Taillard LCG instance generator            c_taillard.c:46-105
 *   min_heads / min_tails                      c_bound_simple.c:277-322
 *   machine pairs (all i<j), lags              c_bound_johnson.c:48-107
 *   Johnson schedules per pair                 c_bound_johnson.c:109-178
 */
#include "bb_gpu.h"
#include "taillard_table.inc"

#include <limits.h>
#include <stdlib.h>
#include <string.h>

int bb_taillard_nb_jobs(int inst)
{
  if (inst <= 30) return 20;
  if (inst <= 60) return 50;
  if (inst <= 90) return 100;
  if (inst <= 110) return 200;
  return 500;
}

int bb_taillard_nb_machines(int inst)
{
  /* classes of ten: 20x{5,10,20} 50x{5,10,20} 100x{5,10,20} 200x{10,20} 500x20 */
  static const unsigned char per_class[12] = {5, 10, 20, 5, 10, 20, 5, 10, 20, 10, 20, 20};
  int c = (inst - 1) / 10;
  return per_class[c < 0 ? 0 : (c > 11 ? 11 : c)];
}

int bb_taillard_best_ub(int inst) { return bb_taillard_table[inst - 1][1]; }
long bb_taillard_seed(int inst) { return bb_taillard_table[inst - 1][0]; }

/* Taillard's generator: x <- 16807 x mod (2^31 - 1) by Schrage's method; the uniform in [0,1) is formed by a
 * single-precision division (that is what the reference does, and it decides the last digit of some values). */
void bb_taillard_processing_times(int inst, int *ptm)
{
  const int n = bb_taillard_nb_jobs(inst), m = bb_taillard_nb_machines(inst);
  long x = bb_taillard_seed(inst);
  for (int idx = 0; idx < n * m; idx++) /* machine-major: idx = machine * n + job */
  {
    long hi = x / 127773L;
    x = 16807L * (x - hi * 127773L) - hi * 2836L;
    if (x < 0)
      x += 2147483647L;
    float u = (float)x / (float)2147483647L;
    ptm[idx] = (int)(1 + (long)((double)u * 99.0));
  }
}

static void compute_heads_tails(bb_pfsp_instance *ins)
{
  const int n = ins->nb_jobs, m = ins->nb_machines;
  for (int k = 0; k < m; k++)
  {
    int best_head = k == 0 ? 0 : INT_MAX, best_tail = k == m - 1 ? 0 : INT_MAX;
    for (int j = 0; j < n; j++)
    {
      int before = 0, after = 0;
      for (int q = 0; q < k; q++)
        before += ins->p_times[q * n + j];
      for (int q = k + 1; q < m; q++)
        after += ins->p_times[q * n + j];
      if (k > 0 && before < best_head)
        best_head = before;
      if (k < m - 1 && after < best_tail)
        best_tail = after;
    }
    ins->min_heads[k] = best_head;
    ins->min_tails[k] = best_tail;
  }
}

/* Johnson's rule for the two-machine problem with keys a = p[m1]+lag, b = p[m2]+lag:
 * jobs with a < b first, by ascending a; then the others by descending b. Ties keep job-id order (the reference's
 * libc merge sort is stable). Position of a job = how many jobs precede it. */
static int precedes(int ai, int bi, int i, int aj, int bj, int j)
{
  const int gi = ai < bi ? 0 : 1, gj = aj < bj ? 0 : 1;
  if (gi != gj)
    return gi < gj;
  const int ki = gi == 0 ? ai : -bi, kj = gj == 0 ? aj : -bj;
  if (ki != kj)
    return ki < kj;
  return i < j;
}

static void compute_pair_tables(bb_pfsp_instance *ins)
{
  const int n = ins->nb_jobs, m = ins->nb_machines;
  int s = 0;
  int *ka = malloc(sizeof(int) * n), *kb = malloc(sizeof(int) * n);
  for (int m1 = 0; m1 + 1 < m; m1++)
    for (int m2 = m1 + 1; m2 < m; m2++, s++)
    {
      ins->machine_pairs_1[s] = m1;
      ins->machine_pairs_2[s] = m2;
      ins->machine_pair_order[s] = s;
      for (int j = 0; j < n; j++)
      {
        int between = 0;
        for (int q = m1 + 1; q < m2; q++)
          between += ins->p_times[q * n + j];
        ins->lags[s * n + j] = between;
        ka[j] = ins->p_times[m1 * n + j] + between;
        kb[j] = ins->p_times[m2 * n + j] + between;
      }
      for (int j = 0; j < n; j++)
      {
        int pos = 0;
        for (int i = 0; i < n; i++)
          if (i != j && precedes(ka[i], kb[i], i, ka[j], kb[j], j))
            pos++;
        ins->johnson_schedules[s * n + pos] = j;
      }
    }
  free(ka);
  free(kb);
}

bb_pfsp_instance *bb_pfsp_instance_from_ptimes(int nb_jobs, int nb_machines, const int *p_times)
{
  if (nb_jobs < 1 || nb_machines < 1 || !p_times)
    return NULL;
  bb_pfsp_instance *ins = calloc(1, sizeof(*ins));
  if (!ins)
    return NULL;
  const int n = nb_jobs, m = nb_machines, P = m * (m - 1) / 2, Pa = P > 0 ? P : 1;
  ins->nb_jobs = n;
  ins->nb_machines = m;
  ins->nb_machine_pairs = P;
  ins->best_known = INT_MAX;
  ins->inst = 0;
  ins->p_times = malloc(sizeof(int) * n * m);
  ins->min_heads = malloc(sizeof(int) * m);
  ins->min_tails = malloc(sizeof(int) * m);
  ins->machine_pairs_1 = malloc(sizeof(int) * Pa);
  ins->machine_pairs_2 = malloc(sizeof(int) * Pa);
  ins->machine_pair_order = malloc(sizeof(int) * Pa);
  ins->lags = malloc(sizeof(int) * Pa * n);
  ins->johnson_schedules = malloc(sizeof(int) * Pa * n);
  memcpy(ins->p_times, p_times, sizeof(int) * n * m);
  compute_heads_tails(ins);
  compute_pair_tables(ins);
  return ins;
}

bb_pfsp_instance *bb_pfsp_instance_taillard(int inst)
{
  if (inst < 1 || inst > 120)
    return NULL;
  const int n = bb_taillard_nb_jobs(inst), m = bb_taillard_nb_machines(inst);
  int *p = malloc(sizeof(int) * n * m);
  bb_taillard_processing_times(inst, p);
  bb_pfsp_instance *ins = bb_pfsp_instance_from_ptimes(n, m, p);
  free(p);
  if (ins)
  {
    ins->best_known = bb_taillard_best_ub(inst);
    ins->inst = inst;
  }
  return ins;
}

void bb_pfsp_instance_free(bb_pfsp_instance *ins)
{
  if (!ins)
    return;
  free(ins->p_times);
  free(ins->min_heads);
  free(ins->min_tails);
  free(ins->machine_pairs_1);
  free(ins->machine_pairs_2);
  free(ins->machine_pair_order);
  free(ins->lags);
  free(ins->johnson_schedules);
  free(ins);
}
