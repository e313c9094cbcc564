/*
 * TEST INFRASTRUCTURE ONLY. Harness that exposes the UNMODIFIED reference CPU library (compiled in place from
 * $(REF)/pfsp by oracle/Makefile into oracle/_ref/libref_pfsp.so) to ctypes. Nothing here computes a bound:
 * it only builds the reference's own structs the way pfsp/pfsp_c.c:33-42 does and forwards arrays.
 * Used to (1) validate oracle_pfsp.c, (2) generate tests/golden, (3) time the CPU baseline.
 */
#include <stdlib.h>
#include <string.h>
#include <limits.h>

#include "lib/c_taillard.h"
#include "lib/c_bound_simple.h"
#include "lib/c_bound_johnson.h"
#include "lib/PFSP_node.h"
#include "lib/PFSP_lib.h"
#include "lib/Pool_atom.h"

typedef struct
{
  int inst;
  lb1_bound_data *l1;
  lb2_bound_data *l2;
} ref_inst;

void *ref_pfsp_open(int inst)
{
  ref_inst *r = malloc(sizeof(ref_inst));
  r->inst = inst;
  int jobs = taillard_get_nb_jobs(inst);
  int machines = taillard_get_nb_machines(inst);
  r->l1 = new_bound_data(jobs, machines);
  taillard_get_processing_times(r->l1->p_times, inst);
  fill_min_heads_tails(r->l1);
  r->l2 = new_johnson_bd_data(r->l1);
  fill_machine_pairs(r->l2);
  fill_lags(r->l1->p_times, r->l2);
  fill_johnson_schedules(r->l1->p_times, r->l2);
  return r;
}

void ref_pfsp_close(void *h)
{
  ref_inst *r = h;
  free_bound_data(r->l1);
  free_johnson_bd_data(r->l2);
  free(r);
}

void ref_pfsp_dims(void *h, int *jobs, int *machines, int *pairs, int *best_ub)
{
  ref_inst *r = h;
  *jobs = r->l1->nb_jobs;
  *machines = r->l1->nb_machines;
  *pairs = r->l2->nb_machine_pairs;
  *best_ub = taillard_get_best_ub(r->inst);
}

void ref_pfsp_tables(void *h, int *p_times, int *min_heads, int *min_tails, int *lags, int *johnson,
                     int *mp1, int *mp2, int *order)
{
  ref_inst *r = h;
  int n = r->l1->nb_jobs, m = r->l1->nb_machines, P = r->l2->nb_machine_pairs;
  memcpy(p_times, r->l1->p_times, sizeof(int) * n * m);
  memcpy(min_heads, r->l1->min_heads, sizeof(int) * m);
  memcpy(min_tails, r->l1->min_tails, sizeof(int) * m);
  memcpy(lags, r->l2->lags, sizeof(int) * P * n);
  memcpy(johnson, r->l2->johnson_schedules, sizeof(int) * P * n);
  memcpy(mp1, r->l2->machine_pairs_1, sizeof(int) * P);
  memcpy(mp2, r->l2->machine_pairs_2, sizeof(int) * P);
  memcpy(order, r->l2->machine_pair_order, sizeof(int) * P);
}

int ref_lb1_bound(void *h, const int *prmu, int limit1, int limit2)
{
  ref_inst *r = h;
  return lb1_bound(r->l1, prmu, limit1, limit2);
}

void ref_lb1_children_bounds(void *h, const int *prmu, int limit1, int limit2, int *lb_begin)
{
  ref_inst *r = h;
  lb1_children_bounds(r->l1, prmu, limit1, limit2, lb_begin);
}

int ref_lb2_bound(void *h, const int *prmu, int limit1, int limit2, int best)
{
  ref_inst *r = h;
  return lb2_bound(r->l1, r->l2, prmu, limit1, limit2, best);
}

int ref_eval_solution(void *h, const int *prmu)
{
  ref_inst *r = h;
  return eval_solution(r->l1, prmu);
}

/* The reference sequential search loop (pfsp/pfsp_c.c:55-63) with an optional budget on popped parents
 * (budget <= 0: run to completion). Returns the number of parents decomposed. */
unsigned long long ref_pfsp_search(void *h, int lb, int *best, unsigned long long *tree,
                                   unsigned long long *sol, long long budget)
{
  ref_inst *r = h;
  int jobs = r->l1->nb_jobs;
  Node root;
  initRoot(&root, jobs);
  SinglePool_atom pool;
  initSinglePool_atom(&pool);
  pushBack(&pool, root);
  unsigned long long parents = 0;
  while (1)
  {
    int hasWork = 0;
    Node parent = popBackFree(&pool, &hasWork);
    if (!hasWork)
      break;
    decompose(jobs, lb, best, r->l1, r->l2, parent, tree, sol, &pool);
    parents++;
    if (budget > 0 && (long long)parents >= budget)
      break;
  }
  deleteSinglePool_atom(&pool);
  return parents;
}

/* One reference decompose() of an arbitrary parent; children returned as flat int16 [count][2+MAX_JOBS]
 * records (depth, limit1, prmu[MAX_JOBS]) in the reference's push order. */
int ref_pfsp_decompose(void *h, int lb, const short *prmu, int depth, int *best, unsigned long long *tree,
                       unsigned long long *sol, short *children_out)
{
  ref_inst *r = h;
  int jobs = r->l1->nb_jobs;
  Node parent;
  memset(&parent, 0, sizeof(parent));
  parent.depth = (int16_t)depth;
  parent.limit1 = (int16_t)(depth - 1);
  for (int i = 0; i < jobs; i++)
    parent.prmu[i] = prmu[i];
  SinglePool_atom pool;
  initSinglePool_atom(&pool);
  decompose(jobs, lb, best, r->l1, r->l2, parent, tree, sol, &pool);
  int cnt = pool.size;
  for (int c = 0; c < cnt; c++)
  {
    Node *nd = &pool.elements[pool.front + c];
    short *o = children_out + (size_t)c * (2 + MAX_JOBS);
    o[0] = nd->depth;
    o[1] = nd->limit1;
    for (int i = 0; i < MAX_JOBS; i++)
      o[2 + i] = i < jobs ? nd->prmu[i] : 0;
  }
  deleteSinglePool_atom(&pool);
  return cnt;
}

/* The reference host prune+branch (pfsp/lib/PFSP_lib.h:51-95) on flat Node arrays. */
int ref_generate_children(const short *parents_flat, int size, int jobs, int *bounds, unsigned long long *tree,
                          unsigned long long *sol, int *best, short *children_flat)
{
  Node *parents = malloc(sizeof(Node) * (size > 0 ? size : 1));
  Node *children = malloc(sizeof(Node) * (size_t)(size > 0 ? size : 1) * jobs);
  for (int i = 0; i < size; i++)
  {
    const short *p = parents_flat + (size_t)i * (2 + MAX_JOBS);
    parents[i].depth = p[0];
    parents[i].limit1 = p[1];
    for (int k = 0; k < MAX_JOBS; k++)
      parents[i].prmu[k] = p[2 + k];
  }
  int index = 0;
  generate_children(parents, children, size, jobs, bounds, tree, sol, best, &index);
  for (int c = 0; c < index; c++)
  {
    short *o = children_flat + (size_t)c * (2 + MAX_JOBS);
    o[0] = children[c].depth;
    o[1] = children[c].limit1;
    for (int k = 0; k < MAX_JOBS; k++)
      o[2 + k] = children[c].prmu[k];
  }
  free(parents);
  free(children);
  return index;
}

int ref_sizeof_node(void) { return (int)sizeof(Node); }
int ref_max_jobs(void) { return MAX_JOBS; }
int ref_max_machines(void) { return MAX_MACHINES; }
