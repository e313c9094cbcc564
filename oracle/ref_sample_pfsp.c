/*
 * TEST/BENCH INFRASTRUCTURE ONLY. Bounded-sample CPU throughput of the UNMODIFIED reference PFSP path
 * (decompose() + SinglePool_atom, compiled in place from $(REF)/pfsp by oracle/Makefile) on all host threads.
 *
 * Why this exists: pfsp_omp_c.out can only run an instance to completion; ta021/LB2 takes hours on a CPU.
 * This driver follows the reference's own 3-step structure (pfsp/pfsp_omp_c.c:101-108 warm-up, :128-331 one
 * private pool per thread, depth-first pop-back + decompose) but stops every thread after `seconds` of wall time
 * and reports explored nodes / elapsed. With -u 1 the incumbent never changes, so every thread walks a piece of
 * the same deterministic tree the GPU engine walks.
 *
 * usage: ref_sample_pfsp.out <inst> <lb> <threads> <seconds>
 * prints one line: inst lb threads tree sol elapsed_s nodes_per_s finished(0/1)
 */
#include <stdio.h>
#include <stdlib.h>
#include <limits.h>
#include <omp.h>

#include "lib/c_taillard.h"
#include "lib/c_bound_simple.h"
#include "lib/c_bound_johnson.h"
#include "lib/PFSP_node.h"
#include "lib/PFSP_lib.h"
#include "lib/Pool_atom.h"

int main(int argc, char **argv)
{
  if (argc < 5)
  {
    fprintf(stderr, "usage: %s inst lb threads seconds\n", argv[0]);
    return 2;
  }
  int inst = atoi(argv[1]), lb = atoi(argv[2]), T = atoi(argv[3]);
  double seconds = atof(argv[4]);
  if (T < 1)
    T = omp_get_num_procs();
  int jobs = taillard_get_nb_jobs(inst), machines = taillard_get_nb_machines(inst);
  if (jobs > MAX_JOBS)
  {
    fprintf(stderr, "instance needs MAX_JOBS=%d\n", jobs);
    return 2;
  }
  lb1_bound_data *l1 = new_bound_data(jobs, machines);
  taillard_get_processing_times(l1->p_times, inst);
  fill_min_heads_tails(l1);
  lb2_bound_data *l2 = new_johnson_bd_data(l1);
  fill_machine_pairs(l2);
  fill_lags(l1->p_times, l2);
  fill_johnson_schedules(l1->p_times, l2);

  int best = taillard_get_best_ub(inst);
  unsigned long long tree = 0, sol = 0;
  Node root;
  initRoot(&root, jobs);
  SinglePool_atom pool;
  initSinglePool_atom(&pool);
  pushBack(&pool, root);

  double t0 = omp_get_wtime();
  const int m = 25; /* reference default -m */
  while (pool.size < T * m)
  {
    int hasWork = 0;
    Node parent = popFrontFree(&pool, &hasWork);
    if (!hasWork)
      break;
    decompose(jobs, lb, &best, l1, l2, parent, &tree, &sol, &pool);
  }
  int finished_threads = 0;
#pragma omp parallel num_threads(T) reduction(+ : tree, sol, finished_threads)
  {
    int id = omp_get_thread_num();
    SinglePool_atom loc;
    initSinglePool_atom(&loc);
    /* strided share of the warm-up pool, node i -> thread i mod T (Pool_atom.c:14-36 does the same split) */
    for (int i = id; i < pool.size; i += T)
      pushBackFree(&loc, pool.elements[pool.front + i]);
    int best_l = best;
    unsigned long long n = 0;
    int done = 0;
    while (1)
    {
      int hasWork = 0;
      Node parent = popBackFree(&loc, &hasWork);
      if (!hasWork)
      {
        done = 1;
        break;
      }
      decompose(jobs, lb, &best_l, l1, l2, parent, &tree, &sol, &loc);
      if ((++n & 1023) == 0 && omp_get_wtime() - t0 > seconds)
        break;
    }
    finished_threads += done;
    deleteSinglePool_atom(&loc);
  }
  double el = omp_get_wtime() - t0;
  printf("%d %d %d %llu %llu %.6f %.1f %d\n", inst, lb, T, tree, sol, el, (double)tree / el, finished_threads == T);
  return 0;
}
