/*
 * pfsp_multigpu.c -> pfsp_multigpu_cuda.out : C host driver with the command line, stdout blocks and multigpu.csv
 * row of the reference's pfsp/pfsp_multigpu_cuda.c (main :513-588, parse_parameters pfsp/lib/PFSP_lib.c:173-320,
 * print_settings / print_results :133-170, print_results_file_multi_gpu pfsp/lib/PFSP_statistic.c:61-113), on top
 * of the C ABI of libbbgpu.so. The whole search - warm-up, parallel phase, drain - runs on the GPUs; there is no
 * CPU bounding path, so -C (CPU helper threads) is accepted and reported but no helper threads are started, and -T
 * (their batch size) is parsed and ignored, as every binary of the reference ignores the flags it does not use.
 */
#include <getopt.h>
#include <limits.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "bb_gpu.h"

static void parse_parameters(int argc, char *argv[], int *inst, int *lb, int *ub, int *m, int *M, int *T, int *D,
                             int *C, int *ws, int *L, double *perc)
{
  static struct option long_options[] = {{"inst", required_argument, NULL, 'i'}, {"lb", required_argument, NULL, 'l'},
                                         {"ub", required_argument, NULL, 'u'},   {"m", required_argument, NULL, 'm'},
                                         {"M", required_argument, NULL, 'M'},    {"T", required_argument, NULL, 'T'},
                                         {"D", required_argument, NULL, 'D'},    {"C", required_argument, NULL, 'C'},
                                         {"ws", required_argument, NULL, 'w'},   {"L", required_argument, NULL, 'L'},
                                         {"perc", required_argument, NULL, 'p'}, {NULL, 0, NULL, 0}};
  /* defaults: PFSP_lib.c:175-185 */
  *inst = 14, *lb = 1, *ub = 1, *m = 25, *M = 50000, *T = 5000, *D = 1, *C = 1, *ws = 1, *L = 1, *perc = 0.5;
  int opt, idx = 0;
  while ((opt = getopt_long(argc, argv, "i:l:u:m:M:T:D:C:w:L:p:", long_options, &idx)) != -1)
  {
    const int v = optarg ? atoi(optarg) : 0;
    const char *bad = NULL;
    switch (opt)
    {
    case 'i': if (v < 1 || v > 120) bad = "unsupported Taillard's instance"; else *inst = v; break;
    case 'l': if (v < 0 || v > 2) bad = "unsupported lower bound function"; else *lb = v; break;
    case 'u': if (v != 0 && v != 1) bad = "unsupported upper bound initialization"; else *ub = v; break;
    case 'm': if (v < 1) bad = "unsupported minimal pool for GPU initialization"; else *m = v; break;
    case 'M': if (v < *m) bad = "unsupported maximal pool for GPU initialization"; else *M = v; break;
    case 'T': if (v < *m) bad = "unsupported maximal pool for CPU multi-core"; else *T = v; break;
    case 'D': if (v < 0) bad = "unsupported number of GPU(s)"; else *D = v; break;
    case 'C': if (v < 0) bad = "unsupported number of CPU Core(s)"; else *C = v; break;
    case 'w': if (v < 0 || v > 1) bad = "unsupported Intra-node Work Stealing option"; else *ws = v; break;
    case 'L': if (v < 0 || v > 1) bad = "unsupported distributed dynamic load balancing option"; else *L = v; break;
    case 'p': if (v <= 0 || v > 100) bad = "unsupported WS percentage for popFrontBulkFree"; else *perc = v / 100.0; break;
    default:
      fprintf(stderr,
              "Usage: %s --inst <value> --lb <value> --ub <value> --m <value> --M <value> --T <value> --D <value> "
              "--C <value> --w <value> --L <value> --perc <value>\n",
              argv[0]);
      exit(EXIT_FAILURE);
    }
    if (bad)
    {
      fprintf(stderr, "Error: %s\n", bad);
      exit(EXIT_FAILURE);
    }
  }
}

static void print_settings(int inst, int machines, int jobs, int ub, int lb, int D, int C, int ws)
{
  static const char *lbname[] = {"lb1_d", "lb1", "lb2"};
  printf("\n=================================================\n");
  printf("Multi-core Multi-GPU C+OpenMP+CUDA (%d GPU(s) - [%d] Multi-core - [%d] Work Stealing)\n\n", D, C, ws);
  printf("Resolution of PFSP Taillard's instance: ta%d (m = %d, n = %d)\n", inst, machines, jobs);
  printf("Initial upper bound: %s\n", ub == 0 ? "inf" : "opt");
  printf("Lower bound function: %s\n", lbname[lb]);
  printf("Branching rule: fwd\n");
  printf("=================================================\n");
}

static void print_phase(const char *title, unsigned long long tree, unsigned long long sol, double t)
{
  printf("\n%s\n", title);
  printf("Size of the explored tree: %llu\n", tree);
  printf("Number of explored solutions: %llu\n", sol);
  printf("Elapsed time: %f [s]\n", t);
}

static void print_results(int optimum, unsigned long long tree, unsigned long long sol, double timer)
{
  printf("\n=================================================\n");
  printf("Size of the explored tree: %llu\n", tree);
  printf("Number of explored solutions: %llu\n", sol);
  printf("Optimal makespan: %d\n", optimum);
  printf("Elapsed time: %.4f [s]\n", timer);
  printf("=================================================\n");
}

static void ull_array(FILE *f, const unsigned long long *a, int n)
{
  fprintf(f, "\"[");
  for (int i = 0; i < n; i++)
    fprintf(f, i + 1 < n ? "%llu," : "%llu", a[i]);
  fprintf(f, "]\",");
}

static void dbl_array(FILE *f, const double *a, int n)
{
  fprintf(f, "\"[");
  for (int i = 0; i < n; i++)
    fprintf(f, i + 1 < n ? "%.4f," : "%.4f", a[i]);
  fprintf(f, "]\",");
}

/* one row of multigpu.csv in the current directory, schema of PFSP_statistic.c:82-106 */
static void append_csv(int inst, int lb, int D, int C, int ws, int optimum, int m, int M, int T,
                       const bb_search_result *res, const bb_stats *per)
{
  FILE *f = fopen("multigpu.csv", "a");
  if (!f)
    return;
  fseek(f, 0, SEEK_END);
  if (ftell(f) == 0)
    fprintf(f, "instance_id,D,C,lower_bound,work_stealing,optimum,m,M,T,total_time,total_tree,total_sol,"
               "exp_tree_gpu,exp_sol_gpu,gen_child_gpu,steals_gpu,success_steals_gpu,termination_gpu,"
               "gpu_memcpy_time,gpu_malloc_time,gpu_kernel_time,gpu_gen_child_time,pool_ops_time,gpu_idle_time,"
               "termination_time\n");
  fprintf(f, "%d,%d,%d,%d,%d,%d,%d,%d,%d,%.4f,%llu,%llu,", inst, D, C, lb, ws, optimum, m, M, T, res->elapsed_s,
          res->tree, res->sol);
  unsigned long long u[6][64];
  double d[7][64];
  memset(u, 0, sizeof(u));
  memset(d, 0, sizeof(d));
  for (int i = 0; i < D; i++)
  {
    u[0][i] = per[i].tree;      /* explored on this GPU */
    u[1][i] = per[i].sol;
    u[2][i] = per[i].tree;      /* children generated = survivors appended on the device */
    u[3][i] = per[i].steals;
    u[4][i] = per[i].steals_ok;
    u[5][i] = 1;                /* every worker leaves through the termination protocol once */
    d[1][i] = res->setup_s;     /* gpu_malloc_time: engine creation (pool cudaMalloc + table upload) */
    d[2][i] = per[i].kernel_ms * 1e-3;
    d[5][i] = res->search_s - per[i].kernel_ms * 1e-3 > 0 ? res->search_s - per[i].kernel_ms * 1e-3 : 0;
    /* gpu_memcpy_time, gpu_gen_child_time, pool_ops_time, termination_time: no such host phases here - 0 */
  }
  for (int k = 0; k < 6; k++)
    ull_array(f, u[k], D);
  for (int k = 0; k < 7; k++)
    dbl_array(f, d[k], D);
  fprintf(f, "\n");
  fclose(f);
}

int main(int argc, char *argv[])
{
  int inst, lb, ub, m, M, T, D, C, ws, LB;
  double perc;
  parse_parameters(argc, argv, &inst, &lb, &ub, &m, &M, &T, &D, &C, &ws, &LB, &perc);
  if (C < 0 || C > 1)
  {
    printf("C is set to %d. Invalid option for this version.\nChoose 0 to unable and 1 to enable multi-core. "
           "Mapping automatically done.\n", C);
    exit(1);
  }
  if (D < 1)
    D = 1; /* the reference divides by the device count and cannot run without a GPU either */
  if (D > 64)
  {
    printf("Execution Terminated. More GPU devices requested than the ones available\n");
    exit(1);
  }
  bb_pfsp_instance *ins = bb_pfsp_instance_taillard(inst);
  if (!ins)
  {
    fprintf(stderr, "Error: %s\n", bb_last_error());
    return EXIT_FAILURE;
  }
  print_settings(inst, ins->nb_machines, ins->nb_jobs, ub, lb, D, C, ws);
  if (ins->nb_jobs > BB_MAX_JOBS_WIDE || ins->nb_machines > BB_MAX_MACHINES)
  {
    /* the reference needs a rebuild with a larger MAX_JOBS for these classes (pfsp/README.md:52) and silently
     * overflows otherwise; refuse instead */
    fprintf(stderr, "Error: ta%d has %d jobs x %d machines; this build supports at most %d x %d\n", inst,
            ins->nb_jobs, ins->nb_machines, BB_MAX_JOBS_WIDE, BB_MAX_MACHINES);
    return EXIT_FAILURE;
  }
  const int optimum0 = ub == 1 ? ins->best_known : INT_MAX;
  bb_config cfg;
  bb_config_default(&cfg);
  cfg.m = m;
  cfg.M = M;
  bb_search_result res;
  bb_stats per[64];
  memset(per, 0, sizeof(per));
  const int rc = bb_pfsp_search(ins, lb, optimum0, D, ws, &cfg, &res, per);
  if (rc != BB_OK)
  {
    if (strstr(bb_last_error(), "GPUs requested"))
      printf("Execution Terminated. More GPU devices requested than the ones available\n");
    fprintf(stderr, "Error: %s\n", bb_last_error());
    return EXIT_FAILURE;
  }
  print_phase("Initial search on CPU completed", res.warm_tree, res.warm_sol, res.warm_s);
  print_phase("Search on Parallel GPU completed", res.tree, res.sol, res.search_s);
  /* the reference prints the CUMULATIVE counters in all three blocks (pfsp_multigpu_cuda.c:122-125, 478-481, 505-508);
   * there is no CPU drain step here, so this block repeats the totals with a zero phase time */
  print_phase("Final on CPU completed", res.tree, res.sol, 0.0);
  printf("\nExploration terminated.\n");
  print_results(res.best, res.tree, res.sol, res.elapsed_s);
  append_csv(inst, lb, D, C, ws, res.best, m, M, T, &res, per);
  bb_pfsp_instance_free(ins);
  return 0;
}
