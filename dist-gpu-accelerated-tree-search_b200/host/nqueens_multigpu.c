/*
 * nqueens_multigpu.c -> nqueens_multigpu_cuda.out : C host driver with the command line and stdout blocks of the
 * reference's nqueens/nqueens_multigpu_cuda.cu (parse_parameters :25-89, print_settings :91-98, print_results
 * :100-108, phase blocks :241-244, 345-348, 364-369), on top of the C ABI of libbbgpu.so. Unlike the reference the
 * GPUs share work dynamically (peer-memory stealing); the counts do not depend on it.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "bb_gpu.h"

static void parse_parameters(int argc, char *argv[], int *N, int *G, int *m, int *M, int *D)
{
  *N = 14, *G = 1, *m = 25, *M = 50000, *D = 1;
  int opt;
  while ((opt = getopt(argc, argv, "N:g:m:M:D:")) != -1)
  {
    const int v = optarg ? atoi(optarg) : 0;
    if (opt != '?' && v <= 0)
    {
      printf("All parameters must be positive integers.\n");
      exit(EXIT_FAILURE);
    }
    switch (opt)
    {
    case 'N': *N = v; break;
    case 'g': *G = v; break;
    case 'm': *m = v; break;
    case 'M':
      if (v < *m)
      {
        fprintf(stderr, "Error: M must be a positive integer, greater or equal to m.\n");
        exit(EXIT_FAILURE);
      }
      *M = v;
      break;
    case 'D': *D = v; break;
    default:
      fprintf(stderr, "Usage: %s -N value -g value -m value -M value -D value\n", argv[0]);
      exit(EXIT_FAILURE);
    }
  }
}

static void print_phase(const char *title, unsigned long long tree, unsigned long long sol, double t)
{
  printf("\n%s\n", title);
  printf("Size of the explored tree: %llu\n", tree);
  printf("Number of explored solutions: %llu\n", sol);
  printf("Elapsed time: %f [s]\n", t);
}

int main(int argc, char *argv[])
{
  int N, G, m, M, D;
  parse_parameters(argc, argv, &N, &G, &m, &M, &D);
  printf("\n=================================================\n");
  printf("Multi-GPU C+OpenMP+CUDA (%d GPUs)\n\n", D);
  printf("Resolution of the %d-Queens instance\n", N);
  printf("  with %d safety check(s) per evaluation\n", G);
  printf("=================================================\n");
  if (N > BB_MAX_QUEENS)
  {
    /* the reference does not check this and overflows Node.board[MAX_QUEENS] (NQueens_node.h:11) */
    fprintf(stderr, "Error: N must not exceed %d.\n", BB_MAX_QUEENS);
    return EXIT_FAILURE;
  }
  bb_config cfg;
  bb_config_default(&cfg);
  cfg.m = m;
  cfg.M = M;
  bb_search_result res;
  bb_stats per[64];
  memset(per, 0, sizeof(per));
  if (D > 64 || bb_nq_search(N, G, D, 1, &cfg, &res, per) != BB_OK)
  {
    fprintf(stderr, "Error: %s\n", D > 64 ? "too many GPUs requested" : bb_last_error());
    return EXIT_FAILURE;
  }
  print_phase("Initial search on CPU completed", res.warm_tree, res.warm_sol, res.warm_s);
  print_phase("Search on GPU completed", res.tree, res.sol, res.search_s);
  print_phase("Search on CPU completed", res.tree, res.sol, 0.0);
  printf("\nExploration terminated.\n");
  printf("\n=================================================\n");
  printf("Size of the explored tree: %llu\n", res.tree);
  printf("Number of explored solutions: %llu\n", res.sol);
  printf("Elapsed time: %.4f [s]\n", res.elapsed_s);
  printf("=================================================\n");
  return 0;
}
