/*
 * oracle_nqueens.c - CPU restatement of the reference N-Queens path.
 * TEST INFRASTRUCTURE ONLY (see oracle.h). Parity: PINNED against oracle/_ref/libref_nqueens.so, nqueens_c.out
 * and OEIS A000170 solution counts (tests/golden/nqueens.json).
 */
#include "oracle.h"

#include <stdlib.h>
#include <string.h>

#define NQ_REC (1 + ORC_MAX_QUEENS) /* flat node: depth, board[MAX_QUEENS] = the reference's 21-byte Node */

/* nqueens/nqueens_c.c:80-96. The board is a permutation, so only the two diagonals can clash. The comparisons
 * are made on int-promoted values (row_pos - d may be negative and then never matches a uint8 row). The check is
 * repeated G times (-g, "safety checks per evaluation"); the result does not depend on G. */
int orc_nq_is_safe(int G, const uint8_t *board, int queen_num, int row_pos)
{
  int safe = 1;
  for (int g = 0; g < G; g++)
    for (int i = 0; i < queen_num; i++)
    {
      const int other = board[i], d = queen_num - i;
      if (other == row_pos - d || other == row_pos + d)
        safe = 0;
    }
  return safe;
}

/* kernel K4, nqueens/nqueens_multigpu_cuda.cu:158-186: one label per (parent, slot k >= depth) */
void orc_nq_labels(int N, int G, const uint8_t *parents, int size, uint8_t *labels)
{
  for (int i = 0; i < size; i++)
  {
    const uint8_t *nd = parents + (size_t)i * NQ_REC;
    const int depth = nd[0];
    for (int k = depth; k < N; k++)
      labels[k + (size_t)i * N] = (uint8_t)orc_nq_is_safe(G, nd + 1, depth, nd[1 + k]);
  }
}

/* host branch step, nqueens_multigpu_cuda.cu:189-210: a parent at depth N is a solution; every safe slot
 * becomes a child (leaves at depth N are pushed and counted in `tree`). */
int orc_nq_generate_children(int N, const uint8_t *parents, int size, const uint8_t *labels,
                             unsigned long long *tree, unsigned long long *sol, uint8_t *children)
{
  int out = 0;
  for (int i = 0; i < size; i++)
  {
    const uint8_t *nd = parents + (size_t)i * NQ_REC;
    const int depth = nd[0];
    if (depth == N)
      (*sol)++;
    for (int j = depth; j < N; j++)
      if (labels[j + (size_t)i * N] == 1)
      {
        uint8_t *ch = children + (size_t)out * NQ_REC;
        memcpy(ch, nd, NQ_REC);
        ch[0] = (uint8_t)(depth + 1);
        uint8_t tmp = ch[1 + depth];
        ch[1 + depth] = ch[1 + j];
        ch[1 + j] = tmp;
        out++;
        (*tree)++;
      }
  }
  return out;
}

/* sequential DFS, nqueens_c.c:99-117 + :134-140 */
void orc_nq_search(int N, int G, long long budget, orc_run_stats *out)
{
  memset(out, 0, sizeof(*out));
  size_t cap = 1024, size = 0;
  uint8_t *pool = malloc(cap * NQ_REC);
  uint8_t nd[NQ_REC];
  memset(nd, 0, sizeof(nd));
  for (int i = 0; i < N; i++)
    nd[1 + i] = (uint8_t)i; /* NQueens_node.c:7-13 */
  memcpy(pool, nd, NQ_REC);
  size = 1;
  while (size > 0)
  {
    size--;
    memcpy(nd, pool + size * NQ_REC, NQ_REC);
    const int depth = nd[0];
    out->parents++;
    if (depth < 64)
      out->depth_hist[depth]++;
    if (depth == N)
      out->sol++;
    for (int j = depth; j < N; j++)
    {
      out->evals++;
      if (!orc_nq_is_safe(G, nd + 1, depth, nd[1 + j]))
        continue;
      if (size + 1 >= cap)
      {
        cap *= 2;
        pool = realloc(pool, cap * NQ_REC);
      }
      uint8_t *ch = pool + size * NQ_REC;
      memcpy(ch, nd, NQ_REC);
      ch[0] = (uint8_t)(depth + 1);
      uint8_t tmp = ch[1 + depth];
      ch[1 + depth] = ch[1 + j];
      ch[1 + j] = tmp;
      size++;
      out->tree++;
    }
    if ((int)size > out->max_pool)
      out->max_pool = (int)size;
    if (budget > 0 && (long long)out->parents >= budget)
      break;
  }
  free(pool);
}
