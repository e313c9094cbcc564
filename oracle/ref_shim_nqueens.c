/*
 * TEST INFRASTRUCTURE ONLY. Harness exposing the UNMODIFIED reference N-Queens CPU code
 * ($(REF)/nqueens/nqueens_c.c, compiled in place with -Dmain=ref_nqueens_c_main) to ctypes.
 * isSafe (nqueens_c.c:80-96), decompose (:99-117) and the DFS loop (:134-140) are the reference's own.
 */
#include <stdint.h>
#include <string.h>
#include "lib/NQueens_node.h"
#include "lib/Pool.h"

uint8_t isSafe(const int G, const uint8_t *board, const uint8_t queen_num, const uint8_t row_pos);
void decompose(const int N, const int G, const Node parent, unsigned long long int *tree_loc,
               unsigned long long int *num_sol, SinglePool *pool);

int ref_nq_is_safe(int G, const uint8_t *board, int queen_num, int row_pos)
{
  return isSafe(G, board, (uint8_t)queen_num, (uint8_t)row_pos);
}

/* One reference decompose(); children as flat uint8 [count][1+MAX_QUEENS] (depth, board[]) in push order. */
int ref_nq_decompose(int N, int G, const uint8_t *board, int depth, unsigned long long *tree,
                     unsigned long long *sol, uint8_t *children_out)
{
  Node parent;
  memset(&parent, 0, sizeof(parent));
  parent.depth = (uint8_t)depth;
  memcpy(parent.board, board, N);
  SinglePool pool;
  initSinglePool(&pool);
  decompose(N, G, parent, tree, sol, &pool);
  int cnt = pool.size;
  for (int c = 0; c < cnt; c++)
  {
    Node *nd = &pool.elements[pool.front + c];
    uint8_t *o = children_out + (size_t)c * (1 + MAX_QUEENS);
    o[0] = nd->depth;
    memcpy(o + 1, nd->board, MAX_QUEENS);
  }
  deleteSinglePool(&pool);
  return cnt;
}

/* The reference DFS (nqueens_c.c:134-140) with an optional budget on popped parents. */
unsigned long long ref_nq_search(int N, int G, unsigned long long *tree, unsigned long long *sol, long long budget)
{
  Node root;
  initRoot(&root, N);
  SinglePool pool;
  initSinglePool(&pool);
  pushBack(&pool, root);
  unsigned long long parents = 0;
  while (1)
  {
    int hasWork = 0;
    Node parent = popBack(&pool, &hasWork);
    if (!hasWork)
      break;
    decompose(N, G, parent, tree, sol, &pool);
    parents++;
    if (budget > 0 && (long long)parents >= budget)
      break;
  }
  deleteSinglePool(&pool);
  return parents;
}

int ref_nq_sizeof_node(void) { return (int)sizeof(Node); }
int ref_nq_max_queens(void) { return MAX_QUEENS; }
