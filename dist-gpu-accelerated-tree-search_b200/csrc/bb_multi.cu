/*
 * bb_multi.cu - whole searches: one host thread per GPU around the per-GPU engines (include/bb_gpu.h,
 * bb_pfsp_search / bb_nq_search). Replaces pfsp_search() (pfsp/pfsp_multigpu_cuda.c:55-511) and nqueens_search()
 * (nqueens/nqueens_multigpu_cuda.cu:213-372).
 */
#include "bb_internal.cuh"

#include <chrono>
#include <climits>
#include <cstring>

enum
{
  KIND_PFSP = 1,
  KIND_NQ = 2
};

static double now_s()
{
  using namespace std::chrono;
  return duration<double>(steady_clock::now().time_since_epoch()).count();
}

static int search_single(int kind, const bb_pfsp_instance *ins, int lb, int best0, int N, int G,
                         const bb_config *cfg, bb_search_result *res, bb_stats *per_gpu)
{
  const double t0 = now_s();
  bb_engine *e = kind == KIND_PFSP ? bb_pfsp_engine_create(ins, lb, cfg) : bb_nq_engine_create(N, G, cfg);
  if (!e)
    return BB_ERR_CUDA;
  bb_stats st;
  memset(&st, 0, sizeof(st));
  int rc = bb_engine_reset(e, best0);
  if (!rc)
    rc = bb_engine_push_root(e);
  if (!rc)
    rc = bb_engine_run(e, 0, &st);
  bb_engine_destroy(e);
  if (rc)
    return rc;
  memset(res, 0, sizeof(*res));
  res->tree = st.tree;
  res->sol = st.sol;
  res->best = st.best;
  res->elapsed_s = now_s() - t0;
  res->kernel_ms_max = st.kernel_ms;
  if (per_gpu)
    per_gpu[0] = st;
  return BB_OK;
}

int bb_multi_search(int kind, const bb_pfsp_instance *ins, int lb, int best0, int N, int G, int D, int ws,
                    const bb_config *cfg, bb_search_result *res, bb_stats *per_gpu)
{
  (void)ws;
  if (D == 1)
    return search_single(kind, ins, lb, best0, N, G, cfg, res, per_gpu);
  return bb_fail(BB_ERR_ARG, "multi-GPU search (D=%d) is not available in this build", D);
}
