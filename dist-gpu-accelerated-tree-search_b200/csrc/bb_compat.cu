/*
 * bb_compat.cu - the drop-in boundary: the exact extern "C" symbols the reference's C drivers call
 * (pfsp/lib/PFSP_gpu_lib.cuh:13-29), re-implemented for sm_100a, plus nq_evaluate_gpu for the N-Queens kernel the
 * reference keeps inside its driver translation unit.
 *
 * Inputs/outputs are the reference's own device buffers (44-byte AoS Node, int32 tables, nodeIndex/sumOffSets,
 * bounds[]), so pfsp_multigpu_cuda.c links and runs unchanged and bounds[] is bit-identical, including the value
 * LB2 returns when its early exit fires. Differences from the reference kernels: instance tables are staged in
 * shared memory once per CTA (the reference re-reads them from global memory through by-value structs), LB2 carries
 * the two-machine recurrence as a single fused add-max per job, and every launch is error-checked
 * (bb_last_cuda_error()).
 */
#include "bb_internal.cuh"

#include <cstdio>

namespace
{

constexpr int CB = 256; /* threads per CTA */

struct compat_smem
{
  int *p;     /* [m*n] */
  int *heads; /* [m] */
  int *tails; /* [m] */
  int *jsched, *lags, *m1, *m2, *order;
};

__device__ inline void stage(int *dst, const int *src, int count)
{
  for (int i = threadIdx.x; i < count; i += blockDim.x)
    dst[i] = src[i];
}

__device__ inline compat_smem carve(int *base, int n, int m, int P, bool lb2)
{
  compat_smem s;
  s.p = base;
  s.heads = s.p + n * m;
  s.tails = s.heads + m;
  s.jsched = s.tails + m;
  s.lags = s.jsched + (lb2 ? P * n : 0);
  s.m1 = s.lags + (lb2 ? P * n : 0);
  s.m2 = s.m1 + (lb2 ? P : 0);
  s.order = s.m2 + (lb2 ? P : 0);
  return s;
}

/* front[] after scheduling prmu[0..limit1] (schedule_front, bounds_gpu.cu:32-50) */
__device__ inline void front_of(const compat_smem &s, int n, int m, const int *prmu, int limit1, int *front)
{
  if (limit1 == -1)
  {
    for (int i = 0; i < m; i++)
      front[i] = s.heads[i];
    return;
  }
  for (int i = 0; i < m; i++)
    front[i] = 0;
  for (int q = 0; q <= limit1; q++)
  {
    const int job = prmu[q];
    front[0] += s.p[job];
    for (int i = 1; i < m; i++)
      front[i] = max(front[i - 1], front[i]) + s.p[i * n + job];
  }
}

__device__ inline void load_node(const Node *nd, int *prmu, int &depth, int &limit1)
{
  /* 44-byte node = 11 words; read it with 32-bit loads */
  const uint32_t *w = reinterpret_cast<const uint32_t *>(nd);
  const uint32_t w0 = w[0];
  depth = (int16_t)(w0 & 0xffff);
  limit1 = (int16_t)(w0 >> 16);
#pragma unroll
  for (int q = 0; q < BB_MAX_JOBS / 2; q++)
  {
    const uint32_t v = w[1 + q];
    prmu[2 * q] = (int16_t)(v & 0xffff);
    prmu[2 * q + 1] = (int16_t)(v >> 16);
  }
}

/* K1: one thread per child, LB1 (PFSP_gpu_lib.cu:43-65 -> bounds_gpu.cu:174-191) */
__global__ void __launch_bounds__(CB) k_compat_lb1(int n, int m, int size, const Node *parents, lb1_bound_data d,
                                                   int *bounds, const int *sumOffSets, const int *nodeIndex)
{
  extern __shared__ int sm[];
  compat_smem s = carve(sm, n, m, 0, false);
  stage(s.p, d.p_times, n * m);
  stage(s.heads, d.min_heads, m);
  stage(s.tails, d.min_tails, m);
  __syncthreads();
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= size)
    return;
  const int parentId = nodeIndex[tid];
  int prmu[BB_MAX_JOBS], depth, limit1;
  load_node(parents + parentId, prmu, depth, limit1);
  int k = tid + depth;
  if (parentId != 0)
    k -= sumOffSets[parentId - 1];
  const int t = prmu[depth];
  prmu[depth] = prmu[k];
  prmu[k] = t;
  int front[BB_MAX_MACHINES], remain[BB_MAX_MACHINES];
  front_of(s, n, m, prmu, limit1 + 1, front);
  for (int i = 0; i < m; i++)
    remain[i] = 0;
  for (int q = limit1 + 2; q < n; q++)
    for (int i = 0; i < m; i++)
      remain[i] += s.p[i * n + prmu[q]];
  int carry = front[0] + remain[0];
  int lb = carry + s.tails[0];
  for (int i = 1; i < m; i++)
  {
    carry = max(carry, front[i] + remain[i]);
    lb = max(lb, carry + s.tails[i]);
  }
  bounds[tid] = lb;
}

/* K2: one thread per parent, LB1_d for all its children (PFSP_gpu_lib.cu:73-102 -> bounds_gpu.cu:193-248) */
__global__ void __launch_bounds__(CB) k_compat_lb1_d(int n, int m, int size, const Node *parents, lb1_bound_data d,
                                                     int *bounds, const int *sumOffSets)
{
  extern __shared__ int sm[];
  compat_smem s = carve(sm, n, m, 0, false);
  stage(s.p, d.p_times, n * m);
  stage(s.heads, d.min_heads, m);
  stage(s.tails, d.min_tails, m);
  __syncthreads();
  const int parentId = blockIdx.x * blockDim.x + threadIdx.x;
  if (parentId >= size)
    return;
  int prmu[BB_MAX_JOBS], depth, limit1;
  load_node(parents + parentId, prmu, depth, limit1);
  int front[BB_MAX_MACHINES], remain[BB_MAX_MACHINES];
  front_of(s, n, m, prmu, limit1, front);
  for (int i = 0; i < m; i++)
    remain[i] = 0;
  for (int q = limit1 + 1; q < n; q++)
    for (int i = 0; i < m; i++)
      remain[i] += s.p[i * n + prmu[q]];
  int out = parentId != 0 ? sumOffSets[parentId - 1] : 0;
  for (int k = limit1 + 1; k < n; k++, out++)
  {
    const int job = prmu[k];
    int lb = front[0] + remain[0] + s.tails[0];
    int done = front[0] + s.p[job];
    for (int i = 1; i < m; i++)
    {
      const int start = max(done, front[i]);
      lb = max(lb, start + remain[i] + s.tails[i]);
      done = start + s.p[i * n + job];
    }
    bounds[out] = lb;
  }
}

/* K3: one thread per child, LB2 with the reference's ordered early exit
 * (PFSP_gpu_lib.cu:105-127 -> bounds_gpu.cu:293-316, 252-291) */
__global__ void __launch_bounds__(CB) k_compat_lb2(int n, int m, int size, int best, const Node *parents,
                                                   lb1_bound_data d1, lb2_bound_data d2, int *bounds,
                                                   const int *sumOffSets, const int *nodeIndex)
{
  extern __shared__ int sm[];
  const int P = d2.nb_machine_pairs;
  compat_smem s = carve(sm, n, m, P, true);
  stage(s.p, d1.p_times, n * m);
  stage(s.heads, d1.min_heads, m);
  stage(s.tails, d1.min_tails, m);
  stage(s.jsched, d2.johnson_schedules, P * n);
  stage(s.lags, d2.lags, P * n);
  stage(s.m1, d2.machine_pairs_1, P);
  stage(s.m2, d2.machine_pairs_2, P);
  stage(s.order, d2.machine_pair_order, P);
  __syncthreads();
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= size)
    return;
  const int parentId = nodeIndex[tid];
  int prmu[BB_MAX_JOBS], depth, limit1;
  load_node(parents + parentId, prmu, depth, limit1);
  int k = tid + depth;
  if (parentId != 0)
    k -= sumOffSets[parentId - 1];
  const int t = prmu[depth];
  prmu[depth] = prmu[k];
  prmu[k] = t;
  int front[BB_MAX_MACHINES];
  front_of(s, n, m, prmu, limit1 + 1, front);
  uint32_t fixed = 0;
  for (int q = 0; q <= limit1 + 1; q++)
    fixed |= 1u << prmu[q];
  int lb = 0;
  for (int l = 0; l < P; l++)
  {
    const int pr = s.order[l];
    const int a = s.m1[pr], b = s.m2[pr];
    int ta = front[a];
    int gap = front[b] - ta; /* t1 - t0 */
    for (int q = 0; q < n; q++)
    {
      const int job = s.jsched[pr * n + q];
      if (!((fixed >> job) & 1u))
      {
        const int pa = s.p[a * n + job], pb = s.p[b * n + job];
        ta += pa;
        gap = __viaddmax_s32(gap, pb - pa, s.lags[pr * n + job] + pb);
      }
    }
    lb = max(lb, max(ta + gap + s.tails[b], ta + s.tails[a]));
    if (lb > best)
      break;
  }
  bounds[tid] = lb;
}

/* K4 (N-Queens): one thread per (parent, slot) (nqueens_multigpu_cuda.cu:158-186) */
__global__ void __launch_bounds__(CB) k_compat_nq(int N, int G, const bb_nq_node *parents, uint8_t *labels, int size)
{
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= size)
    return;
  const int parentId = tid / N, k = tid % N;
  const uint8_t *nd = reinterpret_cast<const uint8_t *>(parents + parentId);
  const int depth = nd[0];
  if (k < depth)
    return; /* the reference leaves these slots unwritten */
  const int cand = nd[1 + k];
  int safe = 1;
  for (int i = 0; i < depth; i++)
  {
    int other = nd[1 + i];
    const int dist = depth - i;
    for (int g = 0; g < G; g++)
    {
      asm volatile("" : "+r"(other));
      safe &= (other != cand - dist) & (other != cand + dist);
    }
  }
  labels[tid] = (uint8_t)safe;
}

inline void check_launch(const char *what) { bb_note_cuda(cudaGetLastError(), what); }

} // namespace

extern "C" void evaluate_gpu(const int jobs, const int lb, const int size, const int nbBlocks, const int parentsSize,
                             int *best, const lb1_bound_data lbound1, const lb2_bound_data lbound2, Node *parents,
                             int *bounds, int *sumOffSets_d, int *nodeIndex_d)
{
  (void)nbBlocks; /* the reference's grid is for BLOCK_SIZE 512; we size our own */
  const int n = jobs, m = lbound1.nb_machines;
  if (n > BB_MAX_JOBS || m > BB_MAX_MACHINES)
  {
    bb_fail(BB_ERR_ARG, "evaluate_gpu: %d jobs x %d machines exceeds %d x %d", n, m, BB_MAX_JOBS, BB_MAX_MACHINES);
    return;
  }
  const size_t sm1 = sizeof(int) * (size_t)(n * m + 2 * m);
  switch (lb)
  {
  case 0:
    if (parentsSize > 0)
      k_compat_lb1_d<<<(parentsSize + CB - 1) / CB, CB, sm1>>>(n, m, parentsSize, parents, lbound1, bounds,
                                                               sumOffSets_d);
    break;
  case 1:
    if (size > 0)
      k_compat_lb1<<<(size + CB - 1) / CB, CB, sm1>>>(n, m, size, parents, lbound1, bounds, sumOffSets_d,
                                                      nodeIndex_d);
    break;
  case 2:
    if (size > 0)
    {
      const int P = lbound2.nb_machine_pairs;
      const size_t sm2 = sm1 + sizeof(int) * (size_t)(2 * P * n + 3 * P);
      static thread_local size_t configured = 0;
      if (sm2 > 48 * 1024 && sm2 > configured)
      {
        bb_note_cuda(cudaFuncSetAttribute(k_compat_lb2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2),
                     "cudaFuncSetAttribute(k_compat_lb2)");
        configured = sm2;
      }
      k_compat_lb2<<<(size + CB - 1) / CB, CB, sm2>>>(n, m, size, *best, parents, lbound1, lbound2, bounds,
                                                      sumOffSets_d, nodeIndex_d);
    }
    break;
  default:
    bb_fail(BB_ERR_ARG, "evaluate_gpu: unknown lower bound %d", lb);
    return;
  }
  check_launch("evaluate_gpu launch");
}

static int *to_device(const int *src, size_t count)
{
  int *d = nullptr;
  cudaError_t err = cudaMalloc((void **)&d, sizeof(int) * (count ? count : 1));
  if (err == cudaSuccess && count)
    err = cudaMemcpy(d, src, sizeof(int) * count, cudaMemcpyHostToDevice);
  bb_note_cuda(err, "table upload");
  return d;
}

extern "C" void lb1_alloc_gpu(lb1_bound_data *lbound1_d, lb1_bound_data *lbound1, int *p_times_d, int *min_heads_d,
                              int *min_tails_d, int jobs, int machines)
{
  (void)p_times_d;
  (void)min_heads_d;
  (void)min_tails_d; /* by-value in the reference too: the caller never sees them (PFSP_gpu_lib.cu:154-172) */
  lbound1_d->p_times = to_device(lbound1->p_times, (size_t)jobs * machines);
  lbound1_d->min_heads = to_device(lbound1->min_heads, machines);
  lbound1_d->min_tails = to_device(lbound1->min_tails, machines);
  lbound1_d->nb_jobs = lbound1->nb_jobs;
  lbound1_d->nb_machines = lbound1->nb_machines;
}

extern "C" void lb2_alloc_gpu(lb2_bound_data *lbound2_d, lb2_bound_data *lbound2, int *johnson_schedule_d,
                              int *lags_d, int *machine_pairs_1_d, int *machine_pairs_2_d,
                              int *machine_pair_order_d, int jobs, int machines)
{
  (void)johnson_schedule_d;
  (void)lags_d;
  (void)machine_pairs_1_d;
  (void)machine_pairs_2_d;
  (void)machine_pair_order_d;
  (void)machines;
  const size_t P = (size_t)lbound2->nb_machine_pairs;
  lbound2_d->johnson_schedules = to_device(lbound2->johnson_schedules, P * jobs);
  lbound2_d->lags = to_device(lbound2->lags, P * jobs);
  lbound2_d->machine_pairs_1 = to_device(lbound2->machine_pairs_1, P);
  lbound2_d->machine_pairs_2 = to_device(lbound2->machine_pairs_2, P);
  lbound2_d->machine_pair_order = to_device(lbound2->machine_pair_order, P);
  lbound2_d->nb_machine_pairs = lbound2->nb_machine_pairs;
  lbound2_d->nb_jobs = lbound2->nb_jobs;
  lbound2_d->nb_machines = lbound2->nb_machines;
}

/* the authors' operation / byte model (PFSP_gpu_lib.cu:213-267), used for the LB2 "useful ops" accounting */
extern "C" int P_of(int M) { return M * (M - 1) / 2; }

extern "C" int flop_lb1(int N, int M, int lim)
{
  const int front = lim * (1 + 2 * (M - 1)); /* schedule_front */
  const int remain = (N - lim) * M;          /* sum_unscheduled */
  const int bound = 2 + 4 * (M - 1);         /* machine_bound_from_parts */
  return front + remain + bound;
}

extern "C" int bytes_per_inv_lb1(int N, int M) { return 4 * (N * M + 1); }

extern "C" int flop_lb2(int N, int M, int lim)
{
  const int front = lim * (1 + 2 * (M - 1));
  const int remain = (N - lim) * M;
  const int per_pair = 4 * (N - lim) + 4;
  return front + remain + P_of(M) * per_pair;
}

extern "C" int bytes_per_inv_lb2(int N, int M, int lim) { return 4 * (N * M + 3 * P_of(M) * (N - lim) + 1); }

extern "C" void nq_evaluate_gpu(const int N, const int G, const bb_nq_node *parents_d, uint8_t *labels_d,
                                const int size)
{
  if (size <= 0)
    return;
  k_compat_nq<<<(size + CB - 1) / CB, CB>>>(N, G, parents_d, labels_d, size);
  check_launch("nq_evaluate_gpu launch");
}
