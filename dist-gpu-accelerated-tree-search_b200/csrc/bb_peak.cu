/*
 * bb_peak.cu - measured integer-pipe ceiling for the LB2 roofline ("compare against a measured
 * IADD3/VIMNMX micro-benchmark peak on the box", because an integer-pipe peak is not a datasheet number).
 *
 * The unit is the LB2 inner step itself: one "slot step" = test one job bit of the child's unscheduled mask
 * (LOP3 -> predicate) + one predicated fused add-max (VIADDMNMX) on one lane - exactly the two instructions the
 * expand kernels issue per (child, machine pair, Johnson slot). The kernel below issues nothing else: four
 * independent chains per thread, register operands, every SM full of warps, so its rate is what the ALU pipe can
 * sustain for this instruction pair. bench.py reports LB2 as slot steps per second against this number.
 */
#include "bb_internal.cuh"

namespace
{

__global__ void __launch_bounds__(256) k_slot_step_peak(int *out, unsigned seed, int a_in, int b_in, int iters)
{
  unsigned u0 = seed ^ (threadIdx.x * 2654435761u), u1 = u0 * 3u + 1u, u2 = u1 * 5u + 7u, u3 = u2 * 9u + 11u;
  int d0 = threadIdx.x, d1 = d0 + 1, d2 = d0 + 2, d3 = d0 + 3;
  /* job bits and table values come out of shared memory with vector loads, as in the real kernel */
  __shared__ uint4 sbit[2];
  __shared__ int4 sa[2], sb[2];
  if (threadIdx.x < 2)
  {
    const unsigned t = threadIdx.x * 4;
    sbit[threadIdx.x] = make_uint4(1u << (t + 1), 1u << (t + 9), 1u << (t + 17), 1u << (t + 25));
    sa[threadIdx.x] = make_int4(a_in + t, a_in - t, a_in + 2 * t, a_in + 1);
    sb[threadIdx.x] = make_int4(b_in - t, b_in + t, b_in - 2 * t, b_in - 1);
  }
  __syncthreads();
  for (int i = 0; i < iters; i++)
  {
#pragma unroll
    for (int h = 0; h < 2; h++)
    {
      const int row = (h + i) & 1; /* loop-dependent row: the loads stay inside the loop */
      const uint4 bt = sbit[row];
      const int4 a = sa[row];
      const int4 b = sb[row];
      if (u0 & bt.x) d0 = __viaddmax_s32(d0, a.x, b.x);
      if (u1 & bt.x) d1 = __viaddmax_s32(d1, a.x, b.x);
      if (u2 & bt.x) d2 = __viaddmax_s32(d2, a.x, b.x);
      if (u3 & bt.x) d3 = __viaddmax_s32(d3, a.x, b.x);
      if (u0 & bt.y) d0 = __viaddmax_s32(d0, a.y, b.y);
      if (u1 & bt.y) d1 = __viaddmax_s32(d1, a.y, b.y);
      if (u2 & bt.y) d2 = __viaddmax_s32(d2, a.y, b.y);
      if (u3 & bt.y) d3 = __viaddmax_s32(d3, a.y, b.y);
      if (u0 & bt.z) d0 = __viaddmax_s32(d0, a.z, b.z);
      if (u1 & bt.z) d1 = __viaddmax_s32(d1, a.z, b.z);
      if (u2 & bt.z) d2 = __viaddmax_s32(d2, a.z, b.z);
      if (u3 & bt.z) d3 = __viaddmax_s32(d3, a.z, b.z);
      if (u0 & bt.w) d0 = __viaddmax_s32(d0, a.w, b.w);
      if (u1 & bt.w) d1 = __viaddmax_s32(d1, a.w, b.w);
      if (u2 & bt.w) d2 = __viaddmax_s32(d2, a.w, b.w);
      if (u3 & bt.w) d3 = __viaddmax_s32(d3, a.w, b.w);
    }
    /* new masks every iteration so that none of the 32 tests can be hoisted (4 extra ALU ops per 64) */
    u0 = (u0 << 1) | (u0 >> 31);
    u1 = (u1 << 3) | (u1 >> 29);
    u2 = (u2 << 5) | (u2 >> 27);
    u3 = (u3 << 7) | (u3 >> 25);
  }
  if (d0 + d1 + d2 + d3 == 0x7fffffff)
    out[0] = d0;
}

} // namespace

extern "C" int bb_measure_slot_step_peak(int device, double *slot_steps_per_s)
{
  if (!slot_steps_per_s)
    return bb_fail(BB_ERR_ARG, "bb_measure_slot_step_peak: null output");
  BB_CUDA_TRY(cudaSetDevice(device));
  int sms = 0;
  BB_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  int *out = nullptr;
  BB_CUDA_TRY(cudaMalloc(&out, sizeof(int)));
  cudaEvent_t e0, e1;
  BB_CUDA_TRY(cudaEventCreate(&e0));
  BB_CUDA_TRY(cudaEventCreate(&e1));
  const int grid = sms * 8, iters = 4096;
  double best = 0;
  for (int rep = 0; rep < 6; rep++)
  {
    BB_CUDA_TRY(cudaEventRecord(e0, 0));
    k_slot_step_peak<<<grid, 256>>>(out, 0x9e3779b9u + rep, 3, -5, iters);
    BB_CUDA_TRY(cudaEventRecord(e1, 0));
    BB_CUDA_TRY(cudaEventSynchronize(e1));
    float ms = 0;
    BB_CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
    const double rate = (double)grid * 256.0 * iters * 32.0 / (ms * 1e-3);
    if (rep > 0 && rate > best)
      best = rate;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  *slot_steps_per_s = best;
  return BB_OK;
}
