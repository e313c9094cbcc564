/*
 * bb_pfsp_packed.cuh - second-generation PFSP expand kernel (sm_100a): two sibling children per lane, carried as
 * packed 16-bit halves through the DPX SIMD instructions (VIMNMX.U16x2, VIADDMNMX.U16x2).
 *
 * Why. ncu on the first-generation kernel (bb_pfsp_kernels.cuh; profiles/r01_ncu_full_v2_pfsp_22_0.txt) showed it
 * instruction-bound: ALU pipe 58 %, about 9.5 instructions per (child, machine) step. Every quantity of the bound -
 * front, remain, processing times, tails, the bound itself - is below 2^16 for the 20-job classes (front+remain+tail
 * <= 39*99 + tail), and two children of the same parent share the parent's front[] and remain[]. So one lane bounds
 * the children (k, k+1) of a parent together:
 *     s2  = max(t2, f2)            VIMNMX.U16x2     start of both jobs on machine i   (f2 = parent front, both halves)
 *     t2  = s2 + p2                IADD             completion                         (p2 = p[i][job_k] | p[i][job_k+1] << 16)
 *     sr2 = s2 + r2                IADD             front + remain of both children    (r2 = parent remain, both halves)
 *     lb2 = max(lb2, sr2 + tail2)  VIADDMNMX.U16x2  LB1_d (c_bound_simple.c:218-244)
 * four ALU instructions per machine for two children; the three operand packings are IMADs on the FMA pipe.
 * LB1 (machine_bound_from_parts, c_bound_simple.c:126-141) needs the running maximum of sr over the machines: note
 * that the child's front+remain on a machine IS sr (the job's time cancels), so LB1 = max_i(max_{k<=i} sr_k + tail_i)
 * - one more VIMNMX. LB2 (LB = 3) continues from the packed sr values with the subset table (see
 * bb_pfsp_kernels.cuh): the two children's table rows are interleaved with PRMT and folded with VIADDMNMX.U16x2.
 *
 * Everything else - warp tiles, pass 1 / pass 2, one cursor reservation per tile, device-side planning - is the
 * structure of the first-generation kernel, and the results are bit-identical to it.
 */
#ifndef BB_PFSP_PACKED_CUH
#define BB_PFSP_PACKED_CUH

#include "bb_pfsp_kernels.cuh"

namespace bbk
{

#ifndef BB_PACKED_MIN_CTAS
#define BB_PACKED_MIN_CTAS 2
#endif

template <int LB, int MM>
__global__ void __launch_bounds__(BB_CTA_THREADS, (LB == 3 ? 2 : BB_PACKED_MIN_CTAS))
pfsp_expand2(bb_ctrl *__restrict__ ctrl, const bb_pool pool, const bb_pfsp_dev T, const int prefilter)
{
  static_assert(LB == 0 || LB == 1 || LB == 3, "packed kernel: LB1_d, LB1, LB2 through the subset table");
  const int nbatch = ctrl->n;
  if (nbatch == 0)
    return;
  constexpr int MU = MM ? MM : BB_MAX_MACHINES;
  constexpr int MP = MM ? ((MM + 1) & ~1) : BB_MAX_MACHINES;
  constexpr int NPL = (6 + MP + 3) / 4;
  const int m = MM ? MM : T.m;
  const int n = T.n;
  const int P = T.P;
  const int cap = pool.capacity;
  const int src = ctrl->src, dst = 1 - src, base = ctrl->base, TS = ctrl->tile;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  extern __shared__ __align__(16) unsigned char smem[];
  const pfsp_smem_layout L = pfsp_layout(m, P, LB, 2);
  uint8_t *p8 = smem + L.off_p8;
  uint32_t *p2tab = reinterpret_cast<uint32_t *>(smem + L.off_p2); /* [m][20][20]: both children's times, one load */
  unsigned char *wbase = smem + L.off_warp + warp * L.warp_bytes;
  uint4 *tile_buf = reinterpret_cast<uint4 *>(wbase + L.off_tile);     /* 2 x [plane][32] */
  int *offs_s = reinterpret_cast<int *>(wbase + L.off_offs);           /* [33] exclusive lane-slot offsets */
  uint16_t *surv_s = reinterpret_cast<uint16_t *>(wbase + L.off_surv); /* survivors of the tile */

  for (int i = threadIdx.x; i < (m * BB_MAX_JOBS) / 4; i += BB_CTA_THREADS)
    reinterpret_cast<uint32_t *>(p8)[i] = reinterpret_cast<const uint32_t *>(T.p8)[i];
  for (int i = threadIdx.x; LB != 3 && i < m * BB_MAX_JOBS * BB_MAX_JOBS; i += BB_CTA_THREADS)
  {
    const int mi = i / (BB_MAX_JOBS * BB_MAX_JOBS), ab = i - mi * (BB_MAX_JOBS * BB_MAX_JOBS);
    const int a = ab / BB_MAX_JOBS, b = ab - a * BB_MAX_JOBS;
    p2tab[i] = (uint32_t)T.p8[mi * BB_MAX_JOBS + a] | ((uint32_t)T.p8[mi * BB_MAX_JOBS + b] << 16);
  }
  __syncthreads();

  const uint4 *__restrict__ splanes = pool.planes[src];
  uint4 *__restrict__ dplanes = pool.planes[dst];

  unsigned long long cnt6[6] = {0, 0, 0, 0, 0, 0}; /* tree, sol, parents, evals, lb2_evals, lb2_pairs */

  const int ntiles = (nbatch + TS - 1) / TS;
  const int tstride = gridDim.x * BB_WARPS_PER_CTA;
  /* coalesced asynchronous copy of one tile's parent planes (lane = parent) into a tile buffer: the NEXT tile is
   * in flight (LDGSTS) while the current one is bounded, so its HBM latency is off the critical path */
  auto fetch_tile = [&](int tile, uint4 *dstbuf) {
    if (tile < ntiles)
    {
      const int first = tile * TS;
      if (lane < min(TS, nbatch - first))
      {
        const size_t node = (size_t)base + first + lane;
#pragma unroll
        for (int q = 0; q < NPL; q++)
        {
          const unsigned sa = (unsigned)__cvta_generic_to_shared(dstbuf + q * 32 + lane);
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(splanes + (size_t)q * cap + node)
                       : "memory");
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  int buf = 0;
  fetch_tile(blockIdx.x * BB_WARPS_PER_CTA + warp, tile_buf);
  for (int tile = blockIdx.x * BB_WARPS_PER_CTA + warp; tile < ntiles; tile += tstride)
  {
    uint4 *tile_s = tile_buf + buf * (NPL * 32);
    fetch_tile(tile + tstride, tile_buf + (buf ^ 1) * (NPL * 32));
    buf ^= 1;
    asm volatile("cp.async.wait_group 1;" ::: "memory"); /* everything but the prefetch just issued has landed */
    __syncwarp();
    const int best = *reinterpret_cast<volatile int *>(&ctrl->best);
    const int first = tile * TS;
    const int cnt = min(TS, nbatch - first);

    int kids = 0;
    if (lane < cnt)
    {
      const uint32_t w5 = reinterpret_cast<const uint32_t *>(&tile_s[32 + lane])[1];
      kids = n - (int)(w5 & 0xffu);
    }
    const int slots = (kids + 1) >> 1; /* lane slots of this parent: two children each */
    const int incl = warp_incl_scan(slots, lane);
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    const int allkids = __reduce_add_sync(0xffffffffu, kids);
    const int myoff = incl - slots; /* first lane slot of the parent this lane loaded (every parent has >= 1) */
    offs_s[lane] = myoff;
    if (lane == 0)
    {
      offs_s[32] = total;
      cnt6[2] += cnt;
      cnt6[3] += allkids;
    }
    __syncwarp();

    /* ---- pass 1: bound two children per lane, remember the survivors ---- */
    int nsurv = 0;
    for (int c0 = 0; c0 < total; c0 += 32)
    {
      const int c = c0 + lane;
      const bool act = c < total;
      /* parent of lane slot c = (number of parents starting at or before c) - 1, from two warp votes instead of a
       * dependent binary search through shared memory */
      const unsigned before = __ballot_sync(0xffffffffu, lane < cnt && myoff <= c0);
      const int rel = myoff - c0;
      const unsigned starts = __reduce_or_sync(0xffffffffu, (lane < cnt && rel > 0 && rel < 32) ? (1u << rel) : 0u);
      const int p = act ? __popc(before) - 1 + __popc(starts & ((2u << lane) - 1u)) : 0;
      /* byte b of plane q of parent p sits at tile + q*512 + p*16 + b */
      const unsigned char *rec = reinterpret_cast<const unsigned char *>(tile_s) + p * 16;
      const uint32_t pw5 = *reinterpret_cast<const uint32_t *>(rec + 512 + 4);
      const int depth = pw5 & 0xffu;
      const int k0 = act ? depth + 2 * (c - offs_s[p]) : depth;
      const bool two = act && (k0 + 1 < n);
      const int k1 = two ? k0 + 1 : k0;
      const int job0 = rec[(k0 >> 4) * 512 + (k0 & 15)];
      const int job1 = rec[(k1 >> 4) * 512 + (k1 & 15)];
      auto rd16 = [&](int idx) -> uint32_t { /* u16 #idx of the record (front starts at #12) */
        return *reinterpret_cast<const uint16_t *>(rec + ((idx * 2) >> 4) * 512 + ((idx * 2) & 15));
      };

      const uint32_t smask = pw5 >> 8, fullmask = (1u << n) - 1u;
      const uint32_t UA = ~(smask | (1u << job0)) & fullmask, UB = ~(smask | (1u << job1)) & fullmask;
      const uint32_t *p2row = p2tab + job0 * BB_MAX_JOBS + job1;
      uint32_t sr[MU]; /* front + remain of both children per machine (LB 3 reads them back) */
      uint32_t lb2 = 0u, t2 = 0u, carry2 = 0u;
#pragma unroll
      for (int i = 0; i < MU; i++)
      {
        sr[i] = 0u;
        if (MM || i < m)
        {
          const uint32_t f2 = rd16(12 + i) * 0x10001u;
          const uint32_t r2 = rd16(12 + MP + i) * 0x10001u;
          const uint32_t p2 = LB != 3 ? p2row[i * (BB_MAX_JOBS * BB_MAX_JOBS)]
                                      : (uint32_t)p8[i * BB_MAX_JOBS + job0] +
                                            ((uint32_t)p8[i * BB_MAX_JOBS + job1] << 16);
          const uint32_t s2 = i == 0 ? f2 : __vmaxu2(t2, f2);
          t2 = s2 + p2;
          sr[i] = s2 + r2;
          if (LB == 1)
          {
            carry2 = __vmaxu2(carry2, sr[i]);
            lb2 = __viaddmax_u16x2(carry2, T.tails2[i], lb2);
          }
          else
            lb2 = __viaddmax_u16x2(sr[i], T.tails2[i], lb2);
        }
      }
      const int lbA = (int)(lb2 & 0xffffu), lbB = (int)(lb2 >> 16);

      const bool leaf = (depth + 1 == n); /* then the parent has one child: `two` is false */
      if (act && leaf)
      {
        /* PFSP_lib.h:64-71: a leaf counts as a solution and may improve the incumbent */
        cnt6[1]++;
        if (lbA < best)
          atomicMin(&ctrl->best, lbA);
      }
      bool keepA = act && !leaf && (lbA < best);
      bool keepB = two && (lbB < best);

      if (LB == 3)
      {
        /* LB2 = max(LB1_d, max_a(sr[a] + G_a(U))): the LB1_d test above is exact as a prefilter */
        const bool candA = prefilter ? keepA : (act && !leaf);
        const bool candB = prefilter ? keepB : two;
        if (candA || candB)
        {
          const uint4 *rowA = reinterpret_cast<const uint4 *>(T.g16 + (size_t)UA * T.grow);
          const uint4 *rowB = reinterpret_cast<const uint4 *>(T.g16 + (size_t)UB * T.grow);
          constexpr int NV = (MU - 1 + 7) / 8; /* uint4 loads per row */
          uint32_t v2 = lb2;
#pragma unroll
          for (int q = 0; q < NV; q++)
          {
            const uint4 ga = __ldg(rowA + q), gb = __ldg(rowB + q);
            const uint32_t xa[4] = {ga.x, ga.y, ga.z, ga.w}, xb[4] = {gb.x, gb.y, gb.z, gb.w};
#pragma unroll
            for (int h = 0; h < 4; h++)
            {
              const int a0 = q * 8 + 2 * h;
              if (a0 < MU - 1 && (MM || a0 < m - 1))
                v2 = __viaddmax_u16x2(sr[a0], __byte_perm(xa[h], xb[h], 0x5410), v2);
              if (a0 + 1 < MU - 1 && (MM || a0 + 1 < m - 1))
                v2 = __viaddmax_u16x2(sr[a0 + 1], __byte_perm(xa[h], xb[h], 0x7632), v2);
            }
          }
          keepA = candA && ((int)(v2 & 0xffffu) < best);
          keepB = candB && ((int)(v2 >> 16) < best);
          const int ne = (candA ? 1 : 0) + (candB ? 1 : 0);
          cnt6[4] += ne;
          cnt6[5] += (unsigned long long)ne * P;
        }
      }

      /* ---- prune: survivors are only recorded here; they are branched and written in pass 2 ---- */
      const unsigned ma = __ballot_sync(0xffffffffu, keepA), mb = __ballot_sync(0xffffffffu, keepB);
      const unsigned lt = (1u << lane) - 1u;
      if (keepA)
        surv_s[nsurv + __popc(ma & lt)] = (uint16_t)((p << 8) | k0);
      if (keepB)
        surv_s[nsurv + __popc(ma) + __popc(mb & lt)] = (uint16_t)((p << 8) | k1);
      nsurv += __popc(ma) + __popc(mb);
    }
    __syncwarp();

    /* ---- pass 2: one reservation on the destination cursor per tile, then every lane writes one survivor ---- */
    if (nsurv)
    {
      unsigned pos = 0;
      if (lane == 0)
      {
        pos = atomicAdd(&ctrl->tail, (unsigned)nsurv);
        cnt6[0] += nsurv;
      }
      /* the cursor value is first needed at the stores below: the child records are built while the atomic is in
       * flight */
      {
        for (int i0 = 0; i0 < nsurv; i0 += 32)
        {
          const int i = i0 + lane;
          const unsigned amask = __ballot_sync(0xffffffffu, i < nsurv); /* lane 0 is always in it */
          if (i < nsurv)
          {
            const int e = surv_s[i];
            const int p = e >> 8, k = e & 0xff;
            const unsigned char *rec = reinterpret_cast<const unsigned char *>(tile_s) + p * 16;
            const uint32_t pw5 = *reinterpret_cast<const uint32_t *>(rec + 512 + 4);
            const int depth = pw5 & 0xffu;
            const int job = rec[(k >> 4) * 512 + (k & 15)];
            const int jobd = rec[(depth >> 4) * 512 + (depth & 15)];
            auto rd16 = [&](int idx) -> int {
              return *reinterpret_cast<const uint16_t *>(rec + ((idx * 2) >> 4) * 512 + ((idx * 2) & 15));
            };
            uint32_t w[NPL * 4];
            const uint4 q0 = tile_s[p];
            w[0] = q0.x;
            w[1] = q0.y;
            w[2] = q0.z;
            w[3] = q0.w;
            w[4] = *reinterpret_cast<const uint32_t *>(rec + 512);
            const uint32_t x = (uint32_t)(job ^ jobd);
#pragma unroll
            for (int wi = 0; wi < 5; wi++)
            {
              uint32_t patch = 0;
              if (wi == (depth >> 2))
                patch ^= x << (8 * (depth & 3));
              if (wi == (k >> 2))
                patch ^= x << (8 * (k & 3));
              w[wi] ^= patch;
            }
            w[5] = (uint32_t)(depth + 1) | (((pw5 >> 8) | (1u << job)) << 8);
            int t = 0;
#pragma unroll
            for (int i2 = 0; i2 < MP; i2 += 2)
            {
              uint32_t f2 = 0, r2 = 0;
#pragma unroll
              for (int h = 0; h < 2; h++)
              {
                const int mi = i2 + h;
                if (mi < MU && (MM || mi < m))
                {
                  const int fi = rd16(12 + mi), ri = rd16(12 + MP + mi);
                  const int pji = p8[mi * BB_MAX_JOBS + job];
                  t = (mi == 0 ? fi : max(t, fi)) + pji;
                  f2 |= (uint32_t)t << (16 * h);
                  r2 |= (uint32_t)(ri - pji) << (16 * h);
                }
              }
              w[6 + i2 / 2] = f2;
              w[6 + MP / 2 + i2 / 2] = r2;
            }
#pragma unroll
            for (int wi = 6 + MP; wi < NPL * 4; wi++)
              w[wi] = 0u;
            const unsigned at = __shfl_sync(amask, pos, 0);
            if (at + (unsigned)nsurv > (unsigned)cap)
              atomicExch(&ctrl->error, 1);
            else
            {
              const size_t slot = (size_t)at + i;
#pragma unroll
              for (int q = 0; q < NPL; q++)
                dplanes[(size_t)q * cap + slot] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
            }
          }
        }
      }
    }
    __syncwarp();
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");

  finish_iteration(ctrl, cnt6, src, base, cap);
}

} // namespace bbk
#endif
