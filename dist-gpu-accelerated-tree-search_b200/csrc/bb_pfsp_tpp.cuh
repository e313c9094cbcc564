/*
 * bb_pfsp_tpp.cuh - third-generation PFSP expand kernel (sm_100a): one LANE owns one PARENT.
 *
 * Why. ncu on the second-generation kernel (bb_pfsp_packed.cuh; profiles/r01_ncu_full_v4_pfsp_21_2.txt) showed a
 * latency-bound kernel at a quarter of the warp slots: 11 issued instructions per (child pair, machine) step, four of
 * them shared-memory loads of the PARENT's front[] / remain[] (every lane looked its parent up in the warp tile
 * again and again), 115 registers, short-scoreboard 1.85 and long-scoreboard 2.14 stalls per issue. The integer ALU
 * pipe of an SM sub-partition issues one warp instruction every second cycle (B300_MICROARCH.md: alu-pipe rt = 2), so
 * what counts is ALU-pipe instructions per step. Here the lane that loaded a parent keeps that parent's front[] and
 * remain[] in REGISTERS (both 16-bit halves hold the same value) and walks its own children two at a time:
 *
 *     p2  = PRMT(row[job0], row[job1])    processing times of both children on machine i (bytes -> 16-bit halves)
 *     s2' = max(s2 + p2, f2[i+1])         VIADDMNMX.U16x2   start of both jobs on machine i+1
 *     sr2 = s2 + r2[i]                    IMAD (FMA pipe)   front + remain of both children on machine i
 *     lb2 = max(lb2, sr2 + tail2[i])      VIADDMNMX.U16x2   LB1_d (c_bound_simple.c:218-244)
 *
 * three ALU-pipe instructions and no shared-memory load per step; the only loads of an iteration are the two
 * jobs' processing-time rows (5 words each, bank-conflict free by construction). LB1 (c_bound_simple.c:126-141) adds
 * the running maximum of sr2, LB2 continues from the sr2 values with the subset table exactly as in
 * bb_pfsp_kernels.cuh (LB = 3). Parents that are neighbours in the pool are at the same depth almost always (the
 * stacks are filled generation by generation), so the lanes of a warp run the same number of child pairs.
 *
 * Pass 2 (branch + compact) is dense as before - one lane per survivor, one cursor reservation per tile - but reads
 * the parent's whole record with 16-byte shared-memory loads and its job's time row, instead of one load per value.
 *
 * Nothing waits for the cursor. The atomicAdd that reserves a tile's slots comes back after 1-2 microseconds (every
 * warp of the grid hits the same address; first capture of this kernel, profiles/r02_ncu_a_pfsp_*.txt: 12 % of the
 * stall samples of the LB2 variant and 36 % of the LB1_d variant sat on its return, another 11 % / 33 % on the load of
 * the incumbent, which shared the cursor's cache line). So a tile's pass 2 runs one tile LATE: the warp issues the
 * reservation for tile t, bounds tile t+1, and only then writes tile t's children - three tile buffers rotate
 * (pass 2 pending / being bounded / being prefetched with cp.async) and two survivor lists. The incumbent is read
 * once per launch and lives in a different cache line than the cursor (bb_ctrl).
 * The processing-time rows are staged in shared memory with one TMA bulk copy per CTA (cp.async.bulk on an mbarrier:
 * the table is contiguous and already in its shared-memory layout). The parent tiles stay on per-lane cp.async:
 * one TMA bulk copy per record plane (512 contiguous bytes in the pool and in the tile) was measured too and is
 * slower at a prefetch distance of one tile (ta022 / LB1_d 39.3 -> 41.8 ms, ta021 / LB2 425 -> 431 ms); a longer
 * distance needs a fourth tile buffer that shared memory does not have at m = 20.
 * Results are bit-identical to the earlier kernels (same child set, same counters).
 */
#ifndef BB_PFSP_TPP_CUH
#define BB_PFSP_TPP_CUH

#include "bb_pfsp_kernels.cuh"

#include <type_traits>

namespace bbk
{

struct tpp_smem_layout
{
  int off_prow, off_mbar, off_warp, warp_bytes, off_tile, off_surv, total;
};

/* machines (compile-time class MM as in pfsp_mm; 0 = runtime m <= 20) -> shared-memory layout of pfsp_expand3 */
__host__ __device__ inline tpp_smem_layout tpp_layout(int m)
{
  tpp_smem_layout L;
  const int mu = pfsp_mm(m) ? pfsp_mm(m) : BB_MAX_MACHINES;
  const int pw = (mu + 3) / 4;
  int o = 0;
  L.off_prow = o; /* [pw][BB_MAX_JOBS] u32: word q of job j = times of j on machines 4q..4q+3, one byte each */
  o += (pw * BB_MAX_JOBS * 4 + 15) & ~15;
  L.off_mbar = o; /* mbarrier of the TMA bulk copy that stages the rows */
  o += 16;
  L.off_warp = o;
  int w = 0;
  L.off_tile = w; /* 3 x [plane][32] uint4: pass 2 pending, being bounded, being prefetched */
  w += 3 * pfsp_planes(m) * BB_TILE_MAX * 16;
  L.off_surv = w; /* 2 x survivors of a tile: (parent lane << 8) | position in prmu */
  w += 2 * BB_TILE_MAX * BB_MAX_JOBS * 2;
  L.warp_bytes = (w + 15) & ~15;
  L.total = o + L.warp_bytes * BB_WARPS_PER_CTA;
  return L;
}

/* prmt.b32 in its default mode: a selector nibble with bit 3 set replicates the SIGN of the chosen byte, which is
 * 0x00 for processing times <= 127 - that is how the two zero bytes of (p0 | p1 << 16) are made in one instruction */
template <unsigned SEL>
__device__ __forceinline__ uint32_t prmt_c(uint32_t a, uint32_t b)
{
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "n"(SEL));
  return d;
}

/* byte B of a and byte B of b -> 16-bit halves */
template <int B>
__device__ __forceinline__ uint32_t pair_bytes(uint32_t a, uint32_t b)
{
  return prmt_c<(unsigned)(B | ((8 + B) << 4) | ((4 + B) << 8) | ((12 + B) << 12))>(a, b);
}

/* a * one + b with `one` a run-time 1: an IMAD, i.e. an add issued on the FMA pipe instead of the ALU pipe */
__device__ __forceinline__ uint32_t add_fma(uint32_t a, uint32_t one, uint32_t b)
{
  uint32_t d;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b));
  return d;
}

template <int LB, int MM>
__global__ void __launch_bounds__(BB_CTA_THREADS, ((LB == 3 || MM == 20 || MM == 0) ? 2 : 3))
pfsp_expand3(bb_ctrl *__restrict__ ctrl, const bb_pool pool, const bb_pfsp_dev T, const int prefilter)
{
  static_assert(LB == 0 || LB == 1 || LB == 3, "lane-per-parent kernel: LB1_d, LB1, LB2 through the subset table");
  constexpr int MU = MM ? MM : BB_MAX_MACHINES;
  constexpr int MP = MM ? ((MM + 1) & ~1) : BB_MAX_MACHINES;
  constexpr int NPL = (6 + MP + 3) / 4;
  constexpr int PW = (MU + 3) / 4;
  constexpr int FW = 6, RW = 6 + MP / 2; /* first record word of front[] / remain[] (two machines per word) */
  const int m = MM ? MM : T.m;
  const int n = T.n;
  const int P = T.P;
  const int cap = pool.capacity;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t one = T.one;

  extern __shared__ __align__(16) unsigned char smem[];
  const tpp_smem_layout L = tpp_layout(m);
  uint32_t *prow = reinterpret_cast<uint32_t *>(smem + L.off_prow);
  unsigned char *wbase = smem + L.off_warp + warp * L.warp_bytes;
  uint4 *tile_buf = reinterpret_cast<uint4 *>(wbase + L.off_tile);

  /* ---- stage the processing-time rows: ONE TMA bulk copy per CTA (cp.async.bulk, 1-D: the table is contiguous and
   * already in its shared-memory layout), completion on an mbarrier ---- */
  {
    const unsigned mbar = (unsigned)__cvta_generic_to_shared(smem + L.off_mbar);
    constexpr unsigned bytes = PW * BB_MAX_JOBS * 4;
    static_assert(bytes % 16 == 0, "bulk copies move multiples of 16 bytes");
    if (threadIdx.x == 0)
    {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar) : "memory");
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                       (unsigned)__cvta_generic_to_shared(prow)),
                   "l"(T.prow), "r"(bytes), "r"(mbar)
                   : "memory");
    }
    __syncthreads(); /* the initialised barrier is visible to every thread before it waits */
    unsigned done = 0;
    for (int spin = 0; spin < (1 << 22) && !done; spin++) /* bounded: a failed copy raises the error flag, not a hang */
      asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}"
                   : "=r"(done)
                   : "r"(mbar)
                   : "memory");
    if (!done && threadIdx.x == 0)
      atomicExch(&ctrl->error, 1);
  }

  /* Everything above touches only this launch's shared memory and a table that never changes, so under
   * programmatic dependent launch it runs while the previous iteration is still finishing. The plan and the pool
   * are the previous iteration's output: wait for it, then let the next launch start its own prologue as soon as
   * SM resources free up. (Both are no-ops for a launch without the attribute.) */
  pdl_wait_then_release();
  const int nbatch = ctrl->n;
  if (nbatch == 0)
    return;
  const int src = ctrl->src, dst = 1 - src, base = ctrl->base, TS = ctrl->tile;

  const uint4 *__restrict__ splanes = pool.planes[src];
  uint4 *__restrict__ dplanes = pool.planes[dst];
  const uint32_t fullmask = (n >= 32) ? 0xffffffffu : ((1u << n) - 1u);

  unsigned long long cnt6[6] = {0, 0, 0, 0, 0, 0}; /* tree, sol, parents, evals, lb2_evals, lb2_pairs */
  const unsigned char *g16_bytes = reinterpret_cast<const unsigned char *>(T.g16);
  const unsigned grow_bytes = (unsigned)T.grow * 2u;
  unsigned nlb2 = 0u; /* this lane's table evaluations of the launch (a 64-bit counter in the pair loop cost 7 % of its instructions) */

  const int ntiles = (nbatch + TS - 1) / TS;
  const int tstride = gridDim.x * BB_WARPS_PER_CTA;
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
  /* pass 2 of one tile: every lane writes one survivor at the reserved position */
  auto write_children = [&](const uint4 *tile_s, const uint16_t *surv, int nsurv, unsigned pos) {
    const unsigned char *tile_b = reinterpret_cast<const unsigned char *>(tile_s);
    const unsigned at = __shfl_sync(0xffffffffu, pos, 0);
    if (at + (unsigned)nsurv > (unsigned)cap)
    {
      if (lane == 0)
        atomicExch(&ctrl->error, 1);
      return;
    }
    for (int i = lane; i < nsurv; i += 32)
    {
      const int e = surv[i];
      const int p = e >> 8, k = e & 0xff;
      uint32_t W[NPL * 4];
#pragma unroll
      for (int q = 0; q < NPL; q++)
      {
        const uint4 v = tile_s[q * 32 + p];
        W[4 * q] = v.x;
        W[4 * q + 1] = v.y;
        W[4 * q + 2] = v.z;
        W[4 * q + 3] = v.w;
      }
      const int pdepth = (int)(W[5] & 0xffu);
      const int job = tile_b[(k >> 4) * 512 + p * 16 + (k & 15)];
      const int jobd = tile_b[(pdepth >> 4) * 512 + p * 16 + (pdepth & 15)];
      uint32_t pw[PW];
#pragma unroll
      for (int q = 0; q < PW; q++)
        pw[q] = prow[q * BB_MAX_JOBS + job];
      /* prmu: swap positions depth and k (PFSP_lib.h:80-82) */
      const uint32_t x = (uint32_t)(job ^ jobd);
#pragma unroll
      for (int wi = 0; wi < 5; wi++)
      {
        uint32_t patch = 0;
        if (wi == (pdepth >> 2))
          patch ^= x << (8 * (pdepth & 3));
        if (wi == (k >> 2))
          patch ^= x << (8 * (k & 3));
        W[wi] ^= patch;
      }
      W[5] = (uint32_t)(pdepth + 1) | (((W[5] >> 8) | (1u << job)) << 8);
      /* child front (completion times of the job after the parent's prefix) and remain (minus the job) */
      uint32_t fo[MP / 2], ro[MP / 2];
      {
        int s = (int)(W[FW] & 0xffffu), tprev = 0;
#pragma unroll
        for (int i = 0; i < MP; i++)
        {
          int t = 0;
          if (i < MU && (MM || i < m))
          {
            const uint32_t word = pw[i >> 2];
            const int pj = (int)((i & 3) == 0   ? prmt_c<0x4440>(word, 0u)
                                 : (i & 3) == 1 ? prmt_c<0x4441>(word, 0u)
                                 : (i & 3) == 2 ? prmt_c<0x4442>(word, 0u)
                                                : prmt_c<0x4443>(word, 0u));
            t = s + pj;
            if (i + 1 < MU)
            {
              const uint32_t fw = W[FW + (i + 1) / 2];
              const int fn = (int)(((i + 1) & 1) ? (fw >> 16) : (fw & 0xffffu));
              s = __viaddmax_s32(s, pj, fn);
            }
          }
          if (i & 1)
            fo[i / 2] = (uint32_t)tprev | ((uint32_t)t << 16);
          tprev = t;
        }
#pragma unroll
        for (int q = 0; q < MP / 2; q++)
        {
          const uint32_t word = pw[q >> 1];
          const uint32_t pp = (q & 1) ? prmt_c<0x4342>(word, 0u) : prmt_c<0x4140>(word, 0u);
          ro[q] = W[RW + q] - pp;
        }
      }
#pragma unroll
      for (int q = 0; q < MP / 2; q++)
      {
        W[FW + q] = fo[q];
        W[RW + q] = ro[q];
      }
#pragma unroll
      for (int wi = 6 + MP; wi < NPL * 4; wi++)
        W[wi] = 0u;
      const size_t slot = (size_t)at + i;
#pragma unroll
      for (int q = 0; q < NPL; q++)
        dplanes[(size_t)q * cap + slot] = make_uint4(W[4 * q], W[4 * q + 1], W[4 * q + 2], W[4 * q + 3]);
    }
  };

  uint16_t *surv_buf = reinterpret_cast<uint16_t *>(wbase + L.off_surv);
  int buf = 0, sbuf = 0;
  /* the tile whose children are still to be written (its cursor reservation is in flight) */
  int pend_nsurv = 0, pend_buf = 0, pend_sbuf = 0;
  unsigned pend_pos = 0;
  /* the incumbent is read ONCE per launch, before any asynchronous copy is in flight: the reference's kernels get
   * `best` by value at launch too (PFSP_gpu_lib.cu:148). A per-tile load, even issued a tile ahead, ended up on the
   * scoreboard slot of the tile prefetch and made the first compare of every tile wait for HBM (a fifth of the LB1_d
   * kernel's stall samples). Leaves still lower ctrl->best at once; the next launch prunes with it. */
  const int best = *reinterpret_cast<volatile int *>(&ctrl->best);
  fetch_tile(blockIdx.x * BB_WARPS_PER_CTA + warp, tile_buf);
  for (int tile = blockIdx.x * BB_WARPS_PER_CTA + warp; tile < ntiles; tile += tstride)
  {
    const uint4 *tile_s = tile_buf + buf * (NPL * 32);
    const unsigned char *tile_b = reinterpret_cast<const unsigned char *>(tile_s);
    uint16_t *surv_s = surv_buf + sbuf * (BB_TILE_MAX * BB_MAX_JOBS);
    const int nbuf = buf == 2 ? 0 : buf + 1;
    fetch_tile(tile + tstride, tile_buf + nbuf * (NPL * 32)); /* the third buffer: neither current nor pending */
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncwarp();
    const int first = tile * TS;
    const int cnt = min(TS, nbatch - first);
    const bool mine = lane < cnt;

    /* ---- this lane's parent: front / remain into registers, both halves alike ---- */
    uint32_t f2[MU], r2[MU];
    int depth = n;
    uint32_t U0 = 0u;
    {
      uint32_t W[NPL * 4];
#pragma unroll
      for (int q = 1; q < NPL; q++)
      {
        const uint4 v = tile_s[q * 32 + lane];
        W[4 * q] = v.x;
        W[4 * q + 1] = v.y;
        W[4 * q + 2] = v.z;
        W[4 * q + 3] = v.w;
      }
      if (mine)
      {
        depth = (int)(W[5] & 0xffu);
        U0 = ~(W[5] >> 8) & fullmask;
      }
#pragma unroll
      for (int i = 0; i < MU; i++)
      {
        f2[i] = (i & 1) ? prmt_c<0x3232>(W[FW + i / 2], 0u) : prmt_c<0x1010>(W[FW + i / 2], 0u);
        r2[i] = (i & 1) ? prmt_c<0x3232>(W[RW + i / 2], 0u) : prmt_c<0x1010>(W[RW + i / 2], 0u);
      }
    }
    const int kids = n - depth; /* 0 for a lane without a parent */
    const int iters = __reduce_max_sync(0xffffffffu, (kids + 1) >> 1);
    const int allkids = __reduce_add_sync(0xffffffffu, kids);
    if (lane == 0)
    {
      cnt6[2] += cnt;
      cnt6[3] += allkids;
    }
    const bool leaf_lane = (kids == 1); /* the only child completes the schedule */

    /* ---- pass 1: bound this parent's children two at a time, remember the survivors by position ---- */
    uint32_t keepmask = 0u;
    /* iterations in which EVERY lane has two children to bound - nearly all, the lanes of a warp being at the same
     * depth - run a copy of the body without per-lane activity and leaf tests */
    const int full_iters = __reduce_min_sync(0xffffffffu, kids >> 1);
    auto pair_iter = [&](const int it, auto fast_tag) {
      constexpr bool FAST = decltype(fast_tag)::value;
      const int k0 = depth + 2 * it;
      const bool act0 = FAST || k0 < n, act1 = FAST || k0 + 1 < n;
      const int kk0 = act0 ? k0 : 0, kk1 = act1 ? k0 + 1 : kk0;
      const int job0 = act0 ? tile_b[(kk0 >> 4) * 512 + lane * 16 + (kk0 & 15)] : 0;
      const int job1 = act1 ? tile_b[(kk1 >> 4) * 512 + lane * 16 + (kk1 & 15)] : job0;
      uint32_t w0[PW], w1[PW];
#pragma unroll
      for (int q = 0; q < PW; q++)
      {
        w0[q] = prow[q * BB_MAX_JOBS + job0];
        w1[q] = prow[q * BB_MAX_JOBS + job1];
      }
      uint32_t sr[MU];
      uint32_t s2 = f2[0], lb2 = 0u, carry2 = 0u;
#pragma unroll
      for (int i = 0; i < MU; i++)
      {
        sr[i] = 0u;
        if (MM || i < m)
        {
          sr[i] = add_fma(s2, one, r2[i]);
          if (LB == 1)
          {
            carry2 = __vmaxu2(carry2, sr[i]);
            lb2 = __viaddmax_u16x2(carry2, T.tails2[i], lb2);
          }
          else
            lb2 = __viaddmax_u16x2(sr[i], T.tails2[i], lb2);
          if (i + 1 < MU)
          {
            const uint32_t a = w0[i >> 2], b = w1[i >> 2];
            const uint32_t p2 = (i & 3) == 0   ? pair_bytes<0>(a, b)
                                : (i & 3) == 1 ? pair_bytes<1>(a, b)
                                : (i & 3) == 2 ? pair_bytes<2>(a, b)
                                               : pair_bytes<3>(a, b);
            s2 = __viaddmax_u16x2(s2, p2, f2[i + 1]);
          }
        }
      }
      const int lbA = (int)(lb2 & 0xffffu), lbB = (int)(lb2 >> 16);
      const bool leaf = !FAST && leaf_lane;
      if (act0 && leaf)
      {
        /* PFSP_lib.h:64-71: a leaf counts as a solution and may improve the incumbent */
        cnt6[1]++;
        if (lbA < best)
          atomicMin(&ctrl->best, lbA);
      }
      bool keepA = act0 && !leaf && (lbA < best);
      bool keepB = act1 && (lbB < best);

      if (LB == 3)
      {
        /* LB2 = max(LB1_d, max_a(sr[a] + G_a(U))): the LB1_d test above is exact as a prefilter */
        const bool candA = prefilter ? keepA : (act0 && !leaf);
        const bool candB = prefilter ? keepB : act1;
        if (__any_sync(0xffffffffu, candA || candB))
        {
          constexpr int NV = (MU - 1 + 7) / 8; /* uint4 loads per row */
          const uint32_t UA = U0 & ~(1u << job0), UB = U0 & ~(1u << job1);
          /* base + set x row bytes: one IMAD.WIDE each on the FMA pipe */
          const uint4 *rowA = reinterpret_cast<const uint4 *>(g16_bytes + (size_t)UA * grow_bytes);
          const uint4 *rowB = reinterpret_cast<const uint4 *>(g16_bytes + (size_t)UB * grow_bytes);
          uint4 ga[NV], gb[NV];
#pragma unroll
          for (int q = 0; q < NV; q++)
          {
            ga[q] = candA ? __ldg(rowA + q) : make_uint4(0u, 0u, 0u, 0u);
            gb[q] = candB ? __ldg(rowB + q) : make_uint4(0u, 0u, 0u, 0u);
          }
          uint32_t v2 = lb2;
#pragma unroll
          for (int q = 0; q < NV; q++)
          {
            const uint32_t xa[4] = {ga[q].x, ga[q].y, ga[q].z, ga[q].w}, xb[4] = {gb[q].x, gb[q].y, gb[q].z, gb[q].w};
#pragma unroll
            for (int h = 0; h < 4; h++)
            {
              const int a0 = q * 8 + 2 * h;
              if (a0 < MU - 1 && (MM || a0 < m - 1))
                v2 = __viaddmax_u16x2(sr[a0], prmt_c<0x5410>(xa[h], xb[h]), v2);
              if (a0 + 1 < MU - 1 && (MM || a0 + 1 < m - 1))
                v2 = __viaddmax_u16x2(sr[a0 + 1], prmt_c<0x7632>(xa[h], xb[h]), v2);
            }
          }
          keepA = candA && ((int)(v2 & 0xffffu) < best);
          keepB = candB && ((int)(v2 >> 16) < best);
          if (candA)
            nlb2++;
          if (candB)
            nlb2++;
        }
      }
      keepmask |= (keepA ? (1u << kk0) : 0u) | (keepB ? (1u << kk1) : 0u);
    };
    {
      int it = 0;
      for (; it < full_iters; it++)
        pair_iter(it, std::true_type{});
      for (; it < iters; it++)
        pair_iter(it, std::false_type{});
    }

    /* ---- survivors of the tile: (lane, position) list in shared memory ---- */
    const int nk = __popc(keepmask);
    const int incl = warp_incl_scan(nk, lane);
    const int nsurv = __shfl_sync(0xffffffffu, incl, 31);
    {
      int at = incl - nk;
      while (keepmask)
      {
        const int k = __ffs(keepmask) - 1;
        keepmask &= keepmask - 1u;
        surv_s[at++] = (uint16_t)((lane << 8) | k);
      }
    }
    __syncwarp();

    /* ---- write the PREVIOUS tile's children (its reservation was issued a whole tile ago), then reserve this
     * tile's slots: the value comes back while the next tile is bounded. (In this order the register holding the
     * reservation is never copied while the atomic is still in flight.) ---- */
    if (pend_nsurv)
      write_children(tile_buf + pend_buf * (NPL * 32), surv_buf + pend_sbuf * (BB_TILE_MAX * BB_MAX_JOBS), pend_nsurv,
                     pend_pos);
    pend_pos = 0;
    if (nsurv && lane == 0)
    {
      pend_pos = atomicAdd(&ctrl->tail, (unsigned)nsurv);
      cnt6[0] += nsurv;
    }
    pend_nsurv = nsurv;
    pend_buf = buf;
    pend_sbuf = sbuf;
    buf = nbuf;
    sbuf ^= 1;
    __syncwarp();
  }
  if (pend_nsurv)
    write_children(tile_buf + pend_buf * (NPL * 32), surv_buf + pend_sbuf * (BB_TILE_MAX * BB_MAX_JOBS), pend_nsurv,
                   pend_pos);
  asm volatile("cp.async.wait_group 0;" ::: "memory");

  cnt6[4] = nlb2; /* children bounded through the table; every one stands for the P machine pairs of lb2_bound */
  cnt6[5] = (unsigned long long)nlb2 * P;
  finish_iteration(ctrl, cnt6, src, base, cap);
}

} // namespace bbk
#endif
