/*
 * bb_pfsp_wide.cuh - fused bound -> prune -> branch -> compact for the 50-, 100- and 200-job Taillard classes
 * (ta031-ta060, ta061-ta090, ta091-ta110).
 *
 * The reference handles instances above 20 jobs by editing MAX_JOBS in pfsp/lib/macro.h:10 and rebuilding
 * (pfsp/README.md:52); its kernels and host code are otherwise the same. Here the engine switches to this kernel and
 * a wider record when an instance has 21..200 jobs (three job-capacity classes, wide_geom<50 / 100 / 200>: the
 * reference's MAX_JOBS; the layout below is the 50-job one, the 100-job record has 25 prmu words, 4 mask words, 13
 * planes, the 200-job record 50 prmu words, 7 mask words, 20 planes; job ids stay 8-bit, which is why the 500-job class
 * is not covered):
 *     words  0..12   prmu[50] as uint8 (first `depth` entries = scheduled prefix)
 *     word   13      depth
 *     words 14..15   scheduled-set bitmask (64 bits)
 *     words 16..25   front[20]  uint16 (cached)
 *     words 26..35   remain[20] uint16 (cached)            -> 36 words = 9 planes of 16 bytes = 144 B
 * Values stay below 2^16: a front is at most (50 + 20 - 1) * 99.
 *
 * Structure: the first-generation kernel of bb_pfsp_kernels.cuh (a warp owns a tile of <= 32 parents in shared memory,
 * one lane per child in pass 1, one lane per survivor and one cursor reservation per tile in pass 2, device-side
 * planning). Bounds: LB 0 = LB1_d (c_bound_simple.c:218-244), LB 1 = LB1 (c_bound_simple.c:126-158), LB 2 = LB2 as
 * the machine-pair loop in the reference's pair order with the gap recurrence D <- max(D + (p_b - p_a), lag + p_b)
 * (c_bound_johnson.c:203-237), pruned as soon as one pair reaches `best` - the same decision as the reference's
 * "break when lb > best, keep iff lb < best". A 2^n subset table does not exist at 50 jobs; the pair tables (up to
 * 190 pairs x 50 slots) are read from global memory: every lane of a warp reads the same slot, so it is one
 * broadcast load served by L1.
 */
#ifndef BB_PFSP_WIDE_CUH
#define BB_PFSP_WIDE_CUH

#include "bb_pfsp_kernels.cuh"

/* record geometry of a job-capacity class NJ (50: ta031-ta060, 100: ta061-ta090, 200: ta091-ta110), in 32-bit words */
template <int NJ>
struct wide_geom
{
  static_assert(NJ == 50 || NJ == 100 || NJ == 200, "job-capacity classes of the wide kernel");
  static constexpr int JOBS = NJ;
  static constexpr int STRIDE = NJ == 50 ? 64 : NJ == 100 ? 128 : 256; /* jobs / Johnson slots per row in the wide tables */
  static constexpr int PRW = (NJ + 3) / 4;           /* words of prmu[]: 13 / 25 / 50 */
  static constexpr int W_DEPTH = PRW;                /* 13 / 25 / 50 */
  static constexpr int W_MASK = PRW + 1;             /* 14 / 26 / 51 */
  static constexpr int MW = (NJ + 31) / 32;          /* words of the scheduled-set bitmask: 2 / 4 / 7 */
  static constexpr int W_FRONT = W_MASK + MW;        /* 16 / 30 / 58 */
  static constexpr int W_REMAIN = W_FRONT + BB_MAX_MACHINES / 2; /* 26 / 40 / 68 */
  static constexpr int NPL = (W_REMAIN + BB_MAX_MACHINES / 2 + 3) / 4; /* planes of 16 bytes: 9 / 13 / 20 */
};
__host__ __device__ inline int wide_class(int n) { return n <= 50 ? 50 : n <= 100 ? 100 : 200; }
__host__ __device__ inline int wide_stride(int nj) { return nj == 50 ? 64 : nj == 100 ? 128 : 256; }
__host__ __device__ inline int wide_planes(int nj)
{
  return nj == 50 ? wide_geom<50>::NPL : nj == 100 ? wide_geom<100>::NPL : wide_geom<200>::NPL;
}

/* instance tables of a wide engine (device global memory) */
struct bb_pfsp_wide_dev
{
  int n, m, P;
  const uint8_t *p8;   /* [m][STRIDE] processing times (STRIDE = wide_geom<NJ>::STRIDE) */
  const int *tails;    /* [m] min_tails */
  const uint8_t *jjob; /* [P][STRIDE] job in Johnson order of the pair (0xff past n) */
  const int *ja;       /* [P][STRIDE] p[m2][job] - p[m1][job] */
  const int *jb;       /* [P][STRIDE] lag + p[m2][job] */
  const int2 *pairs;   /* (m1, m2) in the reference's evaluation order */
};

namespace bbk
{

struct wide_smem_layout
{
  int off_p8, off_tails, off_warp, warp_bytes, off_tile, off_offs, off_surv, off_fc, off_fr, total;
};

__host__ __device__ inline wide_smem_layout wide_layout(int m, int nj)
{
  wide_smem_layout L;
  int o = 0;
  L.off_p8 = o;
  o += m * wide_stride(nj);
  L.off_tails = o;
  o += ((m * 4) + 15) & ~15;
  L.off_warp = o;
  int w = 0;
  L.off_tile = w;
  w += wide_planes(nj) * BB_TILE_MAX * 16;
  L.off_offs = w;
  w += 36 * 4;
  L.off_surv = w; /* (parent << 8) | position, at most 32 x NJ */
  w += BB_TILE_MAX * nj * 2;
  L.off_fc = w; /* [20][32] child front (LB2) */
  w += BB_MAX_MACHINES * 32 * 2;
  L.off_fr = w; /* [20][32] child front + remain (LB2) */
  w += BB_MAX_MACHINES * 32 * 2;
  L.warp_bytes = (w + 15) & ~15;
  L.total = o + L.warp_bytes * BB_WARPS_PER_CTA;
  return L;
}

template <int LB, int NJ>
__global__ void __launch_bounds__(BB_CTA_THREADS, NJ == 50 ? 2 : 1)
pfsp_expand_wide(bb_ctrl *__restrict__ ctrl, const bb_pool pool, const bb_pfsp_wide_dev T, const int prefilter)
{
  const int nbatch = ctrl->n;
  if (nbatch == 0)
    return;
  using G = wide_geom<NJ>;
  constexpr int NPL = G::NPL;
  constexpr int BB_WIDE_STRIDE = G::STRIDE, BB_WIDE_W_DEPTH = G::W_DEPTH, BB_WIDE_W_MASK = G::W_MASK,
                BB_WIDE_W_FRONT = G::W_FRONT, BB_WIDE_W_REMAIN = G::W_REMAIN;
  constexpr int MU = BB_MAX_MACHINES;
  const int m = T.m, n = T.n, P = T.P;
  const int cap = pool.capacity;
  const int src = ctrl->src, dst = 1 - src, base = ctrl->base, TS = ctrl->tile;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  extern __shared__ __align__(16) unsigned char smem[];
  const wide_smem_layout L = wide_layout(m, NJ);
  uint8_t *p8 = smem + L.off_p8;
  int *tails = reinterpret_cast<int *>(smem + L.off_tails);
  unsigned char *wbase = smem + L.off_warp + warp * L.warp_bytes;
  uint4 *tile_s = reinterpret_cast<uint4 *>(wbase + L.off_tile);
  int *offs_s = reinterpret_cast<int *>(wbase + L.off_offs);
  uint16_t *surv_s = reinterpret_cast<uint16_t *>(wbase + L.off_surv);
  uint16_t *fc_s = reinterpret_cast<uint16_t *>(wbase + L.off_fc);
  uint16_t *fr_s = reinterpret_cast<uint16_t *>(wbase + L.off_fr);

  for (int i = threadIdx.x; i < m * BB_WIDE_STRIDE; i += BB_CTA_THREADS)
    p8[i] = T.p8[i];
  for (int i = threadIdx.x; i < m; i += BB_CTA_THREADS)
    tails[i] = T.tails[i];
  __syncthreads();

  const uint4 *__restrict__ splanes = pool.planes[src];
  uint4 *__restrict__ dplanes = pool.planes[dst];
  unsigned long long cnt6[6] = {0, 0, 0, 0, 0, 0}; /* tree, sol, parents, evals, lb2_evals, lb2_pairs */

  const int ntiles = (nbatch + TS - 1) / TS;
  for (int tile = blockIdx.x * BB_WARPS_PER_CTA + warp; tile < ntiles; tile += gridDim.x * BB_WARPS_PER_CTA)
  {
    const int best = *reinterpret_cast<volatile int *>(&ctrl->best);
    const int first = tile * TS;
    const int cnt = min(TS, nbatch - first);
    int kids = 0;
    if (lane < cnt)
    {
      const size_t node = (size_t)base + first + lane;
#pragma unroll
      for (int q = 0; q < NPL; q++)
        tile_s[q * 32 + lane] = splanes[(size_t)q * cap + node];
      const uint32_t wd = reinterpret_cast<const uint32_t *>(&tile_s[(BB_WIDE_W_DEPTH >> 2) * 32 + lane])[BB_WIDE_W_DEPTH & 3];
      kids = n - (int)(wd & 0xffu);
    }
    const int incl = warp_incl_scan(kids, lane);
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    offs_s[lane] = incl - kids;
    if (lane == 0)
    {
      offs_s[32] = total;
      cnt6[2] += cnt;
      cnt6[3] += total;
    }
    __syncwarp();

    /* ---- pass 1: bound every child of the tile, remember the survivors ---- */
    int nsurv = 0;
    for (int c0 = 0; c0 < total; c0 += 32)
    {
      const int c = c0 + lane;
      const bool act = c < total;
      int p = 0;
      if (act)
      {
#pragma unroll
        for (int st = 16; st > 0; st >>= 1)
          if (offs_s[p + st] <= c)
            p += st;
      }
      /* byte b of the record of parent p sits at tile + (b >> 4) * 512 + p * 16 + (b & 15) */
      const unsigned char *rec = reinterpret_cast<const unsigned char *>(tile_s) + p * 16;
      auto rdw = [&](int w) -> uint32_t { return *reinterpret_cast<const uint32_t *>(rec + (w >> 2) * 512 + (w & 3) * 4); };
      auto rd16 = [&](int h) -> int {
        return *reinterpret_cast<const uint16_t *>(rec + ((h * 2) >> 4) * 512 + ((h * 2) & 15));
      };
      const int depth = (int)(rdw(BB_WIDE_W_DEPTH) & 0xffu);
      const int k = act ? depth + (c - offs_s[p]) : depth;
      const int job = rec[(k >> 4) * 512 + (k & 15)];
      uint32_t unsched[G::MW]; /* jobs not scheduled in the CHILD: the parent's set plus this job, complemented */
#pragma unroll
      for (int wq = 0; wq < G::MW; wq++)
        unsched[wq] = ~(rdw(BB_WIDE_W_MASK + wq) | (wq == (job >> 5) ? (1u << (job & 31)) : 0u));

      int fc[MU], rc[MU];
      int lb;
      {
        const int f0 = rd16(2 * BB_WIDE_W_FRONT), r0 = rd16(2 * BB_WIDE_W_REMAIN);
        const int pj0 = p8[job];
        int t = f0 + pj0;
        fc[0] = t;
        rc[0] = r0 - pj0;
        int carry = t + rc[0];
        lb = (LB == 1) ? carry + tails[0] : f0 + r0 + tails[0];
#pragma unroll
        for (int i = 1; i < MU; i++)
        {
          fc[i] = 0;
          rc[i] = 0;
          if (i < m)
          {
            const int fi = rd16(2 * BB_WIDE_W_FRONT + i), ri = rd16(2 * BB_WIDE_W_REMAIN + i);
            const int pji = p8[i * BB_WIDE_STRIDE + job];
            const int s = max(t, fi);
            t = s + pji;
            fc[i] = t;
            rc[i] = ri - pji;
            if (LB == 1)
            {
              carry = max(carry, t + rc[i]);
              lb = max(lb, carry + tails[i]);
            }
            else
              lb = max(lb, s + ri + tails[i]);
          }
        }
      }
      const bool leaf = (depth + 1 == n);
      if (act && leaf)
      {
        /* PFSP_lib.h:64-71: a leaf counts as a solution and may improve the incumbent */
        cnt6[1]++;
        if (lb < best)
          atomicMin(&ctrl->best, lb);
      }
      bool keep = act && !leaf && (lb < best);

      if (LB == 2)
      {
        /* LB2 >= LB1_d (every pair's value contains front+remain+tail of its machines), so `lb >= best` already
         * decides; with prefilter off every inner child walks the pairs like the reference does */
        const bool cand = prefilter ? keep : (act && !leaf);
        keep = cand;
        if (__ballot_sync(0xffffffffu, cand))
        {
#pragma unroll
          for (int i = 0; i < MU; i++)
            if (i < m)
            {
              fc_s[i * 32 + lane] = (uint16_t)fc[i];
              fr_s[i * 32 + lane] = (uint16_t)(fc[i] + rc[i]);
            }
          __syncwarp();
          bool alive = cand;
          int npairs = 0;
          for (int s = 0; s < P; s++)
          {
            if (!__any_sync(0xffffffffu, alive))
              break;
            if (alive)
            {
              const int2 ab = __ldg(T.pairs + s);
              int D = (int)fc_s[ab.y * 32 + lane] - (int)fc_s[ab.x * 32 + lane];
              const uint8_t *jj = T.jjob + s * BB_WIDE_STRIDE;
              const int *va = T.ja + s * BB_WIDE_STRIDE, *vb = T.jb + s * BB_WIDE_STRIDE;
              for (int q = 0; q < n; q++)
              {
                const int jq = __ldg(jj + q);
                uint32_t uw = unsched[0];
#pragma unroll
                for (int wq = 1; wq < G::MW; wq++)
                  if ((jq >> 5) == wq)
                    uw = unsched[wq];
                if ((uw >> (jq & 31)) & 1u)
                  D = add_max(D, __ldg(va + q), __ldg(vb + q));
              }
              const int val = (int)fr_s[ab.x * 32 + lane] + max(D + tails[ab.y], tails[ab.x]);
              npairs++;
              if (val >= best)
              {
                alive = false;
                keep = false;
              }
            }
          }
          if (cand)
          {
            cnt6[4]++;
            cnt6[5] += npairs;
          }
          __syncwarp();
        }
      }

      const unsigned kmask = __ballot_sync(0xffffffffu, keep);
      if (keep)
        surv_s[nsurv + __popc(kmask & ((1u << lane) - 1u))] = (uint16_t)((p << 8) | k);
      nsurv += __popc(kmask);
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
      pos = __shfl_sync(0xffffffffu, pos, 0);
      if (pos + (unsigned)nsurv > (unsigned)cap)
      {
        if (lane == 0)
          atomicExch(&ctrl->error, 1);
      }
      else
        for (int i = lane; i < nsurv; i += 32)
        {
          const int e = surv_s[i];
          const int p = e >> 8, k = e & 0xff;
          const unsigned char *rec = reinterpret_cast<const unsigned char *>(tile_s) + p * 16;
          uint32_t w[NPL * 4];
#pragma unroll
          for (int q = 0; q < NPL; q++)
          {
            const uint4 v = tile_s[q * 32 + p];
            w[4 * q] = v.x;
            w[4 * q + 1] = v.y;
            w[4 * q + 2] = v.z;
            w[4 * q + 3] = v.w;
          }
          const int depth = (int)(w[BB_WIDE_W_DEPTH] & 0xffu);
          const int job = rec[(k >> 4) * 512 + (k & 15)];
          const int jobd = rec[(depth >> 4) * 512 + (depth & 15)];
          /* prmu: swap positions depth and k (PFSP_lib.h:80-82) */
          const uint32_t x = (uint32_t)(job ^ jobd);
#pragma unroll
          for (int wi = 0; wi < G::PRW; wi++)
          {
            uint32_t patch = 0;
            if (wi == (depth >> 2))
              patch ^= x << (8 * (depth & 3));
            if (wi == (k >> 2))
              patch ^= x << (8 * (k & 3));
            w[wi] ^= patch;
          }
          w[BB_WIDE_W_DEPTH] = (uint32_t)(depth + 1);
#pragma unroll
          for (int wq = 0; wq < G::MW; wq++)
            if (wq == (job >> 5))
              w[BB_WIDE_W_MASK + wq] |= 1u << (job & 31);
          /* child front / remain (the same recurrence as in pass 1) */
          int t = 0;
#pragma unroll
          for (int i2 = 0; i2 < MU; i2 += 2)
          {
            const uint32_t fw = w[BB_WIDE_W_FRONT + i2 / 2], rw = w[BB_WIDE_W_REMAIN + i2 / 2];
            uint32_t f2 = 0, r2 = 0;
#pragma unroll
            for (int h = 0; h < 2; h++)
            {
              const int mi = i2 + h;
              if (mi < m)
              {
                const int fi = (int)((fw >> (16 * h)) & 0xffffu), ri = (int)((rw >> (16 * h)) & 0xffffu);
                const int pji = p8[mi * BB_WIDE_STRIDE + job];
                t = (mi == 0 ? fi : max(t, fi)) + pji;
                f2 |= (uint32_t)t << (16 * h);
                r2 |= (uint32_t)(ri - pji) << (16 * h);
              }
            }
            w[BB_WIDE_W_FRONT + i2 / 2] = f2;
            w[BB_WIDE_W_REMAIN + i2 / 2] = r2;
          }
          const size_t slot = (size_t)pos + i;
#pragma unroll
          for (int q = 0; q < NPL; q++)
            dplanes[(size_t)q * cap + slot] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
        }
    }
    __syncwarp();
  }
  finish_iteration(ctrl, cnt6, src, base, cap);
}

} // namespace bbk
#endif
