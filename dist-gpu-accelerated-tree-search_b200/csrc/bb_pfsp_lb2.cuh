/*
 * bb_pfsp_lb2.cuh - staged LB2 expand kernel (sm_100a): bound -> prune -> branch -> compact with the two-machine
 * Johnson bound (lb_makespan, pfsp/lib/c_bound_johnson.c:203-237 / bounds_gpu.cu:252-316).
 *
 * Why a second kernel. The first LB2 kernel (pfsp_expand<2,MM>) walked the machine pairs with one lane per child
 * in the reference's fixed pair order. ncu (profiles/r01_ncu_full_pfsp_29_2.txt) showed it ALU-pipe bound (68 %)
 * with ~9 of 32 lanes alive in the pair loop: most children die after a few pairs while the few survivors of a
 * warp need all P pairs. The prune decision `lb < best` does not depend on the ORDER pairs are tried in (the
 * reference's early exit only changes the value returned for already-pruned children), so this
 * kernel
 *   1. tries pairs in a LEARNED order: every killed child credits the pair that killed it (sampled CTAs, shared
 *      memory histogram); the last CTA of a launch folds the histogram into per-pair scores and re-sorts the
 *      staged tables for the next launch. Pairs that prune most come first;
 *   2. stage A: the first KA pairs are evaluated on the children where they are generated (one lane per child);
 *   3. children still alive are COMPACTED into a per-warp shared-memory queue of self-contained candidate records
 *      (permutation, depth|mask, front[], front+remain[]); whenever 32 are queued, stage B evaluates the
 *      remaining pairs with every lane busy, and survivors go straight to the destination stack
 *      (warp-aggregated append).
 * Results (survivor set, exploredTree, exploredSol, incumbent) are identical to the first kernel and to the
 * reference; only the number of pairs evaluated changes.
 *
 * The recurrence is carried as the gap D = t1 - t0:  D <- max(D + (p1 - p0), lag + p1)  (one VIADDMNMX per
 * unscheduled job), see bb_pfsp_kernels.cuh.
 */
#ifndef BB_PFSP_LB2_CUH
#define BB_PFSP_LB2_CUH

#include "bb_pfsp_kernels.cuh"

#define BB_LB2_WARPS 12
#define BB_LB2_THREADS (BB_LB2_WARPS * 32)
#define BB_LB2_QCAP 64       /* queue slots per warp (power of two, >= 63) */
#define BB_LB2_QWORDS 26     /* words per queued candidate: prmu 5, depth|mask 1, front 10, front+remain 10 */

namespace bbk
{

struct lb2_smem_layout
{
  int off_p8, off_tails, off_jbit, off_ja, off_jb, off_pairs, off_hist, off_sort, off_warp, warp_bytes, off_tile,
      off_offs, off_fc, off_fr, off_q, total;
};

__host__ __device__ inline lb2_smem_layout lb2_layout(int m, int P)
{
  lb2_smem_layout L;
  int o = 0;
  L.off_p8 = o;
  o += ((m * BB_MAX_JOBS) + 15) & ~15;
  L.off_tails = o;
  o += ((m * 4) + 15) & ~15;
  const int slots = (P * BB_MAX_JOBS * 4 + 15) & ~15;
  L.off_jbit = o;
  o += slots;
  L.off_ja = o;
  o += slots;
  L.off_jb = o;
  o += slots;
  L.off_pairs = o;
  o += ((P * 8) + 15) & ~15;
  L.off_hist = o;
  o += (((P + 1) * 4) + 15) & ~15;
  L.off_sort = o;
  o += ((3 * P * 4) + 15) & ~15;
  L.off_warp = o;
  const int mp = pfsp_mp(m);
  int w = 0;
  L.off_tile = w;
  w += pfsp_planes(m) * BB_TILE_MAX * 16;
  L.off_offs = w;
  w += 36 * 4;
  L.off_fc = w;
  w += (mp / 2) * 32 * 4;
  L.off_fr = w;
  w += (mp / 2) * 32 * 4;
  L.off_q = w;
  w += BB_LB2_QWORDS * BB_LB2_QCAP * 4;
  L.warp_bytes = (w + 15) & ~15;
  L.total = o + L.warp_bytes * BB_LB2_WARPS;
  return L;
}

/* value of machine pair s for one child: front/front+remain are read as u16 #i at base[(i >> 1) * RS + (i & 1)] */
template <int RS>
__device__ __forceinline__ int lb2_pair_value(const int s, const uint32_t *__restrict__ jbit,
                                              const int *__restrict__ ja, const int *__restrict__ jb,
                                              const int2 *__restrict__ pairs, const int *__restrict__ tails,
                                              const uint16_t *__restrict__ fcb, const uint16_t *__restrict__ frb,
                                              const uint32_t unsched)
{
  const int2 ab = pairs[s];
  const int fa = fcb[(ab.x >> 1) * RS + (ab.x & 1)];
  const int fb = fcb[(ab.y >> 1) * RS + (ab.y & 1)];
  int D = fb - fa;
  const uint4 *vbit = reinterpret_cast<const uint4 *>(jbit + s * BB_MAX_JOBS);
  const int4 *va = reinterpret_cast<const int4 *>(ja + s * BB_MAX_JOBS);
  const int4 *vb = reinterpret_cast<const int4 *>(jb + s * BB_MAX_JOBS);
#pragma unroll
  for (int q = 0; q < BB_MAX_JOBS / 4; q++)
  {
    const uint4 bits = vbit[q];
    const int4 a4 = va[q];
    const int4 b4 = vb[q];
    if (unsched & bits.x) D = add_max(D, a4.x, b4.x);
    if (unsched & bits.y) D = add_max(D, a4.y, b4.y);
    if (unsched & bits.z) D = add_max(D, a4.z, b4.z);
    if (unsched & bits.w) D = add_max(D, a4.w, b4.w);
  }
  return (int)frb[(ab.x >> 1) * RS + (ab.x & 1)] + max(D + tails[ab.y], tails[ab.x]);
}

/* Last CTA of a launch: fold the kill histogram (indexed by position in the current order) into per-pair scores,
 * sort pairs by score and rewrite the active (ordered) tables for the next launch. */
__device__ inline void lb2_reorder(const bb_pfsp_dev &T, int *sort_s)
{
  const int P = T.P;
  unsigned *hist = T.hist; /* [P + 1]; hist[P] = number of samples since the last re-sort */
  if (hist[P] < 4096u)
    return;
  int *score_s = sort_s, *oldpos_s = sort_s + P, *order_s = sort_s + 2 * P;
  for (int pos = threadIdx.x; pos < P; pos += blockDim.x)
  {
    const int id = T.order[pos];
    const unsigned sc = (T.score[id] >> 1) + hist[pos];
    score_s[id] = (int)(sc > 0x3fffffffu ? 0x3fffffffu : sc);
    oldpos_s[id] = pos;
    hist[pos] = 0u;
  }
  __syncthreads();
  if (threadIdx.x == 0)
    hist[P] = 0u;
  for (int id = threadIdx.x; id < P; id += blockDim.x)
  {
    const int sc = score_s[id], op = oldpos_s[id];
    int rank = 0;
    for (int j = 0; j < P; j++)
      rank += (score_s[j] > sc) || (score_s[j] == sc && oldpos_s[j] < op);
    order_s[rank] = id;
    T.score[id] = (unsigned)sc;
  }
  __syncthreads();
  for (int l = threadIdx.x; l < P; l += blockDim.x)
  {
    T.order[l] = order_s[l];
    T.pairs_act[l] = T.pairs[order_s[l]];
  }
  for (int i = threadIdx.x; i < P * BB_MAX_JOBS; i += blockDim.x)
  {
    const int l = i / BB_MAX_JOBS, q = i - l * BB_MAX_JOBS;
    const int srci = order_s[l] * BB_MAX_JOBS + q;
    T.jbit_act[i] = T.jbit[srci];
    T.ja_act[i] = T.ja[srci];
    T.jb_act[i] = T.jb[srci];
  }
}

template <int MM>
__global__ void __launch_bounds__(BB_LB2_THREADS, 1)
pfsp_expand_lb2(bb_ctrl *__restrict__ ctrl, const bb_pool pool, const bb_pfsp_dev T, const int prefilter,
                const int stage_a)
{
  const int nbatch = ctrl->n;
  if (nbatch == 0)
    return;
  constexpr int MU = MM ? MM : BB_MAX_MACHINES;
  constexpr int MP = MM ? ((MM + 1) & ~1) : BB_MAX_MACHINES;
  constexpr int NPL = (6 + MP + 3) / 4;
  constexpr int QC = BB_LB2_QCAP;
  const int m = MM ? MM : T.m;
  const int n = T.n;
  const int P = T.P;
  const int KA = stage_a < P ? (stage_a & ~1) : P; /* pairs tried before compaction (even: pairs go two at a time) */
  const int cap = pool.capacity;
  const int src = ctrl->src, dst = 1 - src, base = ctrl->base, TS = ctrl->tile;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  extern __shared__ __align__(16) unsigned char smem[];
  const lb2_smem_layout L = lb2_layout(m, P);
  uint8_t *p8 = smem + L.off_p8;
  int *tails = reinterpret_cast<int *>(smem + L.off_tails);
  const uint32_t *jbit = reinterpret_cast<const uint32_t *>(smem + L.off_jbit);
  const int *ja = reinterpret_cast<const int *>(smem + L.off_ja);
  const int *jb = reinterpret_cast<const int *>(smem + L.off_jb);
  const int2 *pairs = reinterpret_cast<const int2 *>(smem + L.off_pairs);
  unsigned *hist_s = reinterpret_cast<unsigned *>(smem + L.off_hist);
  unsigned char *wbase = smem + L.off_warp + warp * L.warp_bytes;
  uint4 *tile_s = reinterpret_cast<uint4 *>(wbase + L.off_tile);    /* [plane][32] */
  int *offs_s = reinterpret_cast<int *>(wbase + L.off_offs);        /* [33] exclusive child offsets */
  uint32_t *fcw_s = reinterpret_cast<uint32_t *>(wbase + L.off_fc); /* [MP/2][32] child front, 2 x u16 per word */
  uint32_t *frw_s = reinterpret_cast<uint32_t *>(wbase + L.off_fr); /* [MP/2][32] child front+remain */
  uint32_t *q_s = reinterpret_cast<uint32_t *>(wbase + L.off_q);    /* [QWORDS][QCAP] queued candidates */

  /* ---- stage the instance tables in their current (learned) pair order ---- */
  for (int i = threadIdx.x; i < (m * BB_MAX_JOBS) / 4; i += BB_LB2_THREADS)
    reinterpret_cast<uint32_t *>(p8)[i] = reinterpret_cast<const uint32_t *>(T.p8)[i];
  for (int i = threadIdx.x; i < m; i += BB_LB2_THREADS)
    tails[i] = T.tails[i];
  {
    const int nv = (P * BB_MAX_JOBS) / 4;
    uint4 *d0 = reinterpret_cast<uint4 *>(smem + L.off_jbit);
    uint4 *d1 = reinterpret_cast<uint4 *>(smem + L.off_ja);
    uint4 *d2 = reinterpret_cast<uint4 *>(smem + L.off_jb);
    for (int i = threadIdx.x; i < nv; i += BB_LB2_THREADS)
    {
      d0[i] = reinterpret_cast<const uint4 *>(T.jbit_act)[i];
      d1[i] = reinterpret_cast<const uint4 *>(T.ja_act)[i];
      d2[i] = reinterpret_cast<const uint4 *>(T.jb_act)[i];
    }
    int2 *dp = reinterpret_cast<int2 *>(smem + L.off_pairs);
    for (int i = threadIdx.x; i < P; i += BB_LB2_THREADS)
      dp[i] = T.pairs_act[i];
    for (int i = threadIdx.x; i <= P; i += BB_LB2_THREADS)
      hist_s[i] = 0u;
  }
  __syncthreads();

  /* one CTA in eight feeds the pair statistics (all of them while the batches are still small) */
  const bool sample = ((blockIdx.x & 7) == 0) || nbatch <= 65536;

  const uint4 *__restrict__ splanes = pool.planes[src];
  uint4 *__restrict__ dplanes = pool.planes[dst];

  unsigned long long cnt6[6] = {0, 0, 0, 0, 0, 0}; /* tree, sol, parents, evals, lb2_evals, lb2_pairs */
  int qhead = 0, qcount = 0;                       /* warp-uniform ring state of the candidate queue */
  int best = *reinterpret_cast<volatile int *>(&ctrl->best);

  /* ---- stage B: the remaining pairs for up to 32 queued candidates, one per lane, then append the survivors ---- */
  auto stage_b = [&](const int take) {
    const int slot = (qhead + lane) & (QC - 1);
    const bool have = lane < take;
    const uint32_t w5 = q_s[5 * QC + slot];
    const uint32_t unsched = ~(w5 >> 8);
    const uint16_t *fcb = reinterpret_cast<const uint16_t *>(q_s + 6 * QC + slot);
    const uint16_t *frb = reinterpret_cast<const uint16_t *>(q_s + (6 + MP / 2) * QC + slot);
    bool alive = have;
    int npairs = 0, killpos = -1;
    for (int s = KA; s < P; s += 2)
    {
      if (!__any_sync(0xffffffffu, alive))
        break;
      if (alive)
      {
        const int v0 = lb2_pair_value<2 * QC>(s, jbit, ja, jb, pairs, tails, fcb, frb, unsched);
        const int s1 = s + 1 < P ? s + 1 : s;
        const int v1 = lb2_pair_value<2 * QC>(s1, jbit, ja, jb, pairs, tails, fcb, frb, unsched);
        npairs += 1 + (s1 != s);
        if (v0 >= best)
        {
          alive = false;
          killpos = s;
        }
        else if (v1 >= best)
        {
          alive = false;
          killpos = s1;
        }
      }
    }
    cnt6[5] += npairs;
    if (sample && killpos >= 0)
      atomicAdd(&hist_s[killpos], 1u);
    const unsigned kmask = __ballot_sync(0xffffffffu, alive);
    if (kmask)
    {
      const int nk = __popc(kmask);
      unsigned pos = 0;
      if (lane == 0)
        pos = atomicAdd(&ctrl->tail, (unsigned)nk);
      pos = __shfl_sync(0xffffffffu, pos, 0);
      if (pos + (unsigned)nk > (unsigned)cap)
      {
        if (lane == 0)
          atomicExch(&ctrl->error, 1);
      }
      else if (alive)
      {
        cnt6[0]++;
        const size_t at = (size_t)pos + __popc(kmask & ((1u << lane) - 1u));
        uint32_t w[NPL * 4];
#pragma unroll
        for (int i = 0; i < 6 + MP / 2; i++)
          w[i] = q_s[i * QC + slot];
#pragma unroll
        for (int i = 0; i < MP / 2; i++) /* remain = (front+remain) - front, two u16 at a time */
          w[6 + MP / 2 + i] = __vsub2(q_s[(6 + MP / 2 + i) * QC + slot], w[6 + i]);
#pragma unroll
        for (int i = 6 + MP; i < NPL * 4; i++)
          w[i] = 0u;
#pragma unroll
        for (int q = 0; q < NPL; q++)
          dplanes[(size_t)q * cap + at] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
      }
    }
    qhead = (qhead + take) & (QC - 1);
    qcount -= take;
    __syncwarp();
  };

  const int ntiles = (nbatch + TS - 1) / TS;
  for (;;)
  {
    /* tiles are handed out dynamically: LB2 work per tile varies with how many children survive */
    int tile = 0;
    if (lane == 0)
      tile = (int)atomicAdd(&ctrl->next_tile, 1u);
    tile = __shfl_sync(0xffffffffu, tile, 0);
    if (tile >= ntiles)
      break;
    best = *reinterpret_cast<volatile int *>(&ctrl->best);
    const int first = tile * TS;
    const int cnt = min(TS, nbatch - first);

    /* ---- coalesced load of the parents' planes into the warp tile ---- */
    int kids = 0;
    if (lane < cnt)
    {
      const size_t node = (size_t)base + first + lane;
#pragma unroll
      for (int q = 0; q < NPL; q++)
        tile_s[q * 32 + lane] = splanes[(size_t)q * cap + node];
      const uint32_t w5 = reinterpret_cast<const uint32_t *>(&tile_s[32 + lane])[1];
      kids = n - (int)(w5 & 0xffu);
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
      const unsigned char *rec = reinterpret_cast<const unsigned char *>(tile_s) + p * 16;
      const uint32_t pw5 = *reinterpret_cast<const uint32_t *>(rec + 512 + 4);
      const int depth = pw5 & 0xffu;
      const int k = act ? depth + (c - offs_s[p]) : depth;
      const int job = rec[(k >> 4) * 512 + (k & 15)];
      const int jobd = rec[(depth >> 4) * 512 + (depth & 15)];

      /* ---- LB1_d on the way (add_front_and_bound, c_bound_simple.c:218-244): the child's front / remain ---- */
      auto rd16 = [&](int idx) -> int {
        return *reinterpret_cast<const uint16_t *>(rec + ((idx * 2) >> 4) * 512 + ((idx * 2) & 15));
      };
      int fc[MU], rc[MU];
      int lb;
      {
        const int f0 = rd16(12), r0 = rd16(12 + MP);
        const int pj0 = p8[job];
        int t = f0 + pj0;
        fc[0] = t;
        rc[0] = r0 - pj0;
        lb = f0 + r0 + tails[0];
#pragma unroll
        for (int i = 1; i < MU; i++)
        {
          fc[i] = 0;
          rc[i] = 0;
          if (MM || i < m)
          {
            const int fi = rd16(12 + i), ri = rd16(12 + MP + i);
            const int pji = p8[i * BB_MAX_JOBS + job];
            const int s = max(t, fi);
            t = s + pji;
            fc[i] = t;
            rc[i] = ri - pji;
            lb = max(lb, s + ri + tails[i]);
          }
        }
      }

      const bool leaf = (depth + 1 == n);
      if (act && leaf)
      {
        cnt6[1]++;
        if (lb < best)
          atomicMin(&ctrl->best, lb);
      }
      /* LB2 >= LB1_d, so `lb >= best` already decides; with the prefilter off every inner child walks the pairs */
      const bool cand = act && !leaf && (prefilter ? (lb < best) : true);
      if (!__ballot_sync(0xffffffffu, cand))
        continue;

      const uint32_t smask = (pw5 >> 8) | (1u << job);
      const uint32_t unsched = ~smask;
      uint32_t fw[MP / 2], frw[MP / 2];
#pragma unroll
      for (int i = 0; i < MP; i += 2)
      {
        const uint32_t f_lo = i < MU ? (uint32_t)fc[i < MU ? i : 0] : 0u;
        const uint32_t f_hi = i + 1 < MU ? (uint32_t)fc[i + 1 < MU ? i + 1 : 0] : 0u;
        const uint32_t r_lo = i < MU ? (uint32_t)rc[i < MU ? i : 0] : 0u;
        const uint32_t r_hi = i + 1 < MU ? (uint32_t)rc[i + 1 < MU ? i + 1 : 0] : 0u;
        fw[i / 2] = f_lo | (f_hi << 16);
        frw[i / 2] = (f_lo + r_lo) | ((f_hi + r_hi) << 16);
        fcw_s[(i / 2) * 32 + lane] = fw[i / 2];
        frw_s[(i / 2) * 32 + lane] = frw[i / 2];
      }
      __syncwarp();

      /* ---- stage A: the KA best pairs, where the children are ---- */
      bool alive = cand;
      int npairs = 0, killpos = -1;
      {
        const uint16_t *fcb = reinterpret_cast<const uint16_t *>(fcw_s + lane);
        const uint16_t *frb = reinterpret_cast<const uint16_t *>(frw_s + lane);
        for (int s = 0; s < KA; s += 2)
        {
          if (!__any_sync(0xffffffffu, alive))
            break;
          if (alive)
          {
            const int v0 = lb2_pair_value<64>(s, jbit, ja, jb, pairs, tails, fcb, frb, unsched);
            const int s1 = s + 1 < P ? s + 1 : s;
            const int v1 = lb2_pair_value<64>(s1, jbit, ja, jb, pairs, tails, fcb, frb, unsched);
            npairs += 1 + (s1 != s);
            if (v0 >= best)
            {
              alive = false;
              killpos = s;
            }
            else if (v1 >= best)
            {
              alive = false;
              killpos = s1;
            }
          }
        }
      }
      if (cand)
      {
        cnt6[4]++;
        cnt6[5] += npairs;
      }
      if (sample && killpos >= 0)
        atomicAdd(&hist_s[killpos], 1u);
      __syncwarp();

      /* ---- compact the children still alive into the warp's queue ---- */
      const unsigned amask = __ballot_sync(0xffffffffu, alive);
      if (amask)
      {
        if (alive)
        {
          const int slot = (qhead + qcount + __popc(amask & ((1u << lane) - 1u))) & (QC - 1);
          const uint4 q0 = tile_s[p];
          uint32_t w[5] = {q0.x, q0.y, q0.z, q0.w, *reinterpret_cast<const uint32_t *>(rec + 512)};
          const uint32_t x = (uint32_t)(job ^ jobd);
#pragma unroll
          for (int wi = 0; wi < 5; wi++)
          {
            uint32_t patch = 0;
            if (wi == (depth >> 2))
              patch ^= x << (8 * (depth & 3));
            if (wi == (k >> 2))
              patch ^= x << (8 * (k & 3));
            q_s[wi * QC + slot] = w[wi] ^ patch;
          }
          q_s[5 * QC + slot] = (uint32_t)(depth + 1) | (smask << 8);
#pragma unroll
          for (int i = 0; i < MP / 2; i++)
          {
            q_s[(6 + i) * QC + slot] = fw[i];
            q_s[(6 + MP / 2 + i) * QC + slot] = frw[i];
          }
        }
        qcount += __popc(amask);
        __syncwarp();
        if (qcount >= 32)
          stage_b(32);
      }
    }
    __syncwarp();
  }
  while (qcount > 0)
    stage_b(qcount < 32 ? qcount : 32);

  /* ---- statistics of the sampled CTAs, counters, commit, plan, and (last CTA) the re-sort of the pair order ---- */
  __syncthreads();
  if (sample)
  {
    unsigned tot = 0;
    for (int i = threadIdx.x; i < P; i += BB_LB2_THREADS)
    {
      const unsigned v = hist_s[i];
      if (v)
      {
        atomicAdd(&T.hist[i], v);
        tot += v;
      }
    }
    if (tot)
      atomicAdd(&T.hist[P], tot);
  }
  __shared__ int last_cta;
  if (threadIdx.x == 0)
    last_cta = 0;
  const bool last = finish_iteration(ctrl, cnt6, src, base, cap);
  if (threadIdx.x == 0 && last)
    last_cta = 1;
  __syncthreads();
  if (last_cta)
    lb2_reorder(T, reinterpret_cast<int *>(smem + L.off_sort));
}

} // namespace bbk
#endif
