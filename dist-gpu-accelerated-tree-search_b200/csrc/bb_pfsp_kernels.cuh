/*
 * bb_pfsp_kernels.cuh - fused bound -> prune -> branch -> compact for PFSP on the device-resident pool (sm_100a).
 *
 * One kernel launch = one pool iteration = what the reference does per batch with
 *   popBackBulk + index prep + 3 H2D copies + evaluate_gpu + D2H + generate_children + pushBackBulk
 *   (pfsp/pfsp_multigpu_cuda.c:227-230, 275-317; pfsp/lib/PFSP_gpu_lib.cu:43-127; pfsp/lib/PFSP_lib.h:51-95).
 *
 * Mapping: a WARP owns a tile of <= 32 consecutive parents. It loads their records with coalesced uint4 reads
 * (lane = parent) into a private shared-memory tile, prefix-sums the child counts with shuffles, and then walks
 * the tile's children 32 at a time with ONE LANE PER CHILD (all lanes busy whatever the depth). The parent's cached
 * front[]/remain[] make a child bound O(m): the child's own front[] falls out of the same loop and is stored with
 * the child, so nothing is ever recomputed from the permutation. Survivors are appended to the destination stack
 * with one warp-aggregated atomicAdd and coalesced uint4 plane writes. The incumbent is a device atomicMin.
 * Instance tables (p_times as bytes, min_tails, Johnson/lag tables) are staged in shared memory once per CTA.
 *
 * No tensor cores: the arithmetic is int32 add / max / compare (DPX VIADDMNMX for the LB2 recurrence).
 *
 * Bounds (value-identical to the reference for the prune decision `lb < best`):
 *   LB = 0  LB1_d   add_front_and_bound, c_bound_simple.c:218-244 / bounds_gpu.cu:115-141
 *   LB = 1  LB1     machine_bound_from_parts on the child, c_bound_simple.c:126-158 / bounds_gpu.cu:94-191
 *   LB = 2  LB2     lb_makespan over all machine pairs, c_bound_johnson.c:203-237 / bounds_gpu.cu:252-316.
 *                   The per-pair recurrence  t0 += p0; t1 = max(t1, t0+lag) + p1  is carried as the gap
 *                   D = t1 - t0:  D <- max(D + (p1-p0), lag + p1), one fused add-max per unscheduled job, and
 *                   t0's final value is front[a] + remain[a], known without the loop. A child is pruned as soon
 *                   as one pair reaches `best`; that is the same decision as the reference's
 *                   "break when lb > best, keep iff lb < best".
 *   LB = 3  LB2 through the SUBSET TABLE (same decision as LB = 2, no pair loop). Unrolling the recurrence gives
 *                   D_final = max( D0 + sum_{j in U} a_j ,  max_{k in U} ( b_k + sum_{j in U after k} a_j ) )
 *                   over the unscheduled set U in the pair's Johnson order. The first term is
 *                   (front+remain)[m2] - (front+remain)[m1]; the second, M_s(U), depends ONLY on the set U and the
 *                   pair - not on the order of the scheduled prefix, so all nodes with the same unscheduled set
 *                   share it. The pair's value becomes
 *                     max( fr[m2] + tail[m2], fr[m1] + tail[m1], fr[m1] + M_s(U) + tail[m2] ),  fr = front + remain,
 *                   whose first two terms are LB1_d terms. Hence
 *                     LB2 = max( LB1_d , max_{m1 < m-1} ( fr[m1] + G_m1(U) ) ),  G_m1(U) = max_{m2 > m1} ( M_(m1,m2)(U) + tail[m2] ),
 *                   and G is a table of 2^n x (m-1) uint16 built once per instance on the device (42 MB at 20x20,
 *                   resident in the 126 MB L2). A child bound is then one 48-byte row read and m-1 add/max.
 *   The root (depth 0) is stored with front = 0; the reference's LB1_d uses min_heads there
 *   (c_bound_simple.c:57-61), which gives the same values because a job can never reach machine i before
 *   min_heads[i].
 */
#ifndef BB_PFSP_KERNELS_CUH
#define BB_PFSP_KERNELS_CUH

#include "bb_internal.cuh"

namespace bbk
{

__device__ __forceinline__ int warp_incl_scan(int v, int lane)
{
#pragma unroll
  for (int d = 1; d < 32; d <<= 1)
  {
    int t = __shfl_up_sync(0xffffffffu, v, d);
    if (lane >= d)
      v += t;
  }
  return v;
}

__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v)
{
#pragma unroll
  for (int d = 16; d > 0; d >>= 1)
    v += __shfl_xor_sync(0xffffffffu, v, d);
  return v;
}

/* D <- max(D + a, b): one VIADDMNMX on sm_90+ (DPX) */
__device__ __forceinline__ int add_max(int d, int a, int b) { return __viaddmax_s32(d, a, b); }

/* Kernels are specialised for the Taillard machine counts; any other m <= 20 uses the m = 20 record layout. */
__host__ __device__ inline int pfsp_mm(int m) { return (m == 5 || m == 10 || m == 20) ? m : 0; }
__host__ __device__ inline int pfsp_mp(int m) { return pfsp_mm(m) ? ((m + 1) & ~1) : BB_MAX_MACHINES; }
__host__ __device__ inline int pfsp_planes(int m) { return (6 + pfsp_mp(m) + 3) / 4; }

struct pfsp_smem_layout
{
  int off_p8, off_tails, off_p2, off_jbit, off_ja, off_jb, off_pairs, off_warp, warp_bytes, off_tile, off_offs,
      off_surv, off_fc, off_fr, total;
};

/* tile_buffers: 1 for the first-generation kernel, 2 for the packed kernel (it prefetches the next tile) */
__host__ __device__ inline pfsp_smem_layout pfsp_layout(int m, int P, int lb, int tile_buffers = 1)
{
  pfsp_smem_layout L;
  int o = 0;
  L.off_p8 = o;
  o += ((m * BB_MAX_JOBS) + 15) & ~15;
  L.off_tails = o;
  o += ((m * 4) + 15) & ~15;
  /* packed kernel, one-machine bounds: p[i][a] | p[i][b] << 16 for every job pair (a, b): [m][20][20] u32. Not for
   * the table LB2 (lb 3): its row reads go through L1, which shares the SM's 228 KB with shared memory */
  L.off_p2 = o;
  if (tile_buffers == 2 && lb != 3)
    o += m * BB_MAX_JOBS * BB_MAX_JOBS * 4;
  L.off_jbit = L.off_ja = L.off_jb = L.off_pairs = o;
  if (lb == 2)
  {
    const int slots = (P * BB_MAX_JOBS * 4 + 15) & ~15;
    L.off_jbit = o;
    o += slots;
    L.off_ja = o;
    o += slots;
    L.off_jb = o;
    o += slots;
    L.off_pairs = o;
    o += ((P * 8) + 15) & ~15;
  }
  L.off_warp = o;
  int w = 0;
  L.off_tile = w;
  w += tile_buffers * pfsp_planes(m) * BB_TILE_MAX * 16;
  L.off_offs = w;
  w += 36 * 4;
  L.off_surv = w; /* survivors of the tile: (parent << 8) | slot, at most 32 parents x BB_MAX_JOBS children */
  w += BB_TILE_MAX * BB_MAX_JOBS * 2;
  L.off_fc = L.off_fr = w;
  if (lb == 2)
  {
    const int mp = pfsp_mp(m);
    L.off_fc = w;
    w += mp * 32 * 2;
    L.off_fr = w;
    w += mp * 32 * 2;
  }
  L.warp_bytes = (w + 15) & ~15;
  L.total = o + L.warp_bytes * BB_WARPS_PER_CTA;
  return L;
}

/* Device-side planning of the next iteration: runs in the last CTA to finish (and in the host-triggered plan
 * kernel after seeding / stealing). Depth-first preference: pop from the stack that was just pushed to, unless it
 * cannot fill a batch and the other one is larger. */
__device__ inline void plan_next(bb_ctrl *c, int preferred, int total_warps)
{
  /* a drained stack rewinds to index 0 (donations are copied out of the pool before the thief is told) */
  for (int s = 0; s < 2; s++)
    if (c->size[s] == c->bot[s])
      c->size[s] = c->bot[s] = 0;
  const int avail[2] = {c->size[0] - c->bot[0], c->size[1] - c->bot[1]};
  const int total = avail[0] + avail[1];
  if (total > c->max_pool)
    c->max_pool = total;
  int src = preferred;
  const int sp = avail[preferred], so = avail[1 - preferred];
  if (!(sp >= c->M || sp >= so))
    src = 1 - preferred;
  /* never pop more parents than the destination stack can take children of (capacity guard): the batch shrinks
   * near capacity instead of overflowing; if this direction has no room at all, try the other one. */
  int n = 0;
  for (int attempt = 0; attempt < 2; attempt++)
  {
    const int room = (c->capacity - c->size[1 - src]) / c->maxkids;
    n = avail[src] < c->M ? avail[src] : c->M;
    if (n > room)
      n = room;
    if (n > 0 || avail[1 - src] == 0)
      break;
    src = 1 - src;
  }
  if (n == 0 && total > 0)
    c->error = 1; /* both stacks too full to expand a single parent */
  if (c->error)
    n = 0;
  c->src = src;
  c->n = n;
  c->base = c->size[src] - n;
  c->tail = (unsigned int)c->size[1 - src];
  int tile = (n + total_warps - 1) / total_warps;
  c->tile = tile < 1 ? 1 : (tile > BB_TILE_MAX ? BB_TILE_MAX : tile);
  c->ticket = 0;
  c->next_tile = 0;
  c->next_round = 0;
  c->next_deep = 0;
}

/* programmatic dependent launch (bb_engine.cu launches the chained iterations with
 * cudaLaunchAttributeProgrammaticStreamSerialization): block until the previous grid has completed and its writes are
 * visible, then allow the next grid of the chain to be scheduled */
__device__ __forceinline__ void pdl_wait_then_release()
{
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

/* counters accumulated per thread, reduced per warp -> CTA -> one global atomic each; then the last CTA to
 * finish commits the stack sizes and plans the next iteration. */
/* returns true in thread 0 of the last CTA to finish, false everywhere else */
__device__ __forceinline__ bool finish_iteration(bb_ctrl *ctrl, unsigned long long (&cnt)[6], int src, int base,
                                                 int cap)
{
  bool last = false;
  __shared__ unsigned long long cta_cnt[6];
  const int lane = threadIdx.x & 31;
  if (threadIdx.x < 6)
    cta_cnt[threadIdx.x] = 0ull;
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 6; i++)
  {
    const unsigned long long v = warp_sum_u64(cnt[i]);
    if (lane == 0 && v)
      atomicAdd(&cta_cnt[i], v);
  }
  __syncthreads();
  if (threadIdx.x < 6 && cta_cnt[threadIdx.x])
    atomicAdd(&ctrl->tree + threadIdx.x, cta_cnt[threadIdx.x]); /* tree, sol, parents, evals, lb2_evals, lb2_pairs */
  __syncthreads();
  if (threadIdx.x == 0)
  {
    __threadfence();
    const unsigned t = atomicAdd(&ctrl->ticket, 1u);
    if (t == gridDim.x - 1)
    {
      __threadfence();
      unsigned tl = *reinterpret_cast<volatile unsigned *>(&ctrl->tail);
      if (tl > (unsigned)cap)
        tl = (unsigned)cap;
      ctrl->size[src] = base;
      ctrl->size[1 - src] = (int)tl;
      ctrl->iterations += 1;
      plan_next(ctrl, 1 - src, gridDim.x * (blockDim.x >> 5));
      last = true;
    }
  }
  return last;
}

template <int LB, int MM>
__global__ void __launch_bounds__(BB_CTA_THREADS, (LB >= 2 ? 2 : 3)) /* LB 3 = LB2 through the subset table */
pfsp_expand(bb_ctrl *__restrict__ ctrl, const bb_pool pool, const bb_pfsp_dev T, const int prefilter)
{
  const int nbatch = ctrl->n;
  if (nbatch == 0)
    return;
  constexpr int MU = MM ? MM : BB_MAX_MACHINES;              /* unroll bound of the machine loops */
  constexpr int MP = MM ? ((MM + 1) & ~1) : BB_MAX_MACHINES; /* u16 entries per cached array in the record */
  constexpr int NPL = (6 + MP + 3) / 4;                      /* uint4 planes per record */
  const int m = MM ? MM : T.m;
  const int n = T.n;
  const int P = T.P;
  const int cap = pool.capacity;
  const int src = ctrl->src, dst = 1 - src, base = ctrl->base, TS = ctrl->tile;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  extern __shared__ __align__(16) unsigned char smem[];
  const pfsp_smem_layout L = pfsp_layout(m, P, LB);
  uint8_t *p8 = smem + L.off_p8;
  int *tails = reinterpret_cast<int *>(smem + L.off_tails);
  const uint32_t *jbit = reinterpret_cast<const uint32_t *>(smem + L.off_jbit);
  const int *ja = reinterpret_cast<const int *>(smem + L.off_ja);
  const int *jb = reinterpret_cast<const int *>(smem + L.off_jb);
  const int2 *pairs = reinterpret_cast<const int2 *>(smem + L.off_pairs);
  unsigned char *wbase = smem + L.off_warp + warp * L.warp_bytes;
  uint4 *tile_s = reinterpret_cast<uint4 *>(wbase + L.off_tile);   /* [plane][32] */
  int *offs_s = reinterpret_cast<int *>(wbase + L.off_offs);       /* [33] exclusive child offsets */
  uint16_t *surv_s = reinterpret_cast<uint16_t *>(wbase + L.off_surv); /* survivors of the tile */
  uint16_t *fc_s = reinterpret_cast<uint16_t *>(wbase + L.off_fc); /* [MP][32] child front        (LB2) */
  uint16_t *fr_s = reinterpret_cast<uint16_t *>(wbase + L.off_fr); /* [MP][32] child front+remain (LB2) */

  /* ---- stage the instance tables (once per CTA) ---- */
  for (int i = threadIdx.x; i < (m * BB_MAX_JOBS) / 4; i += BB_CTA_THREADS)
    reinterpret_cast<uint32_t *>(p8)[i] = reinterpret_cast<const uint32_t *>(T.p8)[i];
  for (int i = threadIdx.x; i < m; i += BB_CTA_THREADS)
    tails[i] = T.tails[i];
  if (LB == 2)
  {
    const int nv = (P * BB_MAX_JOBS) / 4;
    uint4 *d0 = reinterpret_cast<uint4 *>(smem + L.off_jbit);
    uint4 *d1 = reinterpret_cast<uint4 *>(smem + L.off_ja);
    uint4 *d2 = reinterpret_cast<uint4 *>(smem + L.off_jb);
    for (int i = threadIdx.x; i < nv; i += BB_CTA_THREADS)
    {
      d0[i] = reinterpret_cast<const uint4 *>(T.jbit)[i];
      d1[i] = reinterpret_cast<const uint4 *>(T.ja)[i];
      d2[i] = reinterpret_cast<const uint4 *>(T.jb)[i];
    }
    int2 *dp = reinterpret_cast<int2 *>(smem + L.off_pairs);
    for (int i = threadIdx.x; i < P; i += BB_CTA_THREADS)
      dp[i] = T.pairs[i];
  }
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

    /* ---- pass 1: bound every child of the tile, remember the survivors ---- */
    int nsurv = 0;
    for (int c0 = 0; c0 < total; c0 += 32)
    {
      const int c = c0 + lane;
      const bool act = c < total;
      /* parent of child c: last p with offs[p] <= c (lanes >= cnt hold offs == total) */
      int p = 0;
      if (act)
      {
#pragma unroll
        for (int st = 16; st > 0; st >>= 1)
          if (offs_s[p + st] <= c)
            p += st;
      }
      /* byte b of plane q of parent p sits at tile + q*512 + p*16 + b */
      const unsigned char *rec = reinterpret_cast<const unsigned char *>(tile_s) + p * 16;
      const uint32_t pw5 = *reinterpret_cast<const uint32_t *>(rec + 512 + 4);
      const int depth = pw5 & 0xffu;
      const int k = act ? depth + (c - offs_s[p]) : depth;
      const int job = rec[(k >> 4) * 512 + (k & 15)];
      const int jobd = rec[(depth >> 4) * 512 + (depth & 15)];

      /* ---- the bound: one pass down the machines; fc/rc = the child's front/remain ---- */
      auto rd16 = [&](int idx) -> int { /* u16 #idx of the record (front starts at #12) */
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
        int carry = t + rc[0];
        lb = (LB == 1) ? carry + tails[0] : f0 + r0 + tails[0];
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
        /* PFSP_lib.h:64-71: a leaf counts as a solution and may improve the incumbent. For a complete
         * schedule all three bounds equal the makespan. */
        cnt6[1]++;
        if (lb < best)
          atomicMin(&ctrl->best, lb);
      }
      bool keep = act && !leaf && (lb < best);

      if (LB == 3)
      {
        /* LB2 = max(LB1_d, max_a(fr[a] + G_a(U))): the LB1_d test above is exact as a prefilter */
        const bool cand = prefilter ? keep : (act && !leaf);
        if (cand)
        {
          const uint32_t U = ~((pw5 >> 8) | (1u << job)) & ((1u << n) - 1u);
          const uint4 *row = reinterpret_cast<const uint4 *>(T.g16 + (size_t)U * T.grow);
          constexpr int NV = (MU - 1 + 7) / 8; /* uint4 loads per row */
          int v = lb;
#pragma unroll
          for (int q = 0; q < NV; q++)
          {
            const uint4 g = __ldg(row + q);
            const uint32_t gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
            for (int h = 0; h < 8; h++)
            {
              const int a = q * 8 + h;
              if (a < MU - 1 && (MM || a < m - 1))
                v = max(v, fc[a] + rc[a] + (int)((gw[h >> 1] >> (16 * (h & 1))) & 0xffffu));
            }
          }
          keep = v < best;
          cnt6[4]++;
          cnt6[5] += P;
        }
      }

      if (LB == 2)
      {
        /* LB2 >= LB1_d (every pair's value contains front+remain+tail of its machines), so `lb >= best`
         * already decides; with prefilter off every inner child walks the pairs like the reference does. */
        const bool cand = prefilter ? keep : (act && !leaf);
        keep = cand;
        if (__ballot_sync(0xffffffffu, cand))
        {
          const uint32_t unsched = ~((pw5 >> 8) | (1u << job));
#pragma unroll
          for (int i = 0; i < MU; i++)
            if (MM || i < m)
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
              const int2 ab = pairs[s];
              int D = (int)fc_s[ab.y * 32 + lane] - (int)fc_s[ab.x * 32 + lane];
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

      /* ---- prune: survivors are only recorded here; they are branched and written in pass 2 ---- */
      const unsigned kmask = __ballot_sync(0xffffffffu, keep);
      if (keep)
        surv_s[nsurv + __popc(kmask & ((1u << lane) - 1u))] = (uint16_t)((p << 8) | k);
      nsurv += __popc(kmask);
    }
    __syncwarp();

    /* ---- pass 2: ONE reservation on the destination cursor per tile (not per 32 children: the cursor is a single
     * address every warp of the grid hits), then branch + compact with every lane writing a survivor: the child's
     * front/remain are recomputed from the parent (survivors are ~1 child in 10), the planes go out as full
     * 512-byte coalesced stores ---- */
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
        for (int i0 = 0; i0 < nsurv; i0 += 32)
        {
          const int i = i0 + lane;
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
            /* child front / remain (the same recurrence as in pass 1) */
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
            const size_t slot = (size_t)pos + i;
#pragma unroll
            for (int q = 0; q < NPL; q++)
              dplanes[(size_t)q * cap + slot] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
          }
        }
    }
    __syncwarp();
  }

  finish_iteration(ctrl, cnt6, src, base, cap);
}

} // namespace bbk
#endif
