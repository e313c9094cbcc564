/*
 * bb_nq_kernels.cuh - fused evaluate -> branch -> compact for N-Queens on the device-resident pool (sm_100a).
 *
 * One launch = one pool iteration = the reference's per-batch popBack loop + H2D + evaluate_gpu + D2H labels +
 * generate_children + pushBack (nqueens/nqueens_multigpu_cuda.cu:288-314, kernel :158-186, branch :189-210).
 *
 * Same warp-tile mapping as the PFSP kernel: a warp loads <= 32 parents (coalesced), prefix-sums N - depth, then
 * evaluates the candidate slots 32 at a time with one lane per (parent, slot).
 *
 * isSafe (nqueens_c.c:80-96) asks whether a queen at (column depth, row r) shares a diagonal with one of the placed
 * queens (column i, row b_i): b_i == r - (depth - i) or b_i == r + (depth - i), i.e. b_i + i == r + depth or
 * b_i - i == r - depth. The lane that loaded a parent folds its placed queens ONCE into two occupancy words
 * (diagonal b+i and anti-diagonal b-i+N-1, < 2N-1 <= 39 bits each); a candidate is then two bit tests instead of
 * a loop over the placed queens that every one of the N-depth candidates of the parent would repeat. The labels
 * are identical to the reference's. -g (the reference's "safety checks per evaluation" knob) repeats the test G
 * times behind an opaque register barrier, so the work still scales with G.
 *
 * Children at depth N are solutions: the reference pushes them, counts them in exploredTree, and counts
 * exploredSol when they are popped (nqueens_c.c:103-105, 112-113). Here they are counted in both at generation and
 * never written to HBM.
 *
 * Record: plane 0 (uint4) = board[0..15]; N <= 15 keeps depth in byte 15; N >= 16 adds TW 32-bit tail words per
 * node holding board[16..N-1] followed by depth at tail byte N-16.
 */
#ifndef BB_NQ_KERNELS_CUH
#define BB_NQ_KERNELS_CUH

#include "bb_internal.cuh"
#include "bb_pfsp_kernels.cuh"

namespace bbk
{

__host__ __device__ inline int nq_tail_words(int N) { return N <= 15 ? 0 : (N - 16 + 1 + 3) / 4; }

template <int TW>
__global__ void __launch_bounds__(BB_CTA_THREADS, 4)
nq_expand(bb_ctrl *__restrict__ ctrl, const bb_pool pool, const int N, const int G)
{
  const int nbatch = ctrl->n;
  if (nbatch == 0)
    return;
  const int cap = pool.capacity;
  const int src = ctrl->src, dst = 1 - src, base = ctrl->base, TS = ctrl->tile;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int dbyte = TW ? 16 + (N - 16) : 15; /* where depth lives in the 16 + 4*TW byte record */

  /* per warp: 32 rows of 32 bytes (board bytes 0..23 used), offsets [33] */
  __shared__ __align__(16) unsigned char rows_s[BB_WARPS_PER_CTA][32 * 32];
  __shared__ int offs_all[BB_WARPS_PER_CTA][36];
  __shared__ uint16_t surv_all[BB_WARPS_PER_CTA][BB_TILE_MAX * BB_MAX_QUEENS]; /* (parent << 8) | slot */
  __shared__ unsigned long long diag_all[BB_WARPS_PER_CTA][2][32]; /* per parent: occupied diagonals / anti-diagonals */
  unsigned char *rows = rows_s[warp];
  int *offs_s = offs_all[warp];
  uint16_t *surv_s = surv_all[warp];
  unsigned long long *d1_s = diag_all[warp][0], *d2_s = diag_all[warp][1];

  const uint4 *__restrict__ splane = pool.planes[src];
  uint4 *__restrict__ dplane = pool.planes[dst];
  const uint32_t *__restrict__ stail = pool.tailw[src];
  uint32_t *__restrict__ dtail = pool.tailw[dst];

  unsigned long long cnt6[6] = {0, 0, 0, 0, 0, 0};

  const int ntiles = (nbatch + TS - 1) / TS;
  for (int tile = blockIdx.x * BB_WARPS_PER_CTA + warp; tile < ntiles; tile += gridDim.x * BB_WARPS_PER_CTA)
  {
    const int first = tile * TS;
    const int cnt = min(TS, nbatch - first);
    int kids = 0;
    if (lane < cnt)
    {
      const size_t node = (size_t)base + first + lane;
      *reinterpret_cast<uint4 *>(rows + lane * 32) = splane[node];
#pragma unroll
      for (int t = 0; t < TW; t++)
        *reinterpret_cast<uint32_t *>(rows + lane * 32 + 16 + 4 * t) = stail[node * TW + t];
      const int depth = rows[lane * 32 + dbyte];
      kids = N - depth;
      unsigned long long d1 = 0ull, d2 = 0ull;
      for (int i = 0; i < depth; i++)
      {
        const int b = rows[lane * 32 + i];
        d1 |= 1ull << (b + i);
        d2 |= 1ull << (b - i + N - 1);
      }
      d1_s[lane] = d1;
      d2_s[lane] = d2;
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

    /* ---- pass 1: evaluate every candidate slot of the tile, remember the safe inner placements ---- */
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
      const unsigned char *row = rows + p * 32;
      const int depth = row[dbyte];
      const int k = act ? depth + (c - offs_s[p]) : depth;
      const int cand = row[k]; /* row position tried for column `depth` */

      /* isSafe: nqueens_c.c:80-96 / nqueens_multigpu_cuda.cu:170-183, as two diagonal-occupancy tests */
      const unsigned long long d1 = d1_s[p], d2 = d2_s[p];
      int safe = 1;
      int r = cand;
      for (int g = 0; g < G; g++)
      {
        asm volatile("" : "+r"(r)); /* keep the G repetitions as real work, as the reference's -g intends */
        safe &= (int)(((d1 >> (r + depth)) | (d2 >> (r - depth + N - 1))) & 1ull) ^ 1;
      }
      const bool ok = act && safe;
      const bool leaf = (depth + 1 == N);
      if (ok)
      {
        cnt6[0]++; /* exploredTree counts every safe placement, leaves included */
        if (leaf)
          cnt6[1]++;
      }
      const bool keep = ok && !leaf;
      const unsigned kmask = __ballot_sync(0xffffffffu, keep);
      if (keep)
        surv_s[nsurv + __popc(kmask & ((1u << lane) - 1u))] = (uint16_t)((p << 8) | k);
      nsurv += __popc(kmask);
    }
    __syncwarp();

    /* ---- pass 2: one reservation on the destination cursor per tile, then every lane writes one child ---- */
    if (nsurv)
    {
      unsigned pos = 0;
      if (lane == 0)
        pos = atomicAdd(&ctrl->tail, (unsigned)nsurv);
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
            const unsigned char *row = rows + p * 32;
            const int depth = row[dbyte];
            const int cand = row[k];
            const size_t slot = (size_t)pos + i;
            uint32_t w[4 + (TW ? TW : 1)];
            const uint4 q0 = *reinterpret_cast<const uint4 *>(row);
            w[0] = q0.x;
            w[1] = q0.y;
            w[2] = q0.z;
            w[3] = q0.w;
#pragma unroll
            for (int t = 0; t < TW; t++)
              w[4 + t] = *reinterpret_cast<const uint32_t *>(row + 16 + 4 * t);
            /* swap board[depth] <-> board[k] (xor patch), then depth+1 (depth byte xor old^new) */
            const uint32_t x = (uint32_t)(cand ^ row[depth]);
            const uint32_t dx = (uint32_t)(depth ^ (depth + 1));
#pragma unroll
            for (int wi = 0; wi < 4 + TW; wi++)
            {
              uint32_t patch = 0;
              if (wi == (depth >> 2))
                patch ^= x << (8 * (depth & 3));
              if (wi == (k >> 2))
                patch ^= x << (8 * (k & 3));
              if (wi == (dbyte >> 2))
                patch ^= dx << (8 * (dbyte & 3));
              w[wi] ^= patch;
            }
            dplane[slot] = make_uint4(w[0], w[1], w[2], w[3]);
#pragma unroll
            for (int t = 0; t < TW; t++)
              dtail[slot * TW + t] = w[4 + t];
          }
        }
    }
    __syncwarp();
  }
  finish_iteration(ctrl, cnt6, src, base, cap);
}

/*
 * Second-generation N-Queens kernel: no per-candidate pass at all. The lane that owns a parent derives, from the
 * placed queens, the SET of safe rows for the next column as one bit mask
 *     free = ~(rows used | rows hit by a diagonal | rows hit by an anti-diagonal) & ((1 << N) - 1)
 * (the same test as isSafe, nqueens_c.c:80-96, for all candidate rows at once: a queen (i, b) attacks row
 * b + (depth - i) and row b - (depth - i) of column depth). popcount(free) is the number of children, so
 * exploredTree / exploredSol need no enumeration, the tile's survivors are known after one warp scan, ONE cursor
 * reservation is made per tile, and pass 2 writes child j of parent p as "the j-th set bit of free_p". -g repeats the
 * derivation of `free` G times behind an opaque register barrier (the work scales with G as the reference's knob
 * intends). Labels, counts and children are the reference's.
 */
template <int TW>
__global__ void __launch_bounds__(BB_CTA_THREADS, 4)
nq_expand2(bb_ctrl *__restrict__ ctrl, const bb_pool pool, const int N, const int G)
{
  const int nbatch = ctrl->n;
  if (nbatch == 0)
    return;
  const int cap = pool.capacity;
  const int src = ctrl->src, dst = 1 - src, base = ctrl->base, TS = ctrl->tile;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int dbyte = TW ? 16 + (N - 16) : 15; /* where depth lives in the 16 + 4*TW byte record */
  const uint32_t full = (N >= 32) ? 0xffffffffu : ((1u << N) - 1u);

  /* per warp, double buffered (the next tile is copied asynchronously while this one is expanded):
   * 32 rows of 32 bytes (board bytes 0..23 used); plus offsets [33] and the free masks [32] */
  __shared__ __align__(16) unsigned char rows_s[BB_WARPS_PER_CTA][2][32 * 32];
  __shared__ int offs_all[BB_WARPS_PER_CTA][36];
  __shared__ uint32_t free_all[BB_WARPS_PER_CTA][32];
  int *offs_s = offs_all[warp];
  uint32_t *free_s = free_all[warp];

  const uint4 *__restrict__ splane = pool.planes[src];
  uint4 *__restrict__ dplane = pool.planes[dst];
  const uint32_t *__restrict__ stail = pool.tailw[src];
  uint32_t *__restrict__ dtail = pool.tailw[dst];

  unsigned long long cnt6[6] = {0, 0, 0, 0, 0, 0};

  const int ntiles = (nbatch + TS - 1) / TS;
  const int tstride = gridDim.x * BB_WARPS_PER_CTA;
  auto fetch_tile = [&](int tile, unsigned char *dstrows) {
    if (tile < ntiles)
    {
      const int first = tile * TS;
      if (lane < min(TS, nbatch - first))
      {
        const size_t node = (size_t)base + first + lane;
        const unsigned sa = (unsigned)__cvta_generic_to_shared(dstrows + lane * 32);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(splane + node) : "memory");
#pragma unroll
        for (int t = 0; t < TW; t++)
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa + 16 + 4 * t), "l"(stail + node * TW + t)
                       : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  int buf = 0;
  fetch_tile(blockIdx.x * BB_WARPS_PER_CTA + warp, rows_s[warp][0]);
  for (int tile = blockIdx.x * BB_WARPS_PER_CTA + warp; tile < ntiles; tile += tstride)
  {
    const unsigned char *rows = rows_s[warp][buf];
    fetch_tile(tile + tstride, rows_s[warp][buf ^ 1]);
    buf ^= 1;
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncwarp();
    const int first = tile * TS;
    const int cnt = min(TS, nbatch - first);

    /* ---- per parent: the safe rows of the next column ---- */
    int kids = 0, cands = 0;
    uint32_t fr = 0u;
    bool leaf = false;
    if (lane < cnt)
    {
      const unsigned char *row = rows + lane * 32;
      const int depth = row[dbyte];
      leaf = (depth + 1 == N);
      cands = N - depth;
      int dd = depth;
      for (int g = 0; g < G; g++)
      {
        asm volatile("" : "+r"(dd)); /* keep the G repetitions as real work, as the reference's -g intends */
        uint32_t used = 0u, hit = 0u; /* hit bit r: row r of column `depth` is on a diagonal of a placed queen */
        for (int i = 0; i < dd; i++)
        {
          const uint32_t b = row[i];
          const uint32_t d = (uint32_t)(dd - i);
          const uint32_t bit = 1u << b;
          uint32_t up; /* row b + d; PTX shl gives 0 for counts >= 32 (rows >= N do not matter) */
          asm("shl.b32 %0, %1, %2;" : "=r"(up) : "r"(1u), "r"(b + d));
          used |= bit;
          hit |= up | (bit >> d); /* row b - d falls off the low end when b < d */
        }
        fr = ~(used | hit) & full;
      }
      kids = __popc(fr);
      free_s[lane] = fr;
    }
    const int wkids = leaf ? 0 : kids; /* depth-N children are counted, not stored (nqueens_c.c:103-105) */
    const int incl = warp_incl_scan(wkids, lane);
    const int nsurv = __shfl_sync(0xffffffffu, incl, 31);
    offs_s[lane] = incl - wkids;
    const int allkids = __reduce_add_sync(0xffffffffu, kids);
    const int allsol = __reduce_add_sync(0xffffffffu, leaf ? kids : 0);
    const int allcand = __reduce_add_sync(0xffffffffu, cands);
    if (lane == 0)
    {
      offs_s[32] = nsurv;
      cnt6[0] += allkids; /* exploredTree counts every safe placement, leaves included */
      cnt6[1] += allsol;
      cnt6[2] += cnt;
      cnt6[3] += allcand;
    }
    __syncwarp();

    /* ---- one reservation per tile, then every lane writes one child ---- */
    if (nsurv)
    {
      unsigned pos = 0;
      if (lane == 0)
        pos = atomicAdd(&ctrl->tail, (unsigned)nsurv);
      for (int i0 = 0; i0 < nsurv; i0 += 32)
      {
        const int i = i0 + lane;
        const unsigned amask = __ballot_sync(0xffffffffu, i < nsurv); /* lane 0 is always in it */
        if (i < nsurv)
        {
          int p = 0;
#pragma unroll
          for (int st = 16; st > 0; st >>= 1)
            if (offs_s[p + st] <= i)
              p += st;
          const int j = i - offs_s[p];
          const unsigned char *row = rows + p * 32;
          const int depth = row[dbyte];
          uint32_t f = free_s[p];
          for (int q = 0; q < j; q++) /* j-th safe row */
            f &= f - 1u;
          const int r = __ffs(f) - 1;
          int k = depth; /* where row r sits in the unplaced part of the permutation */
          while (row[k] != r)
            k++;
          uint32_t w[4 + (TW ? TW : 1)];
          const uint4 q0 = *reinterpret_cast<const uint4 *>(row);
          w[0] = q0.x;
          w[1] = q0.y;
          w[2] = q0.z;
          w[3] = q0.w;
#pragma unroll
          for (int t = 0; t < TW; t++)
            w[4 + t] = *reinterpret_cast<const uint32_t *>(row + 16 + 4 * t);
          /* swap board[depth] <-> board[k] (xor patch), then depth+1 (depth byte xor old^new) */
          const uint32_t x = (uint32_t)(r ^ row[depth]);
          const uint32_t dx = (uint32_t)(depth ^ (depth + 1));
#pragma unroll
          for (int wi = 0; wi < 4 + TW; wi++)
          {
            uint32_t patch = 0;
            if (wi == (depth >> 2))
              patch ^= x << (8 * (depth & 3));
            if (wi == (k >> 2))
              patch ^= x << (8 * (k & 3));
            if (wi == (dbyte >> 2))
              patch ^= dx << (8 * (dbyte & 3));
            w[wi] ^= patch;
          }
          const unsigned at = __shfl_sync(amask, pos, 0);
          if (at + (unsigned)nsurv > (unsigned)cap)
            atomicExch(&ctrl->error, 1);
          else
          {
            const size_t slot = (size_t)at + i;
            dplane[slot] = make_uint4(w[0], w[1], w[2], w[3]);
#pragma unroll
            for (int t = 0; t < TW; t++)
              dtail[slot * TW + t] = w[4 + t];
          }
        }
      }
    }
    __syncwarp();
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  finish_iteration(ctrl, cnt6, src, base, cap);
}

/*
 * Third-generation N-Queens kernel: the attack masks travel with the node.
 *
 * ncu on nq_expand2 (profiles/r01_ncu_full_v4_nq_17_4194304.txt) showed the ALU pipe 68 % busy and DRAM at 12.5 %:
 * every parent rebuilt its diagonal masks from the board in an O(depth) loop, and every child looked for its row in
 * the unplaced part of the permutation byte by byte. Here the record is 32 bytes (two planes):
 *     plane 0          board[0..15]
 *     plane 1  .x      board[16..19]
 *              .y      rows already used (bit r) | depth << 24
 *              .z      rows of column `depth` attacked along the rising diagonals  (b + (depth - i) == r)
 *              .w      rows of column `depth` attacked along the falling diagonals (b - (depth - i) == r)
 * so isSafe (nqueens_c.c:80-96) for ALL rows of the next column is  free = ~(used | up | down)  - three operations
 * - and a child's masks are  used | bit, (up | bit) << 1, (down | bit) >> 1. The lane that owns a parent lists its
 * safe rows (lane, row) in shared memory; pass 2 is dense, one lane per child: the row's position in the unplaced
 * part of the permutation is found with a zero-byte test over the five board words (the child must be the
 * reference's node: board[depth] <-> board[k] swapped, nqueens_c.c:107-111), and the two planes go out as coalesced
 * 16-byte stores. -g repeats the safety test G times behind a register barrier. Counts and children are the
 * reference's; depth-N children are counted, not stored.
 */
__device__ __forceinline__ uint32_t nq_has_byte(uint32_t w, uint32_t rrrr)
{
  const uint32_t x = w ^ rrrr;
  return (x - 0x01010101u) & ~x & 0x80808080u; /* lowest flagged byte is an exact match */
}

__global__ void __launch_bounds__(BB_CTA_THREADS, 4)
nq_expand3(bb_ctrl *__restrict__ ctrl, const bb_pool pool, const int N, const int G)
{
  const int nbatch = ctrl->n;
  if (nbatch == 0)
    return;
  const int cap = pool.capacity;
  const int src = ctrl->src, dst = 1 - src, base = ctrl->base, TS = ctrl->tile;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t full = (N >= 32) ? 0xffffffffu : ((1u << N) - 1u);

  /* per warp: two tile buffers of [plane][32] uint4 (the next tile is copied asynchronously while this one is
   * expanded) and the list of children (parent lane << 8 | row) */
  __shared__ uint4 tile_all[BB_WARPS_PER_CTA][2][2 * 32];
  __shared__ uint16_t surv_all[BB_WARPS_PER_CTA][BB_TILE_MAX * BB_MAX_QUEENS];
  uint16_t *surv_s = surv_all[warp];

  const uint4 *__restrict__ splanes = pool.planes[src];
  uint4 *__restrict__ dplanes = pool.planes[dst];

  unsigned long long cnt6[6] = {0, 0, 0, 0, 0, 0};

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
        for (int q = 0; q < 2; q++)
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
  fetch_tile(blockIdx.x * BB_WARPS_PER_CTA + warp, tile_all[warp][0]);
  for (int tile = blockIdx.x * BB_WARPS_PER_CTA + warp; tile < ntiles; tile += tstride)
  {
    const uint4 *tile_s = tile_all[warp][buf];
    fetch_tile(tile + tstride, tile_all[warp][buf ^ 1]);
    buf ^= 1;
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncwarp();
    const int first = tile * TS;
    const int cnt = min(TS, nbatch - first);

    /* ---- per parent: the safe rows of the next column ---- */
    uint32_t fr = 0u;
    int kids = 0, cands = 0;
    bool leaf = false;
    if (lane < cnt)
    {
      const uint4 q1 = tile_s[32 + lane];
      const int depth = (int)(q1.y >> 24);
      uint32_t used = q1.y & 0x00ffffffu;
      leaf = (depth + 1 == N);
      cands = N - depth;
      for (int g = 0; g < G; g++)
      {
        asm volatile("" : "+r"(used)); /* keep the G repetitions as real work, as the reference's -g intends */
        fr = ~(used | q1.z | q1.w) & full;
      }
      kids = __popc(fr);
    }
    const int wkids = leaf ? 0 : kids; /* depth-N children are counted, not stored (nqueens_c.c:103-105) */
    const int incl = warp_incl_scan(wkids, lane);
    const int nsurv = __shfl_sync(0xffffffffu, incl, 31);
    const int allkids = __reduce_add_sync(0xffffffffu, kids);
    const int allsol = __reduce_add_sync(0xffffffffu, leaf ? kids : 0);
    const int allcand = __reduce_add_sync(0xffffffffu, cands);
    if (lane == 0)
    {
      cnt6[0] += allkids; /* exploredTree counts every safe placement, leaves included */
      cnt6[1] += allsol;
      cnt6[2] += cnt;
      cnt6[3] += allcand;
    }
    if (!leaf)
    {
      int at = incl - wkids;
      uint32_t f = fr;
      while (f)
      {
        const int r = __ffs(f) - 1;
        f &= f - 1u;
        surv_s[at++] = (uint16_t)((lane << 8) | r);
      }
    }
    __syncwarp();

    /* ---- one reservation per tile, then every lane writes one child ---- */
    if (nsurv)
    {
      unsigned pos = 0;
      if (lane == 0)
        pos = atomicAdd(&ctrl->tail, (unsigned)nsurv);
      for (int i0 = 0; i0 < nsurv; i0 += 32)
      {
        const int i = i0 + lane;
        const unsigned amask = __ballot_sync(0xffffffffu, i < nsurv); /* lane 0 is always in it */
        if (i < nsurv)
        {
          const int e = surv_s[i];
          const int p = e >> 8, r = e & 0xff;
          const uint4 q0 = tile_s[p], q1 = tile_s[32 + p];
          uint32_t w[5] = {q0.x, q0.y, q0.z, q0.w, q1.x};
          const int depth = (int)(q1.y >> 24);
          const uint32_t bit = 1u << r;
          /* where row r sits in the unplaced part of the permutation: first byte equal to r (padding bytes beyond
           * N are zero and come later) */
          const uint32_t rrrr = (uint32_t)r * 0x01010101u;
          int k = 0;
          {
            uint32_t z = nq_has_byte(w[4], rrrr);
            k = 16;
#pragma unroll
            for (int wi = 3; wi >= 0; wi--)
            {
              const uint32_t zz = nq_has_byte(w[wi], rrrr);
              if (zz)
              {
                z = zz;
                k = 4 * wi;
              }
            }
            k += (__ffs(z) - 1) >> 3;
          }
          /* swap board[depth] <-> board[k]: byte k takes the old board[depth], byte depth takes r */
          uint32_t bd = 0;
#pragma unroll
          for (int wi = 0; wi < 5; wi++)
            if (wi == (depth >> 2))
              bd = (w[wi] >> (8 * (depth & 3))) & 0xffu;
          const uint32_t x = bd ^ (uint32_t)r;
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
          const uint32_t used2 = ((q1.y & 0x00ffffffu) | bit) | ((uint32_t)(depth + 1) << 24);
          const uint32_t up2 = ((q1.z | bit) << 1) & full;
          const uint32_t down2 = (q1.w | bit) >> 1;
          const unsigned at = __shfl_sync(amask, pos, 0);
          if (at + (unsigned)nsurv > (unsigned)cap)
            atomicExch(&ctrl->error, 1);
          else
          {
            const size_t slot = (size_t)at + i;
            dplanes[slot] = make_uint4(w[0], w[1], w[2], w[3]);
            dplanes[(size_t)cap + slot] = make_uint4(w[4], used2, up2, down2);
          }
        }
      }
    }
    __syncwarp();
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  finish_iteration(ctrl, cnt6, src, base, cap);
}

/*
 * Fourth-generation N-Queens kernel: nq_expand3's record and per-child work, but ONE cursor reservation per CTA.
 *
 * ncu on nq_expand3 (profiles/r02_ncu_a_nq_17.txt): 64 % of the warp-stall samples sit on the return of the
 * per-tile atomicAdd on the destination cursor. The work per node had shrunk so far that the 131 072 same-address
 * atomics of a 4 M-parent launch - served at about one per nanosecond by the L2 - WERE the launch time. With the
 * masks in the record the number of children of a parent is popcount(free): known after reading 16 bytes. So every
 * CTA takes a contiguous slice of the batch and works in rounds of <= NQ4_ROUND parents:
 *   phase 1  all threads: read plane 1 of the slice (coalesced), free mask and child count per parent -> shared
 *            memory, per-tile totals, a scan over the tiles, ONE atomicAdd for the round;
 *   phase 2  warp tiles as in nq_expand3 (planes prefetched with cp.async, children listed per lane, one lane per
 *            child), every child written at round base + tile base + index. No atomics, no waiting.
 * Plane 1 is read twice per parent, the second time from L2 (a launch's plane 1 is 64 MB at M = 4 M).
 */
#define NQ4_ROUND 2048

__global__ void __launch_bounds__(BB_CTA_THREADS, 4)
nq_expand4(bb_ctrl *__restrict__ ctrl, const bb_pool pool, const int N, const int G)
{
  const int nbatch = ctrl->n;
  if (nbatch == 0)
    return;
  const int cap = pool.capacity;
  const int src = ctrl->src, dst = 1 - src, base = ctrl->base;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t full = (N >= 32) ? 0xffffffffu : ((1u << N) - 1u);

  __shared__ uint32_t free_s[NQ4_ROUND];           /* safe rows of the next column, 0 for parents at depth N-1 */
  __shared__ int tile_base[NQ4_ROUND / 32 + 1];    /* exclusive prefix of the tiles' child counts */
  __shared__ unsigned round_base;
  __shared__ uint4 tile_all[BB_WARPS_PER_CTA][2][2 * 32];
  __shared__ uint16_t surv_all[BB_WARPS_PER_CTA][BB_TILE_MAX * BB_MAX_QUEENS];
  __shared__ __align__(16) unsigned char stage_all[BB_WARPS_PER_CTA][32 * 48]; /* per lane: one child being built */
  uint16_t *surv_s = surv_all[warp];

  const uint4 *__restrict__ splanes = pool.planes[src];
  uint4 *__restrict__ dplanes = pool.planes[dst];

  unsigned long long cnt6[6] = {0, 0, 0, 0, 0, 0};

  /* this CTA's slice of the batch: a multiple of 32 parents so that tiles never straddle two CTAs */
  const int per = (((nbatch + (int)gridDim.x - 1) / (int)gridDim.x) + 31) & ~31;
  const long long lo_ll = (long long)blockIdx.x * per;
  const int lo = lo_ll < nbatch ? (int)lo_ll : nbatch;
  const int hi = min(nbatch, lo + per);

  for (int r0 = lo; r0 < hi; r0 += NQ4_ROUND)
  {
    const int rcnt = min(NQ4_ROUND, hi - r0);
    const int rtiles = (rcnt + 31) >> 5;
    /* ---- phase 1: free masks and child counts of the round ---- */
    {
      /* all of a thread's loads of the round are issued before the first is used (the loop used to wait for each
       * one in turn: 29 % of the kernel's stall samples); the counters are kept per thread - finish_iteration sums
       * them over the CTA - so the only warp reduction left is the tile's child total */
      constexpr int P1 = NQ4_ROUND / BB_CTA_THREADS;
      uint4 q[P1];
#pragma unroll
      for (int j = 0; j < P1; j++)
      {
        const int i = threadIdx.x + j * BB_CTA_THREADS;
        q[j] = make_uint4(0u, 0xff000000u, 0u, 0u); /* depth 255: no parent here */
        if (i < rcnt)
          q[j] = __ldg(splanes + (size_t)cap + base + r0 + i);
      }
      int akids = 0, asol = 0, acand = 0;
#pragma unroll
      for (int j = 0; j < P1; j++)
      {
        const int i = threadIdx.x + j * BB_CTA_THREADS;
        const int depth = (int)(q[j].y >> 24);
        const bool valid = depth != 255;
        uint32_t used = q[j].y & 0x00ffffffu;
        const bool leaf = (depth + 1 == N);
        uint32_t fr = 0u;
        for (int g = 0; g < G; g++)
        {
          asm volatile("" : "+r"(used)); /* keep the G repetitions as real work, as the reference's -g intends */
          fr = ~(used | q[j].z | q[j].w) & full;
        }
        if (!valid)
          fr = 0u;
        const int kids = __popc(fr);
        const int wkids = leaf ? 0 : kids; /* depth-N children are counted, not stored (nqueens_c.c:103-105) */
        free_s[i] = leaf ? 0u : fr;
        akids += kids; /* exploredTree counts every safe placement, leaves included */
        asol += leaf ? kids : 0;
        acand += valid ? N - depth : 0;
        const int tot = __reduce_add_sync(0xffffffffu, wkids);
        if (lane == 0)
          tile_base[i >> 5] = tot;
      }
      cnt6[0] += akids;
      cnt6[1] += asol;
      cnt6[3] += acand;
    }
    if (threadIdx.x == 0)
      cnt6[2] += rcnt;
    __syncthreads();
    if (warp == 0)
    {
      /* exclusive scan over <= 128 tile totals: 4 per lane */
      int v[4], sum = 0;
#pragma unroll
      for (int k = 0; k < 4; k++)
      {
        const int t = lane * 4 + k;
        v[k] = t < rtiles ? tile_base[t] : 0;
        sum += v[k];
      }
      const int incl = warp_incl_scan(sum, lane);
      int run = incl - sum;
#pragma unroll
      for (int k = 0; k < 4; k++)
      {
        const int t = lane * 4 + k;
        if (t < rtiles)
          tile_base[t] = run;
        run += v[k];
      }
      const int total = __shfl_sync(0xffffffffu, incl, 31);
      if (lane == 0)
      {
        unsigned at = 0;
        if (total)
        {
          at = atomicAdd(&ctrl->tail, (unsigned)total);
          if (at + (unsigned)total > (unsigned)cap)
          {
            atomicExch(&ctrl->error, 1);
            at = 0xffffffffu; /* nothing of this round is written */
          }
        }
        round_base = at;
      }
    }
    __syncthreads();
    const unsigned rbase = round_base;

    /* ---- phase 2: warp tiles, one lane per child, no atomics ---- */
    auto fetch_tile = [&](int t, uint4 *dstbuf) {
      if (t < rtiles)
      {
        const int first = t * 32;
        if (lane < min(32, rcnt - first))
        {
          const size_t node = (size_t)base + r0 + first + lane;
#pragma unroll
          for (int q = 0; q < 2; q++)
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
    fetch_tile(warp, tile_all[warp][0]);
    for (int t = warp; t < rtiles; t += BB_WARPS_PER_CTA)
    {
      const uint4 *tile_s = tile_all[warp][buf];
      fetch_tile(t + BB_WARPS_PER_CTA, tile_all[warp][buf ^ 1]);
      buf ^= 1;
      asm volatile("cp.async.wait_group 1;" ::: "memory");
      __syncwarp();
      uint32_t f = free_s[t * 32 + lane];
      const int nk = __popc(f);
      const int incl = warp_incl_scan(nk, lane);
      const int nsurv = __shfl_sync(0xffffffffu, incl, 31);
      int at = incl - nk;
      while (f)
      {
        const int r = __ffs(f) - 1;
        f &= f - 1u;
        surv_s[at++] = (uint16_t)((lane << 8) | r);
      }
      __syncwarp();
      if (rbase != 0xffffffffu)
        for (int i = lane; i < nsurv; i += 32)
        {
          const int e = surv_s[i];
          const int p = e >> 8, r = e & 0xff;
          const uint4 q0 = tile_s[p], q1 = tile_s[32 + p];
          uint32_t w[5] = {q0.x, q0.y, q0.z, q0.w, q1.x};
          const int depth = (int)(q1.y >> 24);
          const uint32_t bit = 1u << r;
          /* where row r sits in the unplaced part of the permutation: first byte equal to r (padding bytes beyond
           * N are zero and come later) */
          const uint32_t rrrr = (uint32_t)r * 0x01010101u;
          uint32_t z = nq_has_byte(w[4], rrrr);
          int k = 16;
#pragma unroll
          for (int wi = 3; wi >= 0; wi--)
          {
            const uint32_t zz = nq_has_byte(w[wi], rrrr);
            if (zz)
            {
              z = zz;
              k = 4 * wi;
            }
          }
          k += (__ffs(z) - 1) >> 3;
          /* swap board[depth] <-> board[k] (nqueens_c.c:107-111): byte k takes the old board[depth], byte depth r.
           * The child record is assembled in a per-lane staging slot of shared memory and the two bytes are patched
           * with byte stores: patching them in registers needs the words chosen by depth and k, i.e. a select chain
           * over five words for each (a third of this pass's instructions before). */
          const uint32_t used2 = ((q1.y & 0x00ffffffu) | bit) | ((uint32_t)(depth + 1) << 24);
          const uint32_t up2 = ((q1.z | bit) << 1) & full;
          const uint32_t down2 = (q1.w | bit) >> 1;
          unsigned char *st = stage_all[warp] + lane * 48; /* bytes 0..19 = board, 20.. = the three mask words */
          *reinterpret_cast<uint4 *>(st) = q0;
          *reinterpret_cast<uint4 *>(st + 16) = make_uint4(q1.x, used2, up2, down2);
          const unsigned char bd = st[depth];
          st[depth] = (unsigned char)r;
          st[k] = bd;
          const uint4 c0 = *reinterpret_cast<const uint4 *>(st), c1 = *reinterpret_cast<const uint4 *>(st + 16);
          const size_t slot = (size_t)rbase + tile_base[t] + i;
          dplanes[slot] = c0;
          dplanes[(size_t)cap + slot] = c1;
        }
      __syncwarp();
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads(); /* free_s / tile_base are reused by the next round */
  }
  finish_iteration(ctrl, cnt6, src, base, cap);
}

/*
 * Fifth-generation N-Queens kernel: nq_expand4 without its CTA barriers. ncu on nq_expand4 showed the warps of a CTA
 * waiting for each other three times per round (phase 1 -> scan by warp 0 -> phase 2 -> next round) and for the
 * round's loads all at the same moment. Here every WARP works on its own, in rounds of up to NQ5_TILES tiles
 * taken from a counter (ctrl->next_round): 32 x NQ5_TILES parents read (all loads in flight together), free masks to the warp's
 * shared memory, the round's child total by NQ5_TILES warp reductions, ONE atomicAdd on the cursor per round
 * (16 384 per 4 M-parent launch, an eighth of what made per-tile reservation the limiter), then the round's tiles
 * exactly as nq_expand4's phase 2 with a running offset instead of a tile table. No __syncthreads until the
 * counters are summed; a warp waiting for its loads or its reservation is covered by the 31 others on the SM.
 */
#define NQ5_TILES 8

/*
 * TAIL = 1 (the default engine, bb_config.packed = 3): nodes with at most `tail` columns left are not expanded level
 * by level through the pool. In pass A of a launch whichever lane is free takes the next such parent of the batch and
 * enumerates the whole subtree below it: the three masks of the current node in registers, an explicit stack in the
 * warp's shared memory (one frame = three words: rows used, the two diagonal masks and the rows still to try as four
 * packed 20-bit masks; the depth is the number of rows used; a node without rows left is never pushed). Only the
 * counters leave the SM. The explored tree, the solutions, the parents decomposed and the candidates evaluated are
 * what the level-by-level expansion counts (a node with d queens placed is decomposed: candidates + N - d, tree +
 * safe rows; with d = N - 1 its safe rows are solutions; parents = deep parents + (tree - solutions) below them),
 * so every total is the reference's. The deepest levels hold nearly all of the tree (N = 17: 99 % below depth 7)
 * and no longer cost 64 bytes of HBM traffic per node. Pass B is the round loop for the other parents. The kernel
 * argument G carries tail << 16.
 */
#define NQ5_WBUF_TILE1 1024                                   /* byte offsets in a warp's buffer */
#define NQ5_WBUF_SURV 2048
#define NQ5_WBUF_STAGE (NQ5_WBUF_SURV + BB_TILE_MAX * BB_MAX_QUEENS * 2)
#define NQ5_WBUF_BYTES (NQ5_WBUF_STAGE + 32 * 48)
#define NQ5_MAX_TAIL 12 /* stack levels of 3 words x 32 lanes: tail - 2 <= (NQ5_WBUF_BYTES - NQ5_WBUF_TILE1) / 384 */

template <int TAIL>
__global__ void __launch_bounds__(BB_CTA_THREADS, 4)
nq_expand5(bb_ctrl *__restrict__ ctrl, const bb_pool pool, const int N, const int G_and_tail)
{
  pdl_wait_then_release(); /* the plan and the pool are the previous iteration's output */
  const int nbatch = ctrl->n;
  if (nbatch == 0)
    return;
  const int G = TAIL ? (G_and_tail & 0xffff) : G_and_tail;
  const int cut = TAIL ? N - (G_and_tail >> 16) : N; /* parents at depth >= cut are enumerated to the end in place */
  const int cap = pool.capacity;
  const int src = ctrl->src, dst = 1 - src, base = ctrl->base;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t full = (N >= 32) ? 0xffffffffu : ((1u << N) - 1u);

  __shared__ uint32_t free_all[BB_WARPS_PER_CTA][NQ5_TILES * 32]; /* written and read by the same lane */
  /* per warp: two tile buffers (2 planes x 32 parents), the tile's child list, one staging slot per lane; from the
   * second tile buffer on it doubles as the lanes' DFS stacks [level][3 words][lane] during pass A */
  __shared__ __align__(16) unsigned char wbuf_all[BB_WARPS_PER_CTA][NQ5_WBUF_BYTES];
  static_assert((NQ5_WBUF_BYTES - NQ5_WBUF_TILE1) / 384 >= NQ5_MAX_TAIL - 2, "DFS stack levels");
  static_assert(BB_MAX_QUEENS <= 20, "stack frames pack 20-bit masks");
  uint4(*tile_w)[2 * 32] = reinterpret_cast<uint4(*)[2 * 32]>(wbuf_all[warp]);
  uint32_t *free_s = free_all[warp];
  uint16_t *surv_s = reinterpret_cast<uint16_t *>(wbuf_all[warp] + NQ5_WBUF_SURV);
  unsigned char *stage_w = wbuf_all[warp] + NQ5_WBUF_STAGE;
  uint32_t *stk = reinterpret_cast<uint32_t *>(wbuf_all[warp] + NQ5_WBUF_TILE1);
  __shared__ unsigned long long dq_all[BB_WARPS_PER_CTA]; /* pass A: the warp's chunk of the batch */
  __shared__ int dlock_all[BB_WARPS_PER_CTA];

  const uint4 *__restrict__ splanes = pool.planes[src];
  uint4 *__restrict__ dplanes = pool.planes[dst];

  unsigned long long cnt6[6] = {0, 0, 0, 0, 0, 0};
  int akids = 0, asol = 0, acand = 0, apar = 0;
  uint32_t vmask[5]; /* high bit of every board byte below N */
#pragma unroll
  for (int wi = 0; wi < 5; wi++)
  {
    const int nb = min(4, max(0, N - 4 * wi));
    vmask[wi] = nb == 4 ? 0x80808080u : (0x80808080u & ((1u << (8 * nb)) - 1u));
  }

  if (TAIL)
  {
    /* ---- pass A: the deep parents of the whole batch. Whichever LANE is free takes the next parent from its
     * warp's chunk (a 64-bit word in shared memory: chunk base << 32 | next index, one atomicAdd hands out a parent
     * and names its chunk consistently); a lane that finds the chunk used up refills it from the batch-wide counter.
     * Subtree sizes differ by orders of magnitude: with fixed shares, or with a warp that waits at the end of every
     * round, most lanes wait for the unluckiest one. ---- */
    constexpr unsigned CH = 64;
    if (lane == 0)
    {
      dq_all[warp] = (unsigned long long)(unsigned)nbatch << 32; /* an empty chunk */
      dlock_all[warp] = 0;
    }
    __syncwarp();
    int lvl = 0, d = 0;
    uint32_t used = 0u, up = 0u, down = 0u, fr = 0u;
    bool active = false;
    for (;;)
    {
      if (!active)
      {
        const unsigned long long tk = atomicAdd(&dq_all[warp], 1ull);
        const unsigned cb = (unsigned)(tk >> 32), ci = (unsigned)tk;
        if (cb == 0xffffffffu)
          break; /* the batch is handed out */
        const unsigned ccnt = cb < (unsigned)nbatch ? min(CH, (unsigned)nbatch - cb) : 0u;
        if (ci >= ccnt)
        {
          if (atomicCAS(&dlock_all[warp], 0, 1) == 0)
          {
            const unsigned long long now = *reinterpret_cast<volatile unsigned long long *>(&dq_all[warp]);
            if ((unsigned)(now >> 32) == cb) /* nobody refilled it meanwhile */
            {
              const unsigned nb = atomicAdd(&ctrl->next_deep, CH);
              atomicExch(&dq_all[warp], (unsigned long long)(nb < (unsigned)nbatch ? nb : 0xffffffffu) << 32);
            }
            __threadfence_block();
            atomicExch(&dlock_all[warp], 0);
          }
          continue;
        }
        const uint4 pq = __ldg(splanes + (size_t)cap + base + cb + ci);
        d = (int)(pq.y >> 24);
        if (d < cut)
          continue; /* a shallow parent: its children go to the pool in pass B */
        used = pq.y & 0x00ffffffu;
        up = pq.z;
        down = pq.w;
        /* the parent itself is decomposed here (nqueens_c.c:99-117) */
        for (int g = 0; g < G; g++)
        {
          asm volatile("" : "+r"(used));
          fr = ~(used | up | down) & full;
        }
        const int kids = __popc(fr);
        apar++;
        acand += N - d;
        akids += kids;
        if (d + 1 == N)
        {
          asol += kids;
          fr = 0u;
        }
        lvl = 0;
        active = fr != 0u;
        continue;
      }
      if (fr == 0u)
      {
        if (lvl == 0)
        {
          active = false;
          continue;
        }
        lvl--; /* the nearest ancestor that still has rows to try: frames without any are never pushed */
        /* a frame is three words: the four 20-bit masks packed, the depth is the number of rows used */
        const uint32_t w0 = stk[(lvl * 3 + 0) * 32 + lane], w1 = stk[(lvl * 3 + 1) * 32 + lane];
        down = stk[(lvl * 3 + 2) * 32 + lane];
        used = w0 & 0xfffffu;
        up = w1 & 0xfffffu;
        fr = (w0 >> 20) | ((w1 >> 20) << 12);
        d = __popc(used);
        /* a pushed frame always has a row left, so the lane goes on to place it in this same pass of the loop: a
         * pass that only backtracks would keep the lane out of the instructions below, which the rest of the warp
         * executes anyway */
      }
      /* place the lowest safe row: the child with d + 1 queens is decomposed at once */
      const uint32_t bit = fr & (0u - fr);
      fr ^= bit;
      uint32_t nu = used | bit;
      const uint32_t nup = ((up | bit) << 1) & full, ndn = (down | bit) >> 1;
      uint32_t cfr = 0u;
      for (int g = 0; g < G; g++)
      {
        asm volatile("" : "+r"(nu));
        cfr = ~(nu | nup | ndn) & full;
      }
      const int ckids = __popc(cfr);
      acand += N - (d + 1);
      akids += ckids;
      if (d + 2 == N)
        asol += ckids; /* its safe rows complete the board */
      else if (cfr)
      {
        if (fr) /* nothing to come back to otherwise */
        {
          stk[(lvl * 3 + 0) * 32 + lane] = used | (fr << 20);
          stk[(lvl * 3 + 1) * 32 + lane] = up | ((fr >> 12) << 20);
          stk[(lvl * 3 + 2) * 32 + lane] = down;
          lvl++;
        }
        d++;
        used = nu;
        up = nup;
        down = ndn;
        fr = cfr;
      }
    }
    /* every node below a deep parent that is not a solution was decomposed: parents = roots + (tree - solutions) */
    apar += akids - asol;
    __syncwarp();
  }

  /* rounds are handed out through a counter, and get smaller towards the end of the batch (each warp asks for
   * what it saw left / the number of warps, at most NQ5_TILES tiles, at least one): with equal static slices the
   * launch ended with most SMs waiting for the slowest warps to finish whole rounds */
  const int nw = (int)gridDim.x * BB_WARPS_PER_CTA;
  const int T = (nbatch + 31) >> 5;
  int seen = 0;

  for (;;)
  {
    const int want = min(NQ5_TILES, max(1, (T - seen) / nw));
    int t0 = 0;
    if (lane == 0)
      t0 = (int)atomicAdd(&ctrl->next_round, (unsigned)want);
    t0 = __shfl_sync(0xffffffffu, t0, 0);
    if (t0 >= T)
      break;
    seen = t0 + want;
    const int r0 = t0 * 32;
    const int rcnt = min(want * 32, nbatch - r0);
    const int rtiles = (rcnt + 31) >> 5;
    auto fetch_tile = [&](int t, uint4 *dstbuf) {
      if (t < rtiles)
      {
        const int first = t * 32;
        if (lane < min(32, rcnt - first))
        {
          const size_t node = (size_t)base + r0 + first + lane;
#pragma unroll
          for (int q = 0; q < 2; q++)
          {
            const unsigned sa = (unsigned)__cvta_generic_to_shared(dstbuf + q * 32 + lane);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(splanes + (size_t)q * cap + node)
                         : "memory");
          }
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    fetch_tile(0, tile_w[0]);
    /* ---- the round's free masks and child total ---- */
    int total = 0;
    {
      uint4 q[NQ5_TILES];
#pragma unroll
      for (int j = 0; j < NQ5_TILES; j++)
      {
        const int i = j * 32 + lane;
        q[j] = make_uint4(0u, 0xff000000u, 0u, 0u); /* depth 255: no parent here */
        if (i < rcnt)
          q[j] = __ldg(splanes + (size_t)cap + base + r0 + i);
      }
#pragma unroll
      for (int j = 0; j < NQ5_TILES; j++)
      {
        if (j * 32 >= rcnt) /* warp-uniform */
          break;
        const int depth = (int)(q[j].y >> 24);
        const bool valid = depth != 255 && !(TAIL && depth >= cut); /* deep parents were counted above */
        uint32_t used = q[j].y & 0x00ffffffu;
        const bool leaf = (depth + 1 == N);
        uint32_t fr = 0u;
        for (int g = 0; g < G; g++)
        {
          asm volatile("" : "+r"(used)); /* keep the G repetitions as real work, as the reference's -g intends */
          fr = ~(used | q[j].z | q[j].w) & full;
        }
        if (!valid)
          fr = 0u;
        const int kids = __popc(fr);
        const int wkids = leaf ? 0 : kids; /* depth-N children are counted, not stored (nqueens_c.c:103-105) */
        akids += kids; /* exploredTree counts every safe placement, leaves included */
        asol += leaf ? kids : 0;
        acand += valid ? N - depth : 0;
        apar += valid ? 1 : 0;
        /* the lane's first index in the tile's child list is found here, where NQ5_TILES independent scans overlap,
         * and travels with the mask: rows in bits 0..19, index (< 640) above */
        const int incl = warp_incl_scan(wkids, lane);
        free_s[j * 32 + lane] = (leaf ? 0u : fr) | ((uint32_t)(incl - wkids) << 20);
        total += __shfl_sync(0xffffffffu, incl, 31);
      }
    }
    unsigned rbase = 0;
    if (total)
    {
      if (lane == 0)
      {
        rbase = atomicAdd(&ctrl->tail, (unsigned)total);
        if (rbase + (unsigned)total > (unsigned)cap)
        {
          atomicExch(&ctrl->error, 1);
          rbase = 0xffffffffu; /* nothing of this round is written */
        }
      }
      rbase = __shfl_sync(0xffffffffu, rbase, 0);
    }

    /* ---- the round's tiles, one lane per child, no atomics ---- */
    int buf = 0, off = 0;
    for (int t = 0; t < rtiles; t++)
    {
      const uint4 *tile_s = tile_w[buf];
      fetch_tile(t + 1, tile_w[buf ^ 1]);
      buf ^= 1;
      asm volatile("cp.async.wait_group 1;" ::: "memory");
      __syncwarp();
      const uint32_t packed = free_s[t * 32 + lane];
      uint32_t f = packed & 0xfffffu;
      int at = (int)(packed >> 20);
      const int nsurv = __shfl_sync(0xffffffffu, at + __popc(f), 31);
      while (f)
      {
        const int r = __ffs(f) - 1;
        f &= f - 1u;
        surv_s[at++] = (uint16_t)((lane << 8) | r);
      }
      __syncwarp();
      if (rbase != 0xffffffffu)
        for (int i = lane; i < nsurv; i += 32)
        {
          const int e = surv_s[i];
          const int p = e >> 8, r = e & 0xff;
          const uint4 q0 = tile_s[p], q1 = tile_s[32 + p];
          uint32_t w[5] = {q0.x, q0.y, q0.z, q0.w, q1.x};
          const int depth = (int)(q1.y >> 24);
          const uint32_t bit = 1u << r;
          /* where row r sits in the board: the board holds every row once in bytes 0..N-1 (the padding bytes are
           * masked off), so exactly one word has a matching byte; the flag of byte j of word i is moved to bit
           * 8j + 7 - i and the lowest set bit is the match (the zero-byte test may also flag bytes ABOVE a true
           * match in the same word, never below) */
          const uint32_t rrrr = (uint32_t)r * 0x01010101u;
          uint32_t mm = 0u;
#pragma unroll
          for (int wi = 0; wi < 5; wi++)
          {
            const uint32_t x = w[wi] ^ rrrr;
            mm |= ((x - 0x01010101u) & ~x & vmask[wi]) >> wi;
          }
          const int hb = __ffs(mm) - 1;
          const int k = 4 * (7 - (hb & 7)) + (hb >> 3);
          const uint32_t used2 = ((q1.y & 0x00ffffffu) | bit) | ((uint32_t)(depth + 1) << 24);
          const uint32_t up2 = ((q1.z | bit) << 1) & full;
          const uint32_t down2 = (q1.w | bit) >> 1;
          unsigned char *st = stage_w + lane * 48;
          *reinterpret_cast<uint4 *>(st) = q0;
          *reinterpret_cast<uint4 *>(st + 16) = make_uint4(q1.x, used2, up2, down2);
          const unsigned char bd = st[depth];
          st[depth] = (unsigned char)r;
          st[k] = bd;
          const uint4 c0 = *reinterpret_cast<const uint4 *>(st), c1 = *reinterpret_cast<const uint4 *>(st + 16);
          const size_t slot = (size_t)rbase + off + i;
          dplanes[slot] = c0;
          dplanes[(size_t)cap + slot] = c1;
        }
      off += nsurv;
      __syncwarp();
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
  cnt6[0] = (unsigned long long)akids;
  cnt6[1] = (unsigned long long)asol;
  cnt6[2] = (unsigned long long)apar;
  cnt6[3] = (unsigned long long)acand;
  finish_iteration(ctrl, cnt6, src, base, cap);
}

} // namespace bbk
#endif
