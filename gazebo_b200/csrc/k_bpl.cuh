// k_bpl.cuh — broadphase of a LARGE world (more geoms than one CTA's shared memory can sort and mark: C4, a 20 000-geom
// rubble pile).  Replaces dxHashSpace::collide (collision_space.cpp:427-593, with its n x n / 8-byte `tested` matrix
// on the stack, :486-488), collideAABBs (collision_space_internal.h:48-82) and the filters of
// ODEPhysics::CollisionCallback (ODEPhysics.cc:948-1008) like k_broadphase does for small worlds: sort the geoms by
// the lower x bound of their AABB, sweep forward from every sorted position while the x intervals overlap, give the
// survivors the full test.  Both are exact AABB-overlap filters, so the pair SET is the reference's; the list comes
// out sorted by (i, j), the canonical order of the deterministic row-order mode.
//   k_bpl_tilesort  one CTA per 1024 geoms: bitonic sort of (lower x bound, geom) in shared memory
//   k_bpl_merge     one thread per geom: its final place = its place in its tile + the number of smaller elements in
//                   every other tile (binary searches); writes the sorted arrays and the upper bounds in sorted order
//   k_bpl_span      one thread per sorted position: the end of its x interval (binary search), its followers cut into
//                   work items of GZ_BPL_SEG; intervals of more than 32 items (the ground plane, the pen's walls) go to
//                   the queue
//   k_bpl_sweep     one warp per work item, all SMs: y / z bounds of the followers from a side array in sorted order
//                   (coalesced, four per lane in flight), the full test for the survivors; a pair is counted under its
//                   lower-index geom and logged
//   k_bpl_tail      the queue: all warps share the followers of each queued position (without it the plane's warp
//                   walked all 20 000 geoms alone and WAS the kernel's duration)
//   k_bpl_scan      one CTA: exclusive scan of the counts, n_pairs, capacity flag
//   k_bpl_place     one thread per logged pair: the higher-index partner to its place in the lower geom's list
//                   (first version: a second sweep, 0.19 ms at C4)
//   k_bpl_emit      one warp per geom: puts its partners in ascending order by counting; long lists (the plane's:
//                   thousands) go to k_bpl_emit_big: one CTA each, a bitmap over the geoms and its prefix popcount
#pragma once

namespace gz {

#define GZ_BPL_TILE 1024
#define GZ_BPL_SEG 512      // followers per work item of k_bpl_sweep (a multiple of 32 x GZ_BPL_U)
#define GZ_BPL_NBIG 8       // lists k_bpl_emit can order through a bitmap at the same time
#define GZ_BPL_BIG 256      // ... from this length on

struct BplBuf {
  double *key, *maxx, *tkey;
  double *yz;   // [n][4] lower / upper y and z bounds in sorted order: the sweep's pre-filter reads them coalesced
  int *idx, *cnt, *fill, *tmp, *tidx;
  int *queue;   // [2 * n] (sorted position, first follower not tested yet)
  int *ctl;     // [0] queue length, [1] long lists claimed, [2 ..] their geoms, [2 + GZ_BPL_NBIG] pairs logged, [3 + GZ_BPL_NBIG] work items
  int *bits;    // [GZ_BPL_NBIG][2][(n + 31) / 32 + 1] bitmap and its exclusive prefix popcount
  int npad;
  int *item;    // [item_cap] k_bpl_sweep's work items: sorted position | segment << 27
  int item_cap, big;   // big: GZ_BPL_BIG unless a developer switch says otherwise
};

// total order of the sort: by key, ties by geom index (padding: key +inf, index >= n, sorts last)
__device__ __forceinline__ bool bpl_less(double ka, int ia, double kb, int ib) { return ka < kb || (ka == kb && ia < ib); }

__global__ void __launch_bounds__(GZ_BPL_TILE) k_bpl_tilesort(DevBatch B, const int w, BplBuf Q) {
  __shared__ double sk[GZ_BPL_TILE];
  __shared__ int si[GZ_BPL_TILE];
  const WorldInfo wi = B.world[w];
  const int n = wi.ng, tid = threadIdx.x, g = blockIdx.x * GZ_BPL_TILE + tid;
  const Geom *G = B.geom + wi.geom0;
  double key = (g < n) ? G[g].aabb[0] : GZ_INF;
  if (key != key) key = GZ_INF;   // a NaN bound (a body that left the number line) must not break the total order
  sk[tid] = key;
  si[tid] = g;
  if (g <= n) { Q.cnt[g] = 0; if (g < n) Q.fill[g] = 0; }
  if (g == 0) { Q.ctl[0] = 0; Q.ctl[1] = 0; Q.ctl[2 + GZ_BPL_NBIG] = 0; Q.ctl[3 + GZ_BPL_NBIG] = 0; B.wstep[w].overflow = 0; }   // (k_broadphase's role: the step's capacity mask starts here)
  __syncthreads();
  for (int k = 2; k <= GZ_BPL_TILE; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      const int l = tid ^ j;
      if (l > tid) {
        const double a = sk[tid], b = sk[l];
        const int ia = si[tid], ib = si[l];
        const bool gt = bpl_less(b, ib, a, ia);
        if (gt == ((tid & k) == 0)) { sk[tid] = b; sk[l] = a; si[tid] = ib; si[l] = ia; }
      }
      __syncthreads();
    }
  Q.tkey[g] = sk[tid]; Q.tidx[g] = si[tid];
}

// (four CTAs of 256 per tile: 20 CTAs of 1024 kept 20 of the 148 SMs busy)
__global__ void __launch_bounds__(256) k_bpl_merge(DevBatch B, const int w, BplBuf Q) {
  const WorldInfo wi = B.world[w];
  const int n = wi.ng, tid = (blockIdx.x & 3) * 256 + threadIdx.x, t = blockIdx.x >> 2, ntile = gridDim.x >> 2;
  const Geom *G = B.geom + wi.geom0;
  const double k = Q.tkey[t * GZ_BPL_TILE + tid];
  const int i = Q.tidx[t * GZ_BPL_TILE + tid];
  int rank = tid;
  // binary searches of four tiles at a time, step by step side by side: a search is ten dependent loads from L2
  // (one tile after the other: 70 us at C4's 20 tiles)
  for (int u0 = 0; u0 < ntile; u0 += 4) {
    const double *uk[4];
    const int *ui[4];
    int lo[4];
#pragma unroll
    for (int x = 0; x < 4; x++) {
      const int u = min(u0 + x, ntile - 1);
      uk[x] = Q.tkey + u * GZ_BPL_TILE; ui[x] = Q.tidx + u * GZ_BPL_TILE; lo[x] = 0;
    }
    for (int st = GZ_BPL_TILE >> 1; st > 0; st >>= 1) {
#pragma unroll
      for (int x = 0; x < 4; x++) { const int m = lo[x] + st - 1; if (bpl_less(uk[x][m], ui[x][m], k, i)) lo[x] += st; }
    }
#pragma unroll
    for (int x = 0; x < 4; x++) {   // lo = elements of the tile below (k, i), at most TILE - 1 so far
      if (bpl_less(uk[x][lo[x]], ui[x][lo[x]], k, i)) lo[x]++;
      if (u0 + x < ntile && u0 + x != t) rank += lo[x];
    }
  }
  Q.key[rank] = k; Q.idx[rank] = i;
  if (i < n) {   // (real geoms take the places below n: padding sorts last)
    Q.maxx[rank] = G[i].aabb[1];
    for (int c = 0; c < 4; c++) Q.yz[4 * (size_t)rank + c] = G[i].aabb[2 + c];
  }
}

#define GZ_BPL_U 4   // followers per lane and trip: their keys and bounds are loaded together (a trip is one L2 latency, not four)

// Sorted positions q0 + 32 u + lane (u < GZ_BPL_U) against position p's geom ip: x from the sorted keys, y and z from the
// sorted side array (collideAABBs' own comparisons, collision_space_internal.h:66-72: of the ~800 followers that overlap a
// pile's geom on x, a few dozen overlap it on y and z as well), the full test for the survivors.  A pair that passes is
// counted under its lower geom and logged (in the world's slice of B.pair, which k_bpl_emit fills only later):
// k_bpl_place puts the logged pairs into the per-geom lists once the counts are scanned - the sweep runs once.
// Returns whether all of this lane's positions are still inside p's x interval.
__device__ __forceinline__ bool bpl_follow(const DevBatch &B, const WorldInfo &wi, const Geom *G, const BplBuf &Q, const int ip,
                                           const double mx, const double *pyz, const int q0, const int n) {
  double k[GZ_BPL_U];
  double2 lo2[GZ_BPL_U], hi2[GZ_BPL_U];
#pragma unroll
  for (int u = 0; u < GZ_BPL_U; u++) {
    const int q = min(q0 + 32 * u, n - 1);
    k[u] = Q.key[q];
    lo2[u] = reinterpret_cast<const double2 *>(Q.yz)[2 * (size_t)q];       // y bounds
    hi2[u] = reinterpret_cast<const double2 *>(Q.yz)[2 * (size_t)q + 1];   // z bounds
  }
  bool all_in = true;
#pragma unroll
  for (int u = 0; u < GZ_BPL_U; u++) {
    const int q = q0 + 32 * u;
    const bool in = (q < n) && k[u] <= mx;
    all_in = all_in && in;
    if (!in) continue;
    if (pyz[0] > lo2[u].y || pyz[1] < lo2[u].x || pyz[2] > hi2[u].y || pyz[3] < hi2[u].x) continue;
    const int iq = Q.idx[q];
    const int lo = min(ip, iq), hi = max(ip, iq);
    if (pair_passes(B, G[lo], G[hi])) {
      atomicAdd(&Q.cnt[lo], 1);
      const int at = atomicAdd(&Q.ctl[2 + GZ_BPL_NBIG], 1);
      if (at < wi.pair_cap) { B.pair[2 * (wi.pair0 + at)] = lo; B.pair[2 * (wi.pair0 + at) + 1] = hi; }   // (beyond: k_bpl_scan raises the flag)
    }
  }
  return all_in;
}

// One thread per sorted position: where its x interval ends (binary search: the keys are sorted), and its followers cut
// into work items of GZ_BPL_SEG; an interval of more than 32 items (the ground plane, the pen's walls), or one the item
// table cannot take, goes to the queue of k_bpl_tail.
__global__ void __launch_bounds__(256) k_bpl_span(DevBatch B, const int w, BplBuf Q) {
  const int n = B.world[w].ng;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n - 1; p += gridDim.x * blockDim.x) {
    const double mx = Q.maxx[p];
    int lo = p + 1, hi = n;   // first position in (p, n) whose key is > mx
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (Q.key[mid] <= mx) lo = mid + 1; else hi = mid; }
    const int len = lo - (p + 1);
    if (len <= 0) continue;
    const int s = (len + GZ_BPL_SEG - 1) / GZ_BPL_SEG;
    int at = -1;
    if (s <= 32) { at = atomicAdd(&Q.ctl[3 + GZ_BPL_NBIG], s); if (at + s > Q.item_cap) at = -1; }
    if (at < 0) { const int e = atomicAdd(&Q.ctl[0], 1); Q.queue[2 * e] = p; Q.queue[2 * e + 1] = p + 1; continue; }
    for (int k = 0; k < s; k++) Q.item[at + k] = p | (k << 27);
  }
}

// One warp per work item, all SMs (first version: one warp per position walking its whole interval - the warps of the
// longest intervals were the kernel's duration, 0.18 ms at C4 with the SMs 5 % busy).
__global__ void __launch_bounds__(256) k_bpl_sweep(DevBatch B, const int w, BplBuf Q) {
  const WorldInfo wi = B.world[w];
  const int n = wi.ng, lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarp = (gridDim.x * blockDim.x) >> 5;
  const Geom *G = B.geom + wi.geom0;
  const int nitem = min(Q.ctl[3 + GZ_BPL_NBIG], Q.item_cap);
  for (int it = warp; it < nitem; it += nwarp) {
    const int code = Q.item[it], p = code & ((1 << 27) - 1), k = (unsigned)code >> 27;
    const int ip = Q.idx[p];
    const double mx = Q.maxx[p];
    const double pyz[4] = {Q.yz[4 * (size_t)p], Q.yz[4 * (size_t)p + 1], Q.yz[4 * (size_t)p + 2], Q.yz[4 * (size_t)p + 3]};
    const int qb = p + 1 + k * GZ_BPL_SEG;
    for (int q0 = qb; q0 < min(qb + GZ_BPL_SEG, n); q0 += 32 * GZ_BPL_U) {
      const bool in = bpl_follow(B, wi, G, Q, ip, mx, pyz, q0 + lane, n);
      if (!__all_sync(0xffffffffu, in)) break;   // sorted: nothing further overlaps p on x
    }
  }
}

__global__ void __launch_bounds__(256) k_bpl_tail(DevBatch B, const int w, BplBuf Q) {
  const WorldInfo wi = B.world[w];
  const int n = wi.ng, lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarp = (gridDim.x * blockDim.x) >> 5;
  const Geom *G = B.geom + wi.geom0;
  const int nq = Q.ctl[0];
  for (int e = 0; e < nq; e++) {
    const int p = Q.queue[2 * e], qs = Q.queue[2 * e + 1];
    const int ip = Q.idx[p];
    const double mx = Q.maxx[p];
    const double pyz[4] = {Q.yz[4 * (size_t)p], Q.yz[4 * (size_t)p + 1], Q.yz[4 * (size_t)p + 2], Q.yz[4 * (size_t)p + 3]};
    for (int q0 = qs + 32 * GZ_BPL_U * warp; q0 < n; q0 += 32 * GZ_BPL_U * nwarp) {
      const bool in = bpl_follow(B, wi, G, Q, ip, mx, pyz, q0 + lane, n);
      if (!__all_sync(0xffffffffu, in)) break;   // sorted: this warp's later chunks are outside as well
    }
  }
}

__global__ void __launch_bounds__(1024) k_bpl_scan(DevBatch B, const int w, BplBuf Q) {
  __shared__ int wsum[32];
  const WorldInfo wi = B.world[w];
  const int n = wi.ng, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int carry = 0;   // (the same in every thread)
  for (int i0 = 0; i0 < n; i0 += 1024 * 8) {   // eight counts per thread and trip
    const int ib = i0 + tid * 8;
    int c[8];
#pragma unroll
    for (int x = 0; x < 8; x++) c[x] = (ib + x < n) ? Q.cnt[ib + x] : 0;
    int sum = 0;
#pragma unroll
    for (int x = 0; x < 8; x++) { const int cc = c[x]; c[x] = sum; sum += cc; }
    int v = sum;
    for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += t; }
    if (lane == 31) wsum[warp] = v;
    __syncthreads();
    if (warp == 0) {
      int s = wsum[lane];
      for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, s, d); if (lane >= d) s += t; }
      wsum[lane] = s;
    }
    __syncthreads();
    const int before = carry + (warp > 0 ? wsum[warp - 1] : 0) + v - sum;
#pragma unroll
    for (int x = 0; x < 8; x++) if (ib + x < n) Q.cnt[ib + x] = before + c[x];
    carry += wsum[31];
    __syncthreads();
  }
  if (tid == 0) {
    int total = carry;
    Q.cnt[n] = total;
    WorldStep &ws = B.wstep[w];
    if (total > wi.pair_cap) { ws.overflow |= 1; total = wi.pair_cap; }
    ws.n_pairs = total;
  }
}

// One thread per logged pair: its place in the list of its lower geom (cnt holds the exclusive scan now).
__global__ void __launch_bounds__(256) k_bpl_place(DevBatch B, const int w, BplBuf Q) {
  const WorldInfo wi = B.world[w];
  const int total = min(Q.ctl[2 + GZ_BPL_NBIG], wi.pair_cap);
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const int lo = B.pair[2 * (wi.pair0 + e)], hi = B.pair[2 * (wi.pair0 + e) + 1];
    const int at = Q.cnt[lo] + atomicAdd(&Q.fill[lo], 1);
    if (at < wi.pair_cap) Q.tmp[at] = hi;
  }
}

__global__ void __launch_bounds__(256) k_bpl_emit(DevBatch B, const int w, BplBuf Q) {
  const WorldInfo wi = B.world[w];
  const int n = wi.ng, lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarp = (gridDim.x * blockDim.x) >> 5;
  for (int lo = warp; lo < n; lo += nwarp) {
    const int base = Q.cnt[lo], L = min(Q.cnt[lo + 1], wi.pair_cap) - base;
    if (L >= Q.big) {   // a long list (the plane's: thousands): a whole CTA orders it, k_bpl_emit_big
      int slot = -1;
      if (lane == 0) slot = atomicAdd(&Q.ctl[1], 1);
      slot = __shfl_sync(0xffffffffu, slot, 0);
      if (slot < GZ_BPL_NBIG) { if (lane == 0) Q.ctl[2 + slot] = lo; continue; }
      // (more long lists than bitmaps: counting still gives the right order)
    }
    for (int e = lane; e < L; e += 32) {
      const int hi = Q.tmp[base + e];
      int rank = 0;
      for (int f = 0; f < L; f++) rank += Q.tmp[base + f] < hi;   // partners of a geom are distinct
      B.pair[2 * (wi.pair0 + base + rank)] = wi.geom0 + lo;
      B.pair[2 * (wi.pair0 + base + rank) + 1] = wi.geom0 + hi;
    }
  }
}

// One CTA per long partner list: a bitmap over the geoms, its prefix popcount (block scan), rank = set bits below the
// partner's (partners of a geom are distinct).
__global__ void __launch_bounds__(1024) k_bpl_emit_big(DevBatch B, const int w, BplBuf Q) {
  __shared__ int wsum[32];
  __shared__ int carry_s;
  const int slot = blockIdx.x;
  if (slot >= min(Q.ctl[1], GZ_BPL_NBIG)) return;
  const WorldInfo wi = B.world[w];
  const int n = wi.ng, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwords = (n + 31) / 32 + 1;
  const int lo = Q.ctl[2 + slot];
  const int base = Q.cnt[lo], L = min(Q.cnt[lo + 1], wi.pair_cap) - base;
  int *bits = Q.bits + (size_t)slot * 2 * nwords, *pre = bits + nwords;
  for (int k = tid; k < nwords; k += 1024) bits[k] = 0;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int e = tid; e < L; e += 1024) { const int hi = Q.tmp[base + e]; atomicOr(&bits[hi >> 5], 1 << (hi & 31)); }
  __syncthreads();
  for (int k0 = 0; k0 < nwords; k0 += 1024) {
    const int k = k0 + tid;
    const int c = k < nwords ? __popc(__ldcg(&bits[k])) : 0;
    int v = c;
    for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += t; }
    if (lane == 31) wsum[warp] = v;
    __syncthreads();
    if (warp == 0) {
      int sx = wsum[lane];
      for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, sx, d); if (lane >= d) sx += t; }
      wsum[lane] = sx;
    }
    __syncthreads();
    if (k < nwords) pre[k] = carry_s + (warp > 0 ? wsum[warp - 1] : 0) + v - c;
    __syncthreads();
    if (tid == 0) carry_s += wsum[31];
    __syncthreads();
  }
  for (int e = tid; e < L; e += 1024) {
    const int hi = Q.tmp[base + e];
    const int rank = pre[hi >> 5] + __popc(__ldcg(&bits[hi >> 5]) & ((1u << (hi & 31)) - 1u));
    B.pair[2 * (wi.pair0 + base + rank)] = wi.geom0 + lo;
    B.pair[2 * (wi.pair0 + base + rank) + 1] = wi.geom0 + hi;
  }
}

}  // namespace gz
