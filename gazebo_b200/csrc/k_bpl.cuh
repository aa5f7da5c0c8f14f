// k_bpl.cuh — broadphase of a LARGE world (more geoms than one CTA's shared memory can sort and mark: C4, a 20 000-geom
// rubble pile).  Replaces dxHashSpace::collide (collision_space.cpp:427-593, with its n x n / 8-byte `tested` matrix
// on the stack, :486-488), collideAABBs (collision_space_internal.h:48-82) and the filters of
// ODEPhysics::CollisionCallback (ODEPhysics.cc:948-1008) like k_broadphase does for small worlds: sort the geoms by
// the lower x bound of their AABB, sweep forward from every sorted position while the x intervals overlap, give the
// survivors the full test.  Both are exact AABB-overlap filters, so the pair SET is the reference's; the list comes
// out sorted by (i, j), the canonical order of the deterministic row-order mode.
//   k_bpl_sort   one CTA: bitonic sort of (lower x bound, geom) in global memory, then the upper bounds in sorted order
//   k_bpl_sweep  one warp per sorted position, all SMs; pass 0 counts the pairs of every lower-index geom,
//                pass 1 stores the higher-index partners at the places the scan assigned
//   k_bpl_scan   one CTA: exclusive scan of the counts, n_pairs, capacity flag
//   k_bpl_emit   one warp per geom: ranks its partners by counting (lists are short, tens to a few hundred) and
//                writes the pairs in ascending order
#pragma once

namespace gz {

struct BplBuf { double *key, *maxx; int *idx, *cnt, *fill, *tmp; int npad; };

__global__ void __launch_bounds__(1024) k_bpl_sort(DevBatch B, const int w, BplBuf Q) {
  const WorldInfo wi = B.world[w];
  const int n = wi.ng, tid = threadIdx.x, nt = blockDim.x, npad = Q.npad;
  const Geom *G = B.geom + wi.geom0;
  for (int i = tid; i < npad; i += nt) { Q.key[i] = (i < n) ? G[i].aabb[0] : GZ_INF; Q.idx[i] = i; }
  for (int i = tid; i <= n; i += nt) { Q.cnt[i] = 0; if (i < n) Q.fill[i] = 0; }
  __syncthreads();
  for (int k = 2; k <= npad; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < npad; i += nt) {
        const int l = i ^ j;
        if (l > i) {
          const double a = Q.key[i], b = Q.key[l];
          const int ia = Q.idx[i], ib = Q.idx[l];
          const bool gt = a > b || (a == b && ia > ib);   // padding (index >= n, key +inf) sorts last
          if (gt == ((i & k) == 0)) { Q.key[i] = b; Q.key[l] = a; Q.idx[i] = ib; Q.idx[l] = ia; }
        }
      }
      __syncthreads();
    }
  for (int p = tid; p < n; p += nt) Q.maxx[p] = G[Q.idx[p]].aabb[1];
  if (tid == 0) B.wstep[w].overflow = 0;   // (k_broadphase's role: the step's capacity mask starts here)
}

template <int PASS>
__global__ void __launch_bounds__(256) k_bpl_sweep(DevBatch B, const int w, BplBuf Q) {
  const WorldInfo wi = B.world[w];
  const int n = wi.ng, lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarp = (gridDim.x * blockDim.x) >> 5;
  const Geom *G = B.geom + wi.geom0;
  for (int p = warp; p < n - 1; p += nwarp) {
    const int ip = Q.idx[p];
    const double mx = Q.maxx[p];
    for (int q0 = p + 1; q0 < n; q0 += 32) {
      const int q = q0 + lane;
      const bool in = (q < n) && Q.key[q] <= mx;
      if (in) {
        const int iq = Q.idx[q];
        const int lo = min(ip, iq), hi = max(ip, iq);
        if (pair_passes(B, G[lo], G[hi])) {
          if (PASS == 0) atomicAdd(&Q.cnt[lo], 1);
          else {
            const int at = Q.cnt[lo] + atomicAdd(&Q.fill[lo], 1);   // cnt holds the exclusive scan now
            if (at < wi.pair_cap) Q.tmp[at] = hi;
          }
        }
      }
      if (!__all_sync(0xffffffffu, in)) break;   // sorted: nothing further overlaps p on x
    }
  }
}

__global__ void __launch_bounds__(1024) k_bpl_scan(DevBatch B, const int w, BplBuf Q) {
  __shared__ int wsum[32];
  __shared__ int carry_s;
  const WorldInfo wi = B.world[w];
  const int n = wi.ng, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int i0 = 0; i0 < n; i0 += 1024) {
    const int i = i0 + tid;
    const int c = (i < n) ? Q.cnt[i] : 0;
    int v = c;
    for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += t; }
    if (lane == 31) wsum[warp] = v;
    __syncthreads();
    if (warp == 0) {
      int s = wsum[lane];
      for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, s, d); if (lane >= d) s += t; }
      wsum[lane] = s;
    }
    __syncthreads();
    const int before = carry_s + (warp > 0 ? wsum[warp - 1] : 0) + v - c;
    if (i < n) Q.cnt[i] = before;
    __syncthreads();
    if (tid == 0) carry_s += wsum[31];
    __syncthreads();
  }
  if (tid == 0) {
    int total = carry_s;
    Q.cnt[n] = total;
    WorldStep &ws = B.wstep[w];
    if (total > wi.pair_cap) { ws.overflow |= 1; total = wi.pair_cap; }
    ws.n_pairs = total;
  }
}

__global__ void __launch_bounds__(256) k_bpl_emit(DevBatch B, const int w, BplBuf Q) {
  const WorldInfo wi = B.world[w];
  const int n = wi.ng, lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarp = (gridDim.x * blockDim.x) >> 5;
  for (int lo = warp; lo < n; lo += nwarp) {
    const int base = Q.cnt[lo], L = min(Q.cnt[lo + 1], wi.pair_cap) - base;
    for (int e = lane; e < L; e += 32) {
      const int hi = Q.tmp[base + e];
      int rank = 0;
      for (int f = 0; f < L; f++) rank += Q.tmp[base + f] < hi;   // partners of a geom are distinct
      B.pair[2 * (wi.pair0 + base + rank)] = wi.geom0 + lo;
      B.pair[2 * (wi.pair0 + base + rank) + 1] = wi.geom0 + hi;
    }
  }
}

}  // namespace gz
