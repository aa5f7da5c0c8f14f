// k_large.cuh — the one-CTA-per-world kernels of the step for a LARGE world (C4: 20 000 bodies, ~50 000 candidate
// pairs, ~34 000 contacts): k_contacts and k_assemble1 take 0.9 and 1.4 ms there on one CTA of 1024 threads (latency of
// dependent global loads, 20-130 trips per thread), a quarter and more of the step.  Same work, same results, spread over a
// cooperative grid with the grid barrier of k_pgsl.cuh between the phases; one launch per world.
#pragma once
#include <cooperative_groups.h>

namespace gz {

// k_contacts for a large world: every CTA takes a contiguous slice of the pair list.  Phase 1: the slice's contact
// count; phase 2: the slice's place = the sum of the slices before it, then a block scan over the slice and the
// contacts written in pair order (the reference's order: ODEPhysics::CollisionCallback appends as pairs come).
__global__ void __launch_bounds__(256) k_contacts_large(DevBatch B, const int w, int *part, unsigned *bar) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  unsigned bar_target = *bar;
  grid.sync();
  const WorldInfo wi = B.world[w];
  WorldStep &ws = B.wstep[w];
  const int np = ws.n_pairs, G = gridDim.x, tid = threadIdx.x;
  const int gt = blockIdx.x * blockDim.x + tid, GT = G * blockDim.x;
  const int per = ((np + G - 1) / G + 255) / 256 * 256;   // pairs per CTA: whole tiles
  const int k_lo = min(np, blockIdx.x * per), k_hi = min(np, k_lo + per);
  __shared__ int warp_sums[8];
  __shared__ int carry;
  const int lane = tid & 31, warp = tid >> 5;
  for (int b = gt; b < wi.nb; b += GT) B.bstep[wi.body0 + b].has_contact = 0;
  {
    int s = 0;
    for (int k = k_lo + tid; k < k_hi; k += 256) s += B.pair_nc[wi.pair0 + k];
    s = __reduce_add_sync(0xffffffffu, s);
    if (lane == 0) warp_sums[warp] = s;
    __syncthreads();
    if (tid == 0) { int t = 0; for (int q = 0; q < 8; q++) t += warp_sums[q]; part[blockIdx.x] = t; }
  }
  pgsl_barrier(bar, bar_target);
  {
    int s = 0, tot = 0;
    for (int q = tid; q < G; q += 256) { const int v = __ldcg(&part[q]); tot += v; if (q < (int)blockIdx.x) s += v; }
    s = __reduce_add_sync(0xffffffffu, s); tot = __reduce_add_sync(0xffffffffu, tot);
    __syncthreads();
    if (lane == 0) warp_sums[warp] = s;
    __syncthreads();
    if (tid == 0) { int t = 0; for (int q = 0; q < 8; q++) t += warp_sums[q]; carry = t; }
    __syncthreads();
    if (lane == 0) warp_sums[warp] = tot;
    __syncthreads();
    if (gt == 0) {
      int nc = 0;
      for (int q = 0; q < 8; q++) nc += warp_sums[q];
      if (nc > wi.contact_cap) { ws.overflow |= 2; nc = wi.contact_cap; }
      ws.n_contacts = nc;
    }
    __syncthreads();
  }
  for (int k0 = k_lo; k0 < k_hi; k0 += 256) {
    const int k = k0 + tid;
    const int c = (k < k_hi) ? B.pair_nc[wi.pair0 + k] : 0;
    int v = c;
    for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += t; }
    if (lane == 31) warp_sums[warp] = v;
    __syncthreads();
    if (warp == 0) {
      int s = (lane < 8) ? warp_sums[lane] : 0;
      for (int d = 1; d < 8; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, s, d); if (lane >= d) s += t; }
      if (lane < 8) warp_sums[lane] = s;
    }
    __syncthreads();
    const int off = carry + (v - c) + (warp > 0 ? warp_sums[warp - 1] : 0);
    const int total = warp_sums[7];
    if (c > 0) contacts_of_pair(B, wi, k, c, off);
    __syncthreads();
    if (tid == 0) carry += total;
    __syncthreads();
  }
}

// k_assemble1 for a large world: auto-disable, islands (lock-free union-find over the joints and contacts), which islands are
// stepped, the body prologue, row counts and offsets.
__global__ void __launch_bounds__(256) k_assemble1_large(DevBatch B, const int w, unsigned *bar) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  unsigned bar_target = *bar;
  grid.sync();
  const WorldInfo wi = B.world[w];
  const WorldStep &ws = B.wstep[w];
  const int gt = blockIdx.x * blockDim.x + threadIdx.x, GT = gridDim.x * blockDim.x;
  for (int i = gt; i < wi.nb; i += GT) autodisable_body(B, wi, i);
  pgsl_barrier(bar, bar_target);
  int *const lab0 = &B.bstep[wi.body0].label;
  const size_t ls = sizeof(BodyStep) / sizeof(int);
  auto root = [&](int x) { for (;;) { const int l = __ldcg(lab0 + (size_t)x * ls); if (l == x) return x; x = l; } };
  const int ne = wi.nj + ws.n_contacts;
  // lock-free union-find: an edge links the larger of its two current roots under the smaller one by compare-and-swap
  // (only a body that still is a root is ever re-parented, so no link is lost and ONE pass over the edges suffices;
  // labels only ever decrease, so there are no cycles and a component's root ends up its smallest body).  First version:
  // hooking by atomicMin in passes until nothing changed - two grid barriers a pass, ~8 passes for a pile.
  for (int k = gt; k < ne; k += GT) {
    int b1, b2;
    if (k < wi.nj) { const Joint &j = B.joint[wi.joint0 + k]; b1 = j.b1; b2 = j.b2; }
    else { const Contact &c = B.contact[wi.contact0 + (k - wi.nj)]; b1 = c.b1; b2 = c.b2; }
    if (b2 < 0) continue;
    int r1 = root(b1 - wi.body0), r2 = root(b2 - wi.body0);
    while (r1 != r2) {
      const int hi = max(r1, r2), lo = min(r1, r2);
      const int old = atomicCAS(lab0 + (size_t)hi * ls, hi, lo);
      if (old == hi) break;
      r1 = root(old); r2 = root(lo);   // somebody else re-parented hi first: go on from the new roots
    }
  }
  pgsl_barrier(bar, bar_target);
  // flatten (a label overwritten by its root while another thread walks through it still leads to the same root)
  for (int i = gt; i < wi.nb; i += GT) { const int r = root(i); B.bstep[wi.body0 + i].label = r; }
  pgsl_barrier(bar, bar_target);
  // a component is stepped iff it holds an enabled body; every body in it is re-enabled
  for (int i = gt; i < wi.nb; i += GT) B.bstep[wi.body0 + i].active = 0;
  pgsl_barrier(bar, bar_target);
  for (int i = gt; i < wi.nb; i += GT)
    if (!(B.body[wi.body0 + i].flags & GZ_BODY_DISABLED)) B.bstep[wi.body0 + __ldcg(&B.bstep[wi.body0 + i].label)].active = 1;   // benign race: all write 1
  pgsl_barrier(bar, bar_target);
  for (int i = gt; i < wi.nb; i += GT) {
    BodyStep &bs = B.bstep[wi.body0 + i];
    const int lab = __ldcg(&bs.label);
    const int act = __ldcg(&B.bstep[wi.body0 + lab].active);
    if (lab != i) bs.active = act;
    if (act) B.body[wi.body0 + i].flags &= ~GZ_BODY_DISABLED;
  }
  pgsl_barrier(bar, bar_target);
  for (int i = gt; i < wi.nb; i += GT) bodyprep_body(B, wi.body0 + i);
  if (blockIdx.x == 0) info1_world(B, w);   // (joint rows need a scan over the joints: one CTA; contacts are 3 or 4 rows each)
}

}  // namespace gz
