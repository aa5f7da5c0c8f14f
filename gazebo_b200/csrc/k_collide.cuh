// k_collide.cuh — ODEPhysics::UpdateCollision on the device:
//   k_geom        final geom pose + AABB         (dxGeom::computePosr collision_kernel.cpp:462-475,
//                                                  computeAABB box.cpp:60 sphere.cpp:59 capsule.cpp:60
//                                                  cylinder.cpp:59 plane.cpp:80)
//   k_broadphase  candidate pairs per world       (dxHashSpace::collide collision_space.cpp:427-593,
//                                                  dSpaceCollide2 :778-843, collideAABBs
//                                                  collision_space_internal.h:48-82,
//                                                  ODEPhysics::CollisionCallback ODEPhysics.cc:948-1008)
//   k_narrow      dCollide + Gazebo truncation    (collision_kernel.cpp:293-340, ODEPhysics.cc:1046-1091)
//   k_contacts    ordered compaction + surface    (ODEPhysics.cc:1094-1329, dJointAttach ode.cpp:1417-1465)
//
// Pair order is canonical: (lower geom index, higher geom index), ascending — the
// deterministic row-order mode of north_star; the oracle harness sorts the reference's
// pairs the same way.  The pair SET equals the reference's: the hash space is an exact
// AABB-overlap test between the model spaces followed by geom-vs-geom AABB tests.
#pragma once
#include "k_boxbox.cuh"
#include "../../include/gz_detmath.h"
#include "k_cylbox.cuh"
#include "k_mpr.cuh"

namespace gz {

__global__ void k_geom(DevBatch B) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B.ng) return;
  Geom &g = B.geom[i];
  if (g.type == GZ_GEOM_PLANE) {
    // dxPlane::computeAABB: half-space box for axis-aligned normals (float _dequal overload)
    const double *p = g.dims;
    double a[6] = {-GZ_INF, GZ_INF, -GZ_INF, GZ_INF, -GZ_INF, GZ_INF};
    bool z0 = fabsf((float)p[0] - 0.0f) <= 1e-6f, z1 = fabsf((float)p[1] - 0.0f) <= 1e-6f,
         z2 = fabsf((float)p[2] - 0.0f) <= 1e-6f;
    if (z1 && z2) { a[0] = (p[0] > 0) ? -GZ_INF : -p[3]; a[1] = (p[0] > 0) ? p[3] : GZ_INF; }
    else if (z0 && z2) { a[2] = (p[1] > 0) ? -GZ_INF : -p[3]; a[3] = (p[1] > 0) ? p[3] : GZ_INF; }
    else if (z0 && z1) { a[4] = (p[2] > 0) ? -GZ_INF : -p[3]; a[5] = (p[2] > 0) ? p[3] : GZ_INF; }
    for (int k = 0; k < 6; k++) g.aabb[k] = a[k];
    return;
  }
  double pos[3], R[9];
  if (g.body >= 0) {
    const Body &b = B.body[g.body];
    mul0_331(pos, b.R, g.opos);
    pos[0] += b.pos[0]; pos[1] += b.pos[1]; pos[2] += b.pos[2];
    mul0_333(R, b.R, g.oR);
  } else {
    for (int k = 0; k < 3; k++) pos[k] = g.opos[k];
    for (int k = 0; k < 9; k++) R[k] = g.oR[k];
  }
  double xr, yr, zr;
  if (g.type == GZ_GEOM_BOX) {
    const double *s = g.dims;
    xr = 0.5 * (fabs(R[0] * s[0]) + fabs(R[1] * s[1]) + fabs(R[2] * s[2]));
    yr = 0.5 * (fabs(R[3] * s[0]) + fabs(R[4] * s[1]) + fabs(R[5] * s[2]));
    zr = 0.5 * (fabs(R[6] * s[0]) + fabs(R[7] * s[1]) + fabs(R[8] * s[2]));
  } else if (g.type == GZ_GEOM_SPHERE) {
    xr = yr = zr = g.dims[0];
  } else if (g.type == GZ_GEOM_CAPSULE) {
    double r = g.dims[0], lz = g.dims[1];
    xr = fabs(R[2] * lz) * 0.5 + r;
    yr = fabs(R[5] * lz) * 0.5 + r;
    zr = fabs(R[8] * lz) * 0.5 + r;
  } else {  // cylinder
    double r = g.dims[0], lz = g.dims[1];
    xr = fabs(R[0] * r) + fabs(R[1] * r) + 0.5 * fabs(R[2] * lz);
    yr = fabs(R[3] * r) + fabs(R[4] * r) + 0.5 * fabs(R[5] * lz);
    zr = fabs(R[6] * r) + fabs(R[7] * r) + 0.5 * fabs(R[8] * lz);
  }
  for (int k = 0; k < 3; k++) g.pos[k] = pos[k];
  for (int k = 0; k < 9; k++) g.R[k] = R[k];
  g.aabb[0] = pos[0] - xr; g.aabb[1] = pos[0] + xr;
  g.aabb[2] = pos[1] - yr; g.aabb[3] = pos[1] + yr;
  g.aabb[4] = pos[2] - zr; g.aabb[5] = pos[2] + zr;
}

// Are two bodies joined by a persistent (non-contact) joint?  dAreConnectedExcluding
// (ode.cpp:1610-1618) walks body1's joint list; here the per-body CSR of persistent joints.
__device__ inline bool bodies_connected(const DevBatch &B, int b1, int b2) {
  const Body &a = B.body[b1];
  for (int k = a.joint_begin; k < a.joint_end; k++) {
    const Joint &j = B.joint[B.body_joint[k]];
    if ((j.b1 == b1 && j.b2 == b2) || (j.b1 == b2 && j.b2 == b1)) return true;
  }
  return false;
}

__device__ inline bool pair_passes(const DevBatch &B, const Geom &g1, const Geom &g2) {
  if (g1.space == g2.space) return false;   // same dSimpleSpace: never tested at the top level
  // space-level collideAABBs: category/collide bits of the two spaces
  if (((g1.scat & g2.scol) || (g2.scat & g1.scol)) == 0) return false;
  // geom-level collideAABBs
  if (g1.body == g2.body && g1.body >= 0) return false;
  if (((g1.cat & g2.col) || (g2.cat & g1.col)) == 0) return false;
  const double *a = g1.aabb, *b = g2.aabb;
  if (a[0] > b[1] || a[1] < b[0] || a[2] > b[3] || a[3] < b[2] || a[4] > b[5] || a[5] < b[4]) return false;
  // CollisionCallback: joint-connected bodies, then the both-disabled rule
  int b1 = g1.body, b2 = g2.body;
  if (b1 >= 0 && b2 >= 0 && bodies_connected(B, b1, b2)) return false;
  bool e1 = b1 >= 0 ? !(B.body[b1].flags & GZ_BODY_DISABLED) : false;
  bool e2 = b2 >= 0 ? !(B.body[b2].flags & GZ_BODY_DISABLED) : false;
  if (g1.cat != 2u && g2.cat != 2u &&   // GZ_SENSOR_COLLIDE
      ((b1 >= 0 && b2 >= 0 && !e1 && !e2) || (b2 < 0 && b1 >= 0 && !e1) || (b1 < 0 && b2 >= 0 && !e2)))
    return false;
  return true;
}

// Sort-and-sweep broadphase of one world (n >= B.bp_sap_min geoms), one CTA.  The geoms are sorted
// by the lower x bound of their AABB (bitonic sort in shared memory); one warp per sorted position
// p then walks forward while min_x[q] <= max_x[p] — exactly the pairs that overlap on x, every
// other pair fails collideAABBs' first comparison (collision_space_internal.h:66-72) — and gives
// each survivor the full test.  Hits set bit (lower index, higher index) of an n x n bit matrix
// in shared memory; popcounts per row, a warp prefix scan and an in-order walk of the set bits
// emit the pair list sorted by (i, j): the same list as the all-pairs path, whatever order the
// sweep found the pairs in.  (The reference's own structure here is a multi-resolution hash
// space, collision_space.cpp:427-593; both are exact AABB-overlap filters, so the pair SET is
// the same — the tests compare the lists with the reference's, sorted.)
__device__ void broadphase_sap(const DevBatch &B, const int w, const WorldInfo &wi, int *sh) {
  const int n = wi.ng, tid = threadIdx.x, nt = blockDim.x;
  const int lane = tid & 31, warp = tid >> 5, nwarp = nt >> 5;
  int npad = 32;
  while (npad < n) npad <<= 1;
  const int nw = (n + 31) >> 5;
  double *skey = reinterpret_cast<double *>(sh);          // [npad] lower x bound, sorted
  double *smax = skey + npad;                             // [npad] upper x bound of the same geom
  int *sidx = reinterpret_cast<int *>(smax + npad);       // [npad] world-local geom index
  int *rowcnt = sidx + npad;                              // [n + 1]
  unsigned *bm = reinterpret_cast<unsigned *>(rowcnt + n + 1);   // [n][nw] upper-triangular hit matrix
  const Geom *G = B.geom + wi.geom0;
  for (int i = tid; i < npad; i += nt) { skey[i] = (i < n) ? G[i].aabb[0] : GZ_INF; sidx[i] = i; }
  for (int i = tid; i < n * nw; i += nt) bm[i] = 0u;
  __syncthreads();
  for (int k = 2; k <= npad; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < npad; i += nt) {
        const int l = i ^ j;
        if (l > i) {
          const double a = skey[i], b = skey[l];
          const int ia = sidx[i], ib = sidx[l];
          const bool gt = a > b || (a == b && ia > ib);   // padding (index >= n, key +inf) sorts last
          if (gt == ((i & k) == 0)) { skey[i] = b; skey[l] = a; sidx[i] = ib; sidx[l] = ia; }
        }
      }
      __syncthreads();
    }
  for (int p = tid; p < n; p += nt) smax[p] = G[sidx[p]].aabb[1];
  __syncthreads();
  for (int p = warp; p < n - 1; p += nwarp) {
    const int ip = sidx[p];
    const double mx = smax[p];
    for (int q0 = p + 1; q0 < n; q0 += 32) {
      const int q = q0 + lane;
      const bool in = (q < n) && skey[q] <= mx;
      if (in) {
        const int iq = sidx[q];
        const int lo = min(ip, iq), hi = max(ip, iq);
        if (pair_passes(B, G[lo], G[hi])) atomicOr(&bm[lo * nw + (hi >> 5)], 1u << (hi & 31));
      }
      if (!__all_sync(0xffffffffu, in)) break;   // sorted: nothing further overlaps p on x
    }
  }
  __syncthreads();
  for (int i = tid; i < n; i += nt) {
    int c = 0;
    for (int k = i >> 5; k < nw; k++) c += __popc(bm[i * nw + k]);
    rowcnt[i] = c;
  }
  __syncthreads();
  if (warp == 0) {   // exclusive scan of the row counts, 32 at a time
    int carry = 0;
    for (int i0 = 0; i0 < n; i0 += 32) {
      const int i = i0 + lane;
      const int c = (i < n) ? rowcnt[i] : 0;
      int v = c;
      for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += t; }
      if (i < n) rowcnt[i] = carry + v - c;
      carry += __shfl_sync(0xffffffffu, v, 31);
    }
    if (lane == 0) {
      WorldStep &ws = B.wstep[w];
      ws.overflow = 0;
      if (carry > wi.pair_cap) { ws.overflow |= 1; carry = wi.pair_cap; }
      ws.n_pairs = carry;
    }
  }
  __syncthreads();
  for (int i = warp; i < n; i += nwarp) {
    int base = rowcnt[i];
    for (int k0 = (i >> 5) & ~31; k0 < nw; k0 += 32) {
      const int k = k0 + lane;
      unsigned m = (k < nw) ? bm[i * nw + k] : 0u;
      const int c = __popc(m);
      int v = c;
      for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += t; }
      int at = base + v - c;
      while (m) {
        const int j = 32 * k + __ffs(m) - 1;
        m &= m - 1;
        if (at < wi.pair_cap) {
          B.pair[2 * (wi.pair0 + at)] = wi.geom0 + i;
          B.pair[2 * (wi.pair0 + at) + 1] = wi.geom0 + j;
        }
        at++;
      }
      base += __shfl_sync(0xffffffffu, v, 31);
    }
  }
}

// One CTA per world.  Small worlds: two passes over the (i<j) pair matrix, one warp per row i:
// count, block-wide exclusive scan of the row counts, then ordered emission with
// warp ballots.  Output is sorted by (i, j) by construction.  Worlds with bp_sap_min geoms or
// more take the sort-and-sweep path above.
__global__ void __launch_bounds__(512) k_broadphase(DevBatch B) {
  extern __shared__ __align__(16) int sh[];
  const int w = blockIdx.x;
  const WorldInfo wi = B.world[w];
  const int n = wi.ng;
  if (n >= B.bp_sap_min) { broadphase_sap(B, w, wi, sh); return; }
  int *rowcnt = sh;            // [n+1]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const Geom *G = B.geom + wi.geom0;
  for (int i = warp; i < n; i += nwarp) {
    int c = 0;
    for (int j0 = i + 1; j0 < n; j0 += 32) {
      int j = j0 + lane;
      bool hit = (j < n) && pair_passes(B, G[i], G[j]);
      c += __popc(__ballot_sync(0xffffffffu, hit));
    }
    if (lane == 0) rowcnt[i] = c;
  }
  __syncthreads();
  if (threadIdx.x == 0) {   // n is small (<= a few hundred); serial scan is negligible
    int acc = 0;
    for (int i = 0; i < n; i++) { int c = rowcnt[i]; rowcnt[i] = acc; acc += c; }
    rowcnt[n] = acc;
    WorldStep &ws = B.wstep[w];
    ws.overflow = 0;
    if (acc > wi.pair_cap) { ws.overflow |= 1; acc = wi.pair_cap; }
    ws.n_pairs = acc;
  }
  __syncthreads();
  for (int i = warp; i < n; i += nwarp) {
    int base = rowcnt[i];
    for (int j0 = i + 1; j0 < n; j0 += 32) {
      int j = j0 + lane;
      bool hit = (j < n) && pair_passes(B, G[i], G[j]);
      unsigned m = __ballot_sync(0xffffffffu, hit);
      if (hit) {
        int k = base + __popc(m & ((1u << lane) - 1));
        if (k < wi.pair_cap) {
          B.pair[2 * (wi.pair0 + k)] = wi.geom0 + i;
          B.pair[2 * (wi.pair0 + k) + 1] = wi.geom0 + j;
        }
      }
      base += __popc(m);
    }
  }
}

// dCollide's class-pair table (collision_kernel.cpp:168-269).  Returns contacts with the
// normal convention of dCollide(g1, g2): reversed entries call fn(g2, g1) and flip normals.
// CYL = the scene has cylinders: only then are the cylinder-box and MPR colliders compiled in (they double the
// registers and the stack of k_narrow; scenes of boxes, spheres and capsules run the light instantiation)
template <bool CYL>
__device__ inline int collide_dispatch(const Geom &g1, const Geom &g2, COut &o, int *overflow_word) {
  int t1 = g1.type, t2 = g2.type;
  int n = 0;
  bool rev = false;
  if (t1 == GZ_GEOM_SPHERE && t2 == GZ_GEOM_SPHERE) n = col_sphere_sphere(g1, g2, o);
  else if (t1 == GZ_GEOM_SPHERE && t2 == GZ_GEOM_BOX) n = col_sphere_box(g1, g2, o);
  else if (t1 == GZ_GEOM_BOX && t2 == GZ_GEOM_SPHERE) { n = col_sphere_box(g2, g1, o); rev = true; }
  else if (t1 == GZ_GEOM_SPHERE && t2 == GZ_GEOM_PLANE) n = col_sphere_plane(g1, g2, o);
  else if (t1 == GZ_GEOM_PLANE && t2 == GZ_GEOM_SPHERE) { n = col_sphere_plane(g2, g1, o); rev = true; }
  else if (t1 == GZ_GEOM_BOX && t2 == GZ_GEOM_BOX) n = col_box_box(g1, g2, o);
  else if (t1 == GZ_GEOM_BOX && t2 == GZ_GEOM_PLANE) n = col_box_plane(g1, g2, o);
  else if (t1 == GZ_GEOM_PLANE && t2 == GZ_GEOM_BOX) { n = col_box_plane(g2, g1, o); rev = true; }
  else if (t1 == GZ_GEOM_CAPSULE && t2 == GZ_GEOM_SPHERE) n = col_capsule_sphere(g1, g2, o);
  else if (t1 == GZ_GEOM_SPHERE && t2 == GZ_GEOM_CAPSULE) { n = col_capsule_sphere(g2, g1, o); rev = true; }
  else if (t1 == GZ_GEOM_CAPSULE && t2 == GZ_GEOM_BOX) n = col_capsule_box(g1, g2, o);
  else if (t1 == GZ_GEOM_BOX && t2 == GZ_GEOM_CAPSULE) { n = col_capsule_box(g2, g1, o); rev = true; }
  else if (t1 == GZ_GEOM_CAPSULE && t2 == GZ_GEOM_CAPSULE) n = col_capsule_capsule(g1, g2, o);
  else if (t1 == GZ_GEOM_CAPSULE && t2 == GZ_GEOM_PLANE) n = col_capsule_plane(g1, g2, o);
  else if (t1 == GZ_GEOM_PLANE && t2 == GZ_GEOM_CAPSULE) { n = col_capsule_plane(g2, g1, o); rev = true; }
  else if (t1 == GZ_GEOM_CYLINDER && t2 == GZ_GEOM_SPHERE) n = col_cylinder_sphere(g1, g2, o);
  else if (t1 == GZ_GEOM_SPHERE && t2 == GZ_GEOM_CYLINDER) { n = col_cylinder_sphere(g2, g1, o); rev = true; }
  else if (t1 == GZ_GEOM_CYLINDER && t2 == GZ_GEOM_PLANE) n = col_cylinder_plane(g1, g2, o);
  else if (t1 == GZ_GEOM_PLANE && t2 == GZ_GEOM_CYLINDER) { n = col_cylinder_plane(g2, g1, o); rev = true; }
  else if (CYL && t1 == GZ_GEOM_CYLINDER && t2 == GZ_GEOM_BOX) n = col_cylinder_box(g1, g2, o);
  else if (CYL && t1 == GZ_GEOM_BOX && t2 == GZ_GEOM_CYLINDER) { n = col_cylinder_box(g2, g1, o); rev = true; }
  else if (CYL && t1 == GZ_GEOM_CYLINDER && t2 == GZ_GEOM_CYLINDER) n = col_mpr(g1, false, g2, o);
  else if (CYL && t1 == GZ_GEOM_CAPSULE && t2 == GZ_GEOM_CYLINDER) n = col_mpr(g1, true, g2, o);
  else if (CYL && t1 == GZ_GEOM_CYLINDER && t2 == GZ_GEOM_CAPSULE) { n = col_mpr(g2, true, g1, o); rev = true; }
  else n = 0;
  if (n < 0) { atomicOr(overflow_word, 32); n = 0; }   // MPR portal loop cap (k_mpr.cuh)
  if (rev)
    for (int i = 0; i < n; i++) { o.nrm[i][0] = -o.nrm[i][0]; o.nrm[i][1] = -o.nrm[i][1]; o.nrm[i][2] = -o.nrm[i][2]; }
  return n;
}

// one candidate pair
template <bool CYL>
__device__ __forceinline__ void narrow_pair(DevBatch &B, const int w, const WorldInfo &wi, const int k) {
  const int pi = wi.pair0 + k;
  const Geom &g1 = B.geom[B.pair[2 * pi]];
  const Geom &g2 = B.geom[B.pair[2 * pi + 1]];
  int numc = 0;
  // ODEPhysics::Collide: collide bitmask, maxCollide
  if ((g1.bitmask & g2.bitmask) != 0) {
    unsigned maxCollide = 250;
    if (B.p.max_contacts > 0 && B.p.max_contacts < 250) maxCollide = B.p.max_contacts;
    if ((unsigned)g1.max_contacts < maxCollide) maxCollide = g1.max_contacts;
    if ((unsigned)g2.max_contacts < maxCollide) maxCollide = g2.max_contacts;
    COut o;
    numc = collide_dispatch<CYL>(g1, g2, o, &B.wstep[w].overflow);
    int idx[GZ_PAIR_MAXC];
    for (int i = 0; i < GZ_PAIR_MAXC; i++) idx[i] = i;
    if (maxCollide > 0 && (unsigned)numc > maxCollide) {
      double mx = o.depth[maxCollide - 1];
      for (unsigned i = maxCollide; i < (unsigned)numc; ++i)
        if (o.depth[i] > mx) { mx = o.depth[i]; idx[maxCollide - 1] = i; }
      numc = maxCollide;
    }
    double *out = B.pair_contact + (size_t)pi * (7 * GZ_PAIR_MAXC);
    for (int c = 0; c < numc; c++) {
      int s = idx[c];
      for (int i = 0; i < 3; i++) { out[7 * c + i] = o.pos[s][i]; out[7 * c + 3 + i] = o.nrm[s][i]; }
      out[7 * c + 6] = o.depth[s];
    }
  }
  B.pair_nc[pi] = numc;
}

// grid (gx, n_worlds): one thread per candidate pair, gx blocks striding over the world's pairs (gx is sized for
// the pairs a world of the batch typically has, not for the slab's capacity: idle blocks cost launch time)
template <bool CYL>
__global__ void k_narrow(DevBatch B) {
  const int w = blockIdx.y;
  const WorldInfo wi = B.world[w];
  const int n = B.wstep[w].n_pairs;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) narrow_pair<CYL>(B, w, wi, k);
}

// Hertzian contact (ODEPhysics.cc:1186-1251): patch radius = the larger one, curvatures add, both surfaces must agree on
// use_patch_radius; E* from the two moduli and Poisson's ratios.  
__device__ inline void em_combine(const DevBatch &B, const int gi1, const int gi2, double &em, double &em_radius, int &em_surface) {
  const double *x1 = B.geom_ext + (size_t)gi1 * (GZ_SF_COUNT - GZ_SF_HOT) - GZ_SF_HOT, *x2 = B.geom_ext + (size_t)gi2 * (GZ_SF_COUNT - GZ_SF_HOT) - GZ_SF_HOT;
  const double patch = fmax(x1[GZ_SF_PATCH_RADIUS], x2[GZ_SF_PATCH_RADIUS]);
  double curv1 = 0, curv2 = 0;
  if (x1[GZ_SF_SURFACE_RADIUS] > 0) curv1 = 1 / x1[GZ_SF_SURFACE_RADIUS];
  if (x2[GZ_SF_SURFACE_RADIUS] > 0) curv2 = 1 / x2[GZ_SF_SURFACE_RADIUS];
  const double curv_sum = curv1 + curv2;
  double surface_radius = 0;
  if (curv_sum > 0) surface_radius = 1 / curv_sum;
  const bool use_patch = x1[GZ_SF_USE_PATCH_RADIUS] != 0 && x2[GZ_SF_USE_PATCH_RADIUS] != 0;
  const double nu1 = x1[GZ_SF_POISSON], nu2 = x2[GZ_SF_POISSON];
  const double e1 = x1[GZ_SF_ELASTIC_MODULUS], e2 = x2[GZ_SF_ELASTIC_MODULUS];
  const double es = 1.0 / ((1.0 - nu1 * nu1) / e1 + (1.0 - nu2 * nu2) / e2);
  if (es > 0.0) { em = es; em_radius = use_patch ? patch : surface_radius; em_surface = use_patch ? 0 : 1; }
}

// Torsional friction of a pair (ODEPhysics.cc:1168-1232): mu3 = the smaller one, slip3 = the sum x contacts, the radii as
// for Hertzian contact; the row exists (dContactMu3) when mu3 > 0 and the radius the pair uses is > 0.
__device__ inline void tors_combine(const DevBatch &B, const int gi1, const int gi2, const int numc, double &mu3, double &radius, double &slip3, int &tors) {
  const double *x1 = B.geom_ext + (size_t)gi1 * (GZ_SF_COUNT - GZ_SF_HOT) - GZ_SF_HOT, *x2 = B.geom_ext + (size_t)gi2 * (GZ_SF_COUNT - GZ_SF_HOT) - GZ_SF_HOT;
  mu3 = fmin(x1[GZ_SF_MU3], x2[GZ_SF_MU3]);
  slip3 = x1[GZ_SF_SLIP3] + x2[GZ_SF_SLIP3];
  slip3 *= numc;
  const double patch = fmax(x1[GZ_SF_PATCH_RADIUS], x2[GZ_SF_PATCH_RADIUS]);
  double curv1 = 0, curv2 = 0;
  if (x1[GZ_SF_SURFACE_RADIUS] > 0) curv1 = 1 / x1[GZ_SF_SURFACE_RADIUS];
  if (x2[GZ_SF_SURFACE_RADIUS] > 0) curv2 = 1 / x2[GZ_SF_SURFACE_RADIUS];
  const double curv_sum = curv1 + curv2;
  double surface_radius = 0;
  if (curv_sum > 0) surface_radius = 1 / curv_sum;
  const bool use_patch = x1[GZ_SF_USE_PATCH_RADIUS] != 0 && x2[GZ_SF_USE_PATCH_RADIUS] != 0;
  tors = 0;
  if (mu3 > 0) {
    if (use_patch && patch > 0) { tors = 1; radius = patch; }
    else if (!use_patch && surface_radius > 0) { tors = 2; radius = surface_radius; }
  }
}

// The contacts of candidate pair k of a world (c of them, from k_narrow) at place `off` of the world's contact list,
// with the combined surface parameters.
__device__ __forceinline__ void contacts_of_pair(const DevBatch &B, const WorldInfo &wi, const int k, const int c, const int off) {
  const double h = B.p.step_size;
  const int pi = wi.pair0 + k;
  const int gi1 = B.pair[2 * pi], gi2 = B.pair[2 * pi + 1];
  const Geom &g1 = B.geom[gi1];
  const Geom &g2 = B.geom[gi2];
  // surface combination, ODEPhysics.cc:1108-1258
  double kp = 1.0 / (1.0 / g1.surf[GZ_SF_KP] + 1.0 / g2.surf[GZ_SF_KP]);
  double kd = g1.surf[GZ_SF_KD] + g2.surf[GZ_SF_KD];
  double soft_erp = (h * kp) / (h * kp + kd);
  double soft_cfm = 1.0 / (h * kp + kd);
  // fdir1 as written in ODEPhysics.cc:1120-1161: only surface 1's direction ever reaches the contact (the
  // reference rotates surface 2's fd2 but never assigns it), rotated to the world frame by collision 1's pose
  double fd[3] = {g1.fdir1[0], g1.fdir1[1], g1.fdir1[2]};
  const bool fdset = (fd[0] != 0 || fd[1] != 0 || fd[2] != 0);
  double fdw[3] = {0, 0, 0};
  if (fdset) {
    if (g1.type == GZ_GEOM_PLANE) { fdw[0] = fd[0]; fdw[1] = fd[1]; fdw[2] = fd[2]; }   // a plane has no pose: identity
    else mul0_331(fdw, g1.R, fd);
  }
  double mu = fmin(g1.surf[GZ_SF_MU], g2.surf[GZ_SF_MU]);
  double mu2 = fmin(g1.surf[GZ_SF_MU2], g2.surf[GZ_SF_MU2]);
  double slip1 = g1.surf[GZ_SF_SLIP1] + g2.surf[GZ_SF_SLIP1];
  double slip2 = g1.surf[GZ_SF_SLIP2] + g2.surf[GZ_SF_SLIP2];
  slip1 *= c; slip2 *= c;
  double bounce = fmin(g1.surf[GZ_SF_BOUNCE], g2.surf[GZ_SF_BOUNCE]);
  double bounce_vel = fmin(g1.surf[GZ_SF_BOUNCE_VEL], g2.surf[GZ_SF_BOUNCE_VEL]);
  // Hertzian contact (ODEPhysics.cc:1186-1251): patch radius = the larger one, curvatures add, both surfaces must agree
  // on use_patch_radius; E* from the two moduli and Poisson's ratios, only when both moduli are > 0
  double em = 0.0, em_radius = 0.0;
  int em_surface = 0;
  if (g1.em_on && g2.em_on) em_combine(B, gi1, gi2, em, em_radius, em_surface);
  double mu3 = 0.0, tors_radius = 0.0, slip3 = 0.0;
  int tors = 0;
  if ((g1.tors_on & g2.tors_on & 1) && ((g1.tors_on | g2.tors_on) & 2)) tors_combine(B, gi1, gi2, c, mu3, tors_radius, slip3, tors);
  // dJointAttach(contact, b1, b2): a null first body swaps and sets dJOINT_REVERSE
  int b1 = g1.body, b2 = g2.body, reverse = 0;
  if (b1 < 0) { b1 = b2; b2 = -1; reverse = 1; }
  const double *src = B.pair_contact + (size_t)pi * (7 * GZ_PAIR_MAXC);
  for (int j = 0; j < c; j++) {
    int ci = off + j;
    if (ci >= wi.contact_cap) break;
    Contact &C = B.contact[wi.contact0 + ci];
    for (int i = 0; i < 3; i++) { C.pos[i] = src[7 * j + i]; C.normal[i] = src[7 * j + 3 + i]; C.fdir1[i] = fdw[i]; }
    C.depth = src[7 * j + 6];
    C.mu = mu; C.mu2 = mu2; C.soft_erp = soft_erp; C.soft_cfm = soft_cfm;
    C.slip1 = slip1; C.slip2 = slip2; C.bounce = bounce; C.bounce_vel = bounce_vel;
    C.em = em; C.em_radius = em_radius; C.em_surface = em_surface;
    C.mu3 = mu3; C.tors_radius = tors_radius; C.slip3 = slip3; C.tors = tors;
    C.g1 = gi1; C.g2 = gi2; C.b1 = b1; C.b2 = b2; C.reverse = reverse; C.has_fdir1 = fdset ? 1 : 0;
  }
  if (b1 >= 0) B.bstep[b1].has_contact = 1;
  if (b2 >= 0) B.bstep[b2].has_contact = 1;
}

// One CTA per world: exclusive scan of per-pair contact counts, then one thread per pair
// writes its contacts at the scanned offset.
__global__ void __launch_bounds__(1024) k_contacts(DevBatch B) {
  const int w = blockIdx.x;
  const WorldInfo wi = B.world[w];
  WorldStep &ws = B.wstep[w];
  const int np = ws.n_pairs;
  __shared__ int carry;
  __shared__ int warp_sums[32];
  if (threadIdx.x == 0) carry = 0;
  for (int b = threadIdx.x; b < wi.nb; b += blockDim.x) B.bstep[wi.body0 + b].has_contact = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int k0 = 0; k0 < np; k0 += blockDim.x) {
    int k = k0 + threadIdx.x;
    int c = (k < np) ? B.pair_nc[wi.pair0 + k] : 0;
    // block exclusive scan
    int v = c;
    for (int d = 1; d < 32; d <<= 1) { int t = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += t; }
    if (lane == 31) warp_sums[warp] = v;
    __syncthreads();
    if (warp == 0) {
      int s = (lane < nwarp) ? warp_sums[lane] : 0;
      for (int d = 1; d < 32; d <<= 1) { int t = __shfl_up_sync(0xffffffffu, s, d); if (lane >= d) s += t; }
      warp_sums[lane] = s;
    }
    __syncthreads();
    int off = carry + (v - c) + (warp > 0 ? warp_sums[warp - 1] : 0);
    int total = warp_sums[nwarp - 1];
    if (c > 0) contacts_of_pair(B, wi, k, c, off);
    __syncthreads();
    if (threadIdx.x == 0) carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    int nc = carry;
    if (nc > wi.contact_cap) { ws.overflow |= 2; nc = wi.contact_cap; }
    ws.n_contacts = nc;
  }
}

}  // namespace gz
