// k_ray.cuh — ray queries against the geometry of the batch's worlds: the CPU laser of
// gazebo::physics::ODEMultiRayShape::UpdateRays / UpdateCallback (gazebo/physics/ode/ODEMultiRayShape.cc:82-205; SURVEY
// 8f n3) as a CUDA kernel.  Every ray is tested against every geom of its world that passes collideAABBs' filters
// (category / collide bits of the spaces and of the geoms, AABB overlap; rays ride on no body) with ODE's ray
// colliders (deps/opende/src/ray.cpp: sphere :200-253, box :257-354, capsule :357-469, plane :472-503, cylinder
// :506-703); the nearest hit shortens the ray (contact.depth < length, ODEMultiRayShape.cc:185-202).  Only the hit
// distance is needed, so the colliders are restated for the distance alone, in the reference's operation order: the
// ranges come out bit-identical.  One warp per ray, lanes over the world's geoms, minimum by shuffles.
#pragma once

namespace gz {

struct RayQuery {
  int n;
  const int *world;         // [n]
  const double *start;      // [n][3]
  const double *dir;        // [n][3] any length: normalised as dGeomRaySet does (dNormalize3)
  const double *length;     // [n]
  double *out_range;        // [n]
  int *out_geom;            // [n] world-local geom index or -1
};

struct RayT { double p[3], d[3], len; };

// ray_sphere_helper (ray.cpp:200-236): distance to a sphere; mode 1 = take the exit point
__device__ inline bool ray_sphere_t(const RayT &r, const double *c, const double radius, const int mode, double &t) {
  double q[3] = {r.p[0] - c[0], r.p[1] - c[1], r.p[2] - c[2]};
  const double Bq = dot3(q, r.d);
  const double Cq = dot3(q, q) - radius * radius;
  double k = Bq * Bq - Cq;
  if (k < 0) return false;
  k = sqrt(k);
  double alpha;
  if (mode && Cq >= 0) {
    alpha = -Bq + k;
    if (alpha < 0) return false;
  } else {
    alpha = -Bq - k;
    if (alpha < 0) {
      alpha = -Bq + k;
      if (alpha < 0) return false;
    }
  }
  if (alpha > r.len) return false;
  t = alpha;
  return true;
}

// dCollideRayBox (ray.cpp:257-354): slab test in the box frame, the line mirrored so that its direction is >= 0
__device__ inline bool ray_box_t(const RayT &r, const Geom &g, double &t) {
  double tmp[3] = {r.p[0] - g.pos[0], r.p[1] - g.pos[1], r.p[2] - g.pos[2]}, s[3], v[3];
  mul1_331(s, g.R, tmp);
  mul1_331(v, g.R, r.d);
  for (int i = 0; i < 3; i++)
    if (v[i] < 0) { s[i] = -s[i]; v[i] = -v[i]; }
  const double h[3] = {0.5 * g.dims[0], 0.5 * g.dims[1], 0.5 * g.dims[2]};
  if ((s[0] < -h[0] && v[0] <= 0) || s[0] > h[0] || (s[1] < -h[1] && v[1] <= 0) || s[1] > h[1] ||
      (s[2] < -h[2] && v[2] <= 0) || s[2] > h[2] || (dequal(v[0], 0.0) && dequal(v[1], 0.0) && dequal(v[2], 0.0)))
    return false;
  double lo = -GZ_INF, hi = GZ_INF;
  for (int i = 0; i < 3; i++) {
    if (!dequal(v[i], 0.0)) {
      double k = (-h[i] - s[i]) / v[i];
      if (k > lo) lo = k;
      k = (h[i] - s[i]) / v[i];
      if (k < hi) hi = k;
    }
  }
  if (lo > hi) return false;
  const double alpha = (lo >= 0) ? lo : hi;
  if (alpha < 0 || alpha > r.len) return false;
  t = alpha;
  return true;
}

// dCollideRayCapsule (ray.cpp:357-469)
__device__ inline bool ray_capsule_t(const RayT &r, const Geom &g, double &t) {
  const double radius = g.dims[0], lz2 = g.dims[1] * 0.5;
  const double ax[3] = {g.R[2], g.R[5], g.R[8]};
  double cs[3] = {r.p[0] - g.pos[0], r.p[1] - g.pos[1], r.p[2] - g.pos[2]}, q[3], w[3];
  double k = dot3(ax, cs);   // position of the ray start along the axis
  q[0] = k * ax[0] - cs[0]; q[1] = k * ax[1] - cs[1]; q[2] = k * ax[2] - cs[2];
  const double Cq = dot3(q, q) - radius * radius;   // < 0: the start is within the infinite cylinder
  int inside = 0;
  if (Cq < 0) {
    if (k < -lz2) k = -lz2;
    else if (k > lz2) k = lz2;
    w[0] = g.pos[0] + k * ax[0]; w[1] = g.pos[1] + k * ax[1]; w[2] = g.pos[2] + k * ax[2];
    if ((r.p[0] - w[0]) * (r.p[0] - w[0]) + (r.p[1] - w[1]) * (r.p[1] - w[1]) + (r.p[2] - w[2]) * (r.p[2] - w[2]) < radius * radius) inside = 1;
  }
  if (!inside && Cq < 0) {
    k = (k < 0) ? -lz2 : lz2;   // outside the capsule but within the infinite cylinder: only a cap can be hit
  } else {
    const double uv = dot3(ax, r.d);
    w[0] = uv * ax[0] - r.d[0]; w[1] = uv * ax[1] - r.d[1]; w[2] = uv * ax[2] - r.d[2];
    double A = dot3(w, w);
    const double Bq = 2 * dot3(q, w);
    k = Bq * Bq - 4 * A * Cq;
    if (k < 0) {
      if (!inside) return false;
      k = (uv < 0) ? -lz2 : lz2;   // inside and parallel to the axis: a cap
    } else {
      k = sqrt(k);
      A = 1.0 / (2 * A);
      double alpha = (-Bq - k) * A;
      if (alpha < 0) {
        alpha = (-Bq + k) * A;
        if (alpha < 0) return false;
      }
      if (alpha > r.len) return false;
      // the infinite cylinder is hit: between the caps?
      const double hp[3] = {r.p[0] + alpha * r.d[0], r.p[1] + alpha * r.d[1], r.p[2] + alpha * r.d[2]};
      q[0] = hp[0] - g.pos[0]; q[1] = hp[1] - g.pos[1]; q[2] = hp[2] - g.pos[2];
      k = dot3(q, ax);
      if (k >= -lz2 && k <= lz2) { t = alpha; return true; }
      k = (k < 0) ? -lz2 : lz2;
    }
  }
  const double cap[3] = {g.pos[0] + k * ax[0], g.pos[1] + k * ax[1], g.pos[2] + k * ax[2]};
  return ray_sphere_t(r, cap, radius, inside, t);
}

// dCollideRayPlane (ray.cpp:472-503); plane = dims[0..3] (n, d)
__device__ inline bool ray_plane_t(const RayT &r, const Geom &g, double &t) {
  double alpha = g.dims[3] - dot3(g.dims, r.p);
  const double k = dot3(g.dims, r.d);
  if (dequal(k, 0.0)) return false;   // parallel
  alpha /= k;
  if (alpha < 0 || alpha > r.len) return false;
  t = alpha;
  return true;
}

// dCollideRayCylinder (ray.cpp:506-703): caps first (one of them can be hit), then the barrel
__device__ inline bool ray_cylinder_t(const RayT &r, const Geom &g, double &t) {
  const double radius = g.dims[0], half = g.dims[1] * 0.5;
  double tmp[3] = {r.p[0] - g.pos[0], r.p[1] - g.pos[1], r.p[2] - g.pos[2]}, pos[3], dir[3];
  mul1_331(pos, g.R, tmp);
  mul1_331(dir, g.R, r.d);
  const double r2 = radius * radius;
  const double Cq = pos[0] * pos[0] + pos[1] * pos[1] - r2;
  const bool parallel = dequal(dir[0], 0.0) && dequal(dir[1], 0.0);
  const bool perpendicular = dequal(dir[2], 0.0);
  const bool in_radius = Cq <= 0, in_caps = fabs(pos[2]) <= half;
  const bool check_caps = !perpendicular && (!in_caps || in_radius);
  bool check_cyl = !parallel && (!in_radius || in_caps);
  const bool flip_normals = in_caps && in_radius;
  double tt = -GZ_INF;
  if (check_caps) {
    bool flip = false;
    if ((dir[2] < 0 && flip_normals) || (dir[2] > 0 && !flip_normals)) { flip = true; dir[2] = -dir[2]; pos[2] = -pos[2]; }
    tt = (half - pos[2]) / dir[2];
    if (tt >= 0 && tt <= r.len) {
      const double x = pos[0] + tt * dir[0], y = pos[1] + tt * dir[1];
      if (x * x + y * y <= r2) check_cyl = false;   // inside the cap's circle: a hit
      else tt = -GZ_INF;
    } else tt = -GZ_INF;
    if (flip) { dir[2] = -dir[2]; pos[2] = -pos[2]; }
  }
  if (check_cyl) {
    double A = dir[0] * dir[0] + dir[1] * dir[1];
    const double Bq = 2 * (pos[0] * dir[0] + pos[1] * dir[1]);
    double k = Bq * Bq - 4 * A * Cq;
    if (k >= 0 && (Bq < 0 || Bq * Bq <= k)) {
      k = sqrt(k);
      A = 1.0 / (2 * A);
      if (fabs(Bq) <= k) tt = (-Bq + k) * A;   // the start is inside: second solution
      else tt = (-Bq - k) * A;
      if (tt <= r.len) {
        const double z = pos[2] + tt * dir[2];
        if (!(fabs(z) <= half)) tt = -GZ_INF;   // the barrel is hit outside the caps
      } else tt = -GZ_INF;
    }
  }
  if (tt > 0) { t = tt; return true; }
  return false;
}

__global__ void __launch_bounds__(256) k_rays(DevBatch B, RayQuery Q) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarp = (gridDim.x * blockDim.x) >> 5;
  for (int k = warp; k < Q.n; k += nwarp) {
    const WorldInfo wi = B.world[Q.world[k]];
    RayT r;
    for (int c = 0; c < 3; c++) { r.p[c] = Q.start[3 * k + c]; r.d[c] = Q.dir[3 * k + c]; }
    r.len = Q.length[k];
    safe_normalize3(r.d);   // dGeomRaySet (ray.cpp:119-138)
    // dxRay::computeAABB (ray.cpp:60-93)
    double rb[6];
    for (int c = 0; c < 3; c++) {
      const double e = r.p[c] + r.d[c] * r.len;
      if (r.p[c] < e) { rb[2 * c] = r.p[c]; rb[2 * c + 1] = e; } else { rb[2 * c] = e; rb[2 * c + 1] = r.p[c]; }
    }
    double best = r.len;
    int hit = -1;
    for (int gi = lane; gi < wi.ng; gi += 32) {
      const Geom &g = B.geom[wi.geom0 + gi];
      // collideAABBs between the ray space and the geom's space, then between the ray and the geom (GZ_SENSOR_COLLIDE = 2)
      if (((2u & g.scol) || (g.scat & ~2u)) == 0) continue;
      if (((2u & g.col) || (g.cat & ~2u)) == 0) continue;
      const double *a = g.aabb;
      if (rb[0] > a[1] || rb[1] < a[0] || rb[2] > a[3] || rb[3] < a[2] || rb[4] > a[5] || rb[5] < a[4]) continue;
      double t = 0;
      bool ok;
      if (g.type == GZ_GEOM_SPHERE) ok = ray_sphere_t(r, g.pos, g.dims[0], 0, t);
      else if (g.type == GZ_GEOM_BOX) ok = ray_box_t(r, g, t);
      else if (g.type == GZ_GEOM_CAPSULE) ok = ray_capsule_t(r, g, t);
      else if (g.type == GZ_GEOM_CYLINDER) ok = ray_cylinder_t(r, g, t);
      else ok = ray_plane_t(r, g, t);
      if (ok && t < best) { best = t; hit = gi; }   // (geoms of a lane come in ascending order: the first of equal distances stays)
    }
    for (int d = 16; d > 0; d >>= 1) {
      const double ob = __shfl_xor_sync(0xffffffffu, best, d);
      const int oh = __shfl_xor_sync(0xffffffffu, hit, d);
      if (ob < best || (ob == best && oh >= 0 && (hit < 0 || oh < hit))) { best = ob; hit = oh; }
    }
    if (lane == 0) { Q.out_range[k] = best; Q.out_geom[k] = hit; }
  }
}

}  // namespace gz
