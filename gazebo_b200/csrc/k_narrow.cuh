// k_narrow.cuh — primitive-vs-primitive contact generation (one thread per pair).
//
// Each collider reproduces the CONTACT SET AND ORDER of the reference's collider
// for the same pair (dCollide dispatch deps/opende/src/collision_kernel.cpp:293-340;
// primitives cited per function) because Gazebo's truncation rule
// (gazebo/physics/ode/ODEPhysics.cc:1077-1091) and the PGS row order depend on it.
// Gazebo always asks for 250 contacts, so the colliders' own "maxc" culling
// (box.cpp cullPoints) never triggers and is not needed here.
#pragma once
#include "gz_math.cuh"
#include "gz_batch.h"

namespace gz {

struct COut {          // up to GZ_PAIR_MAXC contacts: position, depth; one normal per contact
  double pos[GZ_PAIR_MAXC][3];
  double nrm[GZ_PAIR_MAXC][3];
  double depth[GZ_PAIR_MAXC];
};

// ---- dCollideSpherePlane  (deps/opende/src/sphere.cpp:226-251) --------------
__device__ inline int col_sphere_plane(const Geom &s, const Geom &pl, COut &o) {
  const double *p = pl.dims;
  double r = s.dims[0];
  double k = dot3(s.pos, p);
  double depth = p[3] - k + r;
  if (depth >= 0) {
    for (int i = 0; i < 3; i++) { o.nrm[0][i] = p[i]; o.pos[0][i] = s.pos[i] - p[i] * r; }
    o.depth[0] = depth;
    return 1;
  }
  return 0;
}

// ---- dCollideSpheres (collision_util.cpp:38-67) via dCollideSphereSphere ----
__device__ inline int col_sphere_sphere(const Geom &a, const Geom &b, COut &o) {
  double r1 = a.dims[0], r2 = b.dims[0];
  double t[3] = {a.pos[0] - b.pos[0], a.pos[1] - b.pos[1], a.pos[2] - b.pos[2]};
  double d = sqrt(t[0] * t[0] + t[1] * t[1] + t[2] * t[2]);
  if (d > (r1 + r2)) return 0;
  if (d <= 0) {
    for (int i = 0; i < 3; i++) o.pos[0][i] = a.pos[i];
    o.nrm[0][0] = 1; o.nrm[0][1] = 0; o.nrm[0][2] = 0;
    o.depth[0] = r1 + r2;
  } else {
    double d1 = 1.0 / d;
    for (int i = 0; i < 3; i++) o.nrm[0][i] = (a.pos[i] - b.pos[i]) * d1;
    double k = 0.5 * (r2 - r1 - d);
    for (int i = 0; i < 3; i++) o.pos[0][i] = a.pos[i] + o.nrm[0][i] * k;
    o.depth[0] = r1 + r2 - d;
  }
  return 1;
}

// ---- dCollideSphereBox (sphere.cpp:130-223) ---------------------------------
__device__ inline int col_sphere_box(const Geom &s, const Geom &bx, COut &o) {
  double l[3], t[3], p[3], q[3], r[3];
  int onborder = 0;
  for (int i = 0; i < 3; i++) p[i] = s.pos[i] - bx.pos[i];
  for (int i = 0; i < 3; i++) {
    l[i] = bx.dims[i] * 0.5;
    t[i] = dot3s(p, 1, bx.R + i, 3);
    if (t[i] < -l[i]) { t[i] = -l[i]; onborder = 1; }
    if (t[i] > l[i]) { t[i] = l[i]; onborder = 1; }
  }
  if (!onborder) {
    double min_distance = l[0] - fabs(t[0]);
    int mini = 0;
    for (int i = 1; i < 3; i++) {
      double fd = l[i] - fabs(t[i]);
      if (fd < min_distance) { min_distance = fd; mini = i; }
    }
    for (int i = 0; i < 3; i++) o.pos[0][i] = s.pos[i];
    double tmp[3] = {0, 0, 0};
    tmp[mini] = (t[mini] > 0) ? 1.0 : -1.0;
    mul0_331(o.nrm[0], bx.R, tmp);
    o.depth[0] = min_distance + s.dims[0];
    return 1;
  }
  mul0_331(q, bx.R, t);
  for (int i = 0; i < 3; i++) r[i] = p[i] - q[i];
  double depth = s.dims[0] - sqrt(dot3(r, r));
  if (depth < 0) return 0;
  for (int i = 0; i < 3; i++) { o.pos[0][i] = q[i] + bx.pos[i]; o.nrm[0][i] = r[i]; }
  safe_normalize3(o.nrm[0]);
  o.depth[0] = depth;
  return 1;
}

// ---- dCollideBoxPlane (box.cpp:746-879), flags = 250 -> maxc = 4 ------------
__device__ inline int col_box_plane(const Geom &bx, const Geom &pl, COut &o) {
  const double *R = bx.R;
  const double *n = pl.dims;
  const double *side = bx.dims;
  double Q[3], A[3], B[3];
  for (int i = 0; i < 3; i++) {
    Q[i] = dot3s(n, 1, R + i, 3);
    A[i] = side[i] * Q[i];
    B[i] = fabs(A[i]);
  }
  double depth = n[3] + 0.5 * (B[0] + B[1] + B[2]) - dot3(n, bx.pos);
  if (depth < 0) return 0;
  double p[3] = {bx.pos[0], bx.pos[1], bx.pos[2]};
  for (int i = 0; i < 3; i++) {
    if (A[i] > 0) { for (int k = 0; k < 3; k++) p[k] -= 0.5 * side[i] * R[3 * k + i]; }
    else          { for (int k = 0; k < 3; k++) p[k] += 0.5 * side[i] * R[3 * k + i]; }
  }
  for (int k = 0; k < 3; k++) o.pos[0][k] = p[k];
  o.depth[0] = depth;
  int ret = 1;
  // second and third points: walk from p along the two sides with the smallest projection
  int s1, s2;   // side used for contact 1, contact 2
  if (B[0] < B[1]) {
    if (B[2] < B[0]) { s1 = 2; s2 = (B[0] < B[1]) ? 0 : 1; }
    else { s1 = 0; s2 = (B[1] < B[2]) ? 1 : 2; }
  } else {
    if (B[2] < B[1]) { s1 = 2; s2 = (B[0] < B[1]) ? 0 : 1; }
    else { s1 = 1; s2 = (B[0] < B[2]) ? 0 : 2; }
  }
  int sides[2] = {s1, s2};
  for (int c = 0; c < 2; c++) {
    int j = sides[c];
    if (depth - B[j] < 0) break;
    if (A[j] > 0) { for (int k = 0; k < 3; k++) o.pos[c + 1][k] = p[k] + side[j] * R[3 * k + j]; }
    else          { for (int k = 0; k < 3; k++) o.pos[c + 1][k] = p[k] - side[j] * R[3 * k + j]; }
    o.depth[c + 1] = depth - B[j];
    ret++;
  }
  if (ret == 3) {
    double d4 = o.depth[1] + o.depth[2] - depth;
    if (d4 > 0) {
      for (int k = 0; k < 3; k++) o.pos[3][k] = o.pos[1][k] + o.pos[2][k] - p[k];
      o.depth[3] = d4;
      ret++;
    }
  }
  for (int i = 0; i < ret; i++) for (int k = 0; k < 3; k++) o.nrm[i][k] = n[k];
  return ret;
}

// ---- dCollideSpheres (collision_util.cpp:38-67) on two points, contact slot i ----
__device__ inline int spheres_pp(const double *p1, double r1, const double *p2, double r2, COut &o, int i) {
  double t[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
  double d = sqrt(t[0] * t[0] + t[1] * t[1] + t[2] * t[2]);
  if (d > (r1 + r2)) return 0;
  if (d <= 0) {
    for (int k = 0; k < 3; k++) o.pos[i][k] = p1[k];
    o.nrm[i][0] = 1; o.nrm[i][1] = 0; o.nrm[i][2] = 0;
    o.depth[i] = r1 + r2;
  } else {
    double d1 = 1.0 / d;
    for (int k = 0; k < 3; k++) o.nrm[i][k] = (p1[k] - p2[k]) * d1;
    double kk = 0.5 * (r2 - r1 - d);
    for (int k = 0; k < 3; k++) o.pos[i][k] = p1[k] + o.nrm[i][k] * kk;
    o.depth[i] = r1 + r2 - d;
  }
  return 1;
}

// ---- dClosestLineSegmentPoints (collision_util.cpp:109-225): closest points of segments a1a2, b1b2,
// by Voronoi features in the reference's order: vertex-vertex, edge-vertex, then edge-edge ----
__device__ inline void closest_segment_points(const double *a1, const double *a2, const double *b1, const double *b2,
                                              double *cp1, double *cp2) {
  double a1a2[3], b1b2[3], a1b1[3], a1b2[3], a2b1[3], a2b2[3], n[3];
  for (int k = 0; k < 3; k++) { a1a2[k] = a2[k] - a1[k]; b1b2[k] = b2[k] - b1[k]; a1b1[k] = b1[k] - a1[k]; }
  const double da1 = dot3(a1a2, a1b1), db1 = dot3(b1b2, a1b1);
  if (da1 <= 0 && db1 >= 0) { for (int k = 0; k < 3; k++) { cp1[k] = a1[k]; cp2[k] = b1[k]; } return; }
  for (int k = 0; k < 3; k++) a1b2[k] = b2[k] - a1[k];
  const double da2 = dot3(a1a2, a1b2), db2 = dot3(b1b2, a1b2);
  if (da2 <= 0 && db2 <= 0) { for (int k = 0; k < 3; k++) { cp1[k] = a1[k]; cp2[k] = b2[k]; } return; }
  for (int k = 0; k < 3; k++) a2b1[k] = b1[k] - a2[k];
  const double da3 = dot3(a1a2, a2b1), db3 = dot3(b1b2, a2b1);
  if (da3 >= 0 && db3 >= 0) { for (int k = 0; k < 3; k++) { cp1[k] = a2[k]; cp2[k] = b1[k]; } return; }
  for (int k = 0; k < 3; k++) a2b2[k] = b2[k] - a2[k];
  const double da4 = dot3(a1a2, a2b2), db4 = dot3(b1b2, a2b2);
  if (da4 >= 0 && db4 <= 0) { for (int k = 0; k < 3; k++) { cp1[k] = a2[k]; cp2[k] = b2[k]; } return; }
  const double la = dot3(a1a2, a1a2);
  if (da1 >= 0 && da3 <= 0) {
    const double kk = da1 / la;
    for (int k = 0; k < 3; k++) n[k] = a1b1[k] - kk * a1a2[k];
    if (dot3(b1b2, n) >= 0) { for (int k = 0; k < 3; k++) { cp1[k] = a1[k] + kk * a1a2[k]; cp2[k] = b1[k]; } return; }
  }
  if (da2 >= 0 && da4 <= 0) {
    const double kk = da2 / la;
    for (int k = 0; k < 3; k++) n[k] = a1b2[k] - kk * a1a2[k];
    if (dot3(b1b2, n) <= 0) { for (int k = 0; k < 3; k++) { cp1[k] = a1[k] + kk * a1a2[k]; cp2[k] = b2[k]; } return; }
  }
  const double lb = dot3(b1b2, b1b2);
  if (db1 <= 0 && db2 >= 0) {
    const double kk = -db1 / lb;
    for (int k = 0; k < 3; k++) n[k] = -a1b1[k] - kk * b1b2[k];
    if (dot3(a1a2, n) >= 0) { for (int k = 0; k < 3; k++) { cp1[k] = a1[k]; cp2[k] = b1[k] + kk * b1b2[k]; } return; }
  }
  if (db3 <= 0 && db4 >= 0) {
    const double kk = -db3 / lb;
    for (int k = 0; k < 3; k++) n[k] = -a2b1[k] - kk * b1b2[k];
    if (dot3(a1a2, n) <= 0) { for (int k = 0; k < 3; k++) { cp1[k] = a2[k]; cp2[k] = b1[k] + kk * b1b2[k]; } return; }
  }
  const double kk = dot3(a1a2, b1b2);
  double det = la * lb - kk * kk;
  if (det <= 0) { for (int k = 0; k < 3; k++) { cp1[k] = a1[k]; cp2[k] = b1[k]; } return; }
  det = 1.0 / det;
  const double alpha = (lb * da1 - kk * db1) * det;
  const double beta = (kk * da1 - la * db1) * det;
  for (int k = 0; k < 3; k++) { cp1[k] = a1[k] + alpha * a1a2[k]; cp2[k] = b1[k] + beta * b1b2[k]; }
}

// ---- dClosestLineBoxPoints (collision_util.cpp:247-390): closest points of the segment p1p2 and the
// box (c, R, side): piecewise-linear root of d|D(t)|^2/dt over t in [0, 1], in box coordinates with
// the line mirrored so that its direction is non-negative ----
__device__ inline void closest_line_box_points(const double *p1, const double *p2, const double *c, const double *R,
                                               const double *side, double *lret, double *bret) {
  double tmp[3], s[3], v[3], sign[3], v2[3], h[3], tanchor[3];
  int region[3];
  for (int k = 0; k < 3; k++) tmp[k] = p1[k] - c[k];
  mul1_331(s, R, tmp);
  for (int k = 0; k < 3; k++) tmp[k] = p2[k] - p1[k];
  mul1_331(v, R, tmp);
  for (int i = 0; i < 3; i++) {
    if (v[i] < 0) { s[i] = -s[i]; v[i] = -v[i]; sign[i] = -1; } else sign[i] = 1;
  }
  for (int i = 0; i < 3; i++) { v2[i] = v[i] * v[i]; h[i] = 0.5 * side[i]; }
  const double tanchor_eps = 1e-307;
  for (int i = 0; i < 3; i++) {
    if (v[i] > tanchor_eps) {
      if (s[i] < -h[i]) { region[i] = -1; tanchor[i] = (-h[i] - s[i]) / v[i]; }
      else { region[i] = (s[i] > h[i]); tanchor[i] = (h[i] - s[i]) / v[i]; }
    } else { region[i] = 0; tanchor[i] = 2; }
  }
  double t = 0, dd2dt = 0;
  for (int i = 0; i < 3; i++) dd2dt -= (region[i] ? v2[i] : 0) * tanchor[i];
  if (!(dd2dt >= 0)) {
    bool found = false;
    do {
      double next_t = 1;
      for (int i = 0; i < 3; i++)
        if (tanchor[i] > t && tanchor[i] < 1 && tanchor[i] < next_t) next_t = tanchor[i];
      double next_dd2dt = 0;
      for (int i = 0; i < 3; i++) next_dd2dt += (region[i] ? v2[i] : 0) * (next_t - tanchor[i]);
      if (next_dd2dt >= 0) {
        const double m = (next_dd2dt - dd2dt) / (next_t - t);
        t -= dd2dt / m;
        found = true;
        break;
      }
      for (int i = 0; i < 3; i++)
        if (dequal(tanchor[i], next_t)) { tanchor[i] = (h[i] - s[i]) / v[i]; region[i]++; }
      t = next_t;
      dd2dt = next_dd2dt;
    } while (t < 1);
    if (!found) t = 1;
  }
  for (int i = 0; i < 3; i++) lret[i] = p1[i] + t * tmp[i];
  for (int i = 0; i < 3; i++) {
    tmp[i] = sign[i] * (s[i] + t * v[i]);
    if (tmp[i] < -h[i]) tmp[i] = -h[i];
    else if (tmp[i] > h[i]) tmp[i] = h[i];
  }
  mul0_331(s, R, tmp);
  for (int i = 0; i < 3; i++) bret[i] = s[i] + c[i];
}

// ---- dCollideCapsuleSphere (capsule.cpp:130-162) ------------------------------
__device__ inline int col_capsule_sphere(const Geom &c, const Geom &sp, COut &o) {
  double alpha = c.R[2] * (sp.pos[0] - c.pos[0]) + c.R[5] * (sp.pos[1] - c.pos[1]) + c.R[8] * (sp.pos[2] - c.pos[2]);
  const double lz2 = c.dims[1] * 0.5;
  if (alpha > lz2) alpha = lz2;
  if (alpha < -lz2) alpha = -lz2;
  double p[3];
  for (int k = 0; k < 3; k++) p[k] = c.pos[k] + alpha * c.R[3 * k + 2];
  return spheres_pp(p, c.dims[0], sp.pos, sp.dims[0], o, 0);
}

// ---- dCollideCapsuleBox (capsule.cpp:165-205) -----------------------------------
__device__ inline int col_capsule_box(const Geom &c, const Geom &bx, COut &o) {
  const double clen = c.dims[1] * 0.5;
  double p1[3], p2[3], pl[3], pb[3];
  for (int k = 0; k < 3; k++) { p1[k] = c.pos[k] + clen * c.R[3 * k + 2]; p2[k] = c.pos[k] - clen * c.R[3 * k + 2]; }
  closest_line_box_points(p1, p2, bx.pos, bx.R, bx.dims, pl, pb);
  return spheres_pp(pl, c.dims[0], pb, 0.0, o, 0);
}

// ---- dCollideCapsuleCapsule (capsule.cpp:208-327): (almost) parallel axes give up to two contacts at
// the ends of the overlap, else the closest points of the two axis segments ----
__device__ inline int col_capsule_capsule(const Geom &c1, const Geom &c2, COut &o) {
  const double tolerance = 1e-5;
  const double lz1 = c1.dims[1] * 0.5, lz2 = c2.dims[1] * 0.5, r1 = c1.dims[0], r2 = c2.dims[0];
  const double *pos1 = c1.pos, *pos2 = c2.pos;
  double axis1[3] = {c1.R[2], c1.R[5], c1.R[8]}, axis2[3] = {c2.R[2], c2.R[5], c2.R[8]};
  double sphere1[3], sphere2[3];
  const double a1a2 = dot3(axis1, axis2);
  const double det = 1.0 - a1a2 * a1a2;
  if (det < tolerance) {
    if (a1a2 < 0) { axis2[0] = -axis2[0]; axis2[1] = -axis2[1]; axis2[2] = -axis2[2]; }
    double q[3];
    for (int i = 0; i < 3; i++) q[i] = pos1[i] - pos2[i];
    const double k = dot3(axis1, q);
    const double a1lo = -lz1, a1hi = lz1, a2lo = -lz2 - k, a2hi = lz2 - k;
    const double lo = (a1lo > a2lo) ? a1lo : a2lo, hi = (a1hi < a2hi) ? a1hi : a2hi;
    if (lo <= hi) {
      if (lo < hi) {   // Gazebo asks for 250 contacts: the two-contact branch is always allowed
        for (int i = 0; i < 3; i++) { sphere1[i] = pos1[i] + lo * axis1[i]; sphere2[i] = pos2[i] + (lo + k) * axis2[i]; }
        if (spheres_pp(sphere1, r1, sphere2, r2, o, 0)) {
          for (int i = 0; i < 3; i++) { sphere1[i] = pos1[i] + hi * axis1[i]; sphere2[i] = pos2[i] + (hi + k) * axis2[i]; }
          if (spheres_pp(sphere1, r1, sphere2, r2, o, 1)) return 2;
        }
      }
      const double alpha1 = (lo + hi) * 0.5, alpha2 = alpha1 + k;
      for (int i = 0; i < 3; i++) { sphere1[i] = pos1[i] + alpha1 * axis1[i]; sphere2[i] = pos2[i] + alpha2 * axis2[i]; }
      return spheres_pp(sphere1, r1, sphere2, r2, o, 0);
    }
  }
  double a1[3], a2[3], b1[3], b2[3];
  for (int i = 0; i < 3; i++) {
    a1[i] = pos1[i] + axis1[i] * lz1; a2[i] = pos1[i] - axis1[i] * lz1;
    b1[i] = pos2[i] + axis2[i] * lz2; b2[i] = pos2[i] - axis2[i] * lz2;
  }
  closest_segment_points(a1, a2, b1, b2, sphere1, sphere2);
  return spheres_pp(sphere1, r1, sphere2, r2, o, 0);
}

// ---- dCollideCapsulePlane (capsule.cpp:330-383) -----------------------------
__device__ inline int col_capsule_plane(const Geom &c, const Geom &pl, COut &o) {
  const double *pp = pl.dims;
  double radius = c.dims[0], lz = c.dims[1];
  double sign = (dot3s(pp, 1, c.R + 2, 3) > 0) ? -1.0 : 1.0;
  double p[3];
  for (int k = 0; k < 3; k++) p[k] = c.pos[k] + c.R[3 * k + 2] * lz * 0.5 * sign;
  double k0 = dot3(p, pp);
  double depth = pp[3] - k0 + radius;
  if (depth < 0) return 0;
  for (int k = 0; k < 3; k++) { o.nrm[0][k] = pp[k]; o.pos[0][k] = p[k] - pp[k] * radius; }
  o.depth[0] = depth;
  int n = 1;
  for (int k = 0; k < 3; k++) p[k] = c.pos[k] - c.R[3 * k + 2] * lz * 0.5 * sign;
  k0 = dot3(p, pp);
  depth = pp[3] - k0 + radius;
  if (depth >= 0) {
    for (int k = 0; k < 3; k++) { o.nrm[1][k] = pp[k]; o.pos[1][k] = p[k] - pp[k] * radius; }
    o.depth[1] = depth;
    n = 2;
  }
  return n;
}

// ---- dCollideCylinderSphere (collision_cylinder_sphere.cpp:53-278): one contact; three cases by where
// the sphere centre projects: beside the barrel, over a cap, or near a rim ----
__device__ inline int col_cylinder_sphere(const Geom &c, const Geom &sp, COut &o) {
  const double toleranz = 0.0000001;
  const double radius = c.dims[0], length = c.dims[1], radius2 = sp.dims[0];
  const double *cylpos = c.pos, *SpherePos = sp.pos;
  const double vDir1[3] = {c.R[2], c.R[5], c.R[8]};
  double G1Pos1[3], G1Pos2[3], C[3];
  double s = length * 0.5;
  for (int k = 0; k < 3; k++) { G1Pos2[k] = vDir1[k] * s + cylpos[k]; G1Pos1[k] = vDir1[k] * -s + cylpos[k]; }
  s = (SpherePos[0] - G1Pos1[0]) * vDir1[0] - (G1Pos1[1] - SpherePos[1]) * vDir1[1] - (G1Pos1[2] - SpherePos[2]) * vDir1[2];
  if (s < (-radius2) || s > (length + radius2)) return 0;
  for (int k = 0; k < 3; k++) C[k] = s * vDir1[k] + G1Pos1[k] - SpherePos[k];
  const double t = sqrt(C[0] * C[0] + C[1] * C[1] + C[2] * C[2]);
  if (t > (radius + radius2)) return 0;
  if (t > radius && (s < 0 || s > length)) {
    // rim
    const double *G = (s <= 0) ? G1Pos1 : G1Pos2;
    const double ds = (s <= 0) ? s : (s - length);
    const double depth = radius2 - sqrt(ds * ds + (t - radius) * (t - radius));
    if (depth < 0) return 0;
    for (int k = 0; k < 3; k++) o.pos[0][k] = C[k] / t * -radius + G[k];
    for (int k = 0; k < 3; k++) o.nrm[0][k] = (o.pos[0][k] - SpherePos[k]) / (radius2 - depth);
    o.depth[0] = depth;
    return 1;
  } else if ((radius - t) <= s && (radius - t) <= (length - s)) {
    // barrel
    const double depth = (radius2 + radius) - t;
    if (depth < 0) return 0;
    if (t > (radius2 + toleranz)) {
      for (int k = 0; k < 3; k++) C[k] /= t;
      for (int k = 0; k < 3; k++) { o.pos[0][k] = C[k] * radius2 + SpherePos[k]; o.nrm[0][k] = C[k]; }
    } else {
      for (int k = 0; k < 3; k++) { o.pos[0][k] = C[k] + SpherePos[k]; o.nrm[0][k] = C[k] / t; }
    }
    o.depth[0] = depth;
    return 1;
  } else {
    // caps
    if (s <= (length * 0.5)) {
      const double depth = s + radius2;
      if (depth < 0) return 0;
      for (int k = 0; k < 3; k++) { o.pos[0][k] = radius2 * vDir1[k] + SpherePos[k]; o.nrm[0][k] = vDir1[k]; }
      o.depth[0] = depth;
    } else {
      const double depth = (radius2 + length - s);
      if (depth < 0) return 0;
      for (int k = 0; k < 3; k++) { o.pos[0][k] = radius2 * -vDir1[k] + SpherePos[k]; o.nrm[0][k] = -vDir1[k]; }
      o.depth[0] = depth;
    }
    return 1;
  }
}

// ---- dCollideCylinderPlane (collision_cylinder_plane.cpp:43-267) ------------
__device__ inline int col_cylinder_plane(const Geom &c, const Geom &pl, COut &o) {
  const double toleranz = 0.0000001;
  double radius = c.dims[0], length = c.dims[1];
  const double *pv = pl.dims;
  double N[3] = {pv[0], pv[1], pv[2]};
  double vDir1[3] = {c.R[2], c.R[5], c.R[8]};
  double s = length * 0.5;
  double G1Pos1[3], G1Pos2[3];
  for (int k = 0; k < 3; k++) { G1Pos2[k] = vDir1[k] * s + c.pos[k]; G1Pos1[k] = vDir1[k] * -s + c.pos[k]; }
  int cnt = 0;
  s = vDir1[0] * N[0] + vDir1[1] * N[1] + vDir1[2] * N[2];
  if (s < 0) s += 1.0; else s -= 1.0;
  if (s < toleranz && s > (-toleranz)) {
    double P[3];
    s = pv[3] - dot3(pv, G1Pos1);
    double t = pv[3] - dot3(pv, G1Pos2);
    if (s >= t) {
      if (s >= 0) { for (int k = 0; k < 3; k++) P[k] = G1Pos1[k]; } else return 0;
    } else {
      if (t >= 0) { for (int k = 0; k < 3; k++) P[k] = G1Pos2[k]; } else return 0;
    }
    double V1[3], V2[3];
    if (vDir1[0] < toleranz && vDir1[0] > (-toleranz)) { V1[0] = vDir1[0] + 1.0; V1[1] = vDir1[1]; V1[2] = vDir1[2]; }
    else { V1[0] = vDir1[0]; V1[1] = vDir1[1] + 1.0; V1[2] = vDir1[2]; }
    cross3(V2, V1, vDir1);
    t = sqrt(V2[0] * V2[0] + V2[1] * V2[1] + V2[2] * V2[2]);
    t = radius / t;
    V2[0] *= t; V2[1] *= t; V2[2] *= t;
    cross3(V1, V2, vDir1);
    for (int c4 = 0; c4 < 4; c4++) {
      const double *V = (c4 < 2) ? V1 : V2;
      double q[3];
      if ((c4 & 1) == 0) { for (int k = 0; k < 3; k++) q[k] = P[k] + V[k]; }
      else               { for (int k = 0; k < 3; k++) q[k] = P[k] - V[k]; }
      double d = pv[3] - dot3(pv, q);
      if (d > 0) {
        for (int k = 0; k < 3; k++) { o.pos[cnt][k] = q[k]; o.nrm[cnt][k] = N[k]; }
        o.depth[cnt] = d;
        cnt++;
      }
    }
  } else {
    double t = dot3(N, vDir1);
    double C[3];
    for (int k = 0; k < 3; k++) C[k] = vDir1[k] * t - N[k];
    s = sqrt(C[0] * C[0] + C[1] * C[1] + C[2] * C[2]);
    s = radius / s;
    C[0] *= s; C[1] *= s; C[2] *= s;
    double q[3];
    for (int k = 0; k < 3; k++) q[k] = C[k] + G1Pos1[k];
    double d = pv[3] - dot3(pv, q);
    if (d >= 0) {
      for (int k = 0; k < 3; k++) { o.pos[cnt][k] = q[k]; o.nrm[cnt][k] = N[k]; }
      o.depth[cnt] = d; cnt++;
    }
    for (int k = 0; k < 3; k++) q[k] = C[k] + G1Pos2[k];
    d = pv[3] - pv[0] * q[0] - pv[1] * q[1] - pv[2] * q[2];
    if (d >= 0) {
      for (int k = 0; k < 3; k++) { o.pos[cnt][k] = q[k]; o.nrm[cnt][k] = N[k]; }
      o.depth[cnt] = d; cnt++;
    }
  }
  return cnt;
}

}  // namespace gz
