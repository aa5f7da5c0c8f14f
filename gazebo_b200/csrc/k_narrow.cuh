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

struct COut {          // up to 8 contacts: position, depth; one normal per contact
  double pos[8][3];
  double nrm[8][3];
  double depth[8];
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
