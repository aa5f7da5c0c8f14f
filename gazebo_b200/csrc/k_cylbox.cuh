// k_cylbox.cuh — cylinder vs box narrowphase, a restatement of dCollideCylinderBox
// (deps/opende/src/collision_cylinder_box.cpp:100-1038; registered for (cylinder, box) at
// collision_kernel.cpp:211, Gazebo's build does not define dLIBCCD_BOX_CYL):
//   * separating-axis search over 3 box axes, the cylinder axis, 3 cross products, 8 box-vertex
//     axes and 2 x 12 box-edge / cap-circle axes, keeping the axis of least overlap (:206-574);
//   * |normal . axis| < 0.9: the cylinder's nearest generator line is clipped to the six box
//     planes, 2 contacts (:576-717);
//   * otherwise: the box face most opposed to the normal is clipped, in the cap's frame, to the cap
//     plane and to the 8 planes of the octagon inscribed... circumscribing the cap circle, up to 13
//     contacts (:720-983).
// Every expression keeps the reference's operand order (the library is compiled with -fmad=false);
// the quirks are kept too: dMatrix3Inv's missing parentheses (collision_util.h:192-216) and the
// in-place normalisation of the candidate axis.  Matrices here are 3 x 3 row-major (R[3 i + j]
// is the reference's m[4 i + j]).
#pragma once

namespace gz {

struct CylBox {
  const double *Rc, *pc, *Rb, *pb;
  double axis[3], radius, size, h[3];
  double vert[8][3];
  double diff[3], normal[3];
  double best_depth, best_rb, best_rc;
  int best_axis;
};

__device__ inline void cb_col(const double *R, int c, double *v) { v[0] = R[c]; v[1] = R[3 + c]; v[2] = R[6 + c]; }

// _cldTestAxis (:206-286); v is normalised in place
__device__ inline int cb_test_axis(CylBox &D, double *v, int iaxis) {
  const double fl = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  if (fl < 1e-5) return 1;
  safe_normalize3(v);
  const double fdot1 = dot3(D.axis, v);
  double frc;
  if (fdot1 > 1.0) frc = D.size * 0.5;
  else if (fdot1 < -1.0) frc = D.size * 0.5;
  else frc = fabs(fdot1 * (D.size * 0.5)) + D.radius * sqrt(1.0 - (fdot1 * fdot1));
  double t[3];
  cb_col(D.Rb, 0, t);
  double frb = fabs(dot3(t, v)) * D.h[0];
  cb_col(D.Rb, 1, t);
  frb += fabs(dot3(t, v)) * D.h[1];
  cb_col(D.Rb, 2, t);
  frb += fabs(dot3(t, v)) * D.h[2];
  const double fd = dot3(D.diff, v);
  double depth = frc + frb;
  if (fabs(fd) > depth) return 0;
  depth -= fabs(fd);
  if (depth < D.best_depth) {
    D.best_depth = depth;
    D.normal[0] = v[0]; D.normal[1] = v[1]; D.normal[2] = v[2];
    D.best_axis = iaxis;
    D.best_rb = frb;
    D.best_rc = frc;
    if (fd > 0) { D.normal[0] = -D.normal[0]; D.normal[1] = -D.normal[1]; D.normal[2] = -D.normal[2]; }
  }
  return 1;
}

// _cldTestEdgeCircleAxis (:289-333)
__device__ inline int cb_test_edge_circle(CylBox &D, const double *centre, const double *v0, const double *v1, int iaxis) {
  double dir[3] = {v1[0] - v0[0], v1[1] - v0[1], v1[2] - v0[2]};
  safe_normalize3(dir);
  const double fdot2 = dot3(dir, D.axis);
  if (fabs(fdot2) < 1e-5) return 1;
  double t[3] = {centre[0] - v0[0], centre[1] - v0[1], centre[2] - v0[2]};
  const double fdot1 = dot3(t, D.axis);
  double pnt[3];
  pnt[0] = v0[0] + dir[0] * (fdot1 / fdot2);
  pnt[1] = v0[1] + dir[1] * (fdot1 / fdot2);
  pnt[2] = v0[2] + dir[2] * (fdot1 / fdot2);
  t[0] = centre[0] - pnt[0]; t[1] = centre[1] - pnt[1]; t[2] = centre[2] - pnt[2];
  double tangent[3], ax[3];
  cross3(tangent, t, D.axis);
  cross3(ax, tangent, dir);
  return cb_test_axis(D, ax, iaxis);
}

// _cldTestSeparatingAxes (:336-574)
__device__ inline int cb_separating_axes(CylBox &D) {
  const double eps = 1e-6;
  double ax[3];
  for (int c = 0; c < 3; c++) {
    cb_col(D.Rb, c, ax);
    if (!cb_test_axis(D, ax, 1 + c)) return 0;
  }
  ax[0] = D.axis[0]; ax[1] = D.axis[1]; ax[2] = D.axis[2];
  if (!cb_test_axis(D, ax, 4)) return 0;
  for (int c = 0; c < 3; c++) {   // cylinder axis x box column c (dCalcVectorCross3_114)
    double col[3];
    cb_col(D.Rb, c, col);
    cross3(ax, D.axis, col);
    if (ax[0] * ax[0] + ax[1] * ax[1] + ax[2] * ax[2] > eps)
      if (!cb_test_axis(D, ax, 5 + c)) return 0;
  }
  for (int i = 0; i < 8; i++) {
    double t1[3] = {D.vert[i][0] - D.pc[0], D.vert[i][1] - D.pc[1], D.vert[i][2] - D.pc[2]}, t2[3];
    cross3(t2, D.axis, t1);
    cross3(ax, D.axis, t2);
    if (ax[0] * ax[0] + ax[1] * ax[1] + ax[2] * ax[2] > eps)
      if (!cb_test_axis(D, ax, 8 + i)) return 0;
  }
  // the 12 box edges against the top, then the bottom cap circle
  const int e0[12] = {1, 1, 2, 2, 4, 4, 0, 5, 5, 2, 4, 6};
  const int e1[12] = {0, 3, 3, 0, 1, 7, 7, 3, 6, 6, 5, 7};
  double vcc[3];
  vcc[0] = D.pc[0] + D.axis[0] * (D.size * 0.5);
  vcc[1] = D.pc[1] + D.axis[1] * (D.size * 0.5);
  vcc[2] = D.pc[2] + D.axis[2] * (D.size * 0.5);
  for (int e = 0; e < 12; e++)
    if (!cb_test_edge_circle(D, vcc, D.vert[e0[e]], D.vert[e1[e]], 16 + e)) return 0;
  vcc[0] = D.pc[0] - D.axis[0] * (D.size * 0.5);
  vcc[1] = D.pc[1] - D.axis[1] * (D.size * 0.5);
  vcc[2] = D.pc[2] - D.axis[2] * (D.size * 0.5);
  for (int e = 0; e < 12; e++)
    if (!cb_test_edge_circle(D, vcc, D.vert[e0[e]], D.vert[e1[e]], 28 + e)) return 0;
  return 1;
}

// dClipEdgeToPlane (collision_util.cpp:473-511)
__device__ inline int cb_clip_edge(double *p0, double *p1, const double *pl) {
  const double d0 = pl[0] * p0[0] + pl[1] * p0[1] + pl[2] * p0[2] + pl[3];
  const double d1 = pl[0] * p1[0] + pl[1] * p1[1] + pl[2] * p1[2] + pl[3];
  if (d0 < 0 && d1 < 0) return 0;
  if (d0 > 0 && d1 > 0) return 1;
  if ((d0 > 0 && d1 < 0) || (d0 < 0 && d1 > 0)) {
    double x[3];
    x[0] = p0[0] - (p0[0] - p1[0]) * d0 / (d0 - d1);
    x[1] = p0[1] - (p0[1] - p1[1]) * d0 / (d0 - d1);
    x[2] = p0[2] - (p0[2] - p1[2]) * d0 / (d0 - d1);
    double *dst = (d0 < 0) ? p0 : p1;
    dst[0] = x[0]; dst[1] = x[1]; dst[2] = x[2];
  }
  return 1;
}

// dClipPolyToPlane (collision_util.cpp:514-560)
__device__ inline int cb_clip_poly(const double (*in)[3], int n_in, double (*out)[3], const double *pl) {
  int n = 0;
  int i0 = n_in - 1;
  for (int i1 = 0; i1 < n_in; i0 = i1, i1++) {
    const double d0 = pl[0] * in[i0][0] + pl[1] * in[i0][1] + pl[2] * in[i0][2] + pl[3];
    const double d1 = pl[0] * in[i1][0] + pl[1] * in[i1][1] + pl[2] * in[i1][2] + pl[3];
    if (d0 >= 0) { out[n][0] = in[i0][0]; out[n][1] = in[i0][1]; out[n][2] = in[i0][2]; n++; }
    if ((d0 > 0 && d1 < 0) || (d0 < 0 && d1 > 0)) {
      out[n][0] = in[i0][0] - (in[i0][0] - in[i1][0]) * d0 / (d0 - d1);
      out[n][1] = in[i0][1] - (in[i0][1] - in[i1][1]) * d0 / (d0 - d1);
      out[n][2] = in[i0][2] - (in[i0][2] - in[i1][2]) * d0 / (d0 - d1);
      n++;
    }
  }
  return n;
}

// dMatrix3Inv (collision_util.h:192-216), operator precedence as written there
__device__ inline void cb_mat_inv(const double *m, double *d) {
  const double det = m[0] * (m[4] * m[8] - m[7] * m[5]) - m[1] * (m[3] * m[8] - m[6] * m[5]) + m[2] * (m[3] * m[7] - m[6] * m[4]);
  if (fabs(det) < 0.0005) {
    d[0] = 1; d[1] = 0; d[2] = 0; d[3] = 0; d[4] = 1; d[5] = 0; d[6] = 0; d[7] = 0; d[8] = 1;
    return;
  }
  d[0] = m[4] * m[8] - m[5] * m[7] / det;
  d[1] = -(m[1] * m[8] - m[7] * m[2]) / det;
  d[2] = m[1] * m[5] - m[4] * m[2] / det;
  d[3] = -(m[3] * m[8] - m[5] * m[6]) / det;
  d[4] = m[0] * m[8] - m[6] * m[2] / det;
  d[5] = -(m[0] * m[5] - m[3] * m[2]) / det;
  d[6] = m[3] * m[7] - m[6] * m[4] / det;
  d[7] = -(m[0] * m[7] - m[6] * m[1]) / det;
  d[8] = m[0] * m[4] - m[1] * m[3] / det;
}

// _cldClipCylinderToBox (:576-717)
__device__ inline int cb_clip_cylinder_to_box(CylBox &D, COut &o) {
  double vN[3];
  const double f1 = dot3(D.axis, D.normal);
  vN[0] = D.normal[0] - D.axis[0] * f1;
  vN[1] = D.normal[1] - D.axis[1] * f1;
  vN[2] = D.normal[2] - D.axis[2] * f1;
  safe_normalize3(vN);
  double ct[3], ep0[3], ep1[3];
  for (int k = 0; k < 3; k++) ct[k] = D.pc[k] + vN[k] * D.radius;
  for (int k = 0; k < 3; k++) { ep0[k] = ct[k] + D.axis[k] * (D.size * 0.5); ep1[k] = ct[k] - D.axis[k] * (D.size * 0.5); }
  for (int k = 0; k < 3; k++) { ep0[k] -= D.pb[k]; ep1[k] -= D.pb[k]; }
  for (int s = 0; s < 2; s++)
    for (int c = 0; c < 3; c++) {
      double pl[4];
      cb_col(D.Rb, c, pl);
      if (s) { pl[0] = -pl[0]; pl[1] = -pl[1]; pl[2] = -pl[2]; }
      pl[3] = D.h[c];
      if (!cb_clip_edge(ep0, ep1, pl)) return 0;
    }
  double d0 = D.best_rb + dot3(ep0, D.normal);
  double d1 = D.best_rb + dot3(ep1, D.normal);
  if (d0 < 0) d0 = 0.0;
  if (d1 < 0) d1 = 0.0;
  for (int k = 0; k < 3; k++) { ep0[k] += D.pb[k]; ep1[k] += D.pb[k]; }
  for (int k = 0; k < 3; k++) {
    o.pos[0][k] = ep0[k]; o.nrm[0][k] = -D.normal[k];
    o.pos[1][k] = ep1[k]; o.nrm[1][k] = -D.normal[k];
  }
  o.depth[0] = d0; o.depth[1] = d1;
  return 2;
}

// _cldClipBoxToCylinder (:720-983)
__device__ inline int cb_clip_box_to_cylinder(CylBox &D, COut &o) {
  double cpos[3], cnrm[3] = {0.0, 0.0, 0.0};
  if (dot3(D.axis, D.normal) > 0.0) {
    for (int k = 0; k < 3; k++) cpos[k] = D.pc[k] + D.axis[k] * (D.size * 0.5);
    cnrm[2] = -1.0;
  } else {
    for (int k = 0; k < 3; k++) cpos[k] = D.pc[k] - D.axis[k] * (D.size * 0.5);
    cnrm[2] = 1.0;
  }
  double inv[9], vNr[3];
  cb_mat_inv(D.Rb, inv);
  mul0_331(vNr, inv, D.normal);
  const double a0 = fabs(vNr[0]), a1 = fabs(vNr[1]), a2 = fabs(vNr[2]);
  int iB0, iB1, iB2;
  if (a1 > a0) {
    if (a0 > a2) { iB0 = 1; iB1 = 0; iB2 = 2; }
    else if (a1 > a2) { iB0 = 1; iB1 = 2; iB2 = 0; }
    else { iB0 = 2; iB1 = 1; iB2 = 0; }
  } else {
    if (a1 > a2) { iB0 = 0; iB1 = 1; iB2 = 2; }
    else if (a0 > a2) { iB0 = 0; iB1 = 2; iB2 = 1; }
    else { iB0 = 2; iB1 = 0; iB2 = 1; }
  }
  double centre[3], t[3];
  cb_col(D.Rb, iB0, t);
  if (vNr[iB0] > 0) { for (int k = 0; k < 3; k++) centre[k] = D.pb[k] - D.h[iB0] * t[k]; }
  else              { for (int k = 0; k < 3; k++) centre[k] = D.pb[k] + D.h[iB0] * t[k]; }
  double quad[4][3], A1[16][3], A2[16][3];
  double x1[3], x2[3];
  cb_col(D.Rb, iB1, x1);
  cb_col(D.Rb, iB2, x2);
  for (int k = 0; k < 3; k++) {
    quad[0][k] = centre[k] + D.h[iB1] * x1[k] - D.h[iB2] * x2[k];
    quad[1][k] = centre[k] - D.h[iB1] * x1[k] - D.h[iB2] * x2[k];
    quad[2][k] = centre[k] - D.h[iB1] * x1[k] + D.h[iB2] * x2[k];
    quad[3][k] = centre[k] + D.h[iB1] * x1[k] + D.h[iB2] * x2[k];
  }
  double cinv[9];
  cb_mat_inv(D.Rc, cinv);
  for (int i = 0; i < 4; i++) {
    t[0] = quad[i][0] - cpos[0]; t[1] = quad[i][1] - cpos[1]; t[2] = quad[i][2] - cpos[2];
    mul0_331(quad[i], cinv, t);
  }
  double pl[4] = {cnrm[0], cnrm[1], cnrm[2], 0.0};
  int n1 = cb_clip_poly(quad, 4, A1, pl), n2 = 0;
  // the 8 side planes of the polygon around the cap circle (:186-198: angles pi/8 + k pi/4, accumulated)
  double ang = 3.14159265358979323846 / 8;
  const double inc = ang * 2.0;
  for (int s = 0; s < 8; s++) {
    pl[0] = -gzdet_cos(ang); pl[1] = -gzdet_sin(ang); pl[2] = 0; pl[3] = D.radius;
    ang += inc;
    if ((s & 1) == 0) n2 = cb_clip_poly(A1, n1, A2, pl);
    else n1 = cb_clip_poly(A2, n2, A1, pl);
  }
  int cnt = 0;
  for (int i = 0; i < n1; i++) {
    double p[3];
    mul0_331(p, D.Rc, A1[i]);
    p[0] += cpos[0]; p[1] += cpos[1]; p[2] += cpos[2];
    t[0] = p[0] - D.pc[0]; t[1] = p[1] - D.pc[1]; t[2] = p[2] - D.pc[2];
    const double depth = D.best_rc - dot3(t, D.normal);
    if (depth > 0.0) {
      for (int k = 0; k < 3; k++) { o.pos[cnt][k] = p[k]; o.nrm[cnt][k] = -D.normal[k]; }
      o.depth[cnt] = depth;
      cnt++;
      if (cnt == GZ_PAIR_MAXC) break;
    }
  }
  return cnt;
}

// dCollideCylinderBox / sCylinderBoxData::PerformCollisionChecking (:985-1038)
__device__ __noinline__ int col_cylinder_box(const Geom &c, const Geom &b, COut &o) {
  CylBox D;
  D.Rc = c.R; D.pc = c.pos; D.Rb = b.R; D.pb = b.pos;
  cb_col(c.R, 2, D.axis);
  D.radius = c.dims[0]; D.size = c.dims[1];
  for (int k = 0; k < 3; k++) D.h[k] = b.dims[k] * 0.5;
  // box vertices (:126-173), sign patterns of x, y, z
  const int sx[8] = {-1, 1, -1, 1, 1, 1, -1, -1};
  const int sy[8] = {1, 1, -1, -1, 1, -1, -1, 1};
  const int sz[8] = {-1, -1, -1, -1, 1, 1, 1, 1};
  for (int i = 0; i < 8; i++) {
    const double v[3] = {sx[i] < 0 ? -D.h[0] : D.h[0], sy[i] < 0 ? -D.h[1] : D.h[1], sz[i] < 0 ? -D.h[2] : D.h[2]};
    double t[3];
    mul0_331(t, b.R, v);
    D.vert[i][0] = t[0] + b.pos[0]; D.vert[i][1] = t[1] + b.pos[1]; D.vert[i][2] = t[2] + b.pos[2];
  }
  for (int k = 0; k < 3; k++) { D.diff[k] = c.pos[k] - b.pos[k]; D.normal[k] = 0.0; }
  D.best_depth = GZ_INF; D.best_rb = 0; D.best_rc = 0; D.best_axis = 0;
  if (!cb_separating_axes(D)) return 0;
  if (D.best_axis == 0) return 0;
  const double fdot = dot3(D.normal, D.axis);
  if (fabs(fdot) < 0.9) return cb_clip_cylinder_to_box(D, o);
  return cb_clip_box_to_cylinder(D, o);
}

}  // namespace gz
