// k_boxbox.cuh — box/box contact generation, one thread per pair.
//
// Follows the contact set, order and arithmetic of dBoxBox / dCollideBoxBox
// (deps/opende/src/box.cpp:332-743): 15-axis separating-axis test with the 1.05
// edge fudge factor, edge-edge closest points (dLineClosestApproach,
// collision_util.cpp:70-107), or incident-face clipping against the reference
// rectangle (intersectRectQuad, box.cpp:188-239).  Gazebo requests 250 contacts,
// so all (<= 8) penetrating points are returned and cullPoints is never reached.
#pragma once
#include "k_narrow.cuh"

namespace gz {

// intersectRectQuad (box.cpp:188-239): clip quad p[8] to |x|<h0, |y|<h1; <= 8 points out.
__device__ inline int rect_quad(const double h[2], const double p[8], double ret[16]) {
  double bufA[16], bufB[16];
  for (int i = 0; i < 8; i++) bufA[i] = p[i];
  double *q = bufA, *r = bufB;
  int nq = 4, nr = 0;
  for (int dir = 0; dir <= 1; dir++) {
    for (int sign = -1; sign <= 1; sign += 2) {
      nr = 0;
      bool full = false;
      for (int i = 0; i < nq && !full; i++) {
        const double *pq = q + 2 * i;
        const double *nextq = (i + 1 < nq) ? pq + 2 : q;
        bool in0 = sign * pq[dir] < h[dir];
        if (in0) {
          r[2 * nr] = pq[0]; r[2 * nr + 1] = pq[1];
          nr++;
          if (nr & 8) { full = true; break; }
        }
        bool in1 = sign * nextq[dir] < h[dir];
        if (in0 ^ in1) {
          r[2 * nr + (1 - dir)] = pq[1 - dir] + (nextq[1 - dir] - pq[1 - dir]) /
                                  (nextq[dir] - pq[dir]) * (sign * h[dir] - pq[dir]);
          r[2 * nr + dir] = sign * h[dir];
          nr++;
          if (nr & 8) { full = true; break; }
        }
      }
      double *tmp = q; q = r; r = tmp;
      nq = nr;
      if (full) goto done;
    }
  }
done:
  for (int i = 0; i < 2 * nr; i++) ret[i] = q[i];
  return nr;
}

__device__ inline int col_box_box(const Geom &g1, const Geom &g2, COut &o) {
  const double fudge_factor = 1.05;
  const double *p1 = g1.pos, *R1 = g1.R, *p2 = g2.pos, *R2 = g2.R;
  double p[3], pp[3], normalC[3] = {0, 0, 0}, normal[3];
  double A[3], B[3], Rm[3][3], Q[3][3];
  for (int i = 0; i < 3; i++) p[i] = p2[i] - p1[i];
  mul1_331(pp, R1, p);
  for (int i = 0; i < 3; i++) { A[i] = g1.dims[i] * 0.5; B[i] = g2.dims[i] * 0.5; }
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      Rm[i][j] = dot3s(R1 + i, 3, R2 + j, 3);
      Q[i][j] = fabs(Rm[i][j]);
    }
  double s = -GZ_INF, s2, l, e1;
  int invert_normal = 0, code = 0;
  const double *normalR = nullptr;   // column base pointer into R1 or R2

  // face axes of box 1
  for (int i = 0; i < 3; i++) {
    e1 = pp[i];
    s2 = fabs(e1) - (A[i] + B[0] * Q[i][0] + B[1] * Q[i][1] + B[2] * Q[i][2]);
    if (s2 > 0) return 0;
    if (s2 > s) { s = s2; normalR = R1 + i; invert_normal = (e1 < 0); code = 1 + i; }
  }
  // face axes of box 2
  for (int j = 0; j < 3; j++) {
    e1 = dot3s(R2 + j, 3, p, 1);
    s2 = fabs(e1) - (A[0] * Q[0][j] + A[1] * Q[1][j] + A[2] * Q[2][j] + B[j]);
    if (s2 > 0) return 0;
    if (s2 > s) { s = s2; normalR = R2 + j; invert_normal = (e1 < 0); code = 4 + j; }
  }
  // edge x edge axes: u_i x v_j, code 7 + 3 i + j
#define GZ_EDGE(expr1, expr2, n1, n2, n3, cc)                                   \
  e1 = (expr1);                                                                 \
  s2 = fabs(e1) - (expr2);                                                      \
  if (s2 > 0) return 0;                                                         \
  l = sqrt((n1) * (n1) + (n2) * (n2) + (n3) * (n3));                            \
  if (l > 0) {                                                                  \
    s2 /= l;                                                                    \
    if (s2 * fudge_factor > s) {                                                \
      s = s2; normalR = nullptr;                                                \
      normalC[0] = (n1) / l; normalC[1] = (n2) / l; normalC[2] = (n3) / l;      \
      invert_normal = (e1 < 0); code = (cc);                                    \
    }                                                                           \
  }
  GZ_EDGE(pp[2] * Rm[1][0] - pp[1] * Rm[2][0], (A[1] * Q[2][0] + A[2] * Q[1][0] + B[1] * Q[0][2] + B[2] * Q[0][1]), 0, -Rm[2][0], Rm[1][0], 7);
  GZ_EDGE(pp[2] * Rm[1][1] - pp[1] * Rm[2][1], (A[1] * Q[2][1] + A[2] * Q[1][1] + B[0] * Q[0][2] + B[2] * Q[0][0]), 0, -Rm[2][1], Rm[1][1], 8);
  GZ_EDGE(pp[2] * Rm[1][2] - pp[1] * Rm[2][2], (A[1] * Q[2][2] + A[2] * Q[1][2] + B[0] * Q[0][1] + B[1] * Q[0][0]), 0, -Rm[2][2], Rm[1][2], 9);
  GZ_EDGE(pp[0] * Rm[2][0] - pp[2] * Rm[0][0], (A[0] * Q[2][0] + A[2] * Q[0][0] + B[1] * Q[1][2] + B[2] * Q[1][1]), Rm[2][0], 0, -Rm[0][0], 10);
  GZ_EDGE(pp[0] * Rm[2][1] - pp[2] * Rm[0][1], (A[0] * Q[2][1] + A[2] * Q[0][1] + B[0] * Q[1][2] + B[2] * Q[1][0]), Rm[2][1], 0, -Rm[0][1], 11);
  GZ_EDGE(pp[0] * Rm[2][2] - pp[2] * Rm[0][2], (A[0] * Q[2][2] + A[2] * Q[0][2] + B[0] * Q[1][1] + B[1] * Q[1][0]), Rm[2][2], 0, -Rm[0][2], 12);
  GZ_EDGE(pp[1] * Rm[0][0] - pp[0] * Rm[1][0], (A[0] * Q[1][0] + A[1] * Q[0][0] + B[1] * Q[2][2] + B[2] * Q[2][1]), -Rm[1][0], Rm[0][0], 0, 13);
  GZ_EDGE(pp[1] * Rm[0][1] - pp[0] * Rm[1][1], (A[0] * Q[1][1] + A[1] * Q[0][1] + B[0] * Q[2][2] + B[2] * Q[2][0]), -Rm[1][1], Rm[0][1], 0, 14);
  GZ_EDGE(pp[1] * Rm[0][2] - pp[0] * Rm[1][2], (A[0] * Q[1][2] + A[1] * Q[0][2] + B[0] * Q[2][1] + B[1] * Q[2][0]), -Rm[1][2], Rm[0][2], 0, 15);
#undef GZ_EDGE
  if (!code) return 0;

  if (normalR) { normal[0] = normalR[0]; normal[1] = normalR[3]; normal[2] = normalR[6]; }
  else mul0_331(normal, R1, normalC);
  if (invert_normal) { normal[0] = -normal[0]; normal[1] = -normal[1]; normal[2] = -normal[2]; }
  const double depth = -s;
  int cnum;

  if (code > 6) {
    // edge-edge: one contact halfway between the closest points of the two edges
    double pa[3], pb[3], ua[3], ub[3];
    for (int i = 0; i < 3; i++) pa[i] = p1[i];
    for (int j = 0; j < 3; j++) {
      double sign = (dot3s(normal, 1, R1 + j, 3) > 0) ? 1.0 : -1.0;
      for (int i = 0; i < 3; i++) pa[i] += sign * A[j] * R1[3 * i + j];
    }
    for (int i = 0; i < 3; i++) pb[i] = p2[i];
    for (int j = 0; j < 3; j++) {
      double sign = (dot3s(normal, 1, R2 + j, 3) > 0) ? -1.0 : 1.0;
      for (int i = 0; i < 3; i++) pb[i] += sign * B[j] * R2[3 * i + j];
    }
    for (int i = 0; i < 3; i++) { ua[i] = R1[(code - 7) / 3 + 3 * i]; ub[i] = R2[(code - 7) % 3 + 3 * i]; }
    // dLineClosestApproach
    double d3[3] = {pb[0] - pa[0], pb[1] - pa[1], pb[2] - pa[2]};
    double uaub = dot3(ua, ub);
    double q1 = dot3(ua, d3);
    double q2 = -dot3(ub, d3);
    double d = 1 - uaub * uaub;
    double alpha, beta;
    if (d <= 0.0001) { alpha = 0; beta = 0; }
    else { d = 1.0 / d; alpha = (q1 + uaub * q2) * d; beta = (uaub * q1 + q2) * d; }
    for (int i = 0; i < 3; i++) pa[i] += ua[i] * alpha;
    for (int i = 0; i < 3; i++) pb[i] += ub[i] * beta;
    for (int i = 0; i < 3; i++) o.pos[0][i] = 0.5 * (pa[i] + pb[i]);
    o.depth[0] = depth;
    cnum = 1;
  } else {
    // face-something: a = reference box, b = incident box
    const double *Ra, *Rb, *pa, *pb, *Sa, *Sb;
    if (code <= 3) { Ra = R1; Rb = R2; pa = p1; pb = p2; Sa = A; Sb = B; }
    else           { Ra = R2; Rb = R1; pa = p2; pb = p1; Sa = B; Sb = A; }
    double normal2[3], nr[3], anr[3];
    if (code <= 3) { normal2[0] = normal[0]; normal2[1] = normal[1]; normal2[2] = normal[2]; }
    else { normal2[0] = -normal[0]; normal2[1] = -normal[1]; normal2[2] = -normal[2]; }
    mul1_331(nr, Rb, normal2);
    for (int i = 0; i < 3; i++) anr[i] = fabs(nr[i]);
    int lanr, a1, a2;
    if (anr[1] > anr[0]) {
      if (anr[1] > anr[2]) { a1 = 0; lanr = 1; a2 = 2; }
      else { a1 = 0; a2 = 1; lanr = 2; }
    } else {
      if (anr[0] > anr[2]) { lanr = 0; a1 = 1; a2 = 2; }
      else { a1 = 0; a2 = 1; lanr = 2; }
    }
    double center[3];
    if (nr[lanr] < 0) { for (int i = 0; i < 3; i++) center[i] = pb[i] - pa[i] + Sb[lanr] * Rb[3 * i + lanr]; }
    else              { for (int i = 0; i < 3; i++) center[i] = pb[i] - pa[i] - Sb[lanr] * Rb[3 * i + lanr]; }
    int codeN = (code <= 3) ? code - 1 : code - 4;
    int code1, code2;
    if (codeN == 0) { code1 = 1; code2 = 2; }
    else if (codeN == 1) { code1 = 0; code2 = 2; }
    else { code1 = 0; code2 = 1; }
    double quad[8];
    double c1 = dot3s(center, 1, Ra + code1, 3);
    double c2 = dot3s(center, 1, Ra + code2, 3);
    double m11 = dot3s(Ra + code1, 3, Rb + a1, 3);
    double m12 = dot3s(Ra + code1, 3, Rb + a2, 3);
    double m21 = dot3s(Ra + code2, 3, Rb + a1, 3);
    double m22 = dot3s(Ra + code2, 3, Rb + a2, 3);
    {
      double k1 = m11 * Sb[a1], k2 = m21 * Sb[a1], k3 = m12 * Sb[a2], k4 = m22 * Sb[a2];
      quad[0] = c1 - k1 - k3; quad[1] = c2 - k2 - k4;
      quad[2] = c1 - k1 + k3; quad[3] = c2 - k2 + k4;
      quad[4] = c1 + k1 + k3; quad[5] = c2 + k2 + k4;
      quad[6] = c1 + k1 - k3; quad[7] = c2 + k2 - k4;
    }
    double rect[2] = {Sa[code1], Sa[code2]};
    double ret[16];
    int n = rect_quad(rect, quad, ret);
    if (n < 1) return 0;
    double det1 = 1.0 / (m11 * m22 - m12 * m21);
    m11 *= det1; m12 *= det1; m21 *= det1; m22 *= det1;
    cnum = 0;
    for (int j = 0; j < n; j++) {
      double k1 = m22 * (ret[j * 2] - c1) - m12 * (ret[j * 2 + 1] - c2);
      double k2 = -m21 * (ret[j * 2] - c1) + m11 * (ret[j * 2 + 1] - c2);
      double pt[3];
      for (int i = 0; i < 3; i++) pt[i] = center[i] + k1 * Rb[3 * i + a1] + k2 * Rb[3 * i + a2];
      double dep = Sa[codeN] - dot3(normal2, pt);
      if (dep >= 0) {
        for (int i = 0; i < 3; i++) o.pos[cnum][i] = pt[i] + pa[i];
        o.depth[cnum] = dep;
        cnum++;
      }
    }
    if (cnum < 1) return 0;
  }
  // dCollideBoxBox negates the normal (box.cpp:731-736)
  for (int c = 0; c < cnum; c++) { o.nrm[c][0] = -normal[0]; o.nrm[c][1] = -normal[1]; o.nrm[c][2] = -normal[2]; }
  return cnum;
}

}  // namespace gz
