// gz_math.cuh — small fixed-size double-precision helpers used by every kernel.
//
// The operation ORDER of each helper follows the reference's inline maths
// (deps/opende/include/gazebo/ode/odemath.h:60-420, src/odemath.cpp:36-187,
// src/rotation.cpp:194-321) so that, compiled with -fmad=false, the device
// produces the same IEEE-754 double results as the CPU build
// (-O2 -ffp-contract=off).  Matrices are 3x3 row-major (the reference pads rows
// to 4; the padding never takes part in arithmetic).
#pragma once
#include <math.h>

#define GZ_HD __host__ __device__ __forceinline__
#define GZ_INF (__longlong_as_double(0x7ff0000000000000LL))

namespace gz {

GZ_HD double dot3(const double *a, const double *b) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}
// a with stride sa, b with stride sb (dCalcVectorDot3_41 and friends)
GZ_HD double dot3s(const double *a, int sa, const double *b, int sb) {
  return a[0] * b[0] + a[sa] * b[sb] + a[2 * sa] * b[2 * sb];
}
GZ_HD void cross3(double *r, const double *a, const double *b) {
  double r0 = a[1] * b[2] - a[2] * b[1];
  double r1 = a[2] * b[0] - a[0] * b[2];
  double r2 = a[0] * b[1] - a[1] * b[0];
  r[0] = r0; r[1] = r1; r[2] = r2;
}
// r = M * v   (dMultiply0_331)
GZ_HD void mul0_331(double *r, const double *M, const double *v) {
  double r0 = dot3(M, v), r1 = dot3(M + 3, v), r2 = dot3(M + 6, v);
  r[0] = r0; r[1] = r1; r[2] = r2;
}
// r = M^T * v (dMultiply1_331)
GZ_HD void mul1_331(double *r, const double *M, const double *v) {
  double r0 = dot3s(M, 3, v, 1), r1 = dot3s(M + 1, 3, v, 1), r2 = dot3s(M + 2, 3, v, 1);
  r[0] = r0; r[1] = r1; r[2] = r2;
}
// R = A * B (dMultiply0_333): R(i,j) = sum_k B(k,j) * A(i,k) with the product written
// as dCalcVectorDot3_41(B + j, A_row_i)  -> B first, then A.
GZ_HD void mul0_333(double *R, const double *A, const double *B) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      R[3 * i + j] = B[j] * A[3 * i] + B[3 + j] * A[3 * i + 1] + B[6 + j] * A[3 * i + 2];
}
// R = A * B^T (dMultiply2_333): R(i,j) = dot(B_row_j, A_row_i)
GZ_HD void mul2_333(double *R, const double *A, const double *B) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      R[3 * i + j] = B[3 * j] * A[3 * i] + B[3 * j + 1] * A[3 * i + 1] + B[3 * j + 2] * A[3 * i + 2];
}
GZ_HD double recip_sqrt(double x) { return 1.0 / sqrt(x); }   // dRecipSqrt (common.h:186-188)

// _dSafeNormalize3 (odemath.cpp:43-85)
GZ_HD int safe_normalize3(double *a) {
  double aa0 = fabs(a[0]), aa1 = fabs(a[1]), aa2 = fabs(a[2]);
  double m;
  if (aa1 > aa0) {
    m = (aa2 > aa1) ? aa2 : aa1;
  } else {
    if (aa2 > aa0) m = aa2;
    else {
      if (aa0 <= 0) { a[0] = 1; a[1] = 0; a[2] = 0; return 0; }
      m = aa0;
    }
  }
  a[0] /= m; a[1] /= m; a[2] /= m;
  double l = recip_sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
  a[0] *= l; a[1] *= l; a[2] *= l;
  return 1;
}
// _dSafeNormalize4 (odemath.cpp:118-138)
GZ_HD void normalize4(double *a) {
  double l = dot3(a, a) + a[3] * a[3];
  if (l > 0) {
    l = recip_sqrt(l);
    a[0] *= l; a[1] *= l; a[2] *= l; a[3] *= l;
  } else { a[0] = 1; a[1] = 0; a[2] = 0; a[3] = 0; }
}
// dPlaneSpace (odemath.cpp:153-181)
GZ_HD void plane_space(const double *n, double *p, double *q) {
  if (fabs(n[2]) > 0.70710678118654752440) {
    double a = n[1] * n[1] + n[2] * n[2];
    double k = recip_sqrt(a);
    p[0] = 0; p[1] = -n[2] * k; p[2] = n[1] * k;
    q[0] = a * k; q[1] = -n[0] * p[2]; q[2] = n[0] * p[1];
  } else {
    double a = n[0] * n[0] + n[1] * n[1];
    double k = recip_sqrt(a);
    p[0] = -n[1] * k; p[1] = n[0] * k; p[2] = 0;
    q[0] = -n[2] * p[1]; q[1] = n[2] * p[0]; q[2] = a * k;
  }
}
// dRfromQ (rotation.cpp:239-260)
GZ_HD void q_to_R(const double *q, double *R) {
  double qq1 = 2 * q[1] * q[1];
  double qq2 = 2 * q[2] * q[2];
  double qq3 = 2 * q[3] * q[3];
  R[0] = 1 - qq2 - qq3;
  R[1] = 2 * (q[1] * q[2] - q[0] * q[3]);
  R[2] = 2 * (q[1] * q[3] + q[0] * q[2]);
  R[3] = 2 * (q[1] * q[2] + q[0] * q[3]);
  R[4] = 1 - qq1 - qq3;
  R[5] = 2 * (q[2] * q[3] - q[0] * q[1]);
  R[6] = 2 * (q[1] * q[3] - q[0] * q[2]);
  R[7] = 2 * (q[2] * q[3] + q[0] * q[1]);
  R[8] = 1 - qq1 - qq2;
}
// dQMultiply0..3 (rotation.cpp:194-231)
GZ_HD void qmul0(double *qa, const double *qb, const double *qc) {
  qa[0] = qb[0] * qc[0] - qb[1] * qc[1] - qb[2] * qc[2] - qb[3] * qc[3];
  qa[1] = qb[0] * qc[1] + qb[1] * qc[0] + qb[2] * qc[3] - qb[3] * qc[2];
  qa[2] = qb[0] * qc[2] + qb[2] * qc[0] + qb[3] * qc[1] - qb[1] * qc[3];
  qa[3] = qb[0] * qc[3] + qb[3] * qc[0] + qb[1] * qc[2] - qb[2] * qc[1];
}
GZ_HD void qmul1(double *qa, const double *qb, const double *qc) {
  qa[0] = qb[0] * qc[0] + qb[1] * qc[1] + qb[2] * qc[2] + qb[3] * qc[3];
  qa[1] = qb[0] * qc[1] - qb[1] * qc[0] - qb[2] * qc[3] + qb[3] * qc[2];
  qa[2] = qb[0] * qc[2] - qb[2] * qc[0] - qb[3] * qc[1] + qb[1] * qc[3];
  qa[3] = qb[0] * qc[3] - qb[3] * qc[0] - qb[1] * qc[2] + qb[2] * qc[1];
}
GZ_HD void qmul2(double *qa, const double *qb, const double *qc) {
  qa[0] = qb[0] * qc[0] + qb[1] * qc[1] + qb[2] * qc[2] + qb[3] * qc[3];
  qa[1] = -qb[0] * qc[1] + qb[1] * qc[0] - qb[2] * qc[3] + qb[3] * qc[2];
  qa[2] = -qb[0] * qc[2] + qb[2] * qc[0] - qb[3] * qc[1] + qb[1] * qc[3];
  qa[3] = -qb[0] * qc[3] + qb[3] * qc[0] - qb[1] * qc[2] + qb[2] * qc[1];
}
GZ_HD void qmul3(double *qa, const double *qb, const double *qc) {
  qa[0] = qb[0] * qc[0] - qb[1] * qc[1] - qb[2] * qc[2] - qb[3] * qc[3];
  qa[1] = -qb[0] * qc[1] - qb[1] * qc[0] + qb[2] * qc[3] - qb[3] * qc[2];
  qa[2] = -qb[0] * qc[2] - qb[2] * qc[0] + qb[3] * qc[1] - qb[1] * qc[3];
  qa[3] = -qb[0] * qc[3] - qb[3] * qc[0] + qb[1] * qc[2] - qb[2] * qc[1];
}
// dDQfromW (rotation.cpp:314-321)
GZ_HD void dq_from_w(double *dq, const double *w, const double *q) {
  dq[0] = 0.5 * (-w[0] * q[1] - w[1] * q[2] - w[2] * q[3]);
  dq[1] = 0.5 * (w[0] * q[0] + w[1] * q[3] - w[2] * q[2]);
  dq[2] = 0.5 * (-w[0] * q[3] + w[1] * q[0] + w[2] * q[1]);
  dq[3] = 0.5 * (w[0] * q[2] - w[1] * q[1] + w[2] * q[0]);
}
// _dequal (odemath.h:485-489): |a-b| <= 1e-6
GZ_HD bool dequal(double a, double b) { return fabs(a - b) <= 1e-6; }

// dNormalizeAnglePositive / dNormalizeAngle / dShortestAngularDistance(Update)
// (include/gazebo/ode/misc.h:89-150)
GZ_HD double norm_angle_pos(double a) {
  const double two_pi = 2.0 * 3.14159265358979323846;
  return fmod(fmod(a, two_pi) + two_pi, two_pi);
}
GZ_HD double norm_angle(double a) {
  double r = norm_angle_pos(a);
  if (r > 3.14159265358979323846) r -= 2.0 * 3.14159265358979323846;
  return r;
}
GZ_HD double shortest_angular_distance(double from, double to) {
  return norm_angle(norm_angle_pos(norm_angle_pos(to) - norm_angle_pos(from)));
}
GZ_HD double shortest_angular_distance_update(double from, double to) {
  double result = norm_angle(norm_angle_pos(norm_angle_pos(to) - norm_angle_pos(from)));
  if (fabs(result) > 0.3) return to;
  return from + result;
}

}  // namespace gz
