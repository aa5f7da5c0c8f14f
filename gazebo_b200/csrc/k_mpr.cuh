// k_mpr.cuh — cylinder-cylinder and capsule-cylinder narrowphase: the reference sends
// these pairs to libccd's Minkowski Portal Refinement (dCollideCylinderCylinder / dCollideCapsuleCylinder,
// deps/opende/src/collision_libccd.cpp:305-344 -> ccdCollide :250-303 -> ccdMPRPenetration,
// deps/libccd/src/mpr.c:111-146), one contact per pair.  Restated here, same operation order:
//   geometry      ccdGeomToObj / ToCyl / ToCap (collision_libccd.cpp:72-114): position, the quaternion of
//                 dGeomGetQuaternion (collision_kernel.cpp:702-717: dQfromR of the final rotation for a
//                 geom with an offset, rotation.cpp:261-311) and its inverse (quat.h ccdQuatInvert)
//   supports      ccdSupportCyl / ccdSupportCap (collision_libccd.cpp:151-199), __ccdSupport (support.c:20-34)
//   portal        discoverPortal, refinePortal, findPenetr, findPos, findPenetrTouch / Segment, expandPortal,
//                 portalDir, portalReachTolerance ... (mpr.c:150-548)
//   distances     ccdVec3PointTriDist2 / PointSegmentDist2 (vec3.c:86-212); CCD_EPS = DBL_EPSILON (vec3.h:62)
// ccd.max_iterations = 500, ccd.mpr_tolerance = 1e-6 (collision_libccd.cpp:269-270).  The two portal loops
// the reference leaves unbounded are capped at GZ_MPR_LOOP_CAP trips here (a GPU thread must terminate);
// a pair that hits the cap reports no contact and raises the step's capacity flag (bit 32, sticky: the next
// synchronising call returns GZCUDA_E_CAPACITY), so the divergence from the reference is never silent.
#pragma once

namespace gz {

#define GZ_CCD_EPS 2.2204460492503131e-16
#define GZ_MPR_LOOP_CAP 4096

struct CcdObj { double pos[3], rot[4], rot_inv[4], radius, height; int cap; };   // rot = (x, y, z, w)
struct CcdSupp { double v[3], v1[3], v2[3]; };

__device__ inline bool ccd_zero(double v) { return fabs(v) < GZ_CCD_EPS; }
__device__ inline bool ccd_eq(double _a, double _b) {
  const double ab = fabs(_a - _b);
  if (fabs(ab) < GZ_CCD_EPS) return true;
  const double a = fabs(_a), b = fabs(_b);
  return (b > a) ? (ab < GZ_CCD_EPS * b) : (ab < GZ_CCD_EPS * a);
}
__device__ inline int ccd_sign(double v) { return ccd_zero(v) ? 0 : (v < 0.0 ? -1 : 1); }
__device__ inline double ccd_dot(const double *a, const double *b) {
  double d = a[0] * b[0];
  d += a[1] * b[1];
  d += a[2] * b[2];
  return d;
}
__device__ inline void ccd_cross(double *d, const double *a, const double *b) {
  const double x = (a[1] * b[2]) - (a[2] * b[1]);
  const double y = (a[2] * b[0]) - (a[0] * b[2]);
  const double z = (a[0] * b[1]) - (a[1] * b[0]);
  d[0] = x; d[1] = y; d[2] = z;
}
__device__ inline void ccd_sub2(double *d, const double *v, const double *w) { d[0] = v[0] - w[0]; d[1] = v[1] - w[1]; d[2] = v[2] - w[2]; }
__device__ inline void ccd_add(double *v, const double *w) { v[0] += w[0]; v[1] += w[1]; v[2] += w[2]; }
__device__ inline void ccd_scale(double *d, double k) { d[0] *= k; d[1] *= k; d[2] *= k; }
__device__ inline void ccd_copy(double *d, const double *s) { d[0] = s[0]; d[1] = s[1]; d[2] = s[2]; }
__device__ inline void ccd_normalize(double *d) { const double k = 1.0 / sqrt(ccd_dot(d, d)); ccd_scale(d, k); }
__device__ inline bool ccd_vec_eq0(const double *a) { return ccd_eq(a[0], 0.0) && ccd_eq(a[1], 0.0) && ccd_eq(a[2], 0.0); }
__device__ inline double ccd_dist2(const double *a, const double *b) { double ab[3]; ccd_sub2(ab, a, b); return ccd_dot(ab, ab); }
__device__ inline void ccd_supp_copy(CcdSupp &d, const CcdSupp &s) { ccd_copy(d.v, s.v); ccd_copy(d.v1, s.v1); ccd_copy(d.v2, s.v2); }

// ccdQuatRotVec (quat.h)
__device__ inline void ccd_quat_rot(double *v, const double *q) {
  const double w = q[3], x = q[0], y = q[1], z = q[2];
  const double ww = w * w, xx = x * x, yy = y * y, zz = z * z;
  const double wx = w * x, wy = w * y, wz = w * z, xy = x * y, xz = x * z, yz = y * z;
  const double vx = ww * v[0] + xx * v[0] - yy * v[0] - zz * v[0] + 2 * ((xy - wz) * v[1] + (xz + wy) * v[2]);
  const double vy = ww * v[1] - xx * v[1] + yy * v[1] - zz * v[1] + 2 * ((xy + wz) * v[0] + (yz - wx) * v[2]);
  const double vz = ww * v[2] - xx * v[2] - yy * v[2] + zz * v[2] + 2 * ((xz - wy) * v[0] + (yz + wx) * v[1]);
  v[0] = vx; v[1] = vy; v[2] = vz;
}

// dQfromR (rotation.cpp:261-311) on a 3 x 3 row-major matrix; q = (w, x, y, z)
__device__ inline void q_from_R(double *q, const double *R) {
  const double tr = R[0] + R[4] + R[8];
  double s;
  if (tr >= 0) {
    s = sqrt(tr + 1);
    q[0] = 0.5 * s;
    s = 0.5 * (1.0 / s);
    q[1] = (R[7] - R[5]) * s;
    q[2] = (R[2] - R[6]) * s;
    q[3] = (R[3] - R[1]) * s;
    return;
  }
  int c;
  if (R[4] > R[0]) c = (R[8] > R[4]) ? 2 : 1;
  else c = (R[8] > R[0]) ? 2 : 0;
  if (c == 0) {
    s = sqrt((R[0] - (R[4] + R[8])) + 1);
    q[1] = 0.5 * s;
    s = 0.5 * (1.0 / s);
    q[2] = (R[1] + R[3]) * s;
    q[3] = (R[6] + R[2]) * s;
    q[0] = (R[7] - R[5]) * s;
  } else if (c == 1) {
    s = sqrt((R[4] - (R[8] + R[0])) + 1);
    q[2] = 0.5 * s;
    s = 0.5 * (1.0 / s);
    q[3] = (R[5] + R[7]) * s;
    q[1] = (R[1] + R[3]) * s;
    q[0] = (R[2] - R[6]) * s;
  } else {
    s = sqrt((R[8] - (R[0] + R[4])) + 1);
    q[3] = 0.5 * s;
    s = 0.5 * (1.0 / s);
    q[1] = (R[6] + R[2]) * s;
    q[2] = (R[5] + R[7]) * s;
    q[0] = (R[3] - R[1]) * s;
  }
}

__device__ inline void ccd_obj(CcdObj &o, const Geom &g, bool cap) {
  double q[4];
  q_from_R(q, g.R);
  ccd_copy(o.pos, g.pos);
  o.rot[0] = q[1]; o.rot[1] = q[2]; o.rot[2] = q[3]; o.rot[3] = q[0];
  for (int k = 0; k < 4; k++) o.rot_inv[k] = o.rot[k];
  double len2 = o.rot[0] * o.rot[0];
  len2 += o.rot[1] * o.rot[1];
  len2 += o.rot[2] * o.rot[2];
  len2 += o.rot[3] * o.rot[3];
  if (!(len2 < GZ_CCD_EPS)) {
    len2 = 1.0 / len2;
    o.rot_inv[0] = -o.rot_inv[0] * len2;
    o.rot_inv[1] = -o.rot_inv[1] * len2;
    o.rot_inv[2] = -o.rot_inv[2] * len2;
    o.rot_inv[3] = o.rot_inv[3] * len2;
  }
  o.radius = g.dims[0];
  o.height = g.dims[1] / 2.;
  o.cap = cap ? 1 : 0;
}

__device__ inline void ccd_support1(const CcdObj &o, const double *_dir, double *v) {
  double dir[3];
  ccd_copy(dir, _dir);
  ccd_quat_rot(dir, o.rot_inv);
  if (o.cap) {   // ccdSupportCap
    double p1[3] = {0.0, 0.0, o.height}, p2[3] = {0.0, 0.0, -o.height};
    ccd_copy(v, dir);
    ccd_scale(v, o.radius);
    ccd_add(p1, v);
    ccd_add(p2, v);
    if (ccd_dot(dir, p1) > ccd_dot(dir, p2)) ccd_copy(v, p1); else ccd_copy(v, p2);
  } else {       // ccdSupportCyl
    double zdist = dir[0] * dir[0] + dir[1] * dir[1];
    zdist = sqrt(zdist);
    if (ccd_zero(zdist)) {
      v[0] = 0.; v[1] = 0.; v[2] = ccd_sign(dir[2]) * o.height;
    } else {
      const double rad = o.radius / zdist;
      v[0] = rad * dir[0]; v[1] = rad * dir[1]; v[2] = ccd_sign(dir[2]) * o.height;
    }
  }
  ccd_quat_rot(v, o.rot);
  ccd_add(v, o.pos);
}

// __ccdSupport (support.c:20-34)
__device__ inline void ccd_support(const CcdObj &o1, const CcdObj &o2, const double *_dir, CcdSupp &s) {
  double dir[3];
  ccd_copy(dir, _dir);
  ccd_support1(o1, dir, s.v1);
  ccd_scale(dir, -1.0);
  ccd_support1(o2, dir, s.v2);
  ccd_sub2(s.v, s.v1, s.v2);
}

__device__ inline void mpr_portal_dir(const CcdSupp *P, double *dir) {
  double v2v1[3], v3v1[3];
  ccd_sub2(v2v1, P[2].v, P[1].v);
  ccd_sub2(v3v1, P[3].v, P[1].v);
  ccd_cross(dir, v2v1, v3v1);
  ccd_normalize(dir);
}

__device__ inline bool mpr_reach_tolerance(const CcdSupp *P, const CcdSupp &v4, const double *dir) {
  const double dv1 = ccd_dot(P[1].v, dir), dv2 = ccd_dot(P[2].v, dir), dv3 = ccd_dot(P[3].v, dir), dv4 = ccd_dot(v4.v, dir);
  double dot1 = dv4 - dv1;
  const double dot2 = dv4 - dv2, dot3 = dv4 - dv3;
  dot1 = fmin(dot1, dot2);
  dot1 = fmin(dot1, dot3);
  return ccd_eq(dot1, 1e-6) || dot1 < 1e-6;
}

__device__ inline void mpr_expand_portal(CcdSupp *P, const CcdSupp &v4) {
  double v4v0[3];
  ccd_cross(v4v0, v4.v, P[0].v);
  double dot = ccd_dot(P[1].v, v4v0);
  if (dot > 0.0) {
    dot = ccd_dot(P[2].v, v4v0);
    if (dot > 0.0) ccd_supp_copy(P[1], v4); else ccd_supp_copy(P[3], v4);
  } else {
    dot = ccd_dot(P[3].v, v4v0);
    if (dot > 0.0) ccd_supp_copy(P[2], v4); else ccd_supp_copy(P[1], v4);
  }
}

// discoverPortal (mpr.c:158-283): -1 no intersection, 0 portal built, 1 origin on v1, 2 origin on the segment v0-v1
__device__ inline int mpr_discover_portal(const CcdObj &o1, const CcdObj &o2, CcdSupp *P) {
  double dir[3], va[3], vb[3];
  ccd_copy(P[0].v1, o1.pos);
  ccd_copy(P[0].v2, o2.pos);
  ccd_sub2(P[0].v, P[0].v1, P[0].v2);
  if (ccd_vec_eq0(P[0].v)) { va[0] = GZ_CCD_EPS * 10.; va[1] = 0.0; va[2] = 0.0; ccd_add(P[0].v, va); }
  ccd_copy(dir, P[0].v);
  ccd_scale(dir, -1.);
  ccd_normalize(dir);
  ccd_support(o1, o2, dir, P[1]);
  double dot = ccd_dot(P[1].v, dir);
  if (ccd_zero(dot) || dot < 0.0) return -1;
  ccd_cross(dir, P[0].v, P[1].v);
  if (ccd_zero(ccd_dot(dir, dir))) return ccd_vec_eq0(P[1].v) ? 1 : 2;
  ccd_normalize(dir);
  ccd_support(o1, o2, dir, P[2]);
  dot = ccd_dot(P[2].v, dir);
  if (ccd_zero(dot) || dot < 0.0) return -1;
  ccd_sub2(va, P[1].v, P[0].v);
  ccd_sub2(vb, P[2].v, P[0].v);
  ccd_cross(dir, va, vb);
  ccd_normalize(dir);
  dot = ccd_dot(dir, P[0].v);
  if (dot > 0.0) {
    CcdSupp t;
    ccd_supp_copy(t, P[1]); ccd_supp_copy(P[1], P[2]); ccd_supp_copy(P[2], t);
    ccd_scale(dir, -1.);
  }
  for (int trip = 0; trip < GZ_MPR_LOOP_CAP; trip++) {
    ccd_support(o1, o2, dir, P[3]);
    dot = ccd_dot(P[3].v, dir);
    if (ccd_zero(dot) || dot < 0.0) return -1;
    bool cont = false;
    ccd_cross(va, P[1].v, P[3].v);
    dot = ccd_dot(va, P[0].v);
    if (dot < 0.0 && !ccd_zero(dot)) { ccd_supp_copy(P[2], P[3]); cont = true; }
    if (!cont) {
      ccd_cross(va, P[3].v, P[2].v);
      dot = ccd_dot(va, P[0].v);
      if (dot < 0.0 && !ccd_zero(dot)) { ccd_supp_copy(P[1], P[3]); cont = true; }
    }
    if (!cont) return 0;
    ccd_sub2(va, P[1].v, P[0].v);
    ccd_sub2(vb, P[2].v, P[0].v);
    ccd_cross(dir, va, vb);
    ccd_normalize(dir);
  }
  return -2;   // GZ_MPR_LOOP_CAP trips: the reference would go on (reported, see col_mpr)
}

// refinePortal (mpr.c:285-317)
__device__ inline int mpr_refine_portal(const CcdObj &o1, const CcdObj &o2, CcdSupp *P) {
  double dir[3];
  CcdSupp v4;
  for (int trip = 0; trip < GZ_MPR_LOOP_CAP; trip++) {
    mpr_portal_dir(P, dir);
    double dot = ccd_dot(dir, P[1].v);
    if (ccd_zero(dot) || dot > 0.0) return 0;   // portalEncapsulesOrigin
    ccd_support(o1, o2, dir, v4);
    dot = ccd_dot(v4.v, dir);
    if (!(ccd_zero(dot) || dot > 0.0) || mpr_reach_tolerance(P, v4, dir)) return -1;
    mpr_expand_portal(P, v4);
  }
  return -2;   // GZ_MPR_LOOP_CAP trips (reported, see col_mpr)
}

// __ccdVec3PointSegmentDist2 with P = origin and a witness (vec3.c:86-140)
__device__ inline double ccd_origin_seg_dist2(const double *x0, const double *b, double *witness) {
  const double P[3] = {0.0, 0.0, 0.0};
  double d[3], a[3], dist;
  ccd_sub2(d, b, x0);
  ccd_sub2(a, x0, P);
  double t = -1. * ccd_dot(a, d);
  t /= ccd_dot(d, d);
  if (t < 0.0 || ccd_zero(t)) {
    dist = ccd_dist2(x0, P);
    ccd_copy(witness, x0);
  } else if (t > 1.0 || ccd_eq(t, 1.0)) {
    dist = ccd_dist2(b, P);
    ccd_copy(witness, b);
  } else {
    ccd_copy(witness, d);
    ccd_scale(witness, t);
    ccd_add(witness, x0);
    dist = ccd_dist2(witness, P);
  }
  return dist;
}

// ccdVec3PointTriDist2 with P = origin and a witness (vec3.c:150-212)
__device__ inline double ccd_origin_tri_dist2(const double *x0, const double *B, const double *C, double *witness) {
  const double P[3] = {0.0, 0.0, 0.0};
  double d1[3], d2[3], a[3];
  ccd_sub2(d1, B, x0);
  ccd_sub2(d2, C, x0);
  ccd_sub2(a, x0, P);
  const double v = ccd_dot(d1, d1), w = ccd_dot(d2, d2), p = ccd_dot(a, d1), q = ccd_dot(a, d2), r = ccd_dot(d1, d2);
  const double s = (q * r - w * p) / (w * v - r * r);
  const double t = (-s * r - q) / w;
  double dist;
  if ((ccd_zero(s) || s > 0.0) && (ccd_eq(s, 1.0) || s < 1.0) && (ccd_zero(t) || t > 0.0) && (ccd_eq(t, 1.0) || t < 1.0) &&
      (ccd_eq(t + s, 1.0) || t + s < 1.0)) {
    ccd_scale(d1, s);
    ccd_scale(d2, t);
    ccd_copy(witness, x0);
    ccd_add(witness, d1);
    ccd_add(witness, d2);
    dist = ccd_dist2(witness, P);
  } else {
    double w2[3];
    dist = ccd_origin_seg_dist2(x0, B, witness);
    double dist2 = ccd_origin_seg_dist2(x0, C, w2);
    if (dist2 < dist) { dist = dist2; ccd_copy(witness, w2); }
    dist2 = ccd_origin_seg_dist2(B, C, w2);
    if (dist2 < dist) { dist = dist2; ccd_copy(witness, w2); }
  }
  return dist;
}

// findPos (mpr.c:410-476)
__device__ inline void mpr_find_pos(const CcdSupp *P, double *pos) {
  double dir[3], vec[3], b[4], sum;
  mpr_portal_dir(P, dir);
  ccd_cross(vec, P[1].v, P[2].v); b[0] = ccd_dot(vec, P[3].v);
  ccd_cross(vec, P[3].v, P[2].v); b[1] = ccd_dot(vec, P[0].v);
  ccd_cross(vec, P[0].v, P[1].v); b[2] = ccd_dot(vec, P[3].v);
  ccd_cross(vec, P[2].v, P[1].v); b[3] = ccd_dot(vec, P[0].v);
  sum = b[0] + b[1] + b[2] + b[3];
  if (ccd_zero(sum) || sum < 0.0) {
    b[0] = 0.;
    ccd_cross(vec, P[2].v, P[3].v); b[1] = ccd_dot(vec, dir);
    ccd_cross(vec, P[3].v, P[1].v); b[2] = ccd_dot(vec, dir);
    ccd_cross(vec, P[1].v, P[2].v); b[3] = ccd_dot(vec, dir);
    sum = b[1] + b[2] + b[3];
  }
  const double inv = 1. / sum;
  double p1[3] = {0.0, 0.0, 0.0}, p2[3] = {0.0, 0.0, 0.0};
  for (int i = 0; i < 4; i++) {
    ccd_copy(vec, P[i].v1); ccd_scale(vec, b[i]); ccd_add(p1, vec);
    ccd_copy(vec, P[i].v2); ccd_scale(vec, b[i]); ccd_add(p2, vec);
  }
  ccd_scale(p1, inv);
  ccd_scale(p2, inv);
  ccd_copy(pos, p1);
  ccd_add(pos, p2);
  ccd_scale(pos, 0.5);
}

// ccdCollide + ccdMPRPenetration: 0 or 1 contact; normal = -dir
__device__ __noinline__ int col_mpr(const Geom &g1, bool cap1, const Geom &g2, COut &o) {
  CcdObj o1, o2;
  ccd_obj(o1, g1, cap1);
  ccd_obj(o2, g2, false);
  CcdSupp P[4];
  double depth, dir[3], pos[3];
  const int res = mpr_discover_portal(o1, o2, P);
  if (res == -2) return -1;   // a portal loop hit its trip cap: no contact AND the step is flagged (k_narrow)
  if (res < 0) return 0;
  if (res == 1) {          // findPenetrTouch
    depth = 0.;
    dir[0] = dir[1] = dir[2] = 0.0;
    ccd_copy(pos, P[1].v1); ccd_add(pos, P[1].v2); ccd_scale(pos, 0.5);
  } else if (res == 2) {   // findPenetrSegment
    ccd_copy(pos, P[1].v1); ccd_add(pos, P[1].v2); ccd_scale(pos, 0.5);
    ccd_copy(dir, P[1].v);
    depth = sqrt(ccd_dot(dir, dir));
    ccd_normalize(dir);
  } else {
    const int rr = mpr_refine_portal(o1, o2, P);
    if (rr == -2) return -1;
    if (rr < 0) return 0;
    // findPenetr (mpr.c:320-355)
    CcdSupp v4;
    double pd[3];
    unsigned long iterations = 0UL;
    while (true) {
      mpr_portal_dir(P, pd);
      ccd_support(o1, o2, pd, v4);
      if (mpr_reach_tolerance(P, v4, pd) || iterations > 500UL) {
        depth = ccd_origin_tri_dist2(P[1].v, P[2].v, P[3].v, dir);
        depth = sqrt(depth);
        ccd_normalize(dir);
        mpr_find_pos(P, pos);
        break;
      }
      mpr_expand_portal(P, v4);
      iterations++;
    }
  }
  ccd_scale(dir, -1.);
  for (int k = 0; k < 3; k++) { o.pos[0][k] = pos[k]; o.nrm[0][k] = dir[k]; }
  o.depth[0] = depth;
  return 1;
}

}  // namespace gz
