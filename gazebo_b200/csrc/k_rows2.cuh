// k_rows2.cuh — getInfo2 for persistent joints and contacts, then the row prologue.
// See the header of k_rows.cuh for the reference line map.
#pragma once
#include "k_rows.cuh"

namespace gz {

__device__ inline void row_init(Row &r, int lb1, int lb2, double cfm, double cvmax, int src) {
  for (int k = 0; k < 12; k++) r.J[k] = 0;
  r.c = 0; r.cfm = cfm; r.lo = -GZ_INF; r.hi = GZ_INF; r.findex = -1; r.c_v_max = cvmax;
  r.b1 = lb1; r.b2 = lb2; r.src = src;
}

// dxJointLimitMotor::addLimot (joint.cpp:573-750), rotational (hinge) or linear (slider, with the
// linear torque decoupling vector ltd when the joint has two bodies).  Returns 1 if a row was written.
// dxJointLimitMotor as plain data (joints/joint.cpp:470-484)
struct Limot { double lostop, histop, vel, fmax, fudge, bounce, normal_cfm, stop_erp, stop_cfm, limit_err; int limit; };
__device__ inline Limot limot_of(const Joint &j) {   // the joint's (first) limit motor: joint_param slots 0..8
  Limot L;
  L.lostop = j.param[GZ_JP_LOSTOP]; L.histop = j.param[GZ_JP_HISTOP]; L.vel = j.param[GZ_JP_VEL]; L.fmax = j.param[GZ_JP_FMAX];
  L.fudge = j.param[GZ_JP_FUDGE]; L.bounce = j.param[GZ_JP_BOUNCE]; L.normal_cfm = j.param[GZ_JP_NORMAL_CFM];
  L.stop_erp = j.param[GZ_JP_STOP_ERP]; L.stop_cfm = j.param[GZ_JP_STOP_CFM]; L.limit_err = j.limit_err; L.limit = j.limit;
  return L;
}

__device__ inline int add_limot(DevBatch &B, Joint &j, Row &r, const double *ax1, double fps, bool rotational, const Limot &M) {
  const double fmax = M.fmax, vel = M.vel;
  const double lostop = M.lostop, histop = M.histop;
  int powered = fmax > 0;
  const int limit = M.limit;
  if (!(powered || limit)) return 0;
  const int o = rotational ? 3 : 0;
  for (int k = 0; k < 3; k++) r.J[o + k] = ax1[k];
  if (j.b2 >= 0) for (int k = 0; k < 3; k++) r.J[6 + o + k] = -ax1[k];
  double ltd[3] = {0, 0, 0};
  if (!rotational && j.b2 >= 0) {
    double c[3];
    for (int k = 0; k < 3; k++) c[k] = 0.5 * (B.body[j.b2].pos[k] - B.body[j.b1].pos[k]);
    cross3(ltd, c, ax1);
    for (int k = 0; k < 3; k++) { r.J[3 + k] = ltd[k]; r.J[9 + k] = ltd[k]; }
  }
  if (limit && dequal(lostop, histop)) powered = 0;
  if (powered) {
    r.cfm = M.normal_cfm;
    if (!limit) {
      r.c = vel; r.lo = -fmax; r.hi = fmax;
    } else {
      double fm = fmax;
      if ((vel > 0) || (dequal(vel, 0.0) && limit == 2)) fm = -fm;
      if ((limit == 1 && vel > 0) || (limit == 2 && vel < 0)) fm *= M.fudge;
      // dBodyAddTorque on both bodies from inside getInfo2.  Several powered hinges at
      // their stops on one body would make the summation order matter; atomics keep it
      // race-free (order then differs from the reference in the last bit only).
      Body &b1 = B.body[j.b1];
      double *acc1 = rotational ? b1.tacc : b1.facc;
      atomicAdd(&acc1[0], -fm * ax1[0]); atomicAdd(&acc1[1], -fm * ax1[1]); atomicAdd(&acc1[2], -fm * ax1[2]);
      if (j.b2 >= 0) {
        Body &b2 = B.body[j.b2];
        double *acc2 = rotational ? b2.tacc : b2.facc;
        atomicAdd(&acc2[0], fm * ax1[0]); atomicAdd(&acc2[1], fm * ax1[1]); atomicAdd(&acc2[2], fm * ax1[2]);
        if (!rotational)
          for (int k = 0; k < 3; k++) { atomicAdd(&b1.tacc[k], -fm * ltd[k]); atomicAdd(&b2.tacc[k], -fm * ltd[k]); }
      }
    }
  }
  if (limit) {
    double k = fps * M.stop_erp;
    r.c = -k * M.limit_err;
    r.cfm = M.stop_cfm;
    if (dequal(lostop, histop)) { r.lo = -GZ_INF; r.hi = GZ_INF; }
    else {
      if (limit == 1) { r.lo = 0; r.hi = GZ_INF; } else { r.lo = -GZ_INF; r.hi = 0; }
      const double bounce = M.bounce;
      if (bounce > 0) {
        double vvel = dot3(rotational ? B.body[j.b1].avel : B.body[j.b1].lvel, ax1);
        if (j.b2 >= 0) vvel -= dot3(rotational ? B.body[j.b2].avel : B.body[j.b2].lvel, ax1);
        if (limit == 1) {
          if (vel < 0) { double newc = -bounce * vvel; if (newc > r.c) r.c = newc; }
        } else {
          if (vvel > 0) { double newc = -bounce * vvel; if (newc < r.c) r.c = newc; }
        }
      }
    }
  }
  return 1;
}

__device__ inline int add_limot(DevBatch &B, Joint &j, Row &r, const double *ax1, double fps, bool rotational) {
  return add_limot(B, j, r, ax1, fps, rotational, limot_of(j));
}

// setBall (joint.cpp:93-134): rows 0..2 keep the anchor points together; k = fps * erp
__device__ inline void set_ball(Row *rows, const Body &b1, const Body *b2, const Joint &j, double k) {
  double a1[3], a2[3];
  rows[0].J[0] = 1; rows[1].J[1] = 1; rows[2].J[2] = 1;
  mul0_331(a1, b1.R, j.anchor1);
  // dSetCrossMatrixMinus(J1a, a1)
  rows[0].J[3 + 1] = a1[2];  rows[0].J[3 + 2] = -a1[1];
  rows[1].J[3 + 0] = -a1[2]; rows[1].J[3 + 2] = a1[0];
  rows[2].J[3 + 0] = a1[1];  rows[2].J[3 + 1] = -a1[0];
  if (b2) {
    rows[0].J[6] = -1; rows[1].J[7] = -1; rows[2].J[8] = -1;
    mul0_331(a2, b2->R, j.anchor2);
    // dSetCrossMatrixPlus(J2a, a2)
    rows[0].J[9 + 1] = -a2[2]; rows[0].J[9 + 2] = a2[1];
    rows[1].J[9 + 0] = a2[2];  rows[1].J[9 + 2] = -a2[0];
    rows[2].J[9 + 0] = -a2[1]; rows[2].J[9 + 1] = a2[0];
    for (int i = 0; i < 3; i++) rows[i].c = k * (a2[i] + b2->pos[i] - a1[i] - b1.pos[i]);
  } else {
    for (int i = 0; i < 3; i++) rows[i].c = k * (j.anchor2[i] - a1[i] - b1.pos[i]);
  }
}

// setBall2 (joint.cpp:142-208): the ball rows aligned with `axis` (row 0 along it, with its own erp: the suspension)
__device__ inline void set_ball2(Row *rows, const Body &b1, const Body &b2, const Joint &j, const double *axis, double k1, double k) {
  double q1[3], q2[3], a1[3], a2[3], t[3];
  plane_space(axis, q1, q2);
  for (int i = 0; i < 3; i++) { rows[0].J[i] = axis[i]; rows[1].J[i] = q1[i]; rows[2].J[i] = q2[i]; }
  mul0_331(a1, b1.R, j.anchor1);
  cross3(t, a1, axis); for (int i = 0; i < 3; i++) rows[0].J[3 + i] = t[i];
  cross3(t, a1, q1);   for (int i = 0; i < 3; i++) rows[1].J[3 + i] = t[i];
  cross3(t, a1, q2);   for (int i = 0; i < 3; i++) rows[2].J[3 + i] = t[i];
  for (int i = 0; i < 3; i++) { rows[0].J[6 + i] = -axis[i]; rows[1].J[6 + i] = -q1[i]; rows[2].J[6 + i] = -q2[i]; }
  mul0_331(a2, b2.R, j.anchor2);
  cross3(t, a2, axis); for (int i = 0; i < 3; i++) rows[0].J[9 + i] = -t[i];
  cross3(t, a2, q1);   for (int i = 0; i < 3; i++) rows[1].J[9 + i] = -t[i];
  cross3(t, a2, q2);   for (int i = 0; i < 3; i++) rows[2].J[9 + i] = -t[i];
  double d[3];
  for (int i = 0; i < 3; i++) { a1[i] += b1.pos[i]; a2[i] += b2.pos[i]; }
  for (int i = 0; i < 3; i++) d[i] = a2[i] - a1[i];
  rows[0].c = k1 * dot3(axis, d);
  rows[1].c = k * dot3(q1, d);
  rows[2].c = k * dot3(q2, d);
}

// setFixedOrientation (joint.cpp:212-265) on rows[0..2]; k = fps * erp
__device__ inline void set_fixed_orientation(Row *rows, const Body &b1, const Body *b2, const Joint &j, double k) {
  rows[0].J[3] = 1; rows[1].J[4] = 1; rows[2].J[5] = 1;
  if (b2) { rows[0].J[9] = -1; rows[1].J[10] = -1; rows[2].J[11] = -1; }
  double qerr[4], e[3];
  if (b2) { double qq[4]; qmul1(qq, b1.q, b2->q); qmul2(qerr, qq, j.qrel); }
  else qmul3(qerr, b1.q, j.qrel);
  if (qerr[0] < 0) { qerr[1] = -qerr[1]; qerr[2] = -qerr[2]; qerr[3] = -qerr[3]; }
  mul0_331(e, b1.R, qerr + 1);
  rows[0].c = 2 * k * e[0]; rows[1].c = 2 * k * e[1]; rows[2].c = 2 * k * e[2];
}

// One thread per persistent joint.
__global__ void k_rows_joint(DevBatch B) {
  int ji = blockIdx.x * blockDim.x + threadIdx.x;
  if (ji >= B.nj) return;
  Joint &j = B.joint[ji];
  if (j.m == 0) return;
  const WorldInfo wi = B.world[j.world];
  Row *rows = B.row + wi.row0 + j.row0;
  const Body &b1 = B.body[j.b1];
  const Body *b2 = j.b2 >= 0 ? &B.body[j.b2] : nullptr;
  const int lb1 = j.b1 - wi.body0, lb2 = j.b2 >= 0 ? j.b2 - wi.body0 : -1;
  const double fps = 1.0 / B.p.step_size;
  double erp = B.p.erp;
  if (j.param[GZ_JP_ERP] >= 0) erp = j.param[GZ_JP_ERP];
  for (int r = 0; r < j.m; r++) row_init(rows[r], lb1, lb2, B.p.cfm, B.p.contact_max_correcting_vel, ji);
  if (j.type == GZ_JOINT_HINGE) {
    if (j.param[GZ_JP_CFM] >= 0) for (int r = 0; r < 5; r++) rows[r].cfm = j.param[GZ_JP_CFM];
    const double k = fps * erp;
    set_ball(rows, b1, b2, j, k);
    double ax1[3], p[3], q[3];
    mul0_331(ax1, b1.R, j.axis1);
    plane_space(ax1, p, q);
    for (int i = 0; i < 3; i++) { rows[3].J[3 + i] = p[i]; rows[4].J[3 + i] = q[i]; }
    if (b2) for (int i = 0; i < 3; i++) { rows[3].J[9 + i] = -p[i]; rows[4].J[9 + i] = -q[i]; }
    double ax2[3], bb[3];
    if (b2) mul0_331(ax2, b2->R, j.axis2);
    else { ax2[0] = j.axis2[0]; ax2[1] = j.axis2[1]; ax2[2] = j.axis2[2]; }
    cross3(bb, ax1, ax2);
    rows[3].c = k * dot3(bb, p);
    rows[4].c = k * dot3(bb, q);
    if (j.m == 6) add_limot(B, j, rows[5], ax1, fps, true);
  } else if (j.type == GZ_JOINT_BALL || j.type == GZ_JOINT_UNIVERSAL) {
    // dxJointBall::getInfo2 (ball.cpp:56-69), dxJointUniversal::getInfo2 (universal.cpp:310-389): setBall, then for the
    // universal joint one row that keeps the axes perpendicular and the limit motors of its two axes
    if (j.param[GZ_JP_CFM] >= 0) for (int r = 0; r < (j.type == GZ_JOINT_UNIVERSAL ? 4 : j.m); r++) rows[r].cfm = j.param[GZ_JP_CFM];
    set_ball(rows, b1, b2, j, fps * erp);
    if (j.type == GZ_JOINT_UNIVERSAL) {
      double ax1[3], ax2[3], t[3], p[3];
      mul0_331(ax1, b1.R, j.axis1);
      if (b2) mul0_331(ax2, b2->R, j.axis2);
      else { ax2[0] = j.axis2[0]; ax2[1] = j.axis2[1]; ax2[2] = j.axis2[2]; }
      const double k = dot3(ax1, ax2);
      for (int i = 0; i < 3; i++) t[i] = ax2[i] - k * ax1[i];
      cross3(p, ax1, t);
      safe_normalize3(p);
      for (int i = 0; i < 3; i++) rows[3].J[3 + i] = p[i];
      if (b2) for (int i = 0; i < 3; i++) rows[3].J[9 + i] = -p[i];
      rows[3].c = fps * erp * -k;
      // the limit / motor of axis 1, then of axis 2 (universal.cpp:384-388)
      int row = 4;
      if (j.limit || j.param[GZ_JP_FMAX] > 0) row += add_limot(B, j, rows[row], ax1, fps, true);
      if (j.limit2 || joint_ext(B, j)[GZ_JP_FMAX2] > 0) {
        Limot M2;
        M2.lostop = joint_ext(B, j)[GZ_JP_LOSTOP2]; M2.histop = joint_ext(B, j)[GZ_JP_HISTOP2]; M2.vel = joint_ext(B, j)[GZ_JP_VEL2]; M2.fmax = joint_ext(B, j)[GZ_JP_FMAX2];
        M2.fudge = joint_ext(B, j)[GZ_JP_FUDGE2]; M2.bounce = joint_ext(B, j)[GZ_JP_BOUNCE2]; M2.normal_cfm = joint_ext(B, j)[GZ_JP_NORMAL_CFM2];
        M2.stop_erp = joint_ext(B, j)[GZ_JP_STOP_ERP2]; M2.stop_cfm = joint_ext(B, j)[GZ_JP_STOP_CFM2]; M2.limit_err = j.limit_err2; M2.limit = j.limit2;
        add_limot(B, j, rows[row], ax2, fps, true, M2);
      }
    }
  } else if (j.type == GZ_JOINT_SCREW) {
    // dxJointScrew::getInfo2 (screw.cpp:113-420): two rows keep the anchors together across the axis, one row couples
    // the translation along the axis with the rotation about it (thread pitch, rad / m), two rows keep the axes aligned
    // (as the hinge's), then the stop of the rotation
    if (j.param[GZ_JP_CFM] >= 0) for (int r = 0; r < 5; r++) rows[r].cfm = j.param[GZ_JP_CFM];
    const double *X = joint_ext(B, j);
    const double pitch = X[GZ_JP_THREAD_PITCH];
    double ax1[3], q[3], p[3], a1[3], a2[3] = {0, 0, 0}, t[3];
    mul0_331(ax1, b1.R, j.axis1);
    if (b2) {
      mul0_331(q, b2->R, &X[GZ_JP_OFFSET]);
      for (int i = 0; i < 3; i++) q[i] = b1.pos[i] - q[i] - b2->pos[i];
    } else {
      for (int i = 0; i < 3; i++) q[i] = b1.pos[i] - X[GZ_JP_OFFSET + i];
    }
    const double lin_disp = dot3(ax1, q);
    const bool coarse = fabs(pitch) > 1.0;
    const double lin_err = coarse ? -(lin_disp - j.cum_angle / pitch) : -(pitch * lin_disp - j.cum_angle);
    plane_space(ax1, p, q);
    for (int i = 0; i < 3; i++) { rows[0].J[i] = p[i]; rows[1].J[i] = q[i]; }
    mul0_331(a1, b1.R, j.anchor1);
    cross3(t, p, a1); for (int i = 0; i < 3; i++) rows[0].J[3 + i] = -t[i];
    cross3(t, q, a1); for (int i = 0; i < 3; i++) rows[1].J[3 + i] = -t[i];
    if (coarse) for (int i = 0; i < 3; i++) { rows[2].J[i] = ax1[i]; rows[2].J[3 + i] = -ax1[i] / pitch; }
    else for (int i = 0; i < 3; i++) { rows[2].J[i] = ax1[i] * pitch; rows[2].J[3 + i] = -ax1[i]; }
    const double k = fps * erp;
    if (b2) {
      for (int i = 0; i < 3; i++) { rows[0].J[6 + i] = -p[i]; rows[1].J[6 + i] = -q[i]; }
      mul0_331(a2, b2->R, j.anchor2);
      cross3(t, p, a2); for (int i = 0; i < 3; i++) rows[0].J[9 + i] = t[i];
      cross3(t, q, a2); for (int i = 0; i < 3; i++) rows[1].J[9 + i] = t[i];
      if (coarse) for (int i = 0; i < 3; i++) { rows[2].J[9 + i] = ax1[i] / pitch; rows[2].J[6 + i] = -ax1[i]; }
      else for (int i = 0; i < 3; i++) { rows[2].J[9 + i] = ax1[i]; rows[2].J[6 + i] = -ax1[i] * pitch; }
      double e[3];
      for (int i = 0; i < 3; i++) e[i] = a2[i] + b2->pos[i] - a1[i] - b1.pos[i];
      rows[0].c = k * dot3(e, p); rows[1].c = k * dot3(e, q);
    } else {
      double e[3];
      for (int i = 0; i < 3; i++) e[i] = j.anchor2[i] - a1[i] - b1.pos[i];
      rows[0].c = k * dot3(e, p); rows[1].c = k * dot3(e, q);
    }
    rows[2].c = k * lin_err;
    for (int i = 0; i < 3; i++) { rows[3].J[3 + i] = p[i]; rows[4].J[3 + i] = q[i]; }
    if (b2) for (int i = 0; i < 3; i++) { rows[3].J[9 + i] = -p[i]; rows[4].J[9 + i] = -q[i]; }
    double ax2[3], bb[3];
    if (b2) mul0_331(ax2, b2->R, j.axis2);
    else { ax2[0] = j.axis2[0]; ax2[1] = j.axis2[1]; ax2[2] = j.axis2[2]; }
    cross3(bb, ax1, ax2);
    rows[3].c = k * dot3(bb, p);
    rows[4].c = k * dot3(bb, q);
    if (j.m == 6) add_limot(B, j, rows[5], ax1, fps, true);
  } else if (j.type == GZ_JOINT_GEARBOX) {
    // dxJointGearbox::getInfo2 (gearbox.cpp:62-103): the two angles (cumulative, each against its reference body), one
    // row on the angular velocities about the two axes; the world's erp (the joint's is not consulted)
    const double *X = joint_ext(B, j);
    const double ratio = X[GZ_JP_RATIO];
    const int r1 = (int)X[GZ_JP_REFBODY1], r2 = (int)X[GZ_JP_REFBODY2];
    double g1[3], g2[3];
    mul0_331(g1, b1.R, j.axis1);
    mul0_331(g2, b2->R, j.axis2);
    const double ang1 = body_angle(B, j.b1, r1 >= 0 ? wi.body0 + r1 : -1, j.axis1, j.qrel);
    const double ang2 = body_angle(B, j.b2, r2 >= 0 ? wi.body0 + r2 : -1, j.axis2, &X[GZ_JP_QREL2]);
    j.cum_angle = shortest_angular_distance_update(j.cum_angle, ang1);
    j.cum_angle2 = shortest_angular_distance_update(j.cum_angle2, ang2);
    const double err = shortest_angular_distance(-ratio * j.cum_angle2, j.cum_angle);
    for (int i = 0; i < 3; i++) { rows[0].J[3 + i] = g1[i]; rows[0].J[9 + i] = ratio * g2[i]; }
    rows[0].c = -(fps * B.p.erp) * err;
  } else if (j.type == GZ_JOINT_HINGE2) {
    // dxJointHinge2::getInfo2 (hinge2.cpp:158-222): ball rows aligned with axis 1 (row 0 = the suspension), one row that
    // keeps the angle between the axes, then the limit / motor of axis 1 and the motor of axis 2
    if (j.param[GZ_JP_CFM] >= 0) for (int r = 0; r < 4; r++) rows[r].cfm = j.param[GZ_JP_CFM];
    double ax1[3], ax2[3], q[3];
    mul0_331(ax1, b1.R, j.axis1);
    mul0_331(ax2, b2->R, j.axis2);
    cross3(q, ax1, ax2);
    const double sn = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]), cs = dot3(ax1, ax2);
    safe_normalize3(q);
    set_ball2(rows, b1, *b2, j, ax1, fps * joint_ext(B, j)[GZ_JP_SUSP_ERP], fps * erp);
    for (int i = 0; i < 3; i++) { rows[3].J[3 + i] = q[i]; rows[3].J[9 + i] = -q[i]; }
    rows[3].c = fps * erp * (joint_ext(B, j)[GZ_JP_C0] * sn - joint_ext(B, j)[GZ_JP_S0] * cs);
    int row = 4;
    if (j.limit || j.param[GZ_JP_FMAX] > 0) row += add_limot(B, j, rows[row], ax1, fps, true);
    if (joint_ext(B, j)[GZ_JP_FMAX2] > 0) {
      Limot M2;   // dxJointLimitMotor::init (joint.cpp:470-484) with dParamVel2 / dParamFMax2 set
      M2.lostop = -GZ_INF; M2.histop = GZ_INF; M2.vel = joint_ext(B, j)[GZ_JP_VEL2]; M2.fmax = joint_ext(B, j)[GZ_JP_FMAX2];
      M2.fudge = 1; M2.bounce = 0; M2.normal_cfm = B.p.cfm; M2.stop_erp = B.p.erp; M2.stop_cfm = B.p.cfm; M2.limit_err = 0; M2.limit = 0;
      add_limot(B, j, rows[row], ax2, fps, true, M2);
    }
    rows[0].cfm = joint_ext(B, j)[GZ_JP_SUSP_CFM];
  } else if (j.type == GZ_JOINT_SLIDER) {
    // dxJointSlider::getInfo2 (slider.cpp:151-238); anchor1 holds dxJointSlider::offset
    if (j.param[GZ_JP_CFM] >= 0) for (int r = 0; r < 5; r++) rows[r].cfm = j.param[GZ_JP_CFM];
    double c[3] = {0, 0, 0};
    if (b2) for (int i = 0; i < 3; i++) c[i] = b2->pos[i] - b1.pos[i];
    set_fixed_orientation(rows, b1, b2, j, fps * erp);   // rows 0..2
    double ax1[3], p[3], q[3];
    mul0_331(ax1, b1.R, j.axis1);
    plane_space(ax1, p, q);
    if (b2) {
      double t[3];
      cross3(t, c, p);
      for (int i = 0; i < 3; i++) { t[i] *= 0.5; rows[3].J[3 + i] = t[i]; rows[3].J[9 + i] = t[i]; }
      cross3(t, c, q);
      for (int i = 0; i < 3; i++) { t[i] *= 0.5; rows[4].J[3 + i] = t[i]; rows[4].J[9 + i] = t[i]; }
      for (int i = 0; i < 3; i++) { rows[3].J[6 + i] = -p[i]; rows[4].J[6 + i] = -q[i]; }
    }
    for (int i = 0; i < 3; i++) { rows[3].J[i] = p[i]; rows[4].J[i] = q[i]; }
    const double k = fps * erp;
    if (b2) {
      double ofs[3];
      mul0_331(ofs, b2->R, j.anchor1);
      for (int i = 0; i < 3; i++) c[i] += ofs[i];
      rows[3].c = k * dot3(p, c);
      rows[4].c = k * dot3(q, c);
    } else {
      double ofs[3];
      for (int i = 0; i < 3; i++) ofs[i] = j.anchor1[i] - b1.pos[i];
      rows[3].c = k * dot3(p, ofs);
      rows[4].c = k * dot3(q, ofs);
    }
    if (j.m == 6) add_limot(B, j, rows[5], ax1, fps, false);
  } else if (j.type == GZ_JOINT_FIXED) {
    if (j.param[GZ_JP_CFM] >= 0) for (int r = 0; r < 6; r++) rows[r].cfm = j.param[GZ_JP_CFM];
    const double k = fps * erp;
    set_fixed_orientation(rows + 3, b1, b2, j, k);
    // position rows (fixed.cpp:82-109); anchor1 holds dxJointFixed::offset
    rows[0].J[0] = 1; rows[1].J[1] = 1; rows[2].J[2] = 1;
    double ofs[3];
    mul0_331(ofs, b1.R, j.anchor1);
    if (b2) {
      // dSetCrossMatrixPlus(J1a, ofs)
      rows[0].J[3 + 1] = -ofs[2]; rows[0].J[3 + 2] = ofs[1];
      rows[1].J[3 + 0] = ofs[2];  rows[1].J[3 + 2] = -ofs[0];
      rows[2].J[3 + 0] = -ofs[1]; rows[2].J[3 + 1] = ofs[0];
      rows[0].J[6] = -1; rows[1].J[7] = -1; rows[2].J[8] = -1;
      for (int i = 0; i < 3; i++) rows[i].c = k * (b2->pos[i] - b1.pos[i] + ofs[i]);
    } else {
      for (int i = 0; i < 3; i++) rows[i].c = k * (j.anchor1[i] - b1.pos[i]);
    }
  }
}

// One thread per contact: dxJointContact::getInfo2 with Gazebo's mode flags
// (Bounce | Mu2 | SoftERP | SoftCFM | Approx1 | Approx3 | Slip1 | Slip2 [| FDir1]).
__device__ __forceinline__ void rows_contact_one(DevBatch &B, const int w, const WorldInfo &wi, const int ci) {
  const Contact &C = B.contact[wi.contact0 + ci];
  Row *rows = B.row + wi.row0 + C.row0;
  const Body &b1 = B.body[C.b1];
  const Body *b2 = C.b2 >= 0 ? &B.body[C.b2] : nullptr;
  const int lb1 = C.b1 - wi.body0, lb2 = C.b2 >= 0 ? C.b2 - wi.body0 : -1;
  const double fps = 1.0 / B.p.step_size;
  for (int r = 0; r < wi.rpc; r++)
    row_init(rows[r], lb1, lb2, B.p.cfm, B.p.contact_max_correcting_vel, ~(wi.contact0 + ci));
  double normal[3];
  if (C.reverse) { normal[0] = -C.normal[0]; normal[1] = -C.normal[1]; normal[2] = -C.normal[2]; }
  else { normal[0] = C.normal[0]; normal[1] = C.normal[1]; normal[2] = C.normal[2]; }
  double c1[3], c2[3] = {0, 0, 0};
  for (int i = 0; i < 3; i++) c1[i] = C.pos[i] - b1.pos[i];
  Row &r0 = rows[0];
  for (int i = 0; i < 3; i++) r0.J[i] = normal[i];
  cross3(r0.J + 3, c1, normal);
  if (b2) {
    for (int i = 0; i < 3; i++) { c2[i] = C.pos[i] - b2->pos[i]; r0.J[6 + i] = -normal[i]; }
    cross3(r0.J + 9, c2, normal);
    r0.J[9] = -r0.J[9]; r0.J[10] = -r0.J[10]; r0.J[11] = -r0.J[11];
  }
  // every body carries contactp (ODELink::UpdateSurface always calls dBodySetMaxVel/MinDepth)
  double min_min_depth = b2 ? fmin(b1.min_depth, b2->min_depth) : b1.min_depth;
  double depth = C.depth - min_min_depth;
  if (depth < 0) depth = 0;
  r0.cfm = C.soft_cfm;
  const double motionN = 0;
  double local_erp = C.soft_erp;
  if (C.em > 0) {
    // dContactEM (contact.cpp:163-216): Hertzian stiffness 4/3 E* sqrt(r x) linearised at this depth; kd recovered from
    // the soft cfm / erp, then erp and cfm recomputed with it
    double patch_radius;
    if (C.em_surface) patch_radius = sqrt(C.em_radius * depth);   // (the surface radius is never negative here: :169-170)
    else patch_radius = C.em_radius;
    const double khertz_sqrtx = 4.0 / 3.0 * C.em * sqrt(patch_radius * depth);
    const double kd = 1.0 / r0.cfm - local_erp / fps;
    const double kph = khertz_sqrtx / fps;
    local_erp = (kph) / (kph + kd);
    r0.cfm = 1.0 / (kph + kd);
  }
  const double k = fps * local_erp;
  const double pushout = k * depth + motionN;
  r0.c = pushout;
  double maxvel = b2 ? fmin(b1.max_vel, b2->max_vel) : b1.max_vel;
  if (maxvel > B.p.contact_max_correcting_vel) maxvel = B.p.contact_max_correcting_vel;
  r0.c_v_max = maxvel;
  {  // dContactBounce
    double outgoing = dot3(r0.J, b1.lvel) + dot3(r0.J + 3, b1.avel);
    if (b2) outgoing += dot3(r0.J + 6, b2->lvel) + dot3(r0.J + 9, b2->avel);
    outgoing -= motionN;
    if (C.bounce_vel >= 0 && (-outgoing) > C.bounce_vel) {
      double newc = -C.bounce * outgoing + motionN;
      if (newc > r0.c) r0.c = newc;
    }
  }
  r0.lo = 0; r0.hi = GZ_INF; r0.findex = -2;
  double mu = C.mu < 0 ? 0 : C.mu, mu2 = C.mu2 < 0 ? 0 : C.mu2;
  double t1[3], t2[3];
  if (C.has_fdir1) {
    t1[0] = C.fdir1[0]; t1[1] = C.fdir1[1]; t1[2] = C.fdir1[2];
    cross3(t2, normal, t1);
    if (dequal(t2[0], 0.0) && dequal(t2[1], 0.0) && dequal(t2[2], 0.0)) plane_space(normal, t1, t2);
  } else {
    plane_space(normal, t1, t2);
  }
  const int abs_normal_row = C.row0;   // findex rebasing (quickstep.cpp:369-374)
  for (int f = 0; f < 2; f++) {
    Row &r = rows[1 + f];
    const double *t = f ? t2 : t1;
    for (int i = 0; i < 3; i++) r.J[i] = t[i];
    cross3(r.J + 3, c1, t);
    if (b2) {
      for (int i = 0; i < 3; i++) r.J[6 + i] = -t[i];
      cross3(r.J + 9, c2, t);
      r.J[9] = -r.J[9]; r.J[10] = -r.J[10]; r.J[11] = -r.J[11];
    }
    double m = f ? mu2 : mu;
    r.lo = -m; r.hi = m;
    r.findex = abs_normal_row;
    r.cfm = f ? C.slip2 : C.slip1;
  }
  if (wi.rpc == 4) {
    // torsional friction (contact.cpp:362-437): a moment about the normal, limits +-(3 pi a mu3 / 16) x the normal force
    // (findex = the normal row: dContactApprox3, ODEPhysics.cc:1099), a = the patch radius or sqrt(R x depth).
    // A pair without it keeps an empty placeholder (no body entries are made for it, k_bsched).
    Row &r = rows[3];
    r.findex = abs_normal_row;
    r.lo = 0; r.hi = 0;
    if (C.tors) {
      for (int i = 0; i < 3; i++) { r.J[i] = 0.0; r.J[3 + i] = normal[i]; }
      if (b2) for (int i = 0; i < 3; i++) { r.J[6 + i] = -0.0; r.J[9 + i] = -normal[i]; }
      const double patch_radius = C.tors == 2 ? sqrt(C.tors_radius * depth) : C.tors_radius;
      const double lim = (3 * 3.14159265358979323846 * patch_radius * C.mu3) / 16;
      r.lo = -lim; r.hi = lim;
      if (C.slip3 > 0) r.cfm = C.slip3;
    }
  }
}

// DYNAMIC_INERTIA (quickstep_util.cpp:170-524; switched by qs.dynamic_inertia_reduction, ODEPhysics.cc:186-194).  For
// rows 3 and 4 of every hinge between two dynamic bodies - the two rotational constraint axes - compare the bodies'
// inertia along the axis S; if the ratio exceeds 200 move dm of it from the large to the small one (MOI -/+= dm S S^T)
// unless a row of either adjusted matrix would lose diagonal dominance (generalised line criterion), then invert both
// matrices analytically.  The reference does this inside its serial getInfo2 loop, so a body shared by several hinges
// is adjusted in joint order: one thread per world walks the world's joints in index order (the canonical order).
__device__ inline void sym_inverse(const double *M, double *inv) {   // :396-419, cofactor by cofactor as written
  const double det = M[0] * (M[8] * M[4] - M[7] * M[5]) - M[3] * (M[8] * M[1] - M[7] * M[2]) + M[6] * (M[5] * M[1] - M[4] * M[2]);
  inv[0] = (M[8] * M[4] - M[7] * M[5]) / det;
  inv[1] = -(M[8] * M[1] - M[7] * M[2]) / det;
  inv[2] = (M[5] * M[1] - M[4] * M[2]) / det;
  inv[3] = inv[1];
  inv[4] = (M[8] * M[0] - M[6] * M[2]) / det;
  inv[5] = -(M[5] * M[0] - M[3] * M[2]) / det;
  inv[6] = inv[2];
  inv[7] = inv[5];
  inv[8] = (M[4] * M[0] - M[3] * M[1]) / det;
}
__global__ void k_dyn_inertia(DevBatch B) {
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= B.n_worlds) return;
  const WorldInfo wi = B.world[w];
  if (B.wstep[w].m == 0) return;
  for (int k = 0; k < wi.nj; k++) {
    const Joint &j = B.joint[wi.joint0 + k];
    if (j.type != GZ_JOINT_HINGE || j.m == 0 || j.b2 < 0) continue;
    double *MOI1 = B.bstep[j.b1].MOI, *MOI2 = B.bstep[j.b2].MOI;
    double *inv1 = B.bstep[j.b1].invMOI, *inv2 = B.bstep[j.b2].invMOI;
    for (int r = 3; r <= 4; r++) {
      const Row &row = B.row[wi.row0 + j.row0 + r];
      double S[3] = {row.J[3], row.J[4], row.J[5]};   // J1a
      safe_normalize3(S);
      // m = S' MOI S (dMultiply0_133 then dCalcVectorDot3)
      double t1[3], t2[3];
      for (int c = 0; c < 3; c++) {
        t1[c] = MOI1[c] * S[0] + MOI1[3 + c] * S[1] + MOI1[6 + c] * S[2];
        t2[c] = MOI2[c] * S[0] + MOI2[3 + c] * S[1] + MOI2[6 + c] * S[2];
      }
      const double m1 = t1[0] * S[0] + t1[1] * S[1] + t1[2] * S[2];
      const double m2 = t2[0] * S[0] + t2[1] * S[1] + t2[2] * S[2];
      double m_large = m1, m_small = m2;
      double *ML = MOI1, *MS = MOI2;
      if (m2 > m1) { m_large = m2; m_small = m1; ML = MOI2; MS = MOI1; }
      const double SS[9] = {S[0] * S[0], S[0] * S[1], S[0] * S[2], S[1] * S[0], S[1] * S[1], S[1] * S[2],
                            S[2] * S[0], S[2] * S[1], S[2] * S[2]};
      const double moi_ratio_max = 200.0;
      if (m_large > moi_ratio_max * m_small) {
        const double dm = (m_large - moi_ratio_max * m_small) / (1 + moi_ratio_max);
        int problem = 0;
        for (int rw = 0; rw < 3; ++rw) {
          const int c1 = (rw + 1) % 3, c2 = (rw + 2) % 3;
          const int id = 3 * rw + rw, i1 = 3 * rw + c1, i2 = 3 * rw + c2;
          const double Md_large = ML[id] - dm * SS[id];
          const double Mod_large = fabs(ML[i1] - dm * SS[i1]) + fabs(ML[i2] - dm * SS[i2]);
          const double Md_small = MS[id] + dm * SS[id];
          const double Mod_small = fabs(MS[i1] + dm * SS[i1]) + fabs(MS[i2] + dm * SS[i2]);
          if (Md_large <= Mod_large || Md_small <= Mod_small) problem = 1;   // gamma = 1
        }
        if (problem == 0)
          for (int i = 0; i < 9; i++) { ML[i] -= dm * SS[i]; MS[i] += dm * SS[i]; }
        sym_inverse(MOI1, inv1);   // (the reference re-inverts both whether or not the matrices changed)
        sym_inverse(MOI2, inv2);
      }
    }
  }
}

// grid (gx, n_worlds): one thread per contact, gx blocks striding over the world's contacts
__global__ void k_rows_contact(DevBatch B) {
  const int w = blockIdx.y;
  const WorldInfo wi = B.world[w];
  if (B.wstep[w].m == 0) return;
  const int n = B.wstep[w].n_contacts;
  for (int ci = blockIdx.x * blockDim.x + threadIdx.x; ci < n; ci += gridDim.x * blockDim.x) rows_contact_one(B, w, wi, ci);
}

}  // namespace gz
