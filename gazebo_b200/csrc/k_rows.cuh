// k_rows.cuh — everything of dWorldQuickStep that comes before the PGS sweeps:
//   k_control    plugin stand-in: PD-tracked sinusoid -> dJointAddHingeTorque (hinge.cpp:390-409)
//   k_island     dInternalHandleAutoDisabling (util.cpp:116-252) + island membership
//                (BuildIslandsAndEstimateStepperMemoryRequirements util.cpp:408-558)
//   k_bodyprep   world-frame inverse inertia, gyroscopic torque, gravity (quickstep.cpp:113-195)
//   k_info1      getInfo1 row counts + row offsets (quickstep.cpp:200-236, hinge.cpp:53-78,
//                contact.cpp:46-76, fixed.cpp:52-57)
//   k_rows_joint / k_rows_contact   getInfo2 (hinge.cpp:81-180, joint.cpp:93-134,212-265,
//                573-750, fixed.cpp:59-110, contact.cpp:79-438) and findex rebasing
//                (quickstep.cpp:369-374)
//   k_rowfinish  rhs / rhs_erp / cfm scaling (quickstep.cpp:403-443), warm start (:473-495),
//                iMJ, Ad, J scaling (quickstep_pgs_lcp.cpp:985-1107)
#pragma once
#include "gz_math.cuh"
#include "gz_batch.h"
#include "../../include/gz_detmath.h"

namespace gz {

// getHingeAngle (joint.cpp:442-465) + getHingeAngleFromRelativeQuat (:391-432)
#define GZ_PI 3.14159265358979323846   // M_PI

// the joint_param slots beyond GZ_JP_HOT (hinge2 / universal joints only): a side array, so that struct Joint stays small
__device__ __forceinline__ const double *joint_ext(const DevBatch &B, const Joint &j) {
  return B.joint_ext + (size_t)(&j - B.joint) * (GZ_JP_COUNT - GZ_JP_HOT) - GZ_JP_HOT;   // indexed by the slot number itself
}

// dxJointHinge2::measureAngle1 (hinge2.cpp:34-52): axis 2 seen from body 1, against v1 / v2
__device__ inline double hinge2_angle1(const DevBatch &B, const Joint &j) {
  double p[3], q[3];
  mul0_331(p, B.body[j.b2].R, j.axis2);
  mul1_331(q, B.body[j.b1].R, p);
  const double x = dot3(&joint_ext(B, j)[GZ_JP_V1], q), y = dot3(&joint_ext(B, j)[GZ_JP_V2], q);
  return -gzdet_atan2(y, x);
}

// getHingeAngleFromRelativeQuat (joint.cpp:391-432)
__device__ inline double angle_from_relative_quat(const double *qrel, const double *axis) {
  double cost2 = qrel[0];
  double sint2 = sqrt(qrel[1] * qrel[1] + qrel[2] * qrel[2] + qrel[3] * qrel[3]);
  double theta = (dot3(qrel + 1, axis) >= 0) ? (2 * gzdet_atan2(sint2, cost2)) : (2 * gzdet_atan2(sint2, -cost2));
  if (theta > 3.14159265358979323846) theta -= 2 * 3.14159265358979323846;
  return -theta;
}

__device__ inline double hinge_angle(const DevBatch &B, const Joint &j) {
  double qrel[4];
  const Body &b1 = B.body[j.b1];
  if (j.b2 >= 0) {
    double qq[4];
    qmul1(qq, b1.q, B.body[j.b2].q);
    qmul2(qrel, qq, j.qrel);
  } else {
    qmul3(qrel, b1.q, j.qrel);
  }
  return angle_from_relative_quat(qrel, j.axis1);
}

// getHingeAngle (joint.cpp:442-465) of a body against a reference body (global index, < 0: the world) - the gearbox joint
__device__ inline double body_angle(const DevBatch &B, const int body, const int ref, const double *axis, const double *q0) {
  double qrel[4];
  if (ref >= 0) { double qq[4]; qmul1(qq, B.body[body].q, B.body[ref].q); qmul2(qrel, qq, q0); }
  else qmul3(qrel, B.body[body].q, q0);
  return angle_from_relative_quat(qrel, axis);
}

// dRFrom2Axes (rotation.cpp:95-136) + dRtoQ / dQfromR (:261-311): the frame whose x axis is a and whose y axis is b made
// perpendicular to a, as a quaternion
__device__ inline void quat_from_2_axes(double *q, const double *a_in, const double *b_in) {
  double ax = a_in[0], ay = a_in[1], az = a_in[2], bx = b_in[0], by = b_in[1], bz = b_in[2];
  double R[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  double l = sqrt(ax * ax + ay * ay + az * az);
  if (l > 0.0) {
    l = 1.0 / l;
    ax *= l; ay *= l; az *= l;
    const double k = ax * bx + ay * by + az * bz;
    bx -= k * ax; by -= k * ay; bz -= k * az;
    l = sqrt(bx * bx + by * by + bz * bz);
    if (l > 0.0) {
      l = 1.0 / l;
      bx *= l; by *= l; bz *= l;
      R[0] = ax; R[3] = ay; R[6] = az;
      R[1] = bx; R[4] = by; R[7] = bz;
      R[2] = -by * az + ay * bz; R[5] = -bz * ax + az * bx; R[8] = -bx * ay + ax * by;
    }
  }
  const double tr = R[0] + R[4] + R[8];
  if (tr >= 0) {
    double s = sqrt(tr + 1);
    q[0] = 0.5 * s;
    s = 0.5 * (1.0 / s);
    q[1] = (R[7] - R[5]) * s; q[2] = (R[2] - R[6]) * s; q[3] = (R[3] - R[1]) * s;
    return;
  }
  int c = 0;   // the largest diagonal element, ties as dQfromR's jumps resolve them
  if (R[4] > R[0]) c = (R[8] > R[4]) ? 2 : 1;
  else if (R[8] > R[0]) c = 2;
  if (c == 0) {
    double s = sqrt((R[0] - (R[4] + R[8])) + 1);
    q[1] = 0.5 * s; s = 0.5 * (1.0 / s);
    q[2] = (R[1] + R[3]) * s; q[3] = (R[6] + R[2]) * s; q[0] = (R[7] - R[5]) * s;
  } else if (c == 1) {
    double s = sqrt((R[4] - (R[8] + R[0])) + 1);
    q[2] = 0.5 * s; s = 0.5 * (1.0 / s);
    q[3] = (R[5] + R[7]) * s; q[1] = (R[1] + R[3]) * s; q[0] = (R[2] - R[6]) * s;
  } else {
    double s = sqrt((R[8] - (R[0] + R[4])) + 1);
    q[3] = 0.5 * s; s = 0.5 * (1.0 / s);
    q[1] = (R[6] + R[2]) * s; q[2] = (R[5] + R[7]) * s; q[0] = (R[3] - R[1]) * s;
  }
}

// dxJointUniversal::getAngles (universal.cpp:75-165): the angles about axis 1 and axis 2
__device__ inline void universal_angles(const DevBatch &B, const Joint &j, double &angle1, double &angle2) {
  const Body &b1 = B.body[j.b1];
  double ax1[3], ax2[3], qcross[4], qq[4], qrel[4], qcross2[4];
  mul0_331(ax1, b1.R, j.axis1);
  if (j.b2 >= 0) mul0_331(ax2, B.body[j.b2].R, j.axis2);
  else { ax2[0] = j.axis2[0]; ax2[1] = j.axis2[1]; ax2[2] = j.axis2[2]; }
  quat_from_2_axes(qcross, ax1, ax2);
  qmul1(qq, b1.q, qcross);
  qmul2(qrel, qq, j.qrel);   // joint_qrel = qrel1
  angle1 = angle_from_relative_quat(qrel, j.axis1);
  // the cross for axis 2 = the first one turned by 180 degrees about the direction between the two axes
  qrel[0] = 0;
  qrel[1] = ax1[0] + ax2[0]; qrel[2] = ax1[1] + ax2[1]; qrel[3] = ax1[2] + ax2[2];
  const double l = 1.0 / sqrt(qrel[1] * qrel[1] + qrel[2] * qrel[2] + qrel[3] * qrel[3]);
  qrel[1] *= l; qrel[2] *= l; qrel[3] *= l;
  qmul0(qcross2, qrel, qcross);
  if (j.b2 >= 0) { qmul1(qq, B.body[j.b2].q, qcross2); qmul2(qrel, qq, &joint_ext(B, j)[GZ_JP_QREL2]); }
  else qmul2(qrel, qcross2, &joint_ext(B, j)[GZ_JP_QREL2]);
  angle2 = -angle_from_relative_quat(qrel, j.axis2);
}

// dJointGetSliderPosition (slider.cpp:40-84); anchor1 holds dxJointSlider::offset
__device__ inline double slider_position(const DevBatch &B, const Joint &j) {
  const Body &b1 = B.body[j.b1];
  double ax1[3], q[3];
  mul0_331(ax1, b1.R, j.axis1);
  if (j.b2 >= 0) {
    const Body &b2 = B.body[j.b2];
    mul0_331(q, b2.R, j.anchor1);
    for (int i = 0; i < 3; i++) q[i] = b1.pos[i] - q[i] - b2.pos[i];
  } else {
    for (int i = 0; i < 3; i++) q[i] = b1.pos[i] - j.anchor1[i];
  }
  return dot3(ax1, q);
}

// One thread per body: adds the PD torque of every hinge touching the body, in joint order
// (the order dBodyAddTorque calls arrive in when the host loops over joints).
__global__ void k_control(DevBatch B) {
  int bi = blockIdx.x * blockDim.x + threadIdx.x;
  if (bi >= B.nb) return;
  Body &b = B.body[bi];
  const double t = (double)B.step_index * B.p.step_size;
  for (int k = b.joint_begin; k < b.joint_end; k++) {
    const Joint &j = B.joint[B.body_joint[k]];
    if (j.type != GZ_JOINT_HINGE || !(j.pd[5] > 0)) continue;
    double target = j.pd[0] * gzdet_sin(2.0 * 3.14159265358979323846 * j.pd[1] * t + j.pd[2]);
    double ang = hinge_angle(B, j);
    double angle = shortest_angular_distance_update(j.cum_angle, ang);   // dJointGetHingeAngle
    const Body &b1 = B.body[j.b1];
    double axis[3];
    mul0_331(axis, b1.R, j.axis1);
    double rate = dot3(axis, b1.avel);                                     // dJointGetHingeAngleRate
    if (j.b2 >= 0) rate -= dot3(axis, B.body[j.b2].avel);
    double tau = j.pd[3] * (target - angle) - j.pd[4] * rate;
    if (tau > j.pd[5]) tau = j.pd[5];
    if (tau < -j.pd[5]) tau = -j.pd[5];
    axis[0] *= tau; axis[1] *= tau; axis[2] *= tau;
    if (j.b1 == bi) { b.tacc[0] += axis[0]; b.tacc[1] += axis[1]; b.tacc[2] += axis[2]; }
    else            { b.tacc[0] += -axis[0]; b.tacc[1] += -axis[1]; b.tacc[2] += -axis[2]; }
  }
}

// Device-side stand-in for a plugin that calls Joint::SetVelocity on hinge joints every update, which in this fork
// is Joint::SetVelocityMaximal (gazebo/physics/Joint.cc:828-894; DiffDrivePlugin::OnUpdate plugins/DiffDrivePlugin.cc:100-122,
// ODEHingeJoint::SetVelocity ODEHingeJoint.cc:153-156): the child link (body 1 of the ODE joint) gets the parent's angular
// velocity plus vel along the joint axis, and the linear velocity of a point rigidly carried by the joint anchor.  One thread
// per world walks its joints in order, as the plugin's calls arrive.  (Link frame = centre of mass in this scene format; the
// anchor offsets are taken with ODE's matrix arithmetic, the same in the oracle harness.)
__global__ void k_velcmd(DevBatch B) {
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= B.n_worlds) return;
  const WorldInfo wi = B.world[w];
  for (int k = 0; k < wi.nj; k++) {
    const Joint &j = B.joint[wi.joint0 + k];
    if (!j.has_vel_cmd || j.type != GZ_JOINT_HINGE) continue;
    Body &c = B.body[j.b1];
    double ang[3] = {0.0, 0.0, 0.0}, lin[3] = {0.0, 0.0, 0.0};
    if (j.b2 >= 0) {
      const Body &p = B.body[j.b2];
      double off[3], cr[3];
      mul0_331(off, p.R, j.anchor2);
      cross3(cr, p.avel, off);
      for (int i = 0; i < 3; i++) { ang[i] = p.avel[i]; lin[i] = p.lvel[i] + cr[i]; }
    }
    double axis[3], a[3], jw[3], off[3], cr[3];
    mul0_331(axis, c.R, j.axis1);
    for (int i = 0; i < 3; i++) ang[i] += j.vel_cmd * axis[i];
    mul0_331(a, c.R, j.anchor1);
    for (int i = 0; i < 3; i++) { jw[i] = c.pos[i] + a[i]; off[i] = c.pos[i] - jw[i]; }
    cross3(cr, ang, off);
    for (int i = 0; i < 3; i++) { c.avel[i] = ang[i]; c.lvel[i] = lin[i] + cr[i]; }
  }
}

// Auto-disable of one body (average_samples = 1: the instantaneous velocity is the average) and its island label's
// start value (dxStepBody's bookkeeping, ode.cpp dInternalHandleAutoDisabling).
__device__ __forceinline__ void autodisable_body(const DevBatch &B, const WorldInfo &wi, const int i) {
  const double h = B.p.step_size;
  Body &b = B.body[wi.body0 + i];
  BodyStep &bs = B.bstep[wi.body0 + i];
  bs.label = i;
  bool has_joint = (b.joint_end > b.joint_begin) || bs.has_contact;
  if (has_joint && (b.flags & (GZ_BODY_AUTO_DISABLE | GZ_BODY_DISABLED)) == GZ_BODY_AUTO_DISABLE) {
    int idle = 1;
    double av_l = dot3(b.lvel, b.lvel);
    if (av_l > B.p.adis_linear_threshold_sq) idle = 0;
    else {
      double av_a = dot3(b.avel, b.avel);
      if (av_a > B.p.adis_angular_threshold_sq) idle = 0;
    }
    if (idle) { b.adis_stepsleft--; b.adis_timeleft -= h; }
    else { b.adis_stepsleft = B.p.adis_idle_steps; b.adis_timeleft = B.p.adis_idle_time; }
    if (b.adis_stepsleft <= 0 && b.adis_timeleft <= 0) {
      b.flags |= GZ_BODY_DISABLED;
      for (int k = 0; k < 3; k++) { b.lvel[k] = 0; b.avel[k] = 0; }
    }
  }
}

// One CTA per world (a phase of k_assemble1 below).
__device__ __forceinline__ void island_world(const DevBatch &B, const int w) {
  const WorldInfo wi = B.world[w];
  const WorldStep &ws = B.wstep[w];
  __shared__ int changed;
  // 1. auto-disable
  for (int i = threadIdx.x; i < wi.nb; i += blockDim.x) autodisable_body(B, wi, i);
  __syncthreads();
  // 2. connected components over two-body joints (persistent + contact): the smallest body index of the component
  // becomes everybody's label.  Small worlds: min-label propagation (a few passes).  Large worlds (a pile of
  // thousands: the propagation needs as many passes as the pile is wide, 3 ms at C4): hook the larger root under the
  // smaller one, then point every body at its root, until no edge joins two roots - a logarithmic number of passes.
  if (B.islands_hook) {
    int *const lab0 = &B.bstep[wi.body0].label;
    const size_t ls = sizeof(BodyStep) / sizeof(int);   // labels are BodyStep::label, one per body
    auto root = [&](int x) { for (;;) { const int l = *(volatile int *)(lab0 + (size_t)x * ls); if (l == x) return x; x = l; } };
    for (;;) {
      if (threadIdx.x == 0) changed = 0;
      __syncthreads();
      for (int k = threadIdx.x; k < wi.nj + ws.n_contacts; k += blockDim.x) {
        int b1, b2;
        if (k < wi.nj) { const Joint &j = B.joint[wi.joint0 + k]; b1 = j.b1; b2 = j.b2; }
        else { const Contact &c = B.contact[wi.contact0 + (k - wi.nj)]; b1 = c.b1; b2 = c.b2; }
        if (b2 < 0) continue;
        const int r1 = root(b1 - wi.body0), r2 = root(b2 - wi.body0);
        if (r1 != r2) { atomicMin(lab0 + (size_t)max(r1, r2) * ls, min(r1, r2)); changed = 1; }
      }
      __syncthreads();
      for (int i = threadIdx.x; i < wi.nb; i += blockDim.x) B.bstep[wi.body0 + i].label = root(i);
      __syncthreads();
      const int c = changed;
      __syncthreads();
      if (!c) break;
    }
  } else
  for (;;) {
    if (threadIdx.x == 0) changed = 0;
    __syncthreads();
    for (int k = threadIdx.x; k < wi.nj + ws.n_contacts; k += blockDim.x) {
      int b1, b2;
      if (k < wi.nj) { const Joint &j = B.joint[wi.joint0 + k]; b1 = j.b1; b2 = j.b2; }
      else { const Contact &c = B.contact[wi.contact0 + (k - wi.nj)]; b1 = c.b1; b2 = c.b2; }
      if (b2 < 0) continue;
      int *l1 = &B.bstep[b1].label, *l2 = &B.bstep[b2].label;
      int a = *(volatile int *)l1, b = *(volatile int *)l2;
      if (a < b) { atomicMin(l2, a); changed = 1; }
      else if (b < a) { atomicMin(l1, b); changed = 1; }
    }
    __syncthreads();
    int c = changed;
    __syncthreads();
    if (!c) break;
  }
  // 3. a component is stepped iff it holds an enabled body; every body in it is re-enabled
  for (int i = threadIdx.x; i < wi.nb; i += blockDim.x) B.bstep[wi.body0 + i].active = 0;
  __syncthreads();
  for (int i = threadIdx.x; i < wi.nb; i += blockDim.x) {
    if (!(B.body[wi.body0 + i].flags & GZ_BODY_DISABLED))
      B.bstep[wi.body0 + B.bstep[wi.body0 + i].label].active = 1;   // benign race: all write 1
  }
  __syncthreads();
  for (int i = threadIdx.x; i < wi.nb; i += blockDim.x) {
    BodyStep &bs = B.bstep[wi.body0 + i];
    int act = B.bstep[wi.body0 + bs.label].active;
    if (bs.label != i) bs.active = act;
    if (act) B.body[wi.body0 + i].flags &= ~GZ_BODY_DISABLED;
  }
}

// One thread per body (active bodies only).
__device__ __forceinline__ void bodyprep_body(const DevBatch &B, const int bi) {
  BodyStep &bs = B.bstep[bi];
  for (int k = 0; k < 6; k++) { bs.caccel[k] = 0; bs.caccel_erp[k] = 0; }
  if (!bs.active) return;
  Body &b = B.body[bi];
  double tmp[9], MOI[9];
  mul2_333(tmp, b.invI, b.R);
  mul0_333(bs.invMOI, b.R, tmp);
  if (B.p.dynamic_inertia) {   // quickstep.cpp:119-139 computes MOI for every body; only DYNAMIC_INERTIA reads it later
    mul2_333(tmp, b.I, b.R);
    mul0_333(bs.MOI, b.R, tmp);
  }
  if (!(b.flags & GZ_BODY_NO_GYRO)) {
    mul2_333(tmp, b.I, b.R);
    mul0_333(MOI, b.R, tmp);
    double t3[3], cr[3];
    mul0_331(t3, MOI, b.avel);
    cross3(cr, b.avel, t3);
    b.tacc[0] = b.tacc[0] - cr[0]; b.tacc[1] = b.tacc[1] - cr[1]; b.tacc[2] = b.tacc[2] - cr[2];
  }
  if (!(b.flags & GZ_BODY_NO_GRAVITY)) {
    for (int k = 0; k < 3; k++)
      if (!dequal(B.p.gravity[k], 0.0)) b.facc[k] += b.mass * B.p.gravity[k];
  }
}

// One CTA per world: row count per joint (with the hinge's cumulative-angle side effect),
// exclusive scan -> Joint::row0; contacts follow with 3 rows each.
__device__ __forceinline__ void info1_world(const DevBatch &B, const int w) {
  const WorldInfo wi = B.world[w];
  WorldStep &ws = B.wstep[w];
  __shared__ int carry;
  __shared__ int warp_sums[32];
  if (threadIdx.x == 0) carry = 0;
  if (w == 0)   // the schedule buckets are rebuilt by k_schedule
    for (int i = threadIdx.x; i < GZ_NBUCKET; i += blockDim.x) B.bk_count[i] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int k0 = 0; k0 < wi.nj; k0 += blockDim.x) {
    int k = k0 + threadIdx.x;
    int m = 0;
    if (k < wi.nj) {
      Joint &j = B.joint[wi.joint0 + k];
      if (B.bstep[j.b1].active) {
        if (j.type == GZ_JOINT_HINGE) {
          m = (j.param[GZ_JP_FMAX] > 0) ? 6 : 5;
          double lo = j.param[GZ_JP_LOSTOP], hi = j.param[GZ_JP_HISTOP];
          if (lo <= hi) {
            double angle = hinge_angle(B, j);
            j.cum_angle = shortest_angular_distance_update(j.cum_angle, angle);
            // dxJointLimitMotor::testRotationalLimit (joint.cpp:553-571)
            if (j.cum_angle <= lo) { j.limit = 1; j.limit_err = j.cum_angle - lo; m = 6; }
            else if (j.cum_angle >= hi) { j.limit = 2; j.limit_err = j.cum_angle - hi; m = 6; }
            else j.limit = 0;
          }
        } else if (j.type == GZ_JOINT_FIXED) {
          m = 6;
        } else if (j.type == GZ_JOINT_BALL) {
          m = 3;
        } else if (j.type == GZ_JOINT_SCREW) {
          // dxJointScrew::getInfo1 (screw.cpp:55-110): five rows, a sixth at a stop of the rotation
          m = 5;
          const double lo = j.param[GZ_JP_LOSTOP], hi = j.param[GZ_JP_HISTOP];
          if (lo <= hi) {
            const double angle = hinge_angle(B, j);
            j.cum_angle = shortest_angular_distance_update(j.cum_angle, angle);
            if (j.cum_angle <= lo) { j.limit = 1; j.limit_err = j.cum_angle - lo; m = 6; }
            else if (j.cum_angle >= hi) { j.limit = 2; j.limit_err = j.cum_angle - hi; m = 6; }
            else j.limit = 0;
          }
        } else if (j.type == GZ_JOINT_GEARBOX) {
          m = 1;   // dxJointGearbox::getInfo1 (gearbox.cpp:54-59)
        } else if (j.type == GZ_JOINT_HINGE2) {
          // dxJointHinge2::getInfo1 (hinge2.cpp:108-127)
          m = 4;
          const double lo = j.param[GZ_JP_LOSTOP], hi = j.param[GZ_JP_HISTOP];
          j.limit = 0;
          if ((lo >= -GZ_PI || hi <= GZ_PI) && lo <= hi) {
            const double angle = hinge2_angle1(B, j);
            if (angle <= lo) { j.limit = 1; j.limit_err = angle - lo; }
            else if (angle >= hi) { j.limit = 2; j.limit_err = angle - hi; }
          }
          if (j.limit || j.param[GZ_JP_FMAX] > 0) m++;
          if (joint_ext(B, j)[GZ_JP_FMAX2] > 0) m++;
        } else if (j.type == GZ_JOINT_UNIVERSAL) {
          // dxJointUniversal::getInfo1 (universal.cpp:268-307)
          m = 4;
          const double lo1 = j.param[GZ_JP_LOSTOP], hi1 = j.param[GZ_JP_HISTOP];
          const double lo2 = joint_ext(B, j)[GZ_JP_LOSTOP2], hi2 = joint_ext(B, j)[GZ_JP_HISTOP2];
          const bool limiting1 = lo1 <= hi1, limiting2 = lo2 <= hi2;
          j.limit = 0; j.limit2 = 0;
          if (limiting1 || limiting2) {
            double angle1, angle2;
            universal_angles(B, j, angle1, angle2);
            j.cum_angle = shortest_angular_distance_update(j.cum_angle, angle1);
            j.cum_angle2 = shortest_angular_distance_update(j.cum_angle2, angle2);
            if (limiting1) {
              if (j.cum_angle <= lo1) { j.limit = 1; j.limit_err = j.cum_angle - lo1; }
              else if (j.cum_angle >= hi1) { j.limit = 2; j.limit_err = j.cum_angle - hi1; }
            }
            if (limiting2) {
              if (j.cum_angle2 <= lo2) { j.limit2 = 1; j.limit_err2 = j.cum_angle2 - lo2; }
              else if (j.cum_angle2 >= hi2) { j.limit2 = 2; j.limit_err2 = j.cum_angle2 - hi2; }
            }
          }
          if (j.limit || j.param[GZ_JP_FMAX] > 0) m++;
          if (j.limit2 || joint_ext(B, j)[GZ_JP_FMAX2] > 0) m++;
        } else if (j.type == GZ_JOINT_SLIDER) {
          // dxJointSlider::getInfo1 (slider.cpp:118-148)
          m = (j.param[GZ_JP_FMAX] > 0) ? 6 : 5;
          const double lo = j.param[GZ_JP_LOSTOP], hi = j.param[GZ_JP_HISTOP];
          j.limit = 0;
          if ((lo > -GZ_INF || hi < GZ_INF) && lo <= hi) {
            const double pos = slider_position(B, j);
            if (pos <= lo) { j.limit = 1; j.limit_err = pos - lo; m = 6; }
            else if (pos >= hi) { j.limit = 2; j.limit_err = pos - hi; m = 6; }
          }
        }
      }
      j.m = m;
    }
    int v = m;
    for (int d = 1; d < 32; d <<= 1) { int t = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += t; }
    if (lane == 31) warp_sums[warp] = v;
    __syncthreads();
    if (warp == 0) {
      int s = (lane < nwarp) ? warp_sums[lane] : 0;
      for (int d = 1; d < 32; d <<= 1) { int t = __shfl_up_sync(0xffffffffu, s, d); if (lane >= d) s += t; }
      warp_sums[lane] = s;
    }
    __syncthreads();
    if (k < wi.nj) B.joint[wi.joint0 + k].row0 = carry + (v - m) + (warp > 0 ? warp_sums[warp - 1] : 0);
    int total = warp_sums[nwarp - 1];
    __syncthreads();
    if (threadIdx.x == 0) carry += total;
    __syncthreads();
  }
  const int mj = carry;
  for (int c = threadIdx.x; c < ws.n_contacts; c += blockDim.x) B.contact[wi.contact0 + c].row0 = mj + wi.rpc * c;
  if (threadIdx.x == 0) {
    int m = mj + wi.rpc * ws.n_contacts;
    if (m > wi.row_cap) { ws.overflow |= 4; m = 0; }
    ws.m = m;
    ws.n_fric = 0;   // reused by k_pgs as a cycle counter (instrumentation)
    ws.pad[0] = 0; ws.pad[1] = 0;
    atomicAdd(&B.acc[0], (unsigned long long)m);
    atomicAdd(&B.acc[1], (unsigned long long)(m ? 2 * ws.n_contacts : 0));
    atomicAdd(&B.acc[2], (unsigned long long)ws.n_contacts);
    atomicAdd(&B.acc[3], 1ull);
  }
}

// One CTA per world: auto-disable + islands, the body prologue of its bodies, row counts and offsets - three
// phases that were three launches (k_island, k_bodyprep, k_info1), separated by block barriers.
// (launch bounds: at most 64 registers - the rare joint types' angle code must not cost the common path its occupancy)
__global__ void __launch_bounds__(1024) k_assemble1(DevBatch B) {
  island_world(B, blockIdx.x);
  __syncthreads();
  const WorldInfo wi = B.world[blockIdx.x];
  for (int i = threadIdx.x; i < wi.nb; i += blockDim.x) bodyprep_body(B, wi.body0 + i);
  __syncthreads();
  info1_world(B, blockIdx.x);
}

}  // namespace gz
