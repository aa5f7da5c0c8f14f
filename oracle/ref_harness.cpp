// ref_harness.cpp — ORACLE / TEST INFRASTRUCTURE, never linked into the product.
//
// Drives the reference's OWN compiled opende (oracle/_ref/liboderef*.so, built by
// oracle/Makefile from /root/reference/deps/opende in place) on worlds described
// by the same flat gz_scene arrays libgzcuda consumes.  Everything numerical on
// the hot path is the reference's code: dSpaceCollide / dCollide / the contact
// joint / dxQuickStepper / PGS_LCP / dxUpdateBodies / dxStepBody.
//
// What is RESTATED here (gzserver itself cannot be built in this image: no
// Boost/protobuf/sdformat), following the reference line by line:
//   * ODEPhysics::ODEPhysics / Load world set-up   gazebo/physics/ode/ODEPhysics.cc:140-243
//   * ODELink::Init / UpdateSurface / UpdateMass    gazebo/physics/ode/ODELink.cc:63-134,357-431
//   * ODEModel space layout (one dSimpleSpace per model) gazebo/physics/ode/ODEModel.cc:23-27
//   * ODEPhysics::UpdateCollision                   gazebo/physics/ode/ODEPhysics.cc:366-404
//   * ODEPhysics::CollisionCallback                 gazebo/physics/ode/ODEPhysics.cc:948-1008
//   * ODEPhysics::Collide                           gazebo/physics/ode/ODEPhysics.cc:1012-1331
//   * ODEPhysics::UpdatePhysics                     gazebo/physics/ode/ODEPhysics.cc:407-453
//
// Two stepping modes (SURVEY.md §8c "Order control"):
//   mode 0 "natural"  : contacts attached in dSpaceCollide callback order, then the
//                       unmodified dWorldQuickStep (its own island DFS and row order).
//                       This is what gzserver does; used for the timed CPU baseline.
//   mode 1 "canonical": deterministic row-order mode.  Pairs are canonicalised to
//                       (lower geom index, higher geom index) and sorted; contact joints
//                       are attached in that order; auto-disable is the reference's
//                       dInternalHandleAutoDisabling; islands are connected components
//                       ordered by lowest body index with bodies in index order and joints
//                       in (persistent by index, then contacts by creation) order; each
//                       island is stepped by the reference's exported dxQuickStepper
//                       (deps/opende/src/quickstep.h:31).  No reference arithmetic is
//                       touched; only the ORDER of inputs is fixed.  The GPU path
//                       implements exactly this order.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <csignal>
#include <execinfo.h>
#include <unistd.h>
#include <cstring>
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>
#include <chrono>
#include <atomic>

#include <gazebo/ode/ode.h>
#include "config.h"
#include "objects.h"
#include "joints/joints.h"
#include "collision_kernel.h"
#include "util.h"
#include "quickstep.h"

#include "gzcuda_scene.h"

#define GZ_SENSOR_COLLIDE 0x00000002u  // gazebo/physics/PhysicsTypes.hh:264
#define MAX_COLLIDE_RETURNS 250         // gazebo/physics/ode/ODEPhysicsPrivate.hh
#define MAX_CONTACT_JOINTS 250

namespace {

struct Surface {
  double s[GZ_SF_COUNT];
  double fdir1[3];
  unsigned collide_bitmask;
  int max_contacts;
};

struct ContactRec { int g1, g2; double pos[3], normal[3], depth; };

struct HingePD { double amp, freq, phase, kp, kd, tmax; };

struct RefWorld {
  dWorldID world = nullptr;
  dSpaceID top = nullptr;
  dJointGroupID contactGroup = nullptr;
  std::vector<dSpaceID> spaces;
  std::vector<dBodyID> bodies;
  std::vector<dGeomID> geoms;
  std::vector<Surface> surf;
  std::vector<dJointID> joints;
  std::vector<int> joint_type;
  std::vector<HingePD> pd;
  std::vector<double> vel_cmd;                // per joint, NaN = none (gzref_set_joint_velocity_cmd)
  std::vector<dJointID> contact_joints;       // created this step, creation order
  bool feedback = false;                      // gzref_enable_feedback: dJointSetFeedback on joints and contacts
  std::vector<dJointFeedback> joint_fb, contact_fb;
  std::vector<std::pair<int,int>> pairs;      // colliders of this step
  std::vector<ContactRec> contacts;           // contacts of this step
  dContactGeom contactCollisions[MAX_COLLIDE_RETURNS];
  int indices[MAX_CONTACT_JOINTS];
  long step_index = 0;
  // stats of the last step (mode 1: accumulated over islands, see get_stats)
  double rms_residual[4] = {0, 0, 0, 0};
  int num_contacts = 0;
};

struct RefBatch;
// Persistent host threads for gzref_step_mt (every thread calls dAllocateODEDataForThread
// ONCE, ODEPhysics.cc:149,362): created at the first multi-threaded step, parked on a condition variable between
// calls, so that a short timed window measures stepping and not thread start-up.
struct StepPool {
  std::vector<std::thread> th;
  std::mutex mu;
  std::condition_variable cv_go, cv_done;
  long gen = 0;
  int nsteps = 0, running = 0;
  bool quit = false;
  std::atomic<size_t> next{0};
};

struct RefBatch {
  gz_params p;
  int mode;
  std::vector<RefWorld*> worlds;
  std::vector<int> body_off, geom_off, joint_off;
  StepPool *pool = nullptr;
};

bool g_ode_inited = false;

void segv_handler(int) {
  void *bt[64];
  int n = backtrace(bt, 64);
  backtrace_symbols_fd(bt, n, 2);
  _exit(139);
}

void body_moved_cb(dBodyID) {}      // ODELink::MoveCallback: copies the pose to Gazebo's dirty list
void body_disabled_cb(dBodyID) {}   // ODELink::DisabledCallback: empty in the reference

int geom_index(dGeomID g) { return (int)(intptr_t)dGeomGetData(g); }

// ---- ODEPhysics::CollisionCallback (ODEPhysics.cc:948-1008) -----------------
void near_callback(void *data, dGeomID o1, dGeomID o2) {
  RefWorld *w = static_cast<RefWorld *>(data);
  dBodyID b1 = dGeomGetBody(o1);
  dBodyID b2 = dGeomGetBody(o2);
  if (b1 && b2 && dAreConnectedExcluding(b1, b2, dJointTypeContact)) return;
  if (dGeomIsSpace(o1) || dGeomIsSpace(o2)) {
    dSpaceCollide2(o1, o2, data, &near_callback);
    return;
  }
  // "Exit if both bodies are not enabled" (no sensor geoms, no contact subscribers here)
  if (dGeomGetCategoryBits(o1) != GZ_SENSOR_COLLIDE &&
      dGeomGetCategoryBits(o2) != GZ_SENSOR_COLLIDE &&
      ((b1 && b2 && !dBodyIsEnabled(b1) && !dBodyIsEnabled(b2)) ||
       (!b2 && b1 && !dBodyIsEnabled(b1)) ||
       (!b1 && b2 && !dBodyIsEnabled(b2))))
    return;
  w->pairs.push_back(std::make_pair(geom_index(o1), geom_index(o2)));  // AddCollider
}

// ---- ODEPhysics::Collide (ODEPhysics.cc:1012-1331) --------------------------
void collide_pair(RefBatch *B, RefWorld *w, int i1, int i2) {
  const Surface &s1 = w->surf[i1];
  const Surface &s2 = w->surf[i2];
  if ((s1.collide_bitmask & s2.collide_bitmask) == 0) return;

  unsigned int numc = 0;
  dContact contact;
  memset(&contact, 0, sizeof(contact));

  unsigned int maxCollide = MAX_CONTACT_JOINTS;
  if (B->p.max_contacts > 0 && B->p.max_contacts < MAX_CONTACT_JOINTS)
    maxCollide = B->p.max_contacts;
  if ((unsigned)s1.max_contacts < maxCollide) maxCollide = s1.max_contacts;
  if ((unsigned)s2.max_contacts < maxCollide) maxCollide = s2.max_contacts;

  dContactGeom *cc = w->contactCollisions;
  numc = dCollide(w->geoms[i1], w->geoms[i2], MAX_COLLIDE_RETURNS, cc, sizeof(cc[0]));
  if (numc == 0) return;

  for (int i = 0; i < MAX_CONTACT_JOINTS; i++) w->indices[i] = i;

  if (maxCollide > 0 && numc > maxCollide) {
    double max = cc[maxCollide - 1].depth;
    for (unsigned int i = maxCollide; i < numc; ++i) {
      if (cc[i].depth > max) {
        max = cc[i].depth;
        w->indices[maxCollide - 1] = i;
      }
    }
    numc = maxCollide;
  }

  contact.surface.mode = dContactBounce | dContactMu2 | dContactSoftERP |
                         dContactSoftCFM | dContactApprox1 | dContactApprox3 |
                         dContactSlip1 | dContactSlip2;

  const double h = B->p.step_size;
  double kp = 1.0 / (1.0 / s1.s[GZ_SF_KP] + 1.0 / s2.s[GZ_SF_KP]);
  double kd = s1.s[GZ_SF_KD] + s2.s[GZ_SF_KD];
  contact.surface.soft_erp = (h * kp) / (h * kp + kd);
  contact.surface.soft_cfm = 1.0 / (h * kp + kd);

  // fdir1, AS WRITTEN in ODEPhysics.cc:1120-1161: fd is surface 1's direction1, rotated to the world frame by
  // collision 1's pose.  Surface 2's direction (fd2) is rotated in the reference when surface 1 has none or the
  // larger mu1, but it is never assigned to fd: it has NO effect on the contact, and dContactFDir1 is set only
  // when surface 1 has a direction.  (Gazebo rotates with the collision's Pose3d quaternion; the ODE geom rotation
  // matrix is used here - equal up to rounding, gzserver cannot be built in this image; both sides of the parity
  // test use this form.)
  double fd[3] = {s1.fdir1[0], s1.fdir1[1], s1.fdir1[2]};
  const bool fdset = (fd[0] != 0 || fd[1] != 0 || fd[2] != 0);
  if (fdset) {
    // (a plane is not placeable in ODE and has no pose in the scene format: its collision sits at the identity)
    static const dReal ident[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    const dReal *R = dGeomGetClass(w->geoms[i1]) == dPlaneClass ? ident : dGeomGetRotation(w->geoms[i1]);
    dVector3 v = {fd[0], fd[1], fd[2], 0}, r;
    dMultiply0_331(r, R, v);
    contact.surface.mode |= dContactFDir1;
    contact.fdir1[0] = r[0]; contact.fdir1[1] = r[1]; contact.fdir1[2] = r[2];
  }

  contact.surface.mu = std::min(s1.s[GZ_SF_MU], s2.s[GZ_SF_MU]);
  contact.surface.mu2 = std::min(s1.s[GZ_SF_MU2], s2.s[GZ_SF_MU2]);
  contact.surface.mu3 = std::min(s1.s[GZ_SF_MU3], s2.s[GZ_SF_MU3]);   // (the scene format's default is 0 = off; Gazebo's MuTorsion default is 1)
  contact.surface.slip1 = s1.s[GZ_SF_SLIP1] + s2.s[GZ_SF_SLIP1];
  contact.surface.slip2 = s1.s[GZ_SF_SLIP2] + s2.s[GZ_SF_SLIP2];
  contact.surface.slip3 = s1.s[GZ_SF_SLIP3] + s2.s[GZ_SF_SLIP3];
  contact.surface.slip1 *= numc;
  contact.surface.slip2 *= numc;
  contact.surface.slip3 *= numc;
  // ODEPhysics.cc:1186-1251: patch radius, curvatures, use_patch_radius; torsional friction (dContactMu3) when mu3 > 0
  // and the radius in use is > 0; Hertzian contact (dContactEM) when both surfaces have an elastic modulus
  contact.surface.patch_radius = std::max(s1.s[GZ_SF_PATCH_RADIUS], s2.s[GZ_SF_PATCH_RADIUS]);
  {
    double curv1 = 0, curv2 = 0;
    if (s1.s[GZ_SF_SURFACE_RADIUS] > 0) curv1 = 1 / s1.s[GZ_SF_SURFACE_RADIUS];
    if (s2.s[GZ_SF_SURFACE_RADIUS] > 0) curv2 = 1 / s2.s[GZ_SF_SURFACE_RADIUS];
    const double curvSum = curv1 + curv2;
    contact.surface.surface_radius = 0;
    if (curvSum > 0) contact.surface.surface_radius = 1 / curvSum;
    contact.surface.use_patch_radius = s1.s[GZ_SF_USE_PATCH_RADIUS] != 0 && s2.s[GZ_SF_USE_PATCH_RADIUS] != 0;
    if (contact.surface.mu3 > 0) {
      if ((contact.surface.use_patch_radius && contact.surface.patch_radius > 0) ||
          (!contact.surface.use_patch_radius && contact.surface.surface_radius > 0)) {
        contact.surface.mode |= dContactMu3;
        if (contact.surface.slip3 > 0) contact.surface.mode |= dContactSlip3;
      }
    }
    const double nu1 = s1.s[GZ_SF_POISSON], nu2 = s2.s[GZ_SF_POISSON];
    const double e1 = s1.s[GZ_SF_ELASTIC_MODULUS], e2 = s2.s[GZ_SF_ELASTIC_MODULUS];
    if (e1 > 0 && e2 > 0) {
      contact.surface.elastic_modulus = 1.0 / ((1.0 - nu1 * nu1) / e1 + (1.0 - nu2 * nu2) / e2);
      if (contact.surface.elastic_modulus > 0.0) contact.surface.mode |= dContactEM;
    }
  }
  contact.surface.bounce = std::min(s1.s[GZ_SF_BOUNCE], s2.s[GZ_SF_BOUNCE]);
  contact.surface.bounce_vel = std::min(s1.s[GZ_SF_BOUNCE_VEL], s2.s[GZ_SF_BOUNCE_VEL]);

  dBodyID b1 = dGeomGetBody(w->geoms[i1]);
  dBodyID b2 = dGeomGetBody(w->geoms[i2]);

  for (unsigned int j = 0; j < numc; ++j) {
    contact.geom = cc[w->indices[j]];
    dJointID cj = dJointCreateContact(w->world, w->contactGroup, &contact);
    dJointAttach(cj, b1, b2);
    w->contact_joints.push_back(cj);
    ContactRec r;
    r.g1 = i1; r.g2 = i2;
    for (int k = 0; k < 3; k++) { r.pos[k] = contact.geom.pos[k]; r.normal[k] = contact.geom.normal[k]; }
    r.depth = contact.geom.depth;
    w->contacts.push_back(r);
  }
}

std::atomic<long long> g_phase_ns[2];   // [0] UpdateCollision, [1] UpdatePhysics: thread-seconds since the last reset

// ---- ODEPhysics::UpdateCollision (ODEPhysics.cc:366-404) --------------------
void update_collision(RefBatch *B, RefWorld *w) {
  dJointGroupEmpty(w->contactGroup);
  w->contact_joints.clear();
  w->pairs.clear();
  w->contacts.clear();
  dSpaceCollide(w->top, w, &near_callback);
  if (B->mode == 1) {
    for (auto &pr : w->pairs)
      if (pr.first > pr.second) std::swap(pr.first, pr.second);
    std::sort(w->pairs.begin(), w->pairs.end());
  }
  for (size_t i = 0; i < w->pairs.size(); ++i)
    collide_pair(B, w, w->pairs[i].first, w->pairs[i].second);
  if (w->feedback) {   // ODEPhysics.cc:1319: dJointSetFeedback(contact_joint, &contactFeedback->feedbacks[j])
    w->contact_fb.assign(w->contact_joints.size(), dJointFeedback());
    for (size_t i = 0; i < w->contact_joints.size(); ++i) {
      memset(&w->contact_fb[i], 0, sizeof(dJointFeedback));
      dJointSetFeedback(w->contact_joints[i], &w->contact_fb[i]);
    }
  }
}

// ---- plugin stand-in: PD-tracked sinusoid on hinge joints -------------------
// Same place in the step as a ModelPlugin on Events::worldUpdateBegin
// (gazebo/physics/World.cc:746) calling Joint::SetForce -> dJointAddHingeTorque
// (gazebo/physics/ode/ODEHingeJoint.cc:159-162).
void apply_controllers(RefBatch *B, RefWorld *w) {
  const double t = (double)w->step_index * B->p.step_size;
  for (size_t j = 0; j < w->joints.size(); ++j) {
    const HingePD &c = w->pd[j];
    if (!(c.tmax > 0) || w->joint_type[j] != GZ_JOINT_HINGE) continue;
    double target = c.amp * sin(2.0 * M_PI * c.freq * t + c.phase);
    double angle = dJointGetHingeAngle(w->joints[j]);
    double rate = dJointGetHingeAngleRate(w->joints[j]);
    double tau = c.kp * (target - angle) - c.kd * rate;
    if (tau > c.tmax) tau = c.tmax;
    if (tau < -c.tmax) tau = -c.tmax;
    dJointAddHingeTorque(w->joints[j], tau);
  }
}

// ---- plugin stand-in: Joint::SetVelocity on hinge joints every update ---------
// Joint::SetVelocityMaximal (gazebo/physics/Joint.cc:828-894) as this fork's DiffDrivePlugin::OnUpdate reaches it
// (plugins/DiffDrivePlugin.cc:100-122, ODEHingeJoint::SetVelocity ODEHingeJoint.cc:153-156): the child link (ODE body 1)
// takes the parent's angular velocity plus vel along the joint axis and the linear velocity of a point carried by the
// anchor.  Link frame = centre of mass in this scene format; anchor offsets by ODE's matrix arithmetic.
void apply_velocity_commands(RefWorld *w) {
  for (size_t j = 0; j < w->joints.size() && j < w->vel_cmd.size(); ++j) {
    const double vel = w->vel_cmd[j];
    if (!(vel == vel) || w->joint_type[j] != GZ_JOINT_HINGE) continue;
    dxJointHinge *hj = static_cast<dxJointHinge *>(w->joints[j]);
    dxBody *c = hj->node[0].body, *p = hj->node[1].body;
    double ang[3] = {0.0, 0.0, 0.0}, lin[3] = {0.0, 0.0, 0.0};
    if (p) {
      dVector3 off, cr;
      dMultiply0_331(off, p->posr.R, hj->anchor2);
      dCalcVectorCross3(cr, p->avel, off);
      for (int i = 0; i < 3; i++) { ang[i] = p->avel[i]; lin[i] = p->lvel[i] + cr[i]; }
    }
    dVector3 axis, a, jw, off, cr, angv;
    dMultiply0_331(axis, c->posr.R, hj->axis1);
    for (int i = 0; i < 3; i++) ang[i] += vel * axis[i];
    dMultiply0_331(a, c->posr.R, hj->anchor1);
    for (int i = 0; i < 3; i++) { jw[i] = c->posr.pos[i] + a[i]; off[i] = c->posr.pos[i] - jw[i]; angv[i] = ang[i]; }
    dCalcVectorCross3(cr, angv, off);
    dBodySetAngularVel(c, ang[0], ang[1], ang[2]);
    dBodySetLinearVel(c, lin[0] + cr[0], lin[1] + cr[1], lin[2] + cr[2]);
  }
}

// ---- canonical-order step: islands by lowest body index ---------------------
int uf_find(std::vector<int> &p, int x) {
  while (p[x] != x) { p[x] = p[p[x]]; x = p[x]; }
  return x;
}

void canonical_quickstep(RefBatch *B, RefWorld *w) {
  const double h = B->p.step_size;
  dxWorld *world = w->world;
  dInternalHandleAutoDisabling(world, h);  // util.cpp:116-252, as BuildIslands does first

  const int nb = (int)w->bodies.size();
  for (int i = 0; i < nb; i++) w->bodies[i]->tag = i;  // scratch: world-local index

  std::vector<dxJoint *> alljoints;
  for (dJointID j : w->joints) alljoints.push_back(j);
  for (dJointID j : w->contact_joints) alljoints.push_back(j);

  std::vector<int> parent(nb);
  for (int i = 0; i < nb; i++) parent[i] = i;
  std::vector<char> jen(alljoints.size(), 0);
  for (size_t k = 0; k < alljoints.size(); k++) {
    dxJoint *j = alljoints[k];
    if (!j->node[0].body) continue;      // unattached: in no island (util.cpp:563-565)
    if (!j->isEnabled()) continue;
    jen[k] = 1;
    if (j->node[1].body) {
      int a = uf_find(parent, j->node[0].body->tag), b = uf_find(parent, j->node[1].body->tag);
      if (a != b) parent[std::max(a, b)] = std::min(a, b);   // root = lowest index
    }
  }
  std::vector<char> active(nb, 0);
  for (int i = 0; i < nb; i++)
    if (!(w->bodies[i]->flags & dxBodyDisabled)) active[uf_find(parent, i)] = 1;
  std::vector<int> root(nb);
  for (int i = 0; i < nb; i++) root[i] = uf_find(parent, i);
  // island of every joint, fixed BEFORE stepping: dxQuickStepper overwrites dxBody::tag
  std::vector<int> jroot(alljoints.size(), -1);
  for (size_t k = 0; k < alljoints.size(); k++)
    if (jen[k]) jroot[k] = root[alljoints[k]->node[0].body->tag];

  // bodies reached from an enabled body are re-enabled (util.cpp:480-484)
  for (int i = 0; i < nb; i++)
    if (active[root[i]]) w->bodies[i]->flags &= ~dxBodyDisabled;

  std::vector<dxBody *> ib;
  std::vector<dxJoint *> ij;
  dxWorldProcessContext *ctx = nullptr;
  for (int r = 0; r < nb; r++) {
    if (root[r] != r || !active[r]) continue;
    ib.clear(); ij.clear();
    for (int i = r; i < nb; i++) if (root[i] == r) ib.push_back(w->bodies[i]);
    for (size_t k = 0; k < alljoints.size(); k++)
      if (jroot[k] == r) ij.push_back(alljoints[k]);
    size_t req = dxEstimateQuickStepMemoryRequirements(ib.data(), (int)ib.size(), ij.data(), (int)ij.size());
    ctx = dxReallocateTemporayWorldProcessContext(ctx, req + 4096, NULL, NULL);
    world->qs.num_contacts = 0;
    for (int k = 0; k < 4; k++) world->qs.rms_constraint_residual[k] = 0;
    dxQuickStepper(ctx, world, ib.data(), (int)ib.size(), ij.data(), (int)ij.size(), h);
    ctx->CleanupContext();
    // The reference overwrites qs.rms_* per island (last island wins,
    // quickstep_pgs_lcp.cpp:831-855); num_contacts is summed here so a whole-world
    // figure can be compared.
    w->num_contacts += world->qs.num_contacts;
  }
  if (ctx) dxFreeWorldProcessContext(ctx);
  for (int k = 0; k < 4; k++) w->rms_residual[k] = world->qs.rms_constraint_residual[k];
}

void step_world(RefBatch *B, RefWorld *w) {
  static const bool trace = getenv("GZREF_TRACE") != nullptr;
  if (trace) fprintf(stderr, "step %ld: control\n", w->step_index);
  apply_velocity_commands(w);
  apply_controllers(B, w);
  if (trace) fprintf(stderr, "  collide\n");
  // the phase split of the CPU path (the reference's own -DTIMING split is #undef'd in quickstep_util.h:37): host
  // clock around UpdateCollision and UpdatePhysics, summed over the worlds and steps of the calling thread
  const auto t0 = std::chrono::steady_clock::now();
  update_collision(B, w);
  const auto t1 = std::chrono::steady_clock::now();
  if (trace) fprintf(stderr, "  step (pairs %zu contacts %zu)\n", w->pairs.size(), w->contacts.size());
  w->num_contacts = 0;
  if (B->mode == 1) {
    canonical_quickstep(B, w);
  } else {
    dWorldQuickStep(w->world, B->p.step_size);   // ODEPhysics::UpdatePhysics, ODEPhysics.cc:416
    for (int k = 0; k < 4; k++) w->rms_residual[k] = w->world->qs.rms_constraint_residual[k];
    w->num_contacts = w->world->qs.num_contacts;
  }
  const auto t2 = std::chrono::steady_clock::now();
  g_phase_ns[0].fetch_add((long long)std::chrono::duration_cast<std::chrono::nanoseconds>(t1 - t0).count(), std::memory_order_relaxed);
  g_phase_ns[1].fetch_add((long long)std::chrono::duration_cast<std::chrono::nanoseconds>(t2 - t1).count(), std::memory_order_relaxed);
  w->step_index++;
}

}  // namespace

// ---- the CPU laser: ODEMultiRayShape::UpdateRays / UpdateCallback (gazebo/physics/ode/ODEMultiRayShape.cc:82-205) and
// ODERayShape (ODERayShape.cc:36-60, 95-122) restated through the ODE API; the ray colliders are the reference's
// (deps/opende/src/ray.cpp) ----
struct RayScan {
  dSpaceID super = nullptr, raySpace = nullptr;
  std::vector<dGeomID> rays;
  std::vector<double> length;
  std::vector<int> hit;
};

void ray_callback(void *data, dGeomID o1, dGeomID o2) {
  RayScan *self = static_cast<RayScan *>(data);
  if (dGeomIsSpace(o1) || dGeomIsSpace(o2)) {
    if (dGeomGetSpace(o1) == self->super || dGeomGetSpace(o2) == self->super) dSpaceCollide2(o1, o2, self, &ray_callback);
    if (dGeomGetSpace(o1) == self->raySpace || dGeomGetSpace(o2) == self->raySpace) dSpaceCollide2(o1, o2, self, &ray_callback);
    return;
  }
  dGeomID ray = nullptr, other = nullptr;
  if (dGeomGetClass(o1) == dRayClass) { ray = o1; other = o2; }
  else if (dGeomGetClass(o2) == dRayClass) { ray = o2; other = o1; }
  if (!ray || dGeomGetClass(other) == dRayClass) return;
  dGeomRaySetParams(ray, 0, 0);
  dGeomRaySetClosestHit(ray, 1);
  dContactGeom contact;
  const int n = dCollide(o1, o2, 1, &contact, sizeof(contact));
  const int k = (int)(intptr_t)dGeomGetData(ray);
  if (n > 0 && contact.depth < self->length[k]) {   // shape->SetLength(contact.depth): RayShape::contactLen only, the ODE ray keeps its length
    self->length[k] = contact.depth;
    self->hit[k] = geom_index(other);
  }
}

extern "C" {

// n rays of ONE world against its current geometry (the rays ride on no body: ODERayShape never calls dGeomSetBody,
// so a laser sees its own link's collisions too).  out_range = the ray's length after the scan, out_geom = the
// world-local geom index of the last shortening hit or -1.
int gzref_ray_query(void *h, int world, int n, const double *start, const double *dir, const double *length,
                    double *out_range, int *out_geom) {
  RefBatch *B = static_cast<RefBatch *>(h);
  if (!B || world < 0 || world >= (int)B->worlds.size()) return -1;
  RefWorld *w = B->worlds[world];
  RayScan S;
  S.super = dSimpleSpaceCreate(0);
  S.raySpace = dSimpleSpaceCreate(S.super);
  dGeomSetCategoryBits((dGeomID)S.raySpace, 0x00000002);       // GZ_SENSOR_COLLIDE
  dGeomSetCollideBits((dGeomID)S.raySpace, ~0x00000002u);
  S.length.assign(length, length + n);
  S.hit.assign(n, -1);
  for (int k = 0; k < n; k++) {
    dGeomID r = dCreateRay(S.raySpace, length[k]);
    dGeomSetCategoryBits(r, 0x00000002);
    dGeomSetCollideBits(r, ~0x00000002u);
    dGeomSetData(r, (void *)(intptr_t)k);
    S.rays.push_back(r);
  }
  for (int k = 0; k < n; k++) {
    dGeomRaySet(S.rays[k], start[3 * k], start[3 * k + 1], start[3 * k + 2], dir[3 * k], dir[3 * k + 1], dir[3 * k + 2]);
    dGeomRaySetLength(S.rays[k], length[k]);
  }
  dSpaceCollide2((dGeomID)S.super, (dGeomID)w->top, &S, &ray_callback);
  for (int k = 0; k < n; k++) { out_range[k] = S.length[k]; out_geom[k] = S.hit[k]; }
  for (dGeomID r : S.rays) dGeomDestroy(r);
  dSpaceDestroy(S.raySpace);
  dSpaceDestroy(S.super);
  return 0;
}

// Thread-nanoseconds spent in UpdateCollision / UpdatePhysics since the last call with reset != 0.
void gzref_phase_ns(long long out[2], int reset) {
  out[0] = g_phase_ns[0].load(); out[1] = g_phase_ns[1].load();
  if (reset) { g_phase_ns[0] = 0; g_phase_ns[1] = 0; }
}

void *gzref_create(const gz_scene *sc, const gz_params *p, int mode) {
  if (!g_ode_inited) { dInitODE2(0); g_ode_inited = true; if (getenv("GZREF_TRACE")) signal(SIGSEGV, segv_handler); }
  dAllocateODEDataForThread(dAllocateMaskAll);
  RefBatch *B = new RefBatch;
  B->p = *p;
  B->mode = mode;
  B->body_off.assign(sc->body_off, sc->body_off + sc->n_worlds + 1);
  B->geom_off.assign(sc->geom_off, sc->geom_off + sc->n_worlds + 1);
  B->joint_off.assign(sc->joint_off, sc->joint_off + sc->n_worlds + 1);
  for (int wi = 0; wi < sc->n_worlds; wi++) {
    RefWorld *w = new RefWorld;
    B->worlds.push_back(w);
    // ODEPhysics::ODEPhysics (ODEPhysics.cc:140-163)
    w->world = dWorldCreate();
    w->top = dHashSpaceCreate(0);
    dHashSpaceSetLevels(w->top, -2, 8);
    w->contactGroup = dJointGroupCreate(0);
    // ODEPhysics::Load (ODEPhysics.cc:175-243)
    dWorldSetQuickStepInertiaRatioReduction(w->world, p->dynamic_inertia != 0);  // SDF default false (physics.sdf:199)
    dWorldSetDamping(w->world, 0.0001, 0.0001);
    dWorldSetContactMaxCorrectingVel(w->world, p->contact_max_correcting_vel);
    dWorldSetContactSurfaceLayer(w->world, p->contact_surface_layer);
    dWorldSetAutoDisableFlag(w->world, 1);
    w->world->adis.idle_time = p->adis_idle_time;
    w->world->adis.linear_average_threshold = p->adis_linear_threshold_sq;
    w->world->adis.angular_average_threshold = p->adis_angular_threshold_sq;
    w->world->adis.idle_steps = p->adis_idle_steps;
    w->world->dampingp.linear_threshold = p->damping_linear_threshold_sq;
    w->world->dampingp.angular_threshold = p->damping_angular_threshold_sq;
    dWorldSetGravity(w->world, p->gravity[0], p->gravity[1], p->gravity[2]);
    dWorldSetCFM(w->world, p->cfm);
    dWorldSetERP(w->world, p->erp);
    dWorldSetQuickStepNumIterations(w->world, p->iters);
    dWorldSetQuickStepW(w->world, p->sor_w);
    w->world->qs.friction_iterations = p->friction_iters;
    w->world->qs.warm_start = p->warm_start;
    w->world->qs.smooth_contacts = p->smooth_contacts;
    w->world->qs.contact_sor_scale = p->contact_sor_scale;
    w->world->qs.friction_model = p->friction_model == 2 ? box_friction : pyramid_friction;

    // bodies: ODELink::Init / UpdateMass / UpdateSurface
    for (int b = sc->body_off[wi]; b < sc->body_off[wi + 1]; b++) {
      dBodyID body = dBodyCreate(w->world);
      const int fl = sc->body_flags[b];
      if (fl & GZ_BODY_AUTO_DISABLE) {
        dBodySetAutoDisableDefaults(body);
        dBodySetAutoDisableFlag(body, 1);
      } else {
        dBodySetAutoDisableFlag(body, 0);
      }
      dBodySetGravityMode(body, (fl & GZ_BODY_NO_GRAVITY) ? 0 : 1);
      if (fl & GZ_BODY_NO_GYRO) dBodySetGyroscopicMode(body, 0);
      dBodySetLinearDamping(body, sc->body_damping[2 * b + 0]);
      dBodySetAngularDamping(body, sc->body_damping[2 * b + 1]);
      dMass m;
      dMassSetZero(&m);
      const double *I = sc->body_inertia + 6 * b;
      dMassSetParameters(&m, sc->body_mass[b], 0, 0, 0, I[0], I[1], I[2], I[3], I[4], I[5]);
      dBodySetMass(body, &m);
      dBodySetPosition(body, sc->body_pos[3 * b], sc->body_pos[3 * b + 1], sc->body_pos[3 * b + 2]);
      dQuaternion q = {sc->body_quat[4 * b], sc->body_quat[4 * b + 1], sc->body_quat[4 * b + 2], sc->body_quat[4 * b + 3]};
      dBodySetQuaternion(body, q);
      dBodySetLinearVel(body, sc->body_lvel[3 * b], sc->body_lvel[3 * b + 1], sc->body_lvel[3 * b + 2]);
      dBodySetAngularVel(body, sc->body_avel[3 * b], sc->body_avel[3 * b + 1], sc->body_avel[3 * b + 2]);
      dBodySetMaxVel(body, sc->body_contactp[2 * b + 0]);
      dBodySetMinDepth(body, sc->body_contactp[2 * b + 1]);
      // ODELink::Init (ODELink.cc:121-125) always installs both callbacks; dBodyCreate leaves
      // disabled_callback uninitialised (ode.cpp), so a body that auto-disables would jump
      // through a garbage pointer without this.
      dBodySetMovedCallback(body, body_moved_cb);
      dBodySetDisabledCallback(body, body_disabled_cb);
      if (fl & GZ_BODY_DISABLED) dBodyDisable(body);
      w->bodies.push_back(body);
    }

    // geoms
    int nspaces = 0;
    for (int g = sc->geom_off[wi]; g < sc->geom_off[wi + 1]; g++)
      nspaces = std::max(nspaces, sc->geom_space[g] + 1);
    for (int s = 0; s < nspaces; s++) w->spaces.push_back(dSimpleSpaceCreate(w->top));  // ODEModel.cc:26
    for (int g = sc->geom_off[wi]; g < sc->geom_off[wi + 1]; g++) {
      const double *d = sc->geom_dims + 4 * g;
      dSpaceID sp = w->spaces[sc->geom_space[g]];
      dGeomID geom = nullptr;
      switch (sc->geom_type[g]) {
        case GZ_GEOM_SPHERE: geom = dCreateSphere(sp, d[0]); break;
        case GZ_GEOM_BOX: geom = dCreateBox(sp, d[0], d[1], d[2]); break;
        case GZ_GEOM_CAPSULE: geom = dCreateCapsule(sp, d[0], d[1]); break;
        case GZ_GEOM_CYLINDER: geom = dCreateCylinder(sp, d[0], d[1]); break;
        case GZ_GEOM_PLANE: geom = dCreatePlane(sp, d[0], d[1], d[2], d[3]); break;
        default: return nullptr;
      }
      const int lg = g - sc->geom_off[wi];
      dGeomSetData(geom, (void *)(intptr_t)lg);
      dQuaternion q = {sc->geom_quat[4 * g], sc->geom_quat[4 * g + 1], sc->geom_quat[4 * g + 2], sc->geom_quat[4 * g + 3]};
      if (sc->geom_type[g] != GZ_GEOM_PLANE) {
        if (sc->geom_body[g] >= 0) {
          dGeomSetBody(geom, w->bodies[sc->geom_body[g]]);
          dGeomSetOffsetPosition(geom, sc->geom_pos[3 * g], sc->geom_pos[3 * g + 1], sc->geom_pos[3 * g + 2]);
          dGeomSetOffsetQuaternion(geom, q);
        } else {
          dGeomSetPosition(geom, sc->geom_pos[3 * g], sc->geom_pos[3 * g + 1], sc->geom_pos[3 * g + 2]);
          dGeomSetQuaternion(geom, q);
        }
      }
      // ODECollision::SetCategoryBits/SetCollideBits set the geom AND its space (ODECollision.cc:161-176)
      dGeomSetCategoryBits(geom, sc->geom_category_bits[g]);
      dGeomSetCategoryBits((dGeomID)sp, sc->geom_category_bits[g]);
      dGeomSetCollideBits(geom, sc->geom_collide_bits[g]);
      dGeomSetCollideBits((dGeomID)sp, sc->geom_collide_bits[g]);
      Surface s;
      for (int k = 0; k < GZ_SF_COUNT; k++) s.s[k] = sc->geom_surf[GZ_SF_COUNT * g + k];
      for (int k = 0; k < 3; k++) s.fdir1[k] = sc->geom_fdir1[3 * g + k];
      s.collide_bitmask = sc->geom_collide_bitmask[g];
      s.max_contacts = sc->geom_max_contacts[g];
      w->geoms.push_back(geom);
      w->surf.push_back(s);
    }

    // persistent joints
    for (int j = sc->joint_off[wi]; j < sc->joint_off[wi + 1]; j++) {
      dBodyID b1 = w->bodies[sc->joint_body1[j]];
      dBodyID b2 = sc->joint_body2[j] >= 0 ? w->bodies[sc->joint_body2[j]] : nullptr;
      const double *P = sc->joint_param + GZ_JP_COUNT * j;
      dJointID jt = nullptr;
      if (sc->joint_type[j] == GZ_JOINT_HINGE) {
        jt = dJointCreateHinge(w->world, 0);
        dJointAttach(jt, b1, b2);
        dxJointHinge *hj = (dxJointHinge *)jt;
        for (int k = 0; k < 3; k++) {
          hj->anchor1[k] = sc->joint_anchor1[3 * j + k];
          hj->anchor2[k] = sc->joint_anchor2[3 * j + k];
          hj->axis1[k] = sc->joint_axis1[3 * j + k];
          hj->axis2[k] = sc->joint_axis2[3 * j + k];
        }
        for (int k = 0; k < 4; k++) hj->qrel[k] = sc->joint_qrel[4 * j + k];
        hj->limot.lostop = P[GZ_JP_LOSTOP];
        hj->limot.histop = P[GZ_JP_HISTOP];
        hj->limot.vel = P[GZ_JP_VEL];
        hj->limot.fmax = P[GZ_JP_FMAX];
        hj->limot.fudge_factor = P[GZ_JP_FUDGE];
        hj->limot.bounce = P[GZ_JP_BOUNCE];
        hj->limot.normal_cfm = P[GZ_JP_NORMAL_CFM];
        hj->limot.stop_erp = P[GZ_JP_STOP_ERP];
        hj->limot.stop_cfm = P[GZ_JP_STOP_CFM];
      } else if (sc->joint_type[j] == GZ_JOINT_FIXED) {
        jt = dJointCreateFixed(w->world, 0);
        dJointAttach(jt, b1, b2);
        dxJointFixed *fj = (dxJointFixed *)jt;
        for (int k = 0; k < 3; k++) fj->offset[k] = sc->joint_anchor1[3 * j + k];
        for (int k = 0; k < 4; k++) fj->qrel[k] = sc->joint_qrel[4 * j + k];
      } else if (sc->joint_type[j] == GZ_JOINT_BALL) {
        jt = dJointCreateBall(w->world, 0);
        dJointAttach(jt, b1, b2);
        dxJointBall *bj = (dxJointBall *)jt;
        for (int k = 0; k < 3; k++) {
          bj->anchor1[k] = sc->joint_anchor1[3 * j + k];
          bj->anchor2[k] = sc->joint_anchor2[3 * j + k];
        }
      } else if (sc->joint_type[j] == GZ_JOINT_UNIVERSAL) {
        jt = dJointCreateUniversal(w->world, 0);
        dJointAttach(jt, b1, b2);
        dxJointUniversal *uj = (dxJointUniversal *)jt;
        for (int k = 0; k < 3; k++) {
          uj->anchor1[k] = sc->joint_anchor1[3 * j + k];
          uj->anchor2[k] = sc->joint_anchor2[3 * j + k];
          uj->axis1[k] = sc->joint_axis1[3 * j + k];
          uj->axis2[k] = sc->joint_axis2[3 * j + k];
        }
        for (int k = 0; k < 4; k++) { uj->qrel1[k] = sc->joint_qrel[4 * j + k]; uj->qrel2[k] = P[GZ_JP_QREL2 + k]; }
        uj->limot1.lostop = P[GZ_JP_LOSTOP]; uj->limot1.histop = P[GZ_JP_HISTOP]; uj->limot1.vel = P[GZ_JP_VEL]; uj->limot1.fmax = P[GZ_JP_FMAX];
        uj->limot1.fudge_factor = P[GZ_JP_FUDGE]; uj->limot1.bounce = P[GZ_JP_BOUNCE]; uj->limot1.normal_cfm = P[GZ_JP_NORMAL_CFM];
        uj->limot1.stop_erp = P[GZ_JP_STOP_ERP]; uj->limot1.stop_cfm = P[GZ_JP_STOP_CFM];
        uj->limot2.lostop = P[GZ_JP_LOSTOP2]; uj->limot2.histop = P[GZ_JP_HISTOP2]; uj->limot2.vel = P[GZ_JP_VEL2]; uj->limot2.fmax = P[GZ_JP_FMAX2];
        uj->limot2.fudge_factor = P[GZ_JP_FUDGE2]; uj->limot2.bounce = P[GZ_JP_BOUNCE2]; uj->limot2.normal_cfm = P[GZ_JP_NORMAL_CFM2];
        uj->limot2.stop_erp = P[GZ_JP_STOP_ERP2]; uj->limot2.stop_cfm = P[GZ_JP_STOP_CFM2];
      } else if (sc->joint_type[j] == GZ_JOINT_SCREW) {
        jt = dJointCreateScrew(w->world, 0);
        dJointAttach(jt, b1, b2);
        dxJointScrew *sj = (dxJointScrew *)jt;
        for (int k = 0; k < 3; k++) {
          sj->anchor1[k] = sc->joint_anchor1[3 * j + k]; sj->anchor2[k] = sc->joint_anchor2[3 * j + k];
          sj->axis1[k] = sc->joint_axis1[3 * j + k]; sj->axis2[k] = sc->joint_axis2[3 * j + k];
          sj->offset[k] = P[GZ_JP_OFFSET + k];
        }
        for (int k = 0; k < 4; k++) sj->qrel[k] = sc->joint_qrel[4 * j + k];
        sj->thread_pitch = P[GZ_JP_THREAD_PITCH];
        sj->limot.lostop = P[GZ_JP_LOSTOP]; sj->limot.histop = P[GZ_JP_HISTOP]; sj->limot.vel = P[GZ_JP_VEL]; sj->limot.fmax = P[GZ_JP_FMAX];
        sj->limot.fudge_factor = P[GZ_JP_FUDGE]; sj->limot.bounce = P[GZ_JP_BOUNCE]; sj->limot.normal_cfm = P[GZ_JP_NORMAL_CFM];
        sj->limot.stop_erp = P[GZ_JP_STOP_ERP]; sj->limot.stop_cfm = P[GZ_JP_STOP_CFM];
      } else if (sc->joint_type[j] == GZ_JOINT_GEARBOX) {
        if (!b2) return nullptr;
        jt = dJointCreateGearbox(w->world, 0);
        dJointAttach(jt, b1, b2);
        dxJointGearbox *gj = (dxJointGearbox *)jt;
        for (int k = 0; k < 3; k++) { gj->axis1[k] = sc->joint_axis1[3 * j + k]; gj->axis2[k] = sc->joint_axis2[3 * j + k]; }
        for (int k = 0; k < 4; k++) { gj->qrel1[k] = sc->joint_qrel[4 * j + k]; gj->qrel2[k] = P[GZ_JP_QREL2 + k]; }
        gj->ratio = P[GZ_JP_RATIO];
        gj->refBody1 = (int)P[GZ_JP_REFBODY1] >= 0 ? w->bodies[(int)P[GZ_JP_REFBODY1]] : nullptr;
        gj->refBody2 = (int)P[GZ_JP_REFBODY2] >= 0 ? w->bodies[(int)P[GZ_JP_REFBODY2]] : nullptr;
      } else if (sc->joint_type[j] == GZ_JOINT_HINGE2) {
        if (!b2) return nullptr;
        jt = dJointCreateHinge2(w->world, 0);
        dJointAttach(jt, b1, b2);
        dxJointHinge2 *h2 = (dxJointHinge2 *)jt;
        for (int k = 0; k < 3; k++) {
          h2->anchor1[k] = sc->joint_anchor1[3 * j + k];
          h2->anchor2[k] = sc->joint_anchor2[3 * j + k];
          h2->axis1[k] = sc->joint_axis1[3 * j + k];
          h2->axis2[k] = sc->joint_axis2[3 * j + k];
          h2->v1[k] = P[GZ_JP_V1 + k];      // what dJointSetHinge2Anchor / Axis1 / Axis2 derive (makeV1andV2, getAxisInfo)
          h2->v2[k] = P[GZ_JP_V2 + k];
        }
        h2->c0 = P[GZ_JP_C0];
        h2->s0 = P[GZ_JP_S0];
        h2->susp_erp = P[GZ_JP_SUSP_ERP];
        h2->susp_cfm = P[GZ_JP_SUSP_CFM];
        h2->limot1.lostop = P[GZ_JP_LOSTOP];
        h2->limot1.histop = P[GZ_JP_HISTOP];
        h2->limot1.vel = P[GZ_JP_VEL];
        h2->limot1.fmax = P[GZ_JP_FMAX];
        h2->limot1.fudge_factor = P[GZ_JP_FUDGE];
        h2->limot1.bounce = P[GZ_JP_BOUNCE];
        h2->limot1.normal_cfm = P[GZ_JP_NORMAL_CFM];
        h2->limot1.stop_erp = P[GZ_JP_STOP_ERP];
        h2->limot1.stop_cfm = P[GZ_JP_STOP_CFM];
        h2->limot2.vel = P[GZ_JP_VEL2];       // dParamVel2 / dParamFMax2 (ODEHinge2Joint.cc:184-186): axis 2 is only ever powered
        h2->limot2.fmax = P[GZ_JP_FMAX2];
      } else if (sc->joint_type[j] == GZ_JOINT_SLIDER) {
        jt = dJointCreateSlider(w->world, 0);
        dJointAttach(jt, b1, b2);
        dxJointSlider *sj = (dxJointSlider *)jt;
        for (int k = 0; k < 3; k++) {
          sj->axis1[k] = sc->joint_axis1[3 * j + k];
          sj->offset[k] = sc->joint_anchor1[3 * j + k];
        }
        for (int k = 0; k < 4; k++) sj->qrel[k] = sc->joint_qrel[4 * j + k];
        sj->limot.lostop = P[GZ_JP_LOSTOP];
        sj->limot.histop = P[GZ_JP_HISTOP];
        sj->limot.vel = P[GZ_JP_VEL];
        sj->limot.fmax = P[GZ_JP_FMAX];
        sj->limot.fudge_factor = P[GZ_JP_FUDGE];
        sj->limot.bounce = P[GZ_JP_BOUNCE];
        sj->limot.normal_cfm = P[GZ_JP_NORMAL_CFM];
        sj->limot.stop_erp = P[GZ_JP_STOP_ERP];
        sj->limot.stop_cfm = P[GZ_JP_STOP_CFM];
      } else {
        return nullptr;
      }
      jt->erp = P[GZ_JP_ERP];
      jt->cfm = P[GZ_JP_CFM];
      w->joints.push_back(jt);
      w->joint_type.push_back(sc->joint_type[j]);
      HingePD z = {0, 0, 0, 0, 0, 0};
      w->pd.push_back(z);
    }
  }
  return B;
}

static void pool_destroy(RefBatch *B);
void gzref_destroy(void *h) {
  RefBatch *B = static_cast<RefBatch *>(h);
  pool_destroy(B);
  for (RefWorld *w : B->worlds) {
    dJointGroupDestroy(w->contactGroup);
    dSpaceDestroy(w->top);
    dWorldDestroy(w->world);
    delete w;
  }
  delete B;
}

// pd: [nj_total][6] = amplitude, frequency(Hz), phase, kp, kd, torque clamp (<=0: off)
void gzref_set_hinge_pd(void *h, const double *pd) {
  RefBatch *B = static_cast<RefBatch *>(h);
  for (size_t wi = 0; wi < B->worlds.size(); wi++) {
    RefWorld *w = B->worlds[wi];
    for (size_t j = 0; j < w->joints.size(); j++) {
      const double *c = pd + 6 * (B->joint_off[wi] + j);
      HingePD v = {c[0], c[1], c[2], c[3], c[4], c[5]};
      w->pd[j] = v;
    }
  }
}

void gzref_set_joint_velocity_cmd(void *h, const double *vel) {
  RefBatch *B = static_cast<RefBatch *>(h);
  for (size_t wi = 0; wi < B->worlds.size(); wi++) {
    RefWorld *w = B->worlds[wi];
    w->vel_cmd.assign(vel + B->joint_off[wi], vel + B->joint_off[wi] + w->joints.size());
  }
}

// Link::AddForce / AddTorque before a step (dBodyAddForce / dBodyAddTorque: world frame, cleared by the step):
// force[nb][3], torque[nb][3], either may be null.
void gzref_add_body_wrench(void *h, const double *force, const double *torque) {
  RefBatch *B = static_cast<RefBatch *>(h);
  for (size_t wi = 0; wi < B->worlds.size(); wi++) {
    RefWorld *w = B->worlds[wi];
    for (size_t k = 0; k < w->bodies.size(); k++) {
      const size_t b = (size_t)B->body_off[wi] + k;
      if (force) dBodyAddForce(w->bodies[k], force[3 * b], force[3 * b + 1], force[3 * b + 2]);
      if (torque) dBodyAddTorque(w->bodies[k], torque[3 * b], torque[3 * b + 1], torque[3 * b + 2]);
    }
  }
}

int gzref_step(void *h, int nsteps) {
  RefBatch *B = static_cast<RefBatch *>(h);
  dAllocateODEDataForThread(dAllocateMaskAll);
  for (RefWorld *w : B->worlds)
    for (int s = 0; s < nsteps; s++) step_world(B, w);
  return 0;
}

// worlds dealt to nthreads persistent host threads (BASELINE.md section 3 procedure); a thread takes the next
// world from a shared counter and steps it nsteps times (worlds are independent, so the order is free)
static void pool_worker(RefBatch *B, StepPool *P) {
  dAllocateODEDataForThread(dAllocateMaskAll);
  long seen = 0;
  for (;;) {
    int nsteps;
    {
      std::unique_lock<std::mutex> lk(P->mu);
      P->cv_go.wait(lk, [&] { return P->quit || P->gen != seen; });
      if (P->quit) break;
      seen = P->gen;
      nsteps = P->nsteps;
    }
    for (;;) {
      const size_t wi = P->next.fetch_add(1);
      if (wi >= B->worlds.size()) break;
      for (int s = 0; s < nsteps; s++) step_world(B, B->worlds[wi]);
    }
    {
      std::lock_guard<std::mutex> lk(P->mu);
      if (--P->running == 0) P->cv_done.notify_all();
    }
  }
  dCleanupODEAllDataForThread();
}

static void pool_destroy(RefBatch *B) {
  StepPool *P = B->pool;
  if (!P) return;
  { std::lock_guard<std::mutex> lk(P->mu); P->quit = true; }
  P->cv_go.notify_all();
  for (auto &t : P->th) t.join();
  delete P;
  B->pool = nullptr;
}

int gzref_step_mt(void *h, int nsteps, int nthreads) {
  RefBatch *B = static_cast<RefBatch *>(h);
  if (nthreads < 1) nthreads = 1;
  if (B->pool && (int)B->pool->th.size() != nthreads) pool_destroy(B);
  if (!B->pool) {
    B->pool = new StepPool;
    for (int t = 0; t < nthreads; t++) B->pool->th.emplace_back(pool_worker, B, B->pool);
  }
  StepPool *P = B->pool;
  {
    std::lock_guard<std::mutex> lk(P->mu);
    P->nsteps = nsteps;
    P->next.store(0);
    P->running = nthreads;
    P->gen++;
  }
  P->cv_go.notify_all();
  {
    std::unique_lock<std::mutex> lk(P->mu);
    P->cv_done.wait(lk, [&] { return P->running == 0; });
  }
  return 0;
}

// ---- hooks for the tests of the host adapter (host/gzcuda_ode.cpp, libgzcuda_ode.so) -----------------------
// The adapter works on LIVE ODE objects; these are the ones this harness built through the public ODE C API.
// out[0] = dxWorld*, out[1] = the top-level dxSpace*, out[2] = user pointer for gzref_surface_cb.
void gzref_world_handles(void *h, int world, void **out) {
  RefWorld *w = static_cast<RefBatch *>(h)->worlds[world];
  out[0] = w->world; out[1] = w->top; out[2] = w;
}
// What Gazebo's glue would answer from ODECollision / ODESurfaceParams (struct gzcuda_surface of
// include/gzcuda_ode.h, restated here so that the harness does not depend on the product's headers beyond the scene).
struct AdapterSurface {
  double mu, mu2, kp, kd, slip1, slip2, bounce, bounce_vel, patch_radius, elastic_modulus, fdir1[3];
  unsigned collide_bitmask; int max_contacts; long order_key;
  double poissons_ratio, surface_radius, mu3; int use_patch_radius;
  double slip3;
};
int gzref_surface_cb(dGeomID g, AdapterSurface *o, void *user) {
  RefWorld *w = static_cast<RefWorld *>(user);
  const int i = geom_index(g);
  if (i < 0 || i >= (int)w->surf.size()) return -1;
  const Surface &s = w->surf[i];
  o->mu = s.s[GZ_SF_MU]; o->mu2 = s.s[GZ_SF_MU2]; o->kp = s.s[GZ_SF_KP]; o->kd = s.s[GZ_SF_KD];
  o->slip1 = s.s[GZ_SF_SLIP1]; o->slip2 = s.s[GZ_SF_SLIP2]; o->bounce = s.s[GZ_SF_BOUNCE]; o->bounce_vel = s.s[GZ_SF_BOUNCE_VEL];
  o->patch_radius = s.s[GZ_SF_PATCH_RADIUS]; o->elastic_modulus = s.s[GZ_SF_ELASTIC_MODULUS];
  o->poissons_ratio = s.s[GZ_SF_POISSON]; o->surface_radius = s.s[GZ_SF_SURFACE_RADIUS];
  o->use_patch_radius = s.s[GZ_SF_USE_PATCH_RADIUS] != 0; o->mu3 = s.s[GZ_SF_MU3]; o->slip3 = s.s[GZ_SF_SLIP3];
  for (int k = 0; k < 3; k++) o->fdir1[k] = s.fdir1[k];
  o->collide_bitmask = s.collide_bitmask; o->max_contacts = s.max_contacts;
  o->order_key = i;   // Gazebo's creation order of the collisions = the scene's geom order
  return 0;
}
// The plugin part of a step (World::Update, World.cc:746: plugins run before UpdateCollision): PD torques through
// dJointAddHingeTorque, joint velocity commands through the bodies' velocities.  The adapter then has to honour what
// these calls left in the dxBody objects.
void gzref_plugin_update(void *h, int world) {
  RefBatch *B = static_cast<RefBatch *>(h);
  RefWorld *w = B->worlds[world];
  dAllocateODEDataForThread(dAllocateMaskAll);
  apply_velocity_commands(w);
  apply_controllers(B, w);
  w->step_index++;
}

// runs only ODEPhysics::UpdateCollision on one world (debug / parity of pair lists)
int gzref_collide(void *h, int world) {
  RefBatch *B = static_cast<RefBatch *>(h);
  dAllocateODEDataForThread(dAllocateMaskAll);
  update_collision(B, B->worlds[world]);
  return (int)B->worlds[world]->pairs.size();
}

int gzref_get_pairs(void *h, int world, int *out, int cap) {
  RefWorld *w = static_cast<RefBatch *>(h)->worlds[world];
  int n = (int)w->pairs.size();
  for (int i = 0; i < n && i < cap; i++) { out[2 * i] = w->pairs[i].first; out[2 * i + 1] = w->pairs[i].second; }
  return n;
}

// geoms [n][2]; data [n][7] = pos(3), normal(3), depth
int gzref_get_contacts(void *h, int world, int *geoms, double *data, int cap) {
  RefWorld *w = static_cast<RefBatch *>(h)->worlds[world];
  int n = (int)w->contacts.size();
  for (int i = 0; i < n && i < cap; i++) {
    const ContactRec &r = w->contacts[i];
    geoms[2 * i] = r.g1; geoms[2 * i + 1] = r.g2;
    for (int k = 0; k < 3; k++) { data[7 * i + k] = r.pos[k]; data[7 * i + 3 + k] = r.normal[k]; }
    data[7 * i + 6] = r.depth;
  }
  return n;
}

void gzref_get_body_state(void *h, double *pos, double *quat, double *lvel, double *avel) {
  RefBatch *B = static_cast<RefBatch *>(h);
  size_t b = 0;
  for (RefWorld *w : B->worlds)
    for (dBodyID body : w->bodies) {
      for (int k = 0; k < 3; k++) {
        pos[3 * b + k] = body->posr.pos[k];
        lvel[3 * b + k] = body->lvel[k];
        avel[3 * b + k] = body->avel[k];
      }
      for (int k = 0; k < 4; k++) quat[4 * b + k] = body->q[k];
      b++;
    }
}

void gzref_get_body_flags(void *h, int *flags) {
  RefBatch *B = static_cast<RefBatch *>(h);
  size_t b = 0;
  for (RefWorld *w : B->worlds)
    for (dBodyID body : w->bodies) flags[b++] = (body->flags & dxBodyDisabled) ? GZ_BODY_DISABLED : 0;
}

// per world: rms_constraint_residual[4] (ODEPhysics::GetParam("constraint_residual"),
// ODEPhysics.cc:1630-1635) and the number of contact joints created this step.
void gzref_get_stats(void *h, double *rms_residual, int *num_contacts, int *num_contact_joints) {
  RefBatch *B = static_cast<RefBatch *>(h);
  for (size_t wi = 0; wi < B->worlds.size(); wi++) {
    RefWorld *w = B->worlds[wi];
    for (int k = 0; k < 4; k++) rms_residual[4 * wi + k] = w->rms_residual[k];
    num_contacts[wi] = w->num_contacts;
    num_contact_joints[wi] = (int)w->contact_joints.size();
  }
}

// joint state that persists across steps: lambda[6], lambda_erp[6], cumulative_angle
void gzref_get_joint_state(void *h, double *lambda, double *lambda_erp, double *cum_angle) {
  RefBatch *B = static_cast<RefBatch *>(h);
  size_t j = 0;
  for (RefWorld *w : B->worlds)
    for (size_t k = 0; k < w->joints.size(); k++) {
      dxJoint *jt = w->joints[k];
      for (int r = 0; r < 6; r++) { lambda[6 * j + r] = jt->lambda[r]; lambda_erp[6 * j + r] = jt->lambda_erp[r]; }
      cum_angle[j] = (w->joint_type[k] == GZ_JOINT_HINGE) ? ((dxJointHinge *)jt)->cumulative_angle :
                     (w->joint_type[k] == GZ_JOINT_UNIVERSAL) ? ((dxJointUniversal *)jt)->cumulative_angle1 :
                     (w->joint_type[k] == GZ_JOINT_GEARBOX) ? ((dxJointGearbox *)jt)->cumulative_angle1 :
                     (w->joint_type[k] == GZ_JOINT_SCREW) ? ((dxJointScrew *)jt)->cumulative_angle : 0.0;
      j++;
    }
}

// ODEJoint::SetProvideFeedback (ODEJoint.cc:1262-1275) on every joint, contact feedback as when a
// contact subscriber exists.  Does not change the dynamics (quickstep keeps a copy of J).
void gzref_enable_feedback(void *h) {
  RefBatch *B = static_cast<RefBatch *>(h);
  for (RefWorld *w : B->worlds) {
    w->feedback = true;
    w->joint_fb.assign(w->joints.size(), dJointFeedback());
    for (size_t j = 0; j < w->joints.size(); ++j) {
      memset(&w->joint_fb[j], 0, sizeof(dJointFeedback));
      dJointSetFeedback(w->joints[j], &w->joint_fb[j]);
    }
  }
}

static void fb_copy(const dJointFeedback &f, double *o) {
  for (int k = 0; k < 3; k++) { o[k] = f.f1[k]; o[3 + k] = f.t1[k]; o[6 + k] = f.f2[k]; o[9 + k] = f.t2[k]; }
}

void gzref_get_joint_feedback(void *h, double *out) {
  RefBatch *B = static_cast<RefBatch *>(h);
  for (size_t wi = 0; wi < B->worlds.size(); ++wi) {
    RefWorld *w = B->worlds[wi];
    for (size_t j = 0; j < w->joint_fb.size(); ++j) fb_copy(w->joint_fb[j], out + 12 * (B->joint_off[wi] + j));
  }
}

int gzref_get_contact_feedback(void *h, int world, double *out, int cap) {
  RefBatch *B = static_cast<RefBatch *>(h);
  RefWorld *w = B->worlds[world];
  int n = (int)w->contact_fb.size();
  for (int i = 0; i < n && i < cap; ++i) fb_copy(w->contact_fb[i], out + 12 * i);
  return n;
}

}  // extern "C"
