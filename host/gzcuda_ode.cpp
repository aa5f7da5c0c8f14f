// gzcuda_ode.cpp — libgzcuda_ode.so: the host adapter that puts libgzcuda behind the reference's own step
// interface (include/gzcuda_ode.h).  Host code is C++ over the C ABI of include/gzcuda.h; it reads and writes the
// reference's live ODE objects through the reference's HEADERS only (deps/opende/src/objects.h,
// collision_kernel.h, collision_std.h, joints/*.h: compiled with -I/root/reference/..., no reference source is
// copied and nothing of oracle/ is linked in).  The few ODE functions it calls (dGeomMoved) resolve at load time
// against the ODE the host process already carries (gzserver's libgazebo_ode; the compiled reference in the tests).
//
// A world is mirrored into a one-world gz_scene the first time it is seen and again after
// gzcuda_topology_changed.  Every call uploads what the host may have changed since (pose, velocity, force /
// torque accumulators, enable flags, limit-motor parameters: the plugin API of SURVEY 8b), steps on the device,
// writes pos / q / R / lvel / avel / flags back into the dxBody objects, marks their geoms moved (dGeomMoved) and
// fires moved_callback as dxStepBody does (util.cpp:350-361; ODELink::MoveCallback ODELink.cc:142-178).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cfloat>
#include <map>
#include <string>
#include <vector>

#include <gazebo/ode/ode.h>
#include "config.h"
#include "objects.h"
#include "joints/joints.h"
#include "collision_kernel.h"
#include "collision_std.h"

#include "gzcuda.h"
#include "gzcuda_ode.h"

namespace {

struct Mirror {
  dxWorld *world = nullptr;
  dxSpace *top = nullptr;            // the top-level space of the last gzcuda_dspace_collide
  gzcuda_batch *batch = nullptr;
  std::vector<dxBody *> bodies;
  std::vector<dxGeom *> geoms;
  std::vector<dxJoint *> joints;
  std::vector<double> jparam;        // [nj][GZ_JP_COUNT] as last uploaded
  std::vector<int> enabled;          // [nb] as last seen on either side
  std::vector<double> pos, quat, lvel, avel, force, torque;
  gz_params params;
  gzcuda_surface_fn surface_fn = nullptr;
  void *surface_user = nullptr;
  int max_contacts = 20;
  bool collided = false;             // UpdateCollision ran for the step UpdatePhysics is about to take
  bool stale = true;
  size_t nb_seen = 0, nj_seen = 0;
};

std::map<dxWorld *, Mirror *> g_mirrors;
int g_device = 0;
thread_local std::string g_error;

int fail(int code, const std::string &msg) { g_error = msg; return code; }
int fail_lib(const char *what) { g_error = std::string(what) + ": " + gzcuda_last_error(); return GZCUDA_E_CUDA; }

Mirror *mirror_of(dxWorld *w) {
  auto it = g_mirrors.find(w);
  if (it != g_mirrors.end()) return it->second;
  Mirror *m = new Mirror;
  m->world = w;
  g_mirrors[w] = m;
  return m;
}

// geoms of a space tree, children before siblings are NOT assumed: plain depth-first in list order
void collect_geoms(dxSpace *sp, std::vector<dxGeom *> &out) {
  for (dxGeom *g = sp->first; g; g = g->next) {
    if (g->type >= dFirstSpaceClass && g->type <= dLastSpaceClass) collect_geoms(static_cast<dxSpace *>(g), out);
    else out.push_back(g);
  }
}

// world parameters, read from dxWorld at every step entry (cheap)
void read_params(const dxWorld *w, double h, int max_contacts, gz_params &p) {
  memset(&p, 0, sizeof(p));
  for (int k = 0; k < 3; k++) p.gravity[k] = w->gravity[k];
  p.step_size = h;
  p.erp = w->global_erp; p.cfm = w->global_cfm;
  p.sor_w = w->qs.w;
  p.iters = w->qs.num_iterations;
  p.friction_iters = w->qs.friction_iterations;
  p.contact_max_correcting_vel = w->contactp.max_vel;
  p.contact_surface_layer = w->contactp.min_depth;
  p.max_contacts = max_contacts;
  p.warm_start = w->qs.warm_start;
  p.smooth_contacts = w->qs.smooth_contacts;
  p.contact_sor_scale = w->qs.contact_sor_scale;
  p.adis_idle_steps = w->adis.idle_steps;
  p.adis_idle_time = w->adis.idle_time;
  p.adis_linear_threshold_sq = w->adis.linear_average_threshold;
  p.adis_angular_threshold_sq = w->adis.angular_average_threshold;
  p.damping_linear_threshold_sq = w->dampingp.linear_threshold;
  p.damping_angular_threshold_sq = w->dampingp.angular_threshold;
  p.friction_model = (int)w->qs.friction_model;
  p.dynamic_inertia = w->qs.dynamic_inertia_reduction ? 1 : 0;
}

}  // namespace

namespace {

void mat3(const dReal *R, double *o) {   // dMatrix3 (3 x 4, row padded) -> 9
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) o[3 * i + j] = R[4 * i + j];
}

// dQtoR (rotation.cpp:133-151), the same operation order as the device's q_to_R: the R written back into a dxBody
// is bit for bit the one the device holds
void q_to_R34(const double *q, dReal *R) {
  const double qq1 = 2 * q[1] * q[1], qq2 = 2 * q[2] * q[2], qq3 = 2 * q[3] * q[3];
  R[0] = 1 - qq2 - qq3;                 R[1] = 2 * (q[1] * q[2] - q[0] * q[3]); R[2] = 2 * (q[1] * q[3] + q[0] * q[2]);
  R[4] = 2 * (q[1] * q[2] + q[0] * q[3]); R[5] = 1 - qq1 - qq3;                 R[6] = 2 * (q[2] * q[3] - q[0] * q[1]);
  R[8] = 2 * (q[1] * q[3] - q[0] * q[2]); R[9] = 2 * (q[2] * q[3] + q[0] * q[1]); R[10] = 1 - qq1 - qq2;
  R[3] = R[7] = R[11] = 0;
}

void limot_params(const dxJointLimitMotor &l, double *P) {
  P[GZ_JP_LOSTOP] = l.lostop; P[GZ_JP_HISTOP] = l.histop; P[GZ_JP_VEL] = l.vel; P[GZ_JP_FMAX] = l.fmax;
  P[GZ_JP_FUDGE] = l.fudge_factor; P[GZ_JP_BOUNCE] = l.bounce; P[GZ_JP_NORMAL_CFM] = l.normal_cfm;
  P[GZ_JP_STOP_ERP] = l.stop_erp; P[GZ_JP_STOP_CFM] = l.stop_cfm;
}

// the joint's parameter slots as the host holds them now; returns the GZ_JOINT_* type or 0 if unsupported
int joint_params(const Mirror *m, dxJoint *j, double *P) {
  for (int k = 0; k < GZ_JP_COUNT; k++) P[k] = 0;
  P[GZ_JP_LOSTOP] = -dInfinity; P[GZ_JP_HISTOP] = dInfinity; P[GZ_JP_FUDGE] = 1;
  P[GZ_JP_ERP] = j->erp; P[GZ_JP_CFM] = j->cfm;
  switch (j->type()) {
    case dJointTypeHinge: limot_params(static_cast<dxJointHinge *>(j)->limot, P); return GZ_JOINT_HINGE;
    case dJointTypeSlider: limot_params(static_cast<dxJointSlider *>(j)->limot, P); return GZ_JOINT_SLIDER;
    case dJointTypeFixed: return GZ_JOINT_FIXED;
    case dJointTypeBall: return GZ_JOINT_BALL;
    case dJointTypeUniversal: {
      const dxJointUniversal *uj = static_cast<dxJointUniversal *>(j);
      limot_params(uj->limot1, P);
      const dxJointLimitMotor &l2 = uj->limot2;
      P[GZ_JP_VEL2] = l2.vel; P[GZ_JP_FMAX2] = l2.fmax; P[GZ_JP_LOSTOP2] = l2.lostop; P[GZ_JP_HISTOP2] = l2.histop;
      P[GZ_JP_FUDGE2] = l2.fudge_factor; P[GZ_JP_BOUNCE2] = l2.bounce; P[GZ_JP_NORMAL_CFM2] = l2.normal_cfm;
      P[GZ_JP_STOP_ERP2] = l2.stop_erp; P[GZ_JP_STOP_CFM2] = l2.stop_cfm;
      for (int k = 0; k < 4; k++) P[GZ_JP_QREL2 + k] = uj->qrel2[k];
      return GZ_JOINT_UNIVERSAL;
    }
    case dJointTypeScrew: {
      const dxJointScrew *sj = static_cast<dxJointScrew *>(j);
      limot_params(sj->limot, P);
      P[GZ_JP_THREAD_PITCH] = sj->thread_pitch;
      for (int k = 0; k < 3; k++) P[GZ_JP_OFFSET + k] = sj->offset[k];
      return GZ_JOINT_SCREW;
    }
    case dJointTypeGearbox: {
      const dxJointGearbox *gj = static_cast<dxJointGearbox *>(j);
      P[GZ_JP_RATIO] = gj->ratio;
      for (int k = 0; k < 4; k++) P[GZ_JP_QREL2 + k] = gj->qrel2[k];
      P[GZ_JP_REFBODY1] = P[GZ_JP_REFBODY2] = -1;
      for (size_t i = 0; i < m->bodies.size(); i++) {   // the reference bodies as indices of the mirror
        if (m->bodies[i] == gj->refBody1) P[GZ_JP_REFBODY1] = (double)i;
        if (m->bodies[i] == gj->refBody2) P[GZ_JP_REFBODY2] = (double)i;
      }
      return GZ_JOINT_GEARBOX;
    }
    case dJointTypeHinge2: {
      const dxJointHinge2 *h2 = static_cast<dxJointHinge2 *>(j);
      limot_params(h2->limot1, P);
      P[GZ_JP_VEL2] = h2->limot2.vel; P[GZ_JP_FMAX2] = h2->limot2.fmax;
      P[GZ_JP_SUSP_ERP] = h2->susp_erp; P[GZ_JP_SUSP_CFM] = h2->susp_cfm;
      for (int k = 0; k < 3; k++) { P[GZ_JP_V1 + k] = h2->v1[k]; P[GZ_JP_V2 + k] = h2->v2[k]; }
      P[GZ_JP_C0] = h2->c0; P[GZ_JP_S0] = h2->s0;
      return GZ_JOINT_HINGE2;
    }
    default: return 0;
  }
}

struct GeomRec { dxGeom *g; gzcuda_surface s; size_t seq; };

int build_mirror(Mirror *m, dxSpace *top, double h) {
  dxWorld *w = m->world;
  if (m->batch) { gzcuda_batch_destroy(m->batch); m->batch = nullptr; }
  m->bodies.clear(); m->geoms.clear(); m->joints.clear();
  // ---- bodies and joints: the world's lists are prepended to, so reversed = creation order
  for (dxBody *b = w->firstbody; b; b = static_cast<dxBody *>(b->next)) m->bodies.push_back(b);
  std::reverse(m->bodies.begin(), m->bodies.end());
  for (dxJoint *j = w->firstjoint; j; j = static_cast<dxJoint *>(j->next))
    if (j->type() != dJointTypeContact) m->joints.push_back(j);
  std::reverse(m->joints.begin(), m->joints.end());
  std::map<dxBody *, int> bidx;
  for (size_t i = 0; i < m->bodies.size(); i++) bidx[m->bodies[i]] = (int)i;
  // ---- geoms: ascending order_key of the glue (Gazebo's creation order), else reversed traversal
  std::vector<dxGeom *> raw;
  if (top) collect_geoms(top, raw);
  std::reverse(raw.begin(), raw.end());
  std::vector<GeomRec> gr(raw.size());
  for (size_t i = 0; i < raw.size(); i++) {
    gzcuda_surface s;
    memset(&s, 0, sizeof(s));
    s.mu = s.mu2 = 1.0; s.kp = 1e12; s.kd = 1.0; s.bounce_vel = 1e5; s.collide_bitmask = 0xFFFF; s.max_contacts = 20;
    s.poissons_ratio = 0.3; s.surface_radius = DBL_MAX; s.use_patch_radius = 1; s.mu3 = 1.0;
    s.order_key = (long)i;
    if (m->surface_fn && m->surface_fn(raw[i], &s, m->surface_user) != 0)
      return fail(GZCUDA_E_INVALID, "surface callback failed for a geom");
    gr[i].g = raw[i]; gr[i].s = s; gr[i].seq = i;
  }
  std::stable_sort(gr.begin(), gr.end(), [](const GeomRec &a, const GeomRec &b) { return a.s.order_key < b.s.order_key; });
  const int nb = (int)m->bodies.size(), ng = (int)gr.size(), nj = (int)m->joints.size();
  // ---- flat scene of one world
  std::vector<double> bpos(3 * nb), bquat(4 * nb), blv(3 * nb), bav(3 * nb), bmass(nb), binert(6 * nb), bdamp(2 * nb), bcp(2 * nb);
  std::vector<int> bflags(nb);
  for (int i = 0; i < nb; i++) {
    const dxBody *b = m->bodies[i];
    for (int k = 0; k < 3; k++) { bpos[3 * i + k] = b->posr.pos[k]; blv[3 * i + k] = b->lvel[k]; bav[3 * i + k] = b->avel[k]; }
    for (int k = 0; k < 4; k++) bquat[4 * i + k] = b->q[k];
    bmass[i] = b->mass.mass;
    const dReal *I = b->mass.I;
    binert[6 * i + 0] = I[0]; binert[6 * i + 1] = I[5]; binert[6 * i + 2] = I[10];
    binert[6 * i + 3] = I[1]; binert[6 * i + 4] = I[2]; binert[6 * i + 5] = I[6];
    int f = 0;
    if (b->flags & dxBodyNoGravity) f |= GZ_BODY_NO_GRAVITY;
    if (b->flags & dxBodyAutoDisable) f |= GZ_BODY_AUTO_DISABLE;
    if (b->flags & dxBodyDisabled) f |= GZ_BODY_DISABLED;
    if (!(b->flags & dxBodyGyroscopic)) f |= GZ_BODY_NO_GYRO;
    if (b->flags & (dxBodyFlagFiniteRotation | dxBodyFlagFiniteRotationAxis | dxBodyMaxAngularSpeed))
      return fail(GZCUDA_E_INVALID, "finite-rotation and max-angular-speed bodies are not supported (Gazebo never sets them)");
    bflags[i] = f;
    bdamp[2 * i] = (b->flags & dxBodyLinearDamping) ? b->dampingp.linear_scale : 0.0;
    bdamp[2 * i + 1] = (b->flags & dxBodyAngularDamping) ? b->dampingp.angular_scale : 0.0;
    const dxContactParameters &cp = b->contactp ? *b->contactp : w->contactp;
    bcp[2 * i] = cp.max_vel; bcp[2 * i + 1] = cp.min_depth;
  }
  std::vector<int> gtype(ng), gbody(ng), gspace(ng), gmaxc(ng);
  std::vector<double> gdims(4 * ng, 0.0), gpos(3 * ng, 0.0), gquat(4 * ng, 0.0), grot(9 * ng, 0.0), gsurf((size_t)GZ_SF_COUNT * ng), gfd(3 * ng);
  std::vector<unsigned> gcat(ng), gcol(ng), gmask(ng);
  std::map<dxSpace *, int> sidx;
  for (int i = 0; i < ng; i++) {
    dxGeom *g = gr[i].g;
    const gzcuda_surface &s = gr[i].s;
    m->geoms.push_back(g);
    if (g->type < dSphereClass || g->type > dPlaneClass)
      return fail(GZCUDA_E_INVALID, "geom class is not one of sphere, box, capsule, cylinder, plane");
    gtype[i] = g->type;   // dSphereClass .. dPlaneClass = GZ_GEOM_SPHERE .. GZ_GEOM_PLANE
    double *d = &gdims[4 * i];
    switch (g->type) {
      case dSphereClass: d[0] = static_cast<dxSphere *>(g)->radius; break;
      case dBoxClass: for (int k = 0; k < 3; k++) d[k] = static_cast<dxBox *>(g)->side[k]; break;
      case dCapsuleClass: d[0] = static_cast<dxCapsule *>(g)->radius; d[1] = static_cast<dxCapsule *>(g)->lz; break;
      case dCylinderClass: d[0] = static_cast<dxCylinder *>(g)->radius; d[1] = static_cast<dxCylinder *>(g)->lz; break;
      default: for (int k = 0; k < 4; k++) d[k] = static_cast<dxPlane *>(g)->p[k]; break;
    }
    gquat[4 * i] = 1.0;
    double *R = &grot[9 * i];
    R[0] = R[4] = R[8] = 1.0;
    gbody[i] = -1;
    if (g->type != dPlaneClass) {
      if (g->body) {
        auto it = bidx.find(g->body);
        if (it == bidx.end()) return fail(GZCUDA_E_INVALID, "a geom is attached to a body of another world");
        gbody[i] = it->second;
        if (g->offset_posr) { for (int k = 0; k < 3; k++) gpos[3 * i + k] = g->offset_posr->pos[k]; mat3(g->offset_posr->R, R); }
      } else if (g->final_posr) {
        for (int k = 0; k < 3; k++) gpos[3 * i + k] = g->final_posr->pos[k];
        mat3(g->final_posr->R, R);
      }
    }
    auto sp = sidx.find(g->parent_space);
    if (g->parent_space == top) gspace[i] = -1 - i;   // directly in the top space: collides with everything (own id, below)
    else if (sp == sidx.end()) { const int id = (int)sidx.size(); sidx[g->parent_space] = id; gspace[i] = id; }
    else gspace[i] = sp->second;
    gcat[i] = (unsigned)g->category_bits; gcol[i] = (unsigned)g->collide_bits;
    gmask[i] = s.collide_bitmask; gmaxc[i] = s.max_contacts;
    double *S = &gsurf[(size_t)GZ_SF_COUNT * i];
    S[GZ_SF_MU] = s.mu; S[GZ_SF_MU2] = s.mu2; S[GZ_SF_KP] = s.kp; S[GZ_SF_KD] = s.kd; S[GZ_SF_SLIP1] = s.slip1;
    S[GZ_SF_SLIP2] = s.slip2; S[GZ_SF_BOUNCE] = s.bounce; S[GZ_SF_BOUNCE_VEL] = s.bounce_vel;
    S[GZ_SF_PATCH_RADIUS] = s.patch_radius; S[GZ_SF_ELASTIC_MODULUS] = s.elastic_modulus;
    S[GZ_SF_POISSON] = s.poissons_ratio; S[GZ_SF_SURFACE_RADIUS] = s.surface_radius;
    S[GZ_SF_USE_PATCH_RADIUS] = s.use_patch_radius ? 1.0 : 0.0; S[GZ_SF_MU3] = s.mu3; S[GZ_SF_SLIP3] = s.slip3;
    for (int k = 0; k < 3; k++) gfd[3 * i + k] = s.fdir1[k];
  }
  {   // geoms that sit directly in the top-level space form "spaces" of their own
    int next = (int)sidx.size();
    for (int i = 0; i < ng; i++) if (gspace[i] < 0) gspace[i] = next++;
  }
  std::vector<int> jtype(nj), jb1(nj), jb2(nj);
  std::vector<double> ja1(3 * nj, 0.0), ja2(3 * nj, 0.0), jx1(3 * nj, 0.0), jx2(3 * nj, 0.0), jq(4 * nj, 0.0);
  m->jparam.assign((size_t)GZ_JP_COUNT * nj, 0.0);
  for (int i = 0; i < nj; i++) {
    dxJoint *j = m->joints[i];
    jtype[i] = joint_params(m, j, &m->jparam[(size_t)GZ_JP_COUNT * i]);
    if (!jtype[i]) return fail(GZCUDA_E_INVALID, "joint type is not one of hinge, slider, fixed, ball, universal, hinge2, gearbox, screw");
    if (!j->node[0].body) return fail(GZCUDA_E_INVALID, "a joint without bodies");
    auto i1 = bidx.find(j->node[0].body);
    if (i1 == bidx.end()) return fail(GZCUDA_E_INVALID, "a joint is attached to a body of another world");
    jb1[i] = i1->second;
    jb2[i] = -1;
    if (j->node[1].body) {
      auto i2 = bidx.find(j->node[1].body);
      if (i2 == bidx.end()) return fail(GZCUDA_E_INVALID, "a joint is attached to a body of another world");
      jb2[i] = i2->second;
    }
    jq[4 * i] = 1.0;
    if (jtype[i] == GZ_JOINT_HINGE) {
      const dxJointHinge *hj = static_cast<dxJointHinge *>(j);
      for (int k = 0; k < 3; k++) { ja1[3 * i + k] = hj->anchor1[k]; ja2[3 * i + k] = hj->anchor2[k]; jx1[3 * i + k] = hj->axis1[k]; jx2[3 * i + k] = hj->axis2[k]; }
      for (int k = 0; k < 4; k++) jq[4 * i + k] = hj->qrel[k];
    } else if (jtype[i] == GZ_JOINT_FIXED) {
      const dxJointFixed *fj = static_cast<dxJointFixed *>(j);
      for (int k = 0; k < 3; k++) ja1[3 * i + k] = fj->offset[k];
      for (int k = 0; k < 4; k++) jq[4 * i + k] = fj->qrel[k];
    } else if (jtype[i] == GZ_JOINT_SCREW) {
      const dxJointScrew *sj = static_cast<dxJointScrew *>(j);
      for (int k = 0; k < 3; k++) { ja1[3 * i + k] = sj->anchor1[k]; ja2[3 * i + k] = sj->anchor2[k]; jx1[3 * i + k] = sj->axis1[k]; jx2[3 * i + k] = sj->axis2[k]; }
      for (int k = 0; k < 4; k++) jq[4 * i + k] = sj->qrel[k];
    } else if (jtype[i] == GZ_JOINT_GEARBOX) {
      const dxJointGearbox *gj = static_cast<dxJointGearbox *>(j);
      for (int k = 0; k < 3; k++) { jx1[3 * i + k] = gj->axis1[k]; jx2[3 * i + k] = gj->axis2[k]; }
      for (int k = 0; k < 4; k++) jq[4 * i + k] = gj->qrel1[k];
    } else if (jtype[i] == GZ_JOINT_UNIVERSAL) {
      const dxJointUniversal *uj = static_cast<dxJointUniversal *>(j);
      for (int k = 0; k < 3; k++) { ja1[3 * i + k] = uj->anchor1[k]; ja2[3 * i + k] = uj->anchor2[k]; jx1[3 * i + k] = uj->axis1[k]; jx2[3 * i + k] = uj->axis2[k]; }
      for (int k = 0; k < 4; k++) jq[4 * i + k] = uj->qrel1[k];
    } else if (jtype[i] == GZ_JOINT_HINGE2) {
      const dxJointHinge2 *h2 = static_cast<dxJointHinge2 *>(j);
      for (int k = 0; k < 3; k++) { ja1[3 * i + k] = h2->anchor1[k]; ja2[3 * i + k] = h2->anchor2[k]; jx1[3 * i + k] = h2->axis1[k]; jx2[3 * i + k] = h2->axis2[k]; }
    } else if (jtype[i] == GZ_JOINT_BALL) {
      const dxJointBall *bj = static_cast<dxJointBall *>(j);
      for (int k = 0; k < 3; k++) { ja1[3 * i + k] = bj->anchor1[k]; ja2[3 * i + k] = bj->anchor2[k]; }
    } else {
      const dxJointSlider *sj = static_cast<dxJointSlider *>(j);
      for (int k = 0; k < 3; k++) { jx1[3 * i + k] = sj->axis1[k]; ja1[3 * i + k] = sj->offset[k]; }
      for (int k = 0; k < 4; k++) jq[4 * i + k] = sj->qrel[k];
    }
  }
  const int boff[2] = {0, nb}, goff[2] = {0, ng}, joff[2] = {0, nj};
  gz_scene sc;
  memset(&sc, 0, sizeof(sc));
  sc.n_worlds = 1; sc.body_off = boff; sc.geom_off = goff; sc.joint_off = joff;
  sc.body_pos = bpos.data(); sc.body_quat = bquat.data(); sc.body_lvel = blv.data(); sc.body_avel = bav.data();
  sc.body_mass = bmass.data(); sc.body_inertia = binert.data(); sc.body_flags = bflags.data();
  sc.body_damping = bdamp.data(); sc.body_contactp = bcp.data();
  sc.geom_type = gtype.data(); sc.geom_body = gbody.data(); sc.geom_space = gspace.data(); sc.geom_dims = gdims.data();
  sc.geom_pos = gpos.data(); sc.geom_quat = gquat.data(); sc.geom_rot = grot.data();
  sc.geom_category_bits = gcat.data(); sc.geom_collide_bits = gcol.data(); sc.geom_collide_bitmask = gmask.data();
  sc.geom_max_contacts = gmaxc.data(); sc.geom_surf = gsurf.data(); sc.geom_fdir1 = gfd.data();
  sc.joint_type = jtype.data(); sc.joint_body1 = jb1.data(); sc.joint_body2 = jb2.data();
  sc.joint_anchor1 = ja1.data(); sc.joint_anchor2 = ja2.data(); sc.joint_axis1 = jx1.data(); sc.joint_axis2 = jx2.data();
  sc.joint_qrel = jq.data(); sc.joint_param = m->jparam.data();
  read_params(w, h, m->max_contacts, m->params);
  m->batch = gzcuda_batch_create(&sc, &m->params, g_device);
  if (!m->batch) return fail_lib("gzcuda_batch_create");
  // the batch normalised the quaternions once more (dBodySetQuaternion's convention): put the live ones in, bit for bit
  if (nb > 0 && gzcuda_batch_set_body_state_raw(m->batch, 0, nb, bpos.data(), bquat.data(), blv.data(), bav.data()) != 0)
    return fail_lib("gzcuda_batch_set_body_state_raw");
  m->enabled.assign(nb, 1);
  for (int i = 0; i < nb; i++) m->enabled[i] = (m->bodies[i]->flags & dxBodyDisabled) ? 0 : 1;
  m->pos.resize(3 * nb); m->quat.resize(4 * nb); m->lvel.resize(3 * nb); m->avel.resize(3 * nb);
  m->force.resize(3 * nb); m->torque.resize(3 * nb);
  m->nb_seen = (size_t)w->nb; m->nj_seen = m->joints.size();
  m->stale = false;
  m->collided = false;
  return 0;
}

}  // namespace

namespace {

// world parameters the host may have changed through dWorldSet* / ODEPhysics::SetParam since the last call
int push_params(Mirror *m, double h) {
  gz_params p;
  read_params(m->world, h, m->max_contacts, p);
  const gz_params &o = m->params;
  struct { const char *key; double now, was; } kv[] = {
    {"iters", (double)p.iters, (double)o.iters}, {"sor", p.sor_w, o.sor_w}, {"cfm", p.cfm, o.cfm}, {"erp", p.erp, o.erp},
    {"contact_max_correcting_vel", p.contact_max_correcting_vel, o.contact_max_correcting_vel},
    {"contact_surface_layer", p.contact_surface_layer, o.contact_surface_layer},
    {"max_contacts", (double)p.max_contacts, (double)o.max_contacts},
    {"friction_iters", (double)p.friction_iters, (double)o.friction_iters},
    {"warm_start_factor", p.warm_start, o.warm_start}, {"contact_residual_smoothing", p.smooth_contacts, o.smooth_contacts},
    {"contact_sor_scale", p.contact_sor_scale, o.contact_sor_scale}, {"max_step_size", p.step_size, o.step_size},
    {"friction_model", (double)p.friction_model, (double)o.friction_model},
    {"dynamic_moi_rescaling", (double)p.dynamic_inertia, (double)o.dynamic_inertia},
    {"gravity_x", p.gravity[0], o.gravity[0]}, {"gravity_y", p.gravity[1], o.gravity[1]}, {"gravity_z", p.gravity[2], o.gravity[2]},
  };
  for (auto &e : kv)
    if (e.now != e.was && gzcuda_batch_set_param(m->batch, e.key, e.now) != 0) return fail_lib(e.key);
  m->params = p;
  return 0;
}

// host -> device: what a plugin may have written into the dxBody / dxJoint objects between two calls
int upload(Mirror *m, bool with_wrench) {
  const int nb = (int)m->bodies.size(), nj = (int)m->joints.size();
  if (nb == 0) return 0;
  bool flags_changed = false, any_wrench = false;
  std::vector<int> en(nb);
  for (int i = 0; i < nb; i++) {
    const dxBody *b = m->bodies[i];
    for (int k = 0; k < 3; k++) {
      m->pos[3 * i + k] = b->posr.pos[k]; m->lvel[3 * i + k] = b->lvel[k]; m->avel[3 * i + k] = b->avel[k];
      m->force[3 * i + k] = b->facc[k]; m->torque[3 * i + k] = b->tacc[k];
      any_wrench = any_wrench || b->facc[k] != 0 || b->tacc[k] != 0;
    }
    for (int k = 0; k < 4; k++) m->quat[4 * i + k] = b->q[k];
    en[i] = (b->flags & dxBodyDisabled) ? 0 : 1;
    flags_changed = flags_changed || en[i] != m->enabled[i];
  }
  // identical values are written back unchanged (raw: no renormalisation), so unmoved bodies stay bit-identical
  if (gzcuda_batch_set_body_state_raw(m->batch, 0, nb, m->pos.data(), m->quat.data(), m->lvel.data(), m->avel.data()) != 0)
    return fail_lib("gzcuda_batch_set_body_state_raw");
  if (flags_changed) {   // dBodyEnable / dBodyDisable (ODELink::SetEnabled)
    if (gzcuda_batch_set_body_enabled(m->batch, 0, nb, en.data()) != 0) return fail_lib("gzcuda_batch_set_body_enabled");
    m->enabled = en;
  }
  if (with_wrench && any_wrench &&   // dBodyAddForce / dBodyAddTorque / dJointAddHingeTorque since the last step
      gzcuda_batch_add_body_wrench(m->batch, m->force.data(), m->torque.data()) != 0)
    return fail_lib("gzcuda_batch_add_body_wrench");
  for (int i = 0; i < nj; i++) {   // dJointSetHingeParam / dJointSetSliderParam, joint erp / cfm
    double P[GZ_JP_COUNT];
    joint_params(m, m->joints[i], P);
    double *old = &m->jparam[(size_t)GZ_JP_COUNT * i];
    for (int k = 0; k < GZ_JP_COUNT; k++)
      if (P[k] != old[k] && !(P[k] != P[k] && old[k] != old[k])) {
        if (gzcuda_batch_set_joint_param(m->batch, i, k, P[k]) != 0) return fail_lib("gzcuda_batch_set_joint_param");
        old[k] = P[k];
      }
  }
  return 0;
}

// device -> host after the step: what dxStepBody leaves in a dxBody, then dGeomMoved and moved_callback
int download(Mirror *m) {
  const int nb = (int)m->bodies.size();
  if (nb == 0) return 0;
  std::vector<int> dis(nb);
  if (gzcuda_batch_get_body_state(m->batch, m->pos.data(), m->quat.data(), m->lvel.data(), m->avel.data()) != 0)
    return fail_lib("gzcuda_batch_get_body_state");
  if (gzcuda_batch_get_body_flags(m->batch, dis.data()) != 0) return fail_lib("gzcuda_batch_get_body_flags");
  for (int i = 0; i < nb; i++) {
    dxBody *b = m->bodies[i];
    const bool was_enabled = !(b->flags & dxBodyDisabled);
    for (int k = 0; k < 3; k++) {
      b->posr.pos[k] = m->pos[3 * i + k]; b->lvel[k] = m->lvel[3 * i + k]; b->avel[k] = m->avel[3 * i + k];
      b->facc[k] = 0; b->tacc[k] = 0;     // the step consumed the accumulators (quickstep_update_bodies.cpp:355-364)
    }
    for (int k = 0; k < 4; k++) b->q[k] = m->quat[4 * i + k];
    q_to_R34(&m->quat[4 * i], b->posr.R);
    if (dis[i]) b->flags |= dxBodyDisabled; else b->flags &= ~dxBodyDisabled;
    m->enabled[i] = dis[i] ? 0 : 1;
    const bool stepped = was_enabled || !dis[i];   // in an island that was stepped (a sleeper it touched was woken)
    if (stepped) {
      for (dxGeom *g = b->geom; g; g = g->body_next) dGeomMoved(g);      // util.cpp:350-353
      if (b->moved_callback) b->moved_callback(b);                        // util.cpp:358-359 -> ODELink::MoveCallback
      if (was_enabled && dis[i] && b->disabled_callback) b->disabled_callback(b);   // util.cpp:250-251
    }
  }
  // what getInfo1 and the solver leave in the joints: the hinge's cumulative angle (hinge.cpp:65-77; read by
  // dJointGetHingeAngle, i.e. ODEHingeJoint::GetAngleImpl) and the lambdas of the warm start (quickstep.cpp:523-541)
  const int nj = (int)m->joints.size();
  if (nj > 0) {
    std::vector<double> lam(6 * (size_t)nj), lam_erp(6 * (size_t)nj), cum(nj);
    if (gzcuda_batch_get_joint_state(m->batch, lam.data(), lam_erp.data(), cum.data()) != 0) return fail_lib("gzcuda_batch_get_joint_state");
    for (int i = 0; i < nj; i++) {
      dxJoint *j = m->joints[i];
      for (int k = 0; k < 6; k++) { j->lambda[k] = lam[6 * (size_t)i + k]; j->lambda_erp[k] = lam_erp[6 * (size_t)i + k]; }
      if (j->type() == dJointTypeHinge) static_cast<dxJointHinge *>(j)->cumulative_angle = cum[i];
      if (j->type() == dJointTypeUniversal) static_cast<dxJointUniversal *>(j)->cumulative_angle1 = cum[i];   // (axis 2's stays on the device)
      if (j->type() == dJointTypeGearbox) static_cast<dxJointGearbox *>(j)->cumulative_angle1 = cum[i];
      if (j->type() == dJointTypeScrew) static_cast<dxJointScrew *>(j)->cumulative_angle = cum[i];
    }
  }
  return 0;
}

Mirror *ready(dxWorld *world, dxSpace *top, double h) {
  Mirror *m = mirror_of(world);
  // bodies or persistent joints added / removed without gzcuda_topology_changed: the counts give most of it away
  size_t nj = 0;
  for (dxJoint *j = world->firstjoint; j; j = static_cast<dxJoint *>(j->next)) if (j->type() != dJointTypeContact) nj++;
  if (top && top != m->top) { m->top = top; m->stale = true; }
  if (!m->batch || m->stale || m->nb_seen != (size_t)world->nb || m->nj_seen != nj) {
    if (build_mirror(m, m->top, h) != 0) return nullptr;
  }
  return m;
}

}  // namespace

extern "C" {

void gzcuda_ode_set_surface_callback(dxWorld *world, gzcuda_surface_fn fn, void *user) {
  Mirror *m = mirror_of(world);
  m->surface_fn = fn; m->surface_user = user; m->stale = true;
}
void gzcuda_ode_set_max_contacts(dxWorld *world, int max_contacts) { mirror_of(world)->max_contacts = max_contacts; }
int gzcuda_ode_set_device(int device) {
  if (device < 0 || device >= gzcuda_device_count()) return fail(GZCUDA_E_INVALID, "device index out of range");
  g_device = device;
  return 0;
}
void gzcuda_topology_changed(dxWorld *world) { mirror_of(world)->stale = true; }
const char *gzcuda_ode_last_error(void) { return g_error.c_str(); }

void gzcuda_dworld_release(dxWorld *world) {
  auto it = g_mirrors.find(world);
  if (it == g_mirrors.end()) return;
  if (it->second->batch) gzcuda_batch_destroy(it->second->batch);
  delete it->second;
  g_mirrors.erase(it);
}

}  // extern "C"

extern "C" {

// ODEPhysics::UpdateCollision (ODEPhysics.cc:366-404): the contact joints of the previous step are gone (the
// device rebuilds its contact set), geometry poses follow the bodies, broadphase + narrowphase + Collide run
// on the device.  The host's contactGroup stays empty: contact joints live on the device only.
int gzcuda_dspace_collide(dxSpace *top, dxWorld *world) {
  if (!world) return fail(GZCUDA_E_INVALID, "null world");
  Mirror *m = ready(world, top, world ? 0.001 : 0.0);
  if (!m) return GZCUDA_E_INVALID;
  if (upload(m, false) != 0) return GZCUDA_E_CUDA;
  if (gzcuda_batch_update_collision(m->batch) != 0) return fail_lib("gzcuda_batch_update_collision");
  m->collided = true;
  return 0;
}

// physicsStepFunc (ODEPhysicsPrivate.hh:65): 1 on success, 0 on failure, like dWorldQuickStep (ode.cpp:1980-1995)
int gzcuda_dworld_step(dxWorld *world, double stepsize) {
  if (!world || !(stepsize > 0)) { fail(GZCUDA_E_INVALID, "null world or non-positive step"); return 0; }
  Mirror *m = ready(world, nullptr, stepsize);
  if (!m) return 0;
  if (push_params(m, stepsize) != 0) return 0;
  if (!m->collided) {
    // dWorldQuickStep without a collision pass steps with no contact joints; the device keeps last step's contact
    // set until the next collision pass, so run one now (World::Update always calls UpdateCollision first)
    if (upload(m, false) != 0) return 0;
    if (gzcuda_batch_update_collision(m->batch) != 0) { fail_lib("gzcuda_batch_update_collision"); return 0; }
  }
  if (upload(m, true) != 0) return 0;
  if (gzcuda_world_step(m->batch, stepsize) != 1) { fail_lib("gzcuda_world_step"); return 0; }
  m->collided = false;
  if (download(m) != 0) return 0;
  // solver statistics, dWorldGetQuickStepRMSConstraintResidual etc. (ODEPhysics.cc:1630-1635)
  double rms[4] = {0, 0, 0, 0};
  int nc = 0, rows = 0;
  if (gzcuda_batch_get_stats(m->batch, rms, &nc, &rows) == 0) {
    for (int k = 0; k < 4; k++) world->qs.rms_constraint_residual[k] = rms[k];
    world->qs.num_contacts = nc;
  }
  return 1;
}

int gzcuda_dworld_get_contacts(dxWorld *world, dxGeom **geoms, double *data, int cap) {
  auto it = g_mirrors.find(world);
  if (it == g_mirrors.end() || !it->second->batch) return fail(GZCUDA_E_INVALID, "world is not mirrored");
  Mirror *m = it->second;
  std::vector<int> gi(2 * (size_t)std::max(cap, 1));
  const int n = gzcuda_batch_get_contacts(m->batch, 0, gi.data(), data, cap);
  if (n < 0) return fail_lib("gzcuda_batch_get_contacts");
  if (geoms)
    for (int i = 0; i < n && i < cap; i++) { geoms[2 * i] = m->geoms[gi[2 * i]]; geoms[2 * i + 1] = m->geoms[gi[2 * i + 1]]; }
  return n;
}

int gzcuda_dworld_ray_query(dxWorld *world, int n, const double *start, const double *dir, const double *length,
                            double *range, dxGeom **hit) {
  auto it = g_mirrors.find(world);
  if (it == g_mirrors.end() || !it->second->batch) return fail(GZCUDA_E_INVALID, "world is not mirrored");
  Mirror *m = it->second;
  if (n <= 0) return 0;
  std::vector<int> w((size_t)n, 0), gi((size_t)n, -1);
  const int rc = gzcuda_batch_ray_query(m->batch, n, w.data(), start, dir, length, range, gi.data());
  if (rc < 0) return fail_lib("gzcuda_batch_ray_query");
  if (hit)
    for (int k = 0; k < n; k++) hit[k] = gi[k] >= 0 ? m->geoms[gi[k]] : nullptr;
  return 0;
}

}  // extern "C"
