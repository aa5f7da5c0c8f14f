/*
 * gzcuda_scene.h — flat (structure-of-arrays) description of a batch of
 * independent Gazebo/ODE worlds.  Plain C, plain pointers and sizes.
 *
 * This is the data a Gazebo-side binding hands to libgzcuda when it mirrors the
 * ODE objects that gazebo::physics::ODEPhysics owns (reference, read-only:
 *   dxBody   deps/opende/src/objects.h:157-186
 *   dxGeom   deps/opende/src/collision_kernel.h:96-183, collision_std.h
 *   dxJoint  deps/opende/src/joints/joint.h, hinge.h, fixed.h
 *   dxWorld / dxQuickStepParameters  deps/opende/src/objects.h:98-127,189-207
 *   surface parameters  gazebo/physics/ode/ODESurfaceParams.cc:28-31).
 * The same struct feeds the CPU oracle harness (oracle/ref_harness.cpp), which
 * turns it into real ODE objects through the reference's C API, so both sides
 * see identical inputs.
 *
 * Indexing: all arrays are concatenated over worlds; world w owns bodies
 * [body_off[w], body_off[w+1]), geoms [geom_off[w], geom_off[w+1]) and joints
 * [joint_off[w], joint_off[w+1]).  geom_body / joint_body1 / joint_body2 hold
 * WORLD-LOCAL body indices (0 .. nb_w-1) or -1 for "no body" (static geom /
 * joint to the world frame).
 */
#ifndef GZCUDA_SCENE_H
#define GZCUDA_SCENE_H

#ifdef __cplusplus
extern "C" {
#endif

/* geom_type — the five primitives north_star names (dSphereClass, dBoxClass,
 * dCapsuleClass, dCylinderClass, dPlaneClass; include/gazebo/ode/collision.h). */
enum { GZ_GEOM_SPHERE = 0, GZ_GEOM_BOX = 1, GZ_GEOM_CAPSULE = 2,
       GZ_GEOM_CYLINDER = 3, GZ_GEOM_PLANE = 4 };

/* joint_type (dJointTypeHinge, dJointTypeFixed, …; contact joints are created
 * by the collide stage every step and never appear in the scene). */
enum { GZ_JOINT_HINGE = 1, GZ_JOINT_FIXED = 2, GZ_JOINT_BALL = 3,
       GZ_JOINT_SLIDER = 4, GZ_JOINT_UNIVERSAL = 5,
       GZ_JOINT_HINGE2 = 6,
       GZ_JOINT_GEARBOX = 7,
       GZ_JOINT_SCREW = 8 };    /* dxJointScrew (SDF screw): hinge + slider coupled by the thread pitch; joint_anchor1/2, joint_axis1/2,
                                * joint_qrel as for a hinge, dxJointScrew::offset and thread_pitch in joint_param */  /* dxJointHinge2 (SDF revolute2): two bodies required; joint_anchor1/2, joint_axis1/2 in the
                                * bodies' frames, the derived members v1, v2, c0, s0 in joint_param (below);
                                * GZ_JOINT_GEARBOX = dxJointGearbox: two bodies required as well */

/* body_flags bits */
enum { GZ_BODY_NO_GRAVITY   = 1,   /* dxBodyNoGravity (ODELink::SetGravityMode(false)) */
       GZ_BODY_AUTO_DISABLE = 2,   /* dxBodyAutoDisable (ODELink.cc:70-77: jointless, sensorless models) */
       GZ_BODY_DISABLED     = 4,   /* dxBodyDisabled: initial state, updated by the step */
       GZ_BODY_NO_GYRO      = 8 }; /* clear dxBodyGyroscopic (set by default, ode.cpp:302) */

/* joint_param[] slots (dxJointLimitMotor, joints/joint.cpp:470-484; erp/cfm joint.cpp:66-67) */
enum { GZ_JP_LOSTOP = 0, GZ_JP_HISTOP = 1, GZ_JP_VEL = 2, GZ_JP_FMAX = 3,
       GZ_JP_FUDGE = 4, GZ_JP_BOUNCE = 5, GZ_JP_NORMAL_CFM = 6,
       GZ_JP_STOP_ERP = 7, GZ_JP_STOP_CFM = 8,
       GZ_JP_ERP = 9,      /* joint erp, < 0 = use world erp */
       GZ_JP_CFM = 10,     /* joint cfm, < 0 = use world cfm */
       /* hinge2 only (joints/hinge2.cpp): the motor of axis 2 (limot2: dParamVel2 / dParamFMax2; axis 2 is never
        * limited, hinge2.cpp:125-127), the suspension (dParamSuspensionERP / CFM), and the members ODE derives when the
        * anchor and the axes are set (makeV1andV2 :224-255, getAxisInfo :148-156) */
       GZ_JP_VEL2 = 12, GZ_JP_FMAX2 = 13, GZ_JP_SUSP_ERP = 14, GZ_JP_SUSP_CFM = 15,
       GZ_JP_V1 = 16,      /* v1[3]: 16..18 */
       GZ_JP_V2 = 19,      /* v2[3]: 19..21 */
       GZ_JP_C0 = 22, GZ_JP_S0 = 23,
       /* universal joints: the rest of the second limit motor (limot2: axis 2 has stops too, universal.cpp:268-307; its
        * vel / fmax are GZ_JP_VEL2 / GZ_JP_FMAX2) and qrel2; joint_qrel holds qrel1 (the relative rotations of the
        * joint's cross against the two bodies, computeInitialRelativeRotations, universal.cpp:716-758) */
       GZ_JP_LOSTOP2 = 24, GZ_JP_HISTOP2 = 25, GZ_JP_FUDGE2 = 26, GZ_JP_BOUNCE2 = 27, GZ_JP_NORMAL_CFM2 = 28,
       GZ_JP_STOP_ERP2 = 29, GZ_JP_STOP_CFM2 = 30,
       GZ_JP_QREL2 = 31,   /* qrel2[4]: 31..34 */
       /* gearbox joints (joints/gearbox.cpp: one row that couples the angles of two bodies about their axes, each measured
        * against a reference body): joint_axis1 / joint_axis2 in the bodies' frames, joint_qrel = qrel1, GZ_JP_QREL2 = qrel2,
        * the gear ratio, and the reference bodies as world-local body indices (-1: the world) */
       GZ_JP_RATIO = 35, GZ_JP_REFBODY1 = 14, GZ_JP_REFBODY2 = 15,   /* (the reference bodies overlay the hinge2-only suspension slots) */
       /* screw joints (joints/screw.cpp): thread pitch (rad / m) and offset[3] (computeOffset, screw.cpp:795-820); the stops
        * of the rotation are the joint's limit motor (slots 0..8); a POWERED screw (fmax > 0) is refused: the reference
        * writes that row behind the joint's five rows (screw.cpp:55-110 never counts it, :419 writes it) */
       GZ_JP_THREAD_PITCH = 35, GZ_JP_OFFSET = 16,   /* offset[3]: 16..18 (overlays hinge2's v1) */
       GZ_JP_COUNT = 36 };

/* geom_surf[] slots (ODESurfaceParams + FrictionPyramid) */
enum { GZ_SF_MU = 0, GZ_SF_MU2 = 1, GZ_SF_KP = 2, GZ_SF_KD = 3,
       GZ_SF_SLIP1 = 4, GZ_SF_SLIP2 = 5, GZ_SF_BOUNCE = 6, GZ_SF_BOUNCE_VEL = 7,
       /* Hertzian contact (FrictionPyramid elastic modulus / Poisson's ratio, ODEPhysics.cc:1186-1251, contact.cpp:163-216):
        * when BOTH surfaces of a pair have an elastic modulus > 0 the normal row's stiffness follows 4/3 E* sqrt(r x)
        * with r = the pair's patch radius (the larger of the two, when both surfaces say use_patch_radius) or
        * sqrt(surface radius x depth) with 1/R = 1/R1 + 1/R2.  Gazebo's defaults: patch radius 0, elastic modulus 0 (off),
        * Poisson's ratio 0.3, surface radius DBL_MAX, use_patch_radius 1 (SurfaceParams.cc:28-32). */
       GZ_SF_PATCH_RADIUS = 8, GZ_SF_ELASTIC_MODULUS = 9, GZ_SF_POISSON = 10, GZ_SF_SURFACE_RADIUS = 11,
       GZ_SF_USE_PATCH_RADIUS = 12,
       /* torsional friction (FrictionPyramid::MuTorsion / slipTorsion, ODEPhysics.cc:1168-1232): a pair gets a fourth
        * contact row (dContactMu3, contact.cpp:362-437: a moment about the normal, limits +-3 pi a mu3 / 16 x the normal
        * force, a = the patch radius or sqrt(surface radius x depth)) when min(mu3) > 0 and the radius it uses is > 0.
        * NOTE Gazebo's own default for MuTorsion is 1 (torsion is then switched by the patch radius); this format's
        * default is 0 = off.  Worlds of more than 256 bodies with mu3 > 0 are refused (GZCUDA_E_INVALID). */
       GZ_SF_MU3 = 13, GZ_SF_SLIP3 = 14,
       GZ_SF_COUNT = 15 };

/* World-level parameters, shared by every world of a batch
 * (ODEPhysics::Load, gazebo/physics/ode/ODEPhysics.cc:175-243; dWorldCreate
 * defaults deps/opende/src/ode.cpp:1625-1688). */
typedef struct gz_params {
  double gravity[3];
  double step_size;                  /* h = max_step_size */
  double erp, cfm;                   /* world->global_erp / global_cfm */
  double sor_w;                      /* qs.w */
  int    iters;                      /* qs.num_iterations (SDF <iters>) */
  int    friction_iters;             /* qs.friction_iterations (default 10) */
  double contact_max_correcting_vel; /* world->contactp.max_vel */
  double contact_surface_layer;      /* world->contactp.min_depth */
  int    max_contacts;               /* ODEPhysics::GetMaxContacts (SDF default 20) */
  double warm_start;                 /* qs.warm_start (0.5) */
  double smooth_contacts;            /* qs.smooth_contacts (0.01) */
  double contact_sor_scale;          /* qs.contact_sor_scale (0.25) */
  /* auto-disable defaults copied into bodies (ODEPhysics.cc:208-216; thresholds are stored squared) */
  int    adis_idle_steps;
  double adis_idle_time;
  double adis_linear_threshold_sq;
  double adis_angular_threshold_sq;
  /* damping thresholds (squared speeds, ode.cpp:1678-1679) */
  double damping_linear_threshold_sq;
  double damping_angular_threshold_sq;
  /* qs.friction_model (ode.cpp:1672, ODEPhysics::SetFrictionModel): 0 = pyramid (default), 2 = box
   * (friction limits +-mu instead of +-mu * lambda_normal); 1 = cone is not supported */
  int    friction_model;
  /* qs.dynamic_inertia_reduction (dWorldSetQuickStepInertiaRatioReduction; SDF <use_dynamic_moi_rescaling>,
   * ODEPhysics.cc:186-194, default false): move inertia between the two bodies of a hinge along its constrained
   * axes when their ratio exceeds 200 (DYNAMIC_INERTIA, quickstep_util.cpp:170-524) */
  int    dynamic_inertia;
} gz_params;

typedef struct gz_scene {
  int n_worlds;
  const int *body_off;      /* [n_worlds+1] */
  const int *geom_off;      /* [n_worlds+1] */
  const int *joint_off;     /* [n_worlds+1] */

  /* ---- bodies (dxBody) ---- */
  const double *body_pos;      /* [nb][3]  centre of mass, world frame */
  const double *body_quat;     /* [nb][4]  (w,x,y,z) */
  const double *body_lvel;     /* [nb][3] */
  const double *body_avel;     /* [nb][3] */
  const double *body_mass;     /* [nb] */
  const double *body_inertia;  /* [nb][6]  Ixx Iyy Izz Ixy Ixz Iyz, body frame, about the CoM */
  const int    *body_flags;    /* [nb] GZ_BODY_* */
  const double *body_damping;  /* [nb][2]  linear_scale, angular_scale; 0 = flag clear (ode.cpp:1069-1077) */
  const double *body_contactp; /* [nb][2]  max_vel, min_depth (dBodySetMaxVel/MinDepth, ODELink.cc:357-392) */

  /* ---- geoms (dxGeom + Gazebo surface) ---- */
  const int    *geom_type;     /* [ng] GZ_GEOM_* */
  const int    *geom_body;     /* [ng] world-local body index, -1 = static */
  const int    *geom_space;    /* [ng] world-local id of the dSimpleSpace holding it (one per model, ODEModel.cc:26) */
  const double *geom_dims;     /* [ng][4] box lx,ly,lz | sphere r | capsule/cylinder r,length | plane a,b,c,d */
  const double *geom_pos;      /* [ng][3] offset from the body frame (dGeomSetOffsetPosition) or world pose if static */
  const double *geom_quat;     /* [ng][4] same, rotation */
  const unsigned *geom_category_bits; /* [ng] */
  const unsigned *geom_collide_bits;  /* [ng] */
  const unsigned *geom_collide_bitmask; /* [ng] Gazebo surface collideBitmask (ODEPhysics.cc:1016-1018) */
  const int    *geom_max_contacts;    /* [ng] Collision::GetMaxContacts */
  const double *geom_surf;     /* [ng][GZ_SF_COUNT] */
  const double *geom_fdir1;    /* [ng][3] friction direction in the geom's link frame, 0 = unset */

  /* ---- persistent joints ---- */
  const int    *joint_type;    /* [nj] GZ_JOINT_* */
  const int    *joint_body1;   /* [nj] world-local, never -1 */
  const int    *joint_body2;   /* [nj] world-local or -1 (attached to the world) */
  const double *joint_anchor1; /* [nj][3] body-1 frame (hinge/ball/universal anchor1; fixed, slider: offset) */
  const double *joint_anchor2; /* [nj][3] body-2 frame (or world frame when body2 = -1) */
  const double *joint_axis1;   /* [nj][3] body-1 frame */
  const double *joint_axis2;   /* [nj][3] body-2 frame (or world) */
  const double *joint_qrel;    /* [nj][4] initial relative rotation body1 -> body2 */
  const double *joint_param;   /* [nj][GZ_JP_COUNT] */

  /* ---- optional (may be NULL) ---- */
  const double *geom_rot;      /* [ng][9] row-major rotation replacing geom_quat: a binding that mirrors live ODE geoms
                                * holds matrices (dxGeom::offset_posr / final_posr), and R -> q -> R is not exact */
} gz_scene;

#ifdef __cplusplus
}
#endif
#endif /* GZCUDA_SCENE_H */
