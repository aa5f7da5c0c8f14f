/*
 * gzcuda_ode.h — C ABI of libgzcuda_ode.so (host/gzcuda_ode.cpp): libgzcuda behind the reference's OWN step
 * interface, on the reference's live ODE objects.  Plain C; the ODE types are the reference's opaque ids.
 *
 *   gzcuda_dworld_step      has the type of ODEPhysicsPrivate::physicsStepFunc, int (*)(dxWorld*, dReal)
 *                           (gazebo/physics/ode/ODEPhysicsPrivate.hh:65); ODEPhysics::SetStepType
 *                           (ODEPhysics.cc:923-937) stores it for <solver><type>gzcuda</type> and
 *                           ODEPhysics::UpdatePhysics calls it (ODEPhysics.cc:416-417)
 *   gzcuda_dspace_collide   replaces the body of ODEPhysics::UpdateCollision (ODEPhysics.cc:366-404):
 *                           dSpaceCollide + CollisionCallback (:948-1008) + Collide (:1012-1331)
 *
 * Return conventions follow the interfaces replaced: gzcuda_dworld_step returns 1 on success and 0 on failure
 * like dWorldQuickStep (ode.cpp:1980-1995); the others return 0 or a negative GZCUDA_E_* code;
 * gzcuda_ode_last_error() holds the text Gazebo would send to gzerr.
 */
#ifndef GZCUDA_ODE_H
#define GZCUDA_ODE_H

#ifdef __cplusplus
extern "C" {
#endif

struct dxWorld;
struct dxSpace;
struct dxGeom;

/* What ODEPhysics::Collide reads from the two ODECollision / ODESurfaceParams objects of a pair
 * (ODEPhysics.cc:1016-1258; ODESurfaceParams.hh, FrictionPyramid).  ODE itself does not store it: the glue
 * answers for one geom at mirror time.  order_key: position of the collision in Gazebo's creation order
 * (Base::GetId): geoms are indexed in ascending order_key, which fixes the deterministic pair / row order. */
typedef struct gzcuda_surface {
  double mu, mu2, kp, kd, slip1, slip2, bounce, bounce_vel;
  double patch_radius, elastic_modulus;   /* FrictionPyramid::PatchRadius / ElasticModulus (Hertzian contact when > 0 on both) */
  double fdir1[3];
  unsigned collide_bitmask;
  int max_contacts;
  long order_key;
  /* FrictionPyramid::PoissonsRatio / SurfaceRadius / UsePatchRadius / MuTorsion; pre-set to Gazebo's defaults
   * (0.3, DBL_MAX, 1, 1) before the callback runs.  Torsional friction is then switched on by a patch radius > 0
   * (or use_patch_radius = 0), as in Gazebo (ODEPhysics.cc:1216-1232). */
  double poissons_ratio, surface_radius, mu3;
  int use_patch_radius;
  double slip3;   /* ODESurfaceParams::slipTorsion */
} gzcuda_surface;
typedef int (*gzcuda_surface_fn)(struct dxGeom *geom, gzcuda_surface *out, void *user);

/* Set before the first step; without a callback every geom gets the SDF defaults (surface.sdf: mu = mu2 = 1,
 * kp 1e12, kd 1, bounce threshold 1e5, max_contacts 20) in space traversal order. */
void gzcuda_ode_set_surface_callback(struct dxWorld *world, gzcuda_surface_fn fn, void *user);
/* ODEPhysics::GetMaxContacts (SDF <max_contacts>, default 20): Gazebo state, not in dxWorld. */
void gzcuda_ode_set_max_contacts(struct dxWorld *world, int max_contacts);
/* CUDA device for the worlds mirrored from now on (ODEPhysics::InitForThread, ODEPhysics.cc:360-363). */
int gzcuda_ode_set_device(int device);

int gzcuda_dspace_collide(struct dxSpace *top, struct dxWorld *world);
int gzcuda_dworld_step(struct dxWorld *world, double stepsize);

/* Bodies, geoms or joints were added or removed (model spawn / delete, World.cc:2194; dJointAttach): the mirror is
 * rebuilt from the live objects at the next call.  Warm-start lambdas restart from dxJoint::lambda. */
void gzcuda_topology_changed(struct dxWorld *world);
/* Contacts of the last gzcuda_dspace_collide, for ContactManager (ODEPhysics.cc:1262-1329): geoms[n][2] (dxGeom*),
 * data[n][7] = pos, normal, depth.  Returns the count. */
int gzcuda_dworld_get_contacts(struct dxWorld *world, struct dxGeom **geoms, double *data, int cap);
/* The CPU laser against the world as of the last gzcuda_dspace_collide / gzcuda_dworld_step: what
 * ODEMultiRayShape::UpdateRays does with dSpaceCollide2(raySpace, worldSpace) (ODEMultiRayShape.cc:82-205) - n rays from
 * start[k] along dir[k] up to length[k]; range[k] = distance to the nearest hit or length[k], hit[k] = the dxGeom hit or
 * NULL (may be NULL itself).  Returns 0 or a negative error. */
int gzcuda_dworld_ray_query(struct dxWorld *world, int n, const double *start, const double *dir, const double *length,
                            double *range, struct dxGeom **hit);
/* Frees the mirror and its device batch (ODEPhysics::Fini). */
void gzcuda_dworld_release(struct dxWorld *world);
const char *gzcuda_ode_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* GZCUDA_ODE_H */
