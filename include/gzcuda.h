/*
 * gzcuda.h — C ABI of libgzcuda, the B200 (sm_100a) implementation of Gazebo's
 * per-step ODE physics path.  Plain C: opaque handle, plain pointers and sizes.
 *
 * Each entry point names the reference interface it replaces (file:line under
 * the ci-group/gazebo tree).  INTEGRATION.md shows the ODEPhysics-side binding.
 * All functions return 0 on success or a negative GZCUDA_E* code; the message is
 * available from gzcuda_last_error().  Nothing throws across the boundary
 * (the reference's step function returns int and never throws either,
 * deps/opende/src/ode.cpp:1980-1995).
 *
 * Threading: like ODEPhysics::UpdateCollision/UpdatePhysics, which run on the
 * world's RunLoop thread under physicsUpdateMutex (ODEPhysics.cc:370,413), a batch
 * must be driven by one host thread at a time.  Different batches (e.g. one per GPU)
 * may be driven from different threads.
 */
#ifndef GZCUDA_H
#define GZCUDA_H

#include "gzcuda_scene.h"

#ifdef __cplusplus
extern "C" {
#endif

#define GZCUDA_OK            0
#define GZCUDA_E_INVALID    -1   /* bad argument / unsupported geom or joint type */
#define GZCUDA_E_CUDA       -2   /* CUDA runtime error (no device, launch failure, ...) */
#define GZCUDA_E_CAPACITY   -3   /* a per-world pair / contact / row slab overflowed */

typedef struct gzcuda_batch gzcuda_batch;

/* Library / device */
int         gzcuda_version(void);
const char *gzcuda_last_error(void);
int         gzcuda_device_count(void);

/* World construction: replaces ODEPhysics::ODEPhysics + Load (ODEPhysics.cc:140-243) and
 * the dBodyCreate / dCreate<Geom> / dJointCreate<Type> calls of ODELink::Init (ODELink.cc:63-134),
 * ODE<Shape>Shape::Init and ODEJoint::Load/Attach (ODEJoint.cc:83-216) for a batch of
 * independent worlds on one device.  All scene arrays are copied. */
gzcuda_batch *gzcuda_batch_create(const gz_scene *scene, const gz_params *params, int device);
void          gzcuda_batch_destroy(gzcuda_batch *b);

/* One world step for every world of the batch, n_steps times, without host interaction:
 * ODEPhysics::UpdateCollision (ODEPhysics.cc:366-404) followed by ODEPhysics::UpdatePhysics
 * (:407-453) = dWorldQuickStep (deps/opende/src/ode.cpp:1980-1995).  Asynchronous on the
 * batch's stream; gzcuda_batch_sync or any getter waits for completion. */
int gzcuda_batch_step(gzcuda_batch *b, int n_steps);
/* Same, with HOST buffers, synchronous: uploads wrench_in[nb][6] (force, torque per body; what
 * plugins wrote with dBodyAddForce/dBodyAddTorque since the last step; NULL = none), runs
 * n_steps, downloads state_out[nb][13] = pos(3) quat(4) lvel(3) avel(3) (what ODELink::MoveCallback
 * reads, ODELink.cc:142-178; NULL = none).  Pinned host memory makes the copies asynchronous. */
int gzcuda_batch_step_io(gzcuda_batch *b, int n_steps, const double *wrench_in, double *state_out);
/* State log at a decimated rate (the WorldState stream behind `gz log` and ~/pose/info: gazebo/physics/World.cc:2757-2850,
 * gazebo/util/LogRecord.cc:193,809): n_steps steps in one call; after every `every`-th step the state of all bodies is
 * copied to frames[frame][nb][13] (position 3, quaternion 4, linear 3 and angular 3 velocity; pinned host memory makes the
 * copies run beside the following steps).  Returns the number of frames written (<= max_frames) or a negative error. */
int gzcuda_batch_step_log(gzcuda_batch *b, int n_steps, int every, double *frames, int max_frames);
/* The two halves, for callers that need Events::beforePhysicsUpdate between them
 * (gazebo/physics/World.cc:755-786). */
int gzcuda_batch_update_collision(gzcuda_batch *b);   /* ODEPhysics::UpdateCollision */
int gzcuda_batch_update_physics(gzcuda_batch *b);     /* ODEPhysics::UpdatePhysics   */
/* Drop-in for the step-function pointer `int (*physicsStepFunc)(dxWorld*, dReal)`
 * (gazebo/physics/ode/ODEPhysicsPrivate.hh:65): returns 1 on success, 0 on failure. */
int gzcuda_world_step(gzcuda_batch *b, double step_size);
int gzcuda_batch_sync(gzcuda_batch *b);

/* Body state, arrays over all bodies of the batch in scene order.
 * dBodyGetPosition/Quaternion/LinearVel/AngularVel read by ODELink::MoveCallback
 * (ODELink.cc:142-178); any pointer may be NULL. */
int gzcuda_batch_get_body_state(gzcuda_batch *b, double *pos, double *quat, double *lvel, double *avel);
/* dBodySetPosition/Quaternion/LinearVel/AngularVel (ODELink::OnPoseChange ODELink.cc:245-278,
 * ODELink::SetLinearVel/SetAngularVel); bodies [first, first+count). */
int gzcuda_batch_set_body_state(gzcuda_batch *b, int first, int count, const double *pos,
                                const double *quat, const double *lvel, const double *avel);
/* The same for a binding that mirrors live dxBody objects (host/gzcuda_ode.cpp): the quaternion is taken bit for
 * bit (dBodySetQuaternion normalised it on the host already, ode.cpp:414-427; normalising twice moves the last bit)
 * and R is rebuilt from it as dQtoR does. */
int gzcuda_batch_set_body_state_raw(gzcuda_batch *b, int first, int count, const double *pos,
                                    const double *quat, const double *lvel, const double *avel);
/* dBodyEnable / dBodyDisable (ode.cpp:869-887; ODELink::SetEnabled ODELink.cc:296-302): enabled[count], non-zero =
 * enable (restarts the auto-disable counters), zero = disable.  Stream-ordered. */
int gzcuda_batch_set_body_enabled(gzcuda_batch *b, int first, int count, const int *enabled);
/* dBodyAddForce / dBodyAddTorque for every body ([nb][3] each, NULL = none). */
int gzcuda_batch_add_body_wrench(gzcuda_batch *b, const double *force, const double *torque);
/* dBodyIsEnabled: flags[nb] receives GZ_BODY_DISABLED or 0. */
int gzcuda_batch_get_body_flags(gzcuda_batch *b, int *flags);

/* dJointSetHingeParam & friends (ODEJoint::SetParam ODEJoint.cc:392-510): slot = GZ_JP_*. */
int gzcuda_batch_set_joint_param(gzcuda_batch *b, int joint, int slot, double value);
/* Persistent joint state: lambda[nj][6], lambda_erp[nj][6] (dxJoint::lambda, joint.h:156-157),
 * cumulative hinge angle (dxJointHinge::cumulative_angle). */
int gzcuda_batch_get_joint_state(gzcuda_batch *b, double *lambda, double *lambda_erp, double *cum_angle);
/* Device-side stand-in for a ModelPlugin that tracks a sinusoid with a PD torque through
 * Joint::SetForce -> dJointAddHingeTorque (ODEHingeJoint.cc:159-162) once per step on
 * Events::worldUpdateBegin: pd[nj][6] = amplitude, frequency (Hz), phase, kp, kd, torque clamp
 * (clamp <= 0 disables the joint's controller). */
int gzcuda_batch_set_hinge_pd(gzcuda_batch *b, const double *pd);
/* Device-side stand-in for a plugin that calls Joint::SetVelocity(0, v) on hinge joints every update (this fork's
 * DiffDrivePlugin::OnUpdate, plugins/DiffDrivePlugin.cc:100-122 -> Joint::SetVelocityMaximal, Joint.cc:828-894): vel[nj] in
 * rad/s, NaN = no command for that joint.  Applied before every step, joints in order.  A host plugin does the same through
 * gzcuda_batch_set_body_state. */
int gzcuda_batch_set_joint_velocity_cmd(gzcuda_batch *b, const double *vel);

/* Solver statistics per world: rms_residual[nw][4] (dWorldGetQuickStepRMSConstraintResidual,
 * ODEPhysics::GetParam("constraint_residual") ODEPhysics.cc:1630-1635), number of contact
 * joints created in the last step, number of rows. Any pointer may be NULL. */
int gzcuda_batch_get_stats(gzcuda_batch *b, double *rms_residual, int *num_contacts, int *num_rows);
/* ODEPhysics::SetParam for the numeric solver keys (ODEPhysics.cc:1420-1581):
 * "iters", "sor", "cfm", "erp", "contact_max_correcting_vel", "contact_surface_layer",
 * "max_contacts", "friction_iters", "warm_start_factor", "contact_residual_smoothing",
 * "contact_sor_scale", "max_step_size", "friction_model", "dynamic_moi_rescaling" (ODEPhysics.cc:186-194:
 * DYNAMIC_INERTIA, quickstep_util.cpp:170-524), and "gravity_x" / "gravity_y" / "gravity_z" (dWorldSetGravity). */
int gzcuda_batch_set_param(gzcuda_batch *b, const char *key, double value);

/* Collision products of the last step (ContactManager contents): candidate pairs of one world
 * as world-local geom indices [n][2] sorted ascending, and its contacts: geoms[n][2] and
 * data[n][7] = position(3), normal(3), depth.  Return the count (or a negative error). */
int gzcuda_batch_get_pairs(gzcuda_batch *b, int world, int *pairs, int cap);
/* Ray queries against the CURRENT geometry of the worlds (poses as of the last step or setter): the CPU laser,
 * gazebo::physics::ODEMultiRayShape::UpdateRays / UpdateCallback (gazebo/physics/ode/ODEMultiRayShape.cc:82-205) over
 * ODE's ray colliders (deps/opende/src/ray.cpp:200-703).  n_rays rays, ray k in world ray_world[k] from start[k] along
 * dir[k] (any length: normalised as dGeomRaySet does) up to length[k]; rays carry GZ_SENSOR_COLLIDE bits and ride on no
 * body, as ODERayShape's do.  out_range[k] = distance to the nearest hit, or length[k]; out_geom[k] = world-local index of
 * the geom hit, or -1 (may be NULL).  Host arrays; one kernel, one warp per ray. */
int gzcuda_batch_ray_query(gzcuda_batch *b, int n_rays, const int *ray_world, const double *start, const double *dir,
                           const double *length, double *out_range, int *out_geom);
int gzcuda_batch_get_contacts(gzcuda_batch *b, int world, int *geoms, double *data, int cap);

/* Constraint wrenches of the last step (dJointFeedback {f1[3], t1[3], f2[3], t2[3]}, world frame:
 * dJointSetFeedback, filled by quickstep_update_bodies.cpp:79-119; ODEJoint::GetForceTorque
 * ODEJoint.cc:1010-1090 and the contact wrenches of ODEPhysics::UpdatePhysics ODEPhysics.cc:422-449
 * read them).  Joints: fb[nj][12] in scene joint order; contacts of one world: fb[n][12] in the
 * order of gzcuda_batch_get_contacts, body 1 = the contact joint's first body after dJointAttach's
 * swap.  A joint or contact whose island was not stepped reports zeros.  Return the count. */
int gzcuda_batch_get_joint_feedback(gzcuda_batch *b, double *fb);
int gzcuda_batch_get_contact_feedback(gzcuda_batch *b, int world, double *fb, int cap);

/* gzcuda_batch_get_joint_feedback for callers that cannot allocate.  This entry point has NO counterpart in
 * the reference: it was added for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * gzcuda_batch_get_joint_feedback (fb[nj][12], host buffer; returns nj), but the device scratch the feedback
 * kernel writes is a static array inside the library instead of a cudaMalloc per call: the call makes no
 * cudaMalloc / cudaFree and uses no allocating library routine.  Limits: at most GZCUDA_NOALLOC_MAX_JOINTS
 * joints in the batch (GZCUDA_E_INVALID beyond that); the array (3 MiB) is held by every GPU that loads the
 * library until the process exits; all calls share it, so calls that overlap on the GPU (different batches
 * on different host threads) race and must be serialised by the caller. */
#define GZCUDA_NOALLOC_MAX_JOINTS 32768
int gzcuda_batch_get_joint_feedback_noalloc(gzcuda_batch *b, double *fb);

/* Instrumentation: device time in ms of the last gzcuda_batch_step call (CUDA events on the
 * batch stream) and the number of kernel launches it issued. */
int gzcuda_batch_last_step_ms(gzcuda_batch *b, float *ms, long *launches);
/* Profiling of the dominant kernel: when on, every k_pgs launch is bracketed by CUDA events;
 * gzcuda_batch_pgs_ms returns the summed duration and launch count of the last step call. */
int gzcuda_batch_set_profiling(gzcuda_batch *b, int on);
int gzcuda_batch_pgs_ms(gzcuda_batch *b, float *total_ms, int *launches);
/* Running totals since creation / last reset: out[0] constraint rows, out[1] friction rows,
 * out[2] contacts, out[3] world-steps (the m and m_fric of SURVEY.md section 8d). */
int gzcuda_batch_get_counters(gzcuda_batch *b, unsigned long long out[4], int reset);
/* Instrumentation of the last step's k_pgs launch, out[nw][3]: loop trips and row updates of the
 * world's longest lane, and the SM cycles its sweeps took. */
int gzcuda_batch_get_pgs_debug(gzcuda_batch *b, int *out);
/* Developer instrumentation of k_pgs (enabled by GZCUDA_PGS_PROF=1 when the batch is created):
 * cycle sums over all warps since the last call: out[0] fetch issue, [1] wait, [2] ring load,
 * [3] row update, [4] active lanes, [5] rounds; k_pgs_chain: [8] cycles of rounds with a run item,
 * [9] such rounds, [10] rows of the longest run of each, [11] cycles of the other rounds, [12] their
 * count, [13] sweep cycles in total, [14] worlds. */
int gzcuda_batch_get_pgs_profile(gzcuda_batch *b, unsigned long long out[16]);
/* Developer instrumentation: start and end of each world's sweeps in the last
 * step, ns of the GPU's global timer, out[nw][2]. */
int gzcuda_batch_get_pgs_timeline(gzcuda_batch *b, unsigned long long *out);
/* Per-stage device time of the last step call, ms: collide, assemble, pgs, integrate. */
int gzcuda_batch_stage_ms(gzcuda_batch *b, float ms[4]);

#ifdef __cplusplus
}
#endif
#endif /* GZCUDA_H */
