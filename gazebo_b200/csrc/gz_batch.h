// gz_batch.h — device-resident state of a batch of independent worlds.
//
// Layout in HBM: one record array per object kind, concatenated
// over worlds, each record a small contiguous struct of doubles/ints so a
// thread-per-object kernel reads whole 32-byte sectors and a CTA-per-world
// kernel reads one contiguous slab.  Per-step products (pairs, contacts,
// constraint rows) live in per-world slabs of fixed capacity.
//
// Mirrors (reference, read-only): dxBody deps/opende/src/objects.h:157-186,
// dxGeom collision_kernel.h:96-183, dxJointHinge joints/hinge.h:31-51,
// dContact include/gazebo/ode/contact.h:56-105, the row arrays of
// dxQuickStepper deps/opende/src/quickstep.cpp:262-325.
#pragma once
#include "../../include/gzcuda_scene.h"

// PGS schedule buckets (k_pgs.cuh): c = log2(lanes per solve) in 0..4, rung of the period ladder in 1..159
#define GZ_NRUNG 160
#define GZ_NBUCKET (10 * GZ_NRUNG)
#define GZ_PAIR_MAXC 16  // contacts one geom pair can yield (box-box: 8, cylinder-box: up to 13)
#define GZ_PGSB_NKEY 512   // period histogram of k_border: periods >= 511 share the last key
#define GZ_ENT 20
#define GZ_JP_HOT 12   // joint_param slots kept in struct Joint; the rest ([GZ_JP_COUNT - GZ_JP_HOT] per joint) in joint_ext
#define GZ_SF_HOT 8    // geom_surf slots kept in struct Geom; the rest ([GZ_SF_COUNT - GZ_SF_HOT] per geom) in geom_ext
#define GZ_PGSC_MAXCTA 1280  // k_pgsc: CTAs of one world; worlds beyond fall back to k_pgsd
#define GZ_PGSC_TAB (GZ_PGSC_MAXCTA + 9)          // doubles per body-owner PGS entry (k_pgsb.cuh)
#define GZ_CHAIN_NCLS 4   // shared-memory size classes of chain-mode worlds (k_pgs_chain.cuh), at most 5

namespace gz {

struct Body {
  double pos[3];
  double q[4];
  double R[9];
  double lvel[3], avel[3];
  double facc[3], tacc[3];
  double mass, invMass;
  double I[9], invI[9];        // body frame
  double lin_damp, ang_damp;   // scale; 0 = flag clear
  double max_vel, min_depth;   // dxBody::contactp
  double adis_timeleft;
  int adis_stepsleft;
  int flags;                   // GZ_BODY_*
  int world;
  int joint_begin, joint_end;  // slice of body_joint[] (persistent joints touching this body, joint order)
};

// per-step scratch of a body
struct BodyStep {
  double invMOI[9];
  double MOI[9];      // world-frame inertia: scratch of k_dyn_inertia (DYNAMIC_INERTIA), written only when it is on
  double caccel[6], caccel_erp[6];
  int label;      // island label = lowest world-local body index of the component
  int active;     // in an island stepped this step
  int has_contact;  // touched by a contact joint this step (dxBody::firstjoint != NULL test)
  int pad;
};

struct Geom {
  double dims[4];
  double opos[3];    // offset in body frame, or world pose if static
  double oR[9];
  double pos[3];     // final pose (dxGeom::final_posr)
  double R[9];
  double aabb[6];
  double surf[GZ_SF_HOT];    // geom_surf slots 0..7; the Hertzian-contact slots live in DevBatch::geom_ext (em_on says whether to look)
  double fdir1[3];
  int type, body, space, world;   // body: GLOBAL body index or -1
  unsigned cat, col, scat, scol;  // geom bits and the bits of its space (ODECollision.cc:161-176)
  unsigned bitmask;
  int max_contacts;
  int em_on, tors_on;  // elastic modulus > 0 (ODEPhysics.cc:1241-1251 needs it on BOTH geoms of a pair); tors_on: bit 0 mu3 > 0 (likewise, :1168), bit 1 patch radius > 0 or use_patch_radius off
};

struct Joint {
  double anchor1[3], anchor2[3], axis1[3], axis2[3], qrel[4];
  double param[GZ_JP_HOT];   // joint_param slots 0..11; the hinge2 / universal-only slots live in DevBatch::joint_ext
  double lambda[6], lambda_erp[6];
  double cum_angle;
  double limit_err;
  double cum_angle2, limit_err2;   // universal joints: axis 2 (limot2)
  double pd[6];      // amplitude, frequency, phase, kp, kd, torque clamp (<=0 off)
  int type, b1, b2;  // GLOBAL body indices, b2 = -1 for the world
  int limit;         // dxJointLimitMotor::limit
  int limit2;        // ... of a universal joint's second axis
  int m;             // rows this step (0 = not in an active island)
  int row0;          // first row (world-local)
  int world;
  int has_vel_cmd;   // Joint::SetVelocity stand-in (k_velcmd): vel_cmd is applied before every step
  double vel_cmd;
};

struct Contact {
  double pos[3], normal[3], depth;
  double mu, mu2, soft_erp, soft_cfm, slip1, slip2, bounce, bounce_vel;
  double fdir1[3];
  double em, em_radius;   // Hertzian contact (dContactEM): E* > 0 = on; em_radius = patch radius, or the surface radius (em_surface)
  int g1, g2;        // GLOBAL geom indices (canonical: g1 < g2)
  int b1, b2;        // GLOBAL body indices after dJointAttach's swap (b1 >= 0)
  int reverse;       // dJOINT_REVERSE
  int has_fdir1;
  int row0, em_surface;   // em_surface: the pair does not use the patch radius (contact.cpp:166-172)
  // torsional friction (dContactMu3, contact.cpp:362-437): tors 0 = the pair has no such row (in a world of four-row
  // contacts its fourth row is a placeholder that is never swept), 1 = patch radius, 2 = sqrt(surface radius x depth)
  double mu3, tors_radius, slip3;   // slip3 > 0: dContactSlip3
  int tors, pad_;
};

// One constraint row as the assembly kernels leave it: 24 doubles = 192 B.  (What the sweeps read is the
// packed record k_rowfinish makes of it, k_pgs.cuh.)
struct Row {
  double J[12];      // unscaled; the record holds J * Ad (quickstep_pgs_lcp.cpp:1091-1107)
  double Ad;
  double lo, hi;
  double lambda, lambda_erp;
  double c, cfm, c_v_max;
  int b1, b2;        // WORLD-LOCAL body indices, b2 = -1
  int findex;        // -1 bilateral, -2 contact normal, >=0 absolute (world-local) row of the normal
  int src;           // joint index (>=0, global) or ~contact index (global) for write-back
  int t1a, t2a, t1f; // k_schedule: stage, index in the run array or -1, item head slot | len << 20 or -1 (k_pgs.cuh)
  int next_f;        // k_schedule: lambda index (slab slot, or order[] position in chain mode)
};

struct WorldInfo {
  int body0, nb, geom0, ng, joint0, nj;
  int pair0, pair_cap, contact0, contact_cap, row0, row_cap;
  int slot0, slot_cap;   // packed-record slab of the world (schedule slots), see k_pgs.cuh
  int rpc, pad_;         // rows per contact: 3, or 4 when a surface of the world has torsional friction (mu3 > 0)
};

// validity tag of the row-slot path's records (k_pgs.cuh): 1 .. 2^24 - 1, never 0
#define GZ_EPOCH_PERIOD 0xFFFFFFL
#ifdef __CUDACC__
__host__ __device__
#endif
static inline unsigned gz_epoch(long step_index) { return (unsigned)(step_index % GZ_EPOCH_PERIOD) + 1u; }

struct WorldStep {
  int n_pairs, n_contacts, m, n_fric;
  int overflow;     // bit0 pairs, bit1 contacts, bit2 rows
  int num_contacts_stat;
  int pad[2];
  double rms_residual[4];
  double rms_dlambda[4];
};

struct DevBatch {
  int n_worlds, nb, ng, nj;
  int max_nb, max_ng, max_nj, max_pair_cap, max_contact_cap, max_row_cap;
  long step_index;
  gz_params p;
  WorldInfo *world;
  WorldStep *wstep;
  Body *body;
  BodyStep *bstep;
  int *body_joint;       // CSR payload: global joint indices per body, ascending
  Geom *geom;
  Joint *joint;
  int *pair;             // [pairs][2] global geom indices
  int *pair_nc;          // contacts of each pair after Gazebo's truncation
  double *pair_contact;  // [pairs][GZ_PAIR_MAXC][7]  pos, normal, depth
  Contact *contact;
  Row *row;
  int *order;            // per world slab: order[k] = world-local row index
  double *prec;          // packed row records (34 doubles each), one slab per PGS warp (k_pgs.cuh)
  long prec_cap;         // records the pool holds
  double *runrec;        // chain mode: the records of run items, per world at row0 (34 doubles each), k_pgs.cuh
  double *lamg;          // [slots][2] lambda / lambda_erp spill for worlds whose lambdas exceed shared memory
  int *wsched;           // per world WSched {R, k, nstage, bucket, idx, ...} (8 ints)
  int *bk_count;         // [GZ_NBUCKET] worlds per schedule bucket (k, R) this step
  int *bk_tab;           // [4][GZ_NBUCKET+1]: first warp / first record of each bucket; the same sorted heavy-first
  int *warp_world;       // [n_worlds + GZ_NBUCKET][16] worlds of each PGS warp this step (-1: none)
  int *scratch;          // per world int scratch [7][max_row_cap]
  unsigned long long *prof;  // [16] k_pgs / k_pgs_chain segment cycle sums (GZCUDA_PGS_PROF=1), see gzcuda_batch_get_pgs_profile
  int pgs_prof;
  int pgs_wpw_max;       // worlds per PGS warp the shared-memory budget allows
  int sch_rcap;          // longest schedule period k_schedule's occupancy table holds
  int sch_rpar;          // longest period k_schedule probes on 32 lanes at once (0: the probe is off)
  int pgs_chain_min;     // rows on the busiest body from which a world is scheduled in chain mode
  int pgs_chain_doubles[GZ_CHAIN_NCLS];   // shared memory (doubles) of k_pgs_chain's launches, ascending
  int pgs_rt, pgs_kmin, pgs_kmax;   // scheduler knobs: target period, lanes-per-solve range
  int bp_sap_min;        // geoms from which a world takes the sort-and-sweep broadphase (k_collide.cuh)
  unsigned long long *acc;   // [4] running totals: rows, friction rows, contacts, world-steps
  // ---- body-owner PGS (k_pgsb.cuh): per-body entry lists in a bump-allocated pool
  int4 *ent_aux;         // [ent_cap] {entry of the row's other body or -1, that body, entries back to the normal row, class} (k_pgsc only, else null)
  int *ent_part;         // [ent_cap] the entry of the row's other body, -1 if none (k_pgsd only, else null)
  int islands_hook;      // islands by hooking + pointer jumping (large worlds) instead of min-label propagation
  int pgs_dataflow;      // large worlds: 1 = k_pgsd (dataflow, no rounds), 0 = k_pgsl (rounds with a grid barrier)
  int pgsc_bodies;       // k_pgsc: bodies per CTA = warps per CTA x pgsc_bpw
  int pgsc_bpw;          // k_pgsc: bodies per warp (<= 16; the other lanes of the warp idle: a warp is ONE instruction stream
                         // that runs its bodies' rows one after the other, so fewer bodies per warp = more streams)
  double *joint_ext;     // [nj][GZ_JP_COUNT - GZ_JP_HOT] the joint_param slots only hinge2 / universal joints read
  int *bs_cntt;          // [nb] k_bsched_large: torsional rows per body (null unless a large world has four-row contacts)
  double *geom_ext;      // [ng][GZ_SF_COUNT - GZ_SF_HOT] patch radius, elastic modulus, Poisson's ratio, surface radius, use_patch_radius, mu3, slip3
  int *cta_e0;           // [n_worlds][GZ_PGSC_TAB] first entry of every CTA of k_pgsc (then the world's end), else null
  int *bs_clist;         // [2 * contacts] k_bsched_large's per-body contact lists (large worlds only, else null)
  int *bs_glob;          // [5][max_nb] k_bsched's per-body tables for worlds too large for shared memory (else null)
  double *xch_g;         // [2][lanes] k_pgsl: partial dot products of a round (double-buffered), then [grid][9] statistics partials
  int pgs_impl;          // 2: body-owner sweeps over a cooperative grid (k_pgsl, worlds of more than 512 bodies), 1: body-owner sweeps (k_pgsb), 0: the round-1 row-slot sweeps (k_pgs / k_pgs_chain)
  double *ent;           // [ent_cap][GZ_ENT] half-records, one per (row, body), grouped per body in order[] sequence
  int *ent_row;          // [ent_cap] world-local row index if this entry is the row's body-1 half (lambda write-back), else -1
  double *ent_ad2;       // [ent_cap] 1 / Ad^2 (statistics)
  double *ent_dl;        // [ent_cap][2] last delta of each entry, for entries swept from the pool (statistics)
  long ent_cap;
  unsigned long long *ent_cursor;   // [2] bump allocators, used alternately by step parity (the idle one is reset)
  int4 *bsched;          // [nb] {first entry (pool index), entries, first friction entry (list index), 0}
  int4 *wsb;             // [n_worlds] {period R, stages, first entry, entries}
  int *rk_count;         // [2][GZ_PGSB_NKEY] worlds per period this step (by step parity, like ent_cursor)
  int *wkey;             // [n_worlds] period key | rank among the worlds of that key << 10
  int *cls_ctr;          // [16] work-item counters of k_pgsb's shared-memory class launches (zeroed by k_border)
  int *wperm;            // [n_worlds] sweep order: worlds by period, longest first (k_border)
  int *sticky;           // [4] OR of every step's overflow masks since the last check, first world, spare
  double *io_in;         // [nb][6]  staged external wrench (gzcuda_batch_step_io)
  double *io_out;        // [nb][13] staged pos(3) quat(4) lvel(3) avel(3)
  double *set_stage;     // [nb][13] staged setter input (gzcuda_batch_set_body_state; NaN = leave as it is)
  double *jstage;        // [nj][6]  staged per-joint setter input (set_hinge_pd, set_joint_velocity_cmd)
  double *fb_stage;      // [max(nj, max_contact_cap)][12] joint / contact feedback staging (the getters allocate nothing)
  int *flag_stage;       // [nb] staged body flags (gzcuda_batch_get_body_flags)
};

}  // namespace gz
