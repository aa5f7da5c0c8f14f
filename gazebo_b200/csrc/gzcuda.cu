// gzcuda.cu — C ABI (include/gzcuda.h) over the sm_100a kernels: batch construction,
// per-step launch sequence, state exchange.  There is NO CPU fallback: every entry
// point that computes needs a CUDA device and fails with GZCUDA_E_CUDA without one.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>
#include <vector>

#include "../../include/gzcuda.h"
#include "k_collide.cuh"
#include "k_bpl.cuh"
#include "k_ray.cuh"
#include "k_pgs.cuh"

using namespace gz;

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) { g_err = msg; return code; }

#define CK(call)                                                                         \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess)                                                               \
      return fail(GZCUDA_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));    \
  } while (0)

#define GZ_PGSB_NCLS 8
struct gzcuda_batch;
extern "C" { static int check_sticky(gzcuda_batch *b); }
struct gzcuda_batch {
  DevBatch d;
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t stream_copy = nullptr;   // gzcuda_batch_step_log: device -> host copies of the state frames, beside the steps
  cudaEvent_t ev_frame[2] = {nullptr, nullptr}, ev_copied[2] = {nullptr, nullptr};
  double *log_frame[2] = {nullptr, nullptr};   // two device staging frames [nb][13]
  cudaStream_t stream_hi = nullptr;   // the long-period part of k_pgs when it is launched in two parts (GZCUDA_PGS_RSPLIT)
  cudaEvent_t ev_join_hi = nullptr;
  int pgs_rsplit = 0;
  cudaStream_t stream_chain[GZ_CHAIN_NCLS] = {};   // the chain-mode PGS launches (one per shared-memory size class) run
  cudaEvent_t ev_fork = nullptr, ev_join[GZ_CHAIN_NCLS] = {};   // beside the regular one: fork / join by events
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, evs[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  cudaStream_t stream_aux = nullptr;      // independent kernels of a step run beside the main stream (fork / join by events)
  cudaEvent_t ev_aux_fork = nullptr, ev_aux_join = nullptr;
  std::vector<void *> allocs;
  int *h_sticky = nullptr;            // pinned copy of DevBatch::sticky, read back at every synchronising call
  std::vector<WorldInfo> h_world;
  std::vector<int> h_joint_type;
  float last_ms = 0;
  long last_launches = 0;
  bool stage_valid = false;
  bool have_pd = false;
  bool have_velcmd = false;
  size_t pgs_smem = 0, pgs_chain_smem[GZ_CHAIN_NCLS] = {};
  bool has_cylinders = false;         // picks the k_narrow instantiation with the cylinder-box and MPR colliders
  size_t bp_smem = 0;                 // k_broadphase: dynamic shared memory and CTA size
  int bp_threads = 128;
  PgsCfg pgs, pgs_chain[GZ_CHAIN_NCLS];
  long row_tot = 0;                   // rows of all worlds' slabs (runrec)
  bool bp_large = false;              // worlds too large for the one-CTA broadphase: k_bpl_* (sort, sweep, scan, emit)
  BplBuf bpl = {};
  int pgsl_grid = 0;                  // CTAs of k_pgsl's cooperative grid
  unsigned long long *pgsd_slots = nullptr;   // k_pgsd's exchange slots
  unsigned long long pgsd_launches = 0;       // launches since the slots were last cleared (tags are 32 bits)
  unsigned long long pgsd_wrap = 0xFFFFFFFFull;   // tags in use before the slots are cleared (test switch GZCUDA_PGSD_WRAP: a small number)
  int pgs_large_mode = 1;             // large worlds: 0 k_pgsl (rounds + grid barrier), 1 k_pgsd (dataflow through L2), 2 k_pgsc (+ cluster shared memory)
  int pgsc_grid = 0, pgsc_cluster = 1, pgsc_glob_every = 4, pgsc_sleep = 0;
  size_t pgsc_smem = 0;
  void *ray_buf = nullptr;            // gzcuda_batch_ray_query's device buffers
  int ray_cap = 0;
  std::vector<std::pair<char *, size_t>> guards;   // GZCUDA_GUARD: the guard regions behind the device buffers
  bool pgsm = false;                  // k_pgsm: a world of up to 2048 bodies on one cluster, on chip (k_pgsd when it does not fit)
  int pgsm_cluster = 1, pgsm_threads = 128;
  size_t pgsm_smem = 0;
  int *pgsm_done = nullptr;
  int bsched_t = 128, bp_t_small = 128;   // (developer switches GZCUDA_BSCHED_T, GZCUDA_BP_T)
  int small_t = 128;                  // threads of the one-CTA-per-world kernels (k_contacts, k_assemble1)
  bool large_kernels = false;         // k_contacts_large / k_assemble1_large (cooperative grid) instead of the one-CTA-per-world kernels
  int large_grid = 0;
  int *large_part = nullptr;
  int bsl_grid = 0;                   // CTAs of k_bsched_large's cooperative grid
  PgsbCfg pgsbc[GZ_PGSB_NCLS];        // the shared-memory classes of the one-warp CTAs, ascending budgets
  size_t pgsbc_smem[GZ_PGSB_NCLS] = {};
  int pgsbc_grid[GZ_PGSB_NCLS] = {};         // persistent warps of each class: what the GPU can hold of them at once
  int pgsb_ncls = 1;
  cudaStream_t pgsb_stream[GZ_PGSB_NCLS] = {};   // one launch per class, side by side
  cudaEvent_t pgsb_join[GZ_PGSB_NCLS] = {};
  PgsbCfg pgsb;                       // body-owner sweeps (k_pgsb.cuh): lane layout, CTA shape, shared memory
  bool pgsb_multi = false;
  int pgsb_threads = 32, pgsb_ctas = 0;
  size_t pgsb_smem = 0, bsched_smem = 0;
  bool profiling = false;             // record an event pair around every k_pgs launch
  bool reg_first = false;             // experiment: k_pgs before (and above) the small chain-mode launches
  std::vector<cudaEvent_t> pgs_ev;    // 2 per step of the last step call
  int pgs_pairs = 0;
};

// ---- host restatement of dInvertPDMatrix for n = 3 (deps/opende/src/matrix.cpp:124-228) ----
static bool invert_pd3(const double A[9], double Ainv[9]) {
  double L[9];
  memcpy(L, A, sizeof(L));
  double recip[3];
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < i; j++) {
      double sum = L[3 * i + j];
      for (int k = 0; k < j; k++) sum -= L[3 * i + k] * L[3 * j + k];
      L[3 * i + j] = sum * recip[j];
    }
    double sum = L[3 * i + i];
    for (int k = 0; k < i; k++) sum -= L[3 * i + k] * L[3 * i + k];
    if (sum <= 0.0) return false;
    double s = sqrt(sum);
    L[3 * i + i] = s;
    recip[i] = 1.0 / s;
  }
  for (int c = 0; c < 3; c++) {
    double X[3] = {0, 0, 0}, y[3];
    X[c] = 1.0;
    for (int i = 0; i < 3; i++) {
      double sum = 0.0;
      for (int k = 0; k < i; k++) sum += L[3 * i + k] * y[k];
      y[i] = (X[i] - sum) / L[3 * i + i];
    }
    for (int i = 2; i >= 0; i--) {
      double sum = 0.0;
      for (int k = i + 1; k < 3; k++) sum += L[3 * k + i] * X[k];
      X[i] = (y[i] - sum) / L[3 * i + i];
    }
    for (int k = 0; k < 3; k++) Ainv[3 * k + c] = X[k];
  }
  return true;
}

static inline int cdiv_host(int a, int b) { return (a + b - 1) / b; }

template <class T>
static int dev_alloc(gzcuda_batch *b, T **p, size_t n) {
  if (n == 0) n = 1;
  // developer switch GZCUDA_GUARD=1 (compute-sanitizer is not available on the GPU boxes): every buffer is followed by
  // a guard of 0x7F bytes - an index read from beyond a buffer is then a wild one and faults at once instead of
  // reading a zero that happens to work, and gzcuda_batch_destroy reports guards that were written to
  static const bool guard = getenv("GZCUDA_GUARD") && atoi(getenv("GZCUDA_GUARD")) != 0;
  const size_t bytes = n * sizeof(T), gb = guard ? 4096 : 0;
  CK(cudaMalloc((void **)p, bytes + gb));
  CK(cudaMemset(*p, 0, bytes));
  if (guard) {
    CK(cudaMemset(reinterpret_cast<char *>(*p) + bytes, 0x7F, gb));
    b->guards.push_back(std::make_pair(reinterpret_cast<char *>(*p) + bytes, gb));
  }
  b->allocs.push_back(*p);
  return 0;
}
template <class T>
static int dev_upload(gzcuda_batch *b, T **p, const std::vector<T> &v) {
  int rc = dev_alloc(b, p, v.size());
  if (rc) return rc;
  if (!v.empty()) CK(cudaMemcpy(*p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  return 0;
}

extern "C" {

int gzcuda_version(void) { return 100; }
const char *gzcuda_last_error(void) { return g_err.c_str(); }
int gzcuda_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

void gzcuda_batch_destroy(gzcuda_batch *b) {
  if (!b) return;
  cudaSetDevice(b->device);
  if (b->stream) cudaStreamSynchronize(b->stream);
  if (!b->guards.empty()) {   // GZCUDA_GUARD: a guard that is no longer all 0x7F was written by somebody
    std::vector<unsigned char> h(4096);
    int bad = 0;
    for (auto &g : b->guards) {
      if (cudaMemcpy(h.data(), g.first, g.second, cudaMemcpyDeviceToHost) != cudaSuccess) { bad = -1; break; }
      for (size_t k = 0; k < g.second; k++) if (h[k] != 0x7F) { bad++; break; }
    }
    if (bad) { fprintf(stderr, "gzcuda guard: %d of %zu buffers were written beyond their end\n", bad, b->guards.size()); fflush(stderr); }
  }
  for (void *p : b->allocs) cudaFree(p);
  if (b->h_sticky) cudaFreeHost(b->h_sticky);
  if (b->ev0) cudaEventDestroy(b->ev0);
  if (b->ev1) cudaEventDestroy(b->ev1);
  for (auto &e : b->evs) if (e) cudaEventDestroy(e);
  for (auto &e : b->pgs_ev) cudaEventDestroy(e);
  if (b->ev_fork) cudaEventDestroy(b->ev_fork);
  for (auto &e : b->ev_join) if (e) cudaEventDestroy(e);
  for (auto &st : b->stream_chain) if (st) cudaStreamDestroy(st);
  for (auto &st : b->pgsb_stream) if (st) cudaStreamDestroy(st);
  for (auto &e : b->pgsb_join) if (e) cudaEventDestroy(e);
  if (b->stream_aux) cudaStreamDestroy(b->stream_aux);
  if (b->ev_aux_fork) cudaEventDestroy(b->ev_aux_fork);
  if (b->ev_aux_join) cudaEventDestroy(b->ev_aux_join);
  if (b->stream_copy) cudaStreamDestroy(b->stream_copy);
  for (auto &e : b->ev_frame) if (e) cudaEventDestroy(e);
  for (auto &e : b->ev_copied) if (e) cudaEventDestroy(e);
  for (auto &f : b->log_frame) if (f) cudaFree(f);
  if (b->stream_hi) cudaStreamDestroy(b->stream_hi);
  if (b->ev_join_hi) cudaEventDestroy(b->ev_join_hi);
  if (b->stream) cudaStreamDestroy(b->stream);
  delete b;
}

// the instantiation of the sweep kernels for (timeline profiling, box friction model)
typedef void (*PgsKernel)(DevBatch, PgsCfg);
typedef void (*PgsChainKernel)(DevBatch, PgsCfg, int);
static PgsKernel pgs_kernel(bool prof, bool box) {
  return prof ? (box ? k_pgs<true, true> : k_pgs<true, false>) : (box ? k_pgs<false, true> : k_pgs<false, false>);
}
static PgsChainKernel pgs_chain_kernel(bool prof, bool box) {
  return prof ? (box ? k_pgs_chain<true, true> : k_pgs_chain<true, false>)
              : (box ? k_pgs_chain<false, true> : k_pgs_chain<false, false>);
}

static int build(gzcuda_batch *b, const gz_scene *sc, const gz_params *p) {
  const int W = sc->n_worlds;
  const int nb = sc->body_off[W], ng = sc->geom_off[W], nj = sc->joint_off[W];
  DevBatch &d = b->d;
  memset(&d, 0, sizeof(d));
  d.n_worlds = W; d.nb = nb; d.ng = ng; d.nj = nj; d.p = *p; d.step_index = 0;
  if (const char *e = getenv("GZCUDA_TEST_STEP_INDEX0")) d.step_index = atol(e);   // test hook: start near the record epochs' wrap

  std::vector<Body> hb(nb);
  std::vector<Geom> hg(ng);
  std::vector<Joint> hj(nj);
  std::vector<double> hjx((size_t)(GZ_JP_COUNT - GZ_JP_HOT) * nj), hgx((size_t)(GZ_SF_COUNT - GZ_SF_HOT) * ng);   // the cold slots (joint_ext, geom_ext)
  std::vector<WorldInfo> hw(W);
  std::vector<int> bj;                  // CSR payload
  std::vector<std::vector<int>> adj(nb);

  for (int w = 0; w < W; w++) {
    for (int j = sc->joint_off[w]; j < sc->joint_off[w + 1]; j++) {
      Joint &J = hj[j];
      memset(&J, 0, sizeof(J));
      J.type = sc->joint_type[j];
      if (J.type < GZ_JOINT_HINGE || J.type > GZ_JOINT_SCREW)
        return fail(GZCUDA_E_INVALID, "unknown joint type");
      if ((J.type == GZ_JOINT_HINGE2 || J.type == GZ_JOINT_GEARBOX) && sc->joint_body2[j] < 0)
        return fail(GZCUDA_E_INVALID, "hinge2 and gearbox joints need two bodies (joints/hinge2.cpp:31, gearbox.cpp:67-68)");
      if (J.type == GZ_JOINT_SCREW && sc->joint_param[GZ_JP_COUNT * j + GZ_JP_FMAX] > 0)
        return fail(GZCUDA_E_INVALID, "a powered screw joint (fmax > 0) is not supported: the reference writes its row behind the joint's rows (screw.cpp:55-110, :419)");
      const int nbw = sc->body_off[w + 1] - sc->body_off[w];
      if (sc->joint_body1[j] < 0 || sc->joint_body1[j] >= nbw || sc->joint_body2[j] >= nbw)
        return fail(GZCUDA_E_INVALID, "joint body index out of range");
      J.b1 = sc->body_off[w] + sc->joint_body1[j];
      J.b2 = sc->joint_body2[j] >= 0 ? sc->body_off[w] + sc->joint_body2[j] : -1;
      for (int k = 0; k < 3; k++) {
        J.anchor1[k] = sc->joint_anchor1[3 * j + k]; J.anchor2[k] = sc->joint_anchor2[3 * j + k];
        J.axis1[k] = sc->joint_axis1[3 * j + k]; J.axis2[k] = sc->joint_axis2[3 * j + k];
      }
      for (int k = 0; k < 4; k++) J.qrel[k] = sc->joint_qrel[4 * j + k];
      for (int k = 0; k < GZ_JP_HOT; k++) J.param[k] = sc->joint_param[GZ_JP_COUNT * j + k];
      for (int k = GZ_JP_HOT; k < GZ_JP_COUNT; k++) hjx[(size_t)(GZ_JP_COUNT - GZ_JP_HOT) * j + k - GZ_JP_HOT] = sc->joint_param[GZ_JP_COUNT * j + k];
      J.world = w;
      adj[J.b1].push_back(j);
      if (J.b2 >= 0) adj[J.b2].push_back(j);
    }
  }
  for (int w = 0; w < W; w++) {
    for (int i = sc->body_off[w]; i < sc->body_off[w + 1]; i++) {
      Body &B = hb[i];
      memset(&B, 0, sizeof(B));
      for (int k = 0; k < 3; k++) {
        B.pos[k] = sc->body_pos[3 * i + k]; B.lvel[k] = sc->body_lvel[3 * i + k]; B.avel[k] = sc->body_avel[3 * i + k];
      }
      for (int k = 0; k < 4; k++) B.q[k] = sc->body_quat[4 * i + k];
      normalize4(B.q);          // dBodySetQuaternion (ode.cpp:414-427)
      q_to_R(B.q, B.R);
      B.mass = sc->body_mass[i];
      if (!(B.mass > 0)) return fail(GZCUDA_E_INVALID, "body mass must be > 0");
      B.invMass = 1.0 / B.mass;
      const double *I = sc->body_inertia + 6 * i;   // dMassSetParameters (mass.cpp)
      B.I[0] = I[0]; B.I[4] = I[1]; B.I[8] = I[2];
      B.I[1] = B.I[3] = I[3]; B.I[2] = B.I[6] = I[4]; B.I[5] = B.I[7] = I[5];
      if (!invert_pd3(B.I, B.invI)) {                // dBodySetMass falls back to identity (ode.cpp:535-538)
        for (int k = 0; k < 9; k++) B.invI[k] = (k % 4 == 0) ? 1.0 : 0.0;
      }
      B.lin_damp = sc->body_damping[2 * i]; B.ang_damp = sc->body_damping[2 * i + 1];
      B.max_vel = sc->body_contactp[2 * i]; B.min_depth = sc->body_contactp[2 * i + 1];
      B.flags = sc->body_flags[i];
      B.adis_stepsleft = p->adis_idle_steps;   // dBodySetAutoDisableDefaults (ode.cpp:925-941)
      B.adis_timeleft = p->adis_idle_time;
      B.world = w;
      B.joint_begin = (int)bj.size();
      for (int j : adj[i]) bj.push_back(j);   // ascending joint index (joints were visited in order)
      B.joint_end = (int)bj.size();
    }
  }
  int max_nb = 0, max_ng = 0, max_nj = 0, max_pc = 0, max_cc = 0, max_rc = 0;
  bool any_tors = false;
  long pair_tot = 0, contact_tot = 0, row_tot = 0, slot_tot = 0;
  for (int w = 0; w < W; w++) {
    const int g0 = sc->geom_off[w], g1 = sc->geom_off[w + 1];
    int nspace = 0;
    for (int g = g0; g < g1; g++) nspace = std::max(nspace, sc->geom_space[g] + 1);
    std::vector<unsigned> scat(nspace, 0), scol(nspace, 0);
    for (int g = g0; g < g1; g++) {   // the last geom added to a space leaves its bits on the space
      scat[sc->geom_space[g]] = sc->geom_category_bits[g];
      scol[sc->geom_space[g]] = sc->geom_collide_bits[g];
    }
    for (int g = g0; g < g1; g++) {
      Geom &G = hg[g];
      memset(&G, 0, sizeof(G));
      G.type = sc->geom_type[g];
      if (G.type < GZ_GEOM_SPHERE || G.type > GZ_GEOM_PLANE) return fail(GZCUDA_E_INVALID, "unknown geom type");
      const int nbw = sc->body_off[w + 1] - sc->body_off[w];
      if (sc->geom_body[g] >= nbw) return fail(GZCUDA_E_INVALID, "geom body index out of range");
      G.body = sc->geom_body[g] >= 0 ? sc->body_off[w] + sc->geom_body[g] : -1;
      G.space = sc->geom_space[g];
      G.world = w;
      for (int k = 0; k < 4; k++) G.dims[k] = sc->geom_dims[4 * g + k];
      if (G.type == GZ_GEOM_PLANE) {   // make_sure_plane_normal_has_unit_length (plane.cpp:49-65)
        double *pp = G.dims;
        double l = pp[0] * pp[0] + pp[1] * pp[1] + pp[2] * pp[2];
        if (l > 0) { l = recip_sqrt(l); pp[0] *= l; pp[1] *= l; pp[2] *= l; pp[3] *= l; }
        else { pp[0] = 1; pp[1] = 0; pp[2] = 0; pp[3] = 0; }
        G.body = -1;
      }
      for (int k = 0; k < 3; k++) { G.opos[k] = sc->geom_pos[3 * g + k]; G.fdir1[k] = sc->geom_fdir1[3 * g + k]; }
      if (sc->geom_rot) for (int k = 0; k < 9; k++) G.oR[k] = sc->geom_rot[9 * g + k];
      else q_to_R(sc->geom_quat + 4 * g, G.oR);   // dGeomSetOffsetQuaternion / dGeomSetQuaternion: dQtoR, no normalisation
      for (int k = 0; k < GZ_SF_HOT; k++) G.surf[k] = sc->geom_surf[GZ_SF_COUNT * g + k];
      for (int k = GZ_SF_HOT; k < GZ_SF_COUNT; k++) hgx[(size_t)(GZ_SF_COUNT - GZ_SF_HOT) * g + k - GZ_SF_HOT] = sc->geom_surf[GZ_SF_COUNT * g + k];
      G.em_on = sc->geom_surf[GZ_SF_COUNT * g + GZ_SF_ELASTIC_MODULUS] > 0 ? 1 : 0;
      {   // bit 0: mu3 > 0 (needed on both surfaces of a pair); bit 1: a radius that can switch the row on (on either)
        const double *sf = sc->geom_surf + (size_t)GZ_SF_COUNT * g;
        G.tors_on = (sf[GZ_SF_MU3] > 0 ? 1 : 0) | ((sf[GZ_SF_PATCH_RADIUS] > 0 || sf[GZ_SF_USE_PATCH_RADIUS] == 0) ? 2 : 0);
      }
      G.cat = sc->geom_category_bits[g]; G.col = sc->geom_collide_bits[g];
      G.scat = scat[G.space]; G.scol = scol[G.space];
      G.bitmask = sc->geom_collide_bitmask[g];
      G.max_contacts = sc->geom_max_contacts[g];
      if (G.type == GZ_GEOM_CYLINDER) b->has_cylinders = true;
    }
    // static upper bound on candidate pairs of this world
    const int n = g1 - g0;
    long np = 0;
    const long np_limit = 32L * n;
    for (int i = g0; i < g1 && np <= np_limit; i++)
      for (int j = i + 1; j < g1; j++) {
        const Geom &a = hg[i], &c = hg[j];
        if (a.space == c.space) continue;
        if (((a.scat & c.scol) || (c.scat & a.scol)) == 0) continue;
        if (a.body == c.body && a.body >= 0) continue;
        if (((a.cat & c.col) || (c.cat & a.col)) == 0) continue;
        np++;
      }
    WorldInfo &wi = hw[w];
    wi.body0 = sc->body_off[w]; wi.nb = sc->body_off[w + 1] - wi.body0;
    wi.geom0 = g0; wi.ng = n;
    wi.joint0 = sc->joint_off[w]; wi.nj = sc->joint_off[w + 1] - wi.joint0;
    wi.pair_cap = (int)std::max(1L, std::min(np, np_limit));
    wi.contact_cap = std::min(8 * wi.pair_cap, std::max(64, 24 * n));
    int n_tors = 0, tors_radius = 0;   // a pair has a torsional row only when BOTH surfaces have mu3 > 0 (the smaller one counts)
    for (int i = g0; i < g1; i++) { n_tors += hg[i].tors_on & 1; tors_radius |= hg[i].tors_on & 2; }
    wi.rpc = (n_tors >= 2 && tors_radius) ? 4 : 3;
    any_tors = any_tors || wi.rpc == 4;
    wi.row_cap = 6 * wi.nj + wi.rpc * wi.contact_cap;
    // test hooks: shrink the slabs so that a step overflows them (tests/test_parity_gpu.py::test_capacity_overflow_is_sticky)
    if (const char *e = getenv("GZCUDA_TEST_PAIR_CAP")) wi.pair_cap = std::max(1, std::min(wi.pair_cap, atoi(e)));
    if (const char *e = getenv("GZCUDA_TEST_CONTACT_CAP")) wi.contact_cap = std::max(1, std::min(wi.contact_cap, atoi(e)));
    if (const char *e = getenv("GZCUDA_TEST_ROW_CAP")) wi.row_cap = std::max(1, std::min(wi.row_cap, atoi(e)));
    wi.pair0 = (int)pair_tot; wi.contact0 = (int)contact_tot; wi.row0 = (int)row_tot;
    wi.slot_cap = 0; wi.slot0 = 0;
    // schedule slots a world can need: about twice its rows (periods are padded to a ladder and to k lanes)
    pair_tot += wi.pair_cap; contact_tot += wi.contact_cap; row_tot += wi.row_cap; slot_tot += 2 * wi.row_cap + 64;
    if (row_tot > 0x7fffffffL || slot_tot > 0x7fffffffL) return fail(GZCUDA_E_INVALID, "batch too large (row slab index overflow)");
    max_nb = std::max(max_nb, wi.nb); max_ng = std::max(max_ng, wi.ng); max_nj = std::max(max_nj, wi.nj);
    max_pc = std::max(max_pc, wi.pair_cap); max_cc = std::max(max_cc, wi.contact_cap);
    max_rc = std::max(max_rc, wi.row_cap);
  }
  d.max_nb = max_nb; d.max_ng = max_ng; d.max_nj = max_nj;
  d.max_pair_cap = max_pc; d.max_contact_cap = max_cc; d.max_row_cap = max_rc;
  if (max_nb >= GZ_ENT_NOPART) return fail(GZCUDA_E_INVALID, "worlds are limited to 65534 bodies (16-bit partner ids in the PGS entries)");
  b->h_world = hw;

  int rc;
  if ((rc = dev_upload(b, &d.world, hw))) return rc;
  if ((rc = dev_alloc(b, &d.wstep, (size_t)W))) return rc;
  if ((rc = dev_upload(b, &d.body, hb))) return rc;
  if ((rc = dev_alloc(b, &d.bstep, (size_t)nb))) return rc;
  if ((rc = dev_upload(b, &d.body_joint, bj))) return rc;
  if ((rc = dev_upload(b, &d.geom, hg))) return rc;
  if ((rc = dev_upload(b, &d.joint, hj))) return rc;
  if ((rc = dev_upload(b, &d.joint_ext, hjx))) return rc;
  if ((rc = dev_upload(b, &d.geom_ext, hgx))) return rc;
  if ((rc = dev_alloc(b, &d.pair, (size_t)2 * pair_tot))) return rc;
  if ((rc = dev_alloc(b, &d.pair_nc, (size_t)pair_tot))) return rc;
  if ((rc = dev_alloc(b, &d.pair_contact, (size_t)7 * GZ_PAIR_MAXC * pair_tot))) return rc;
  if ((rc = dev_alloc(b, &d.contact, (size_t)contact_tot))) return rc;
  if ((rc = dev_alloc(b, &d.row, (size_t)row_tot))) return rc;
  if ((rc = dev_alloc(b, &d.order, (size_t)row_tot))) return rc;
  if ((rc = dev_alloc(b, &d.scratch, (size_t)W * ((size_t)7 * max_rc)))) return rc;
  if ((rc = dev_alloc(b, &d.runrec, (size_t)GZ_REC * row_tot))) return rc;
  b->row_tot = row_tot;
  // record pool of the PGS warps: every world's slots, rounded up to whole warps (16 lane pairs per
  // round), plus one partly filled warp per bucket in use; clamped to a share of device memory
  {
    size_t free_b = 0, total_b = 0;
    CK(cudaMemGetInfo(&free_b, &total_b));
    long want = 2 * slot_tot + 64L * 1024 * 16;
    const long most = (long)(free_b / 3 / (GZ_REC * sizeof(double)));
    d.prec_cap = std::max(16L * GZ_RMAX, std::min(want, most));
    if (d.prec_cap > 0x7fffffffL) d.prec_cap = 0x7fffffffL;
  }
  if ((rc = dev_alloc(b, &d.prec, (size_t)GZ_REC * d.prec_cap))) return rc;
  if ((rc = dev_alloc(b, &d.lamg, (size_t)2 * (3 * (size_t)row_tot + 128 * (size_t)W)))) return rc;   // per world 2 x (3 row_cap + 128)
  if ((rc = dev_alloc(b, &d.wsched, (size_t)8 * W))) return rc;
  if ((rc = dev_alloc(b, &d.bk_count, (size_t)GZ_NBUCKET))) return rc;
  if ((rc = dev_alloc(b, &d.bk_tab, (size_t)4 * (GZ_NBUCKET + 1)))) return rc;
  if ((rc = dev_alloc(b, &d.warp_world, (size_t)16 * (W + GZ_NBUCKET)))) return rc;
  if ((rc = dev_alloc(b, &d.acc, (size_t)4))) return rc;
  if ((rc = dev_alloc(b, &d.prof, (size_t)16 + 2 * (size_t)W))) return rc;
  d.pgs_prof = getenv("GZCUDA_PGS_PROF") ? atoi(getenv("GZCUDA_PGS_PROF")) : 0;
  if ((rc = dev_alloc(b, &d.io_in, (size_t)6 * nb))) return rc;
  if ((rc = dev_alloc(b, &d.io_out, (size_t)13 * nb))) return rc;
  if ((rc = dev_alloc(b, &d.set_stage, (size_t)13 * nb))) return rc;
  if ((rc = dev_alloc(b, &d.jstage, (size_t)6 * nj))) return rc;
  if ((rc = dev_alloc(b, &d.fb_stage, (size_t)12 * std::max(nj, max_cc)))) return rc;
  if ((rc = dev_alloc(b, &d.flag_stage, (size_t)nb))) return rc;
  b->h_joint_type.resize(nj);
  for (int j = 0; j < nj; j++) b->h_joint_type[j] = hj[j].type;
  if ((rc = dev_alloc(b, &d.sticky, (size_t)4))) return rc;
  CK(cudaMallocHost((void **)&b->h_sticky, 4 * sizeof(int)));

  // Body-owner PGS (k_pgsb.cuh): the default for worlds of up to 256 bodies.  Entry pool: a row has one entry
  // per body, so 2 x rows is the worst case; clamped to a quarter of the free memory (k_bsched flags a step
  // that does not fit).
  // <= 256 bodies: k_pgsb (warp or CTA per world); beyond: the body-owner sweeps over the whole GPU (k_pgsd: dataflow
  // through L2 slots, one launch per world).  Batches of 8+ worlds of 257..512 bodies keep round 1's row-slot kernels
  // (all worlds in one launch); GZCUDA_PGS_IMPL=0 forces them for any batch of worlds up to 512 bodies.
  d.pgs_impl = (max_nb <= 256) ? 1 : ((max_nb <= 512 && W >= 8) ? 0 : 2);
  // one-CTA-per-world kernels: a world of a few bodies keeps a few lanes busy, and the smaller the CTA the more worlds
  // are resident at once (measured at C2: 32 threads 1.305 ms/step, 128 threads 1.327)
  b->small_t = max_nb <= 16 ? 32 : (max_nb <= 48 ? 64 : 128);
  if (const char *e = getenv("GZCUDA_SMALL_T")) { const int v = atoi(e); if (v == 32 || v == 64 || v == 128 || v == 256) b->small_t = v; }
  if (const char *e = getenv("GZCUDA_BSCHED_T")) { const int v = atoi(e); if (v == 32 || v == 64 || v == 128) b->bsched_t = v; }
  if (const char *e = getenv("GZCUDA_BP_T")) { const int v = atoi(e); if (v == 32 || v == 64 || v == 128) b->bp_t_small = v; }
  d.islands_hook = max_nb > 512;
  if (const char *e = getenv("GZCUDA_ISLANDS_HOOK")) d.islands_hook = atoi(e) != 0;   // developer / test switch
  if (const char *e = getenv("GZCUDA_PGS_IMPL")) {
    const int v = atoi(e);
    if (v == 0 && max_nb <= 512) d.pgs_impl = 0;
    if (v == 2) d.pgs_impl = 2;
  }
  if (any_tors && d.pgs_impl == 0)   // (the round-1 row-owner path, batches of 257..512-body worlds, does not know the four-row contact layout)
    d.pgs_impl = 2;
  if (d.pgs_impl == 2) {
    // one lane per (body, solve) of the largest world, all CTAs resident (cooperative launch)
    int n_sm = 148, per_sm = 0;
    CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, b->device));
    // developer switches: GZCUDA_PGSL_MODE = 0 round-synchronous sweeps (k_pgsl), 1 dataflow through L2 (k_pgsd, the
    // default), 2 dataflow with inboxes in cluster shared memory (k_pgsc: measured, not faster - DESIGN.md §4)
    b->pgs_large_mode = 1;
    if (const char *em = getenv("GZCUDA_PGSL_MODE")) b->pgs_large_mode = std::max(0, std::min(2, atoi(em)));
    if (const char *eb = getenv("GZCUDA_PGSL_BARRIER")) if (atoi(eb) == 1) b->pgs_large_mode = 0;
    d.pgs_dataflow = b->pgs_large_mode >= 1;
    if (d.pgs_dataflow) { CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_pgsd<false>, 256, 0)); }
    else CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_pgsl<false>, 256, 0));
    b->pgsl_grid = cdiv_host(2 * std::max(max_nb, 1), 256);
    int bsl_per_sm = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bsl_per_sm, k_bsched_large, 256, 0));
    b->bsl_grid = std::max(1, std::min(n_sm * std::min(bsl_per_sm, 2), cdiv_host((int)std::max(row_tot, 1L), 256)));
    if (b->pgsl_grid > n_sm * per_sm) return fail(GZCUDA_E_INVALID, "world too large for one cooperative grid");
    if ((rc = dev_alloc(b, &d.xch_g, (size_t)2 * b->pgsl_grid * 256 + (size_t)9 * b->pgsl_grid + 8))) return rc;
    CK(cudaMemset(d.xch_g, 0, ((size_t)2 * b->pgsl_grid * 256 + (size_t)9 * b->pgsl_grid + 8) * sizeof(double)));
    if ((rc = dev_alloc(b, &d.bs_clist, (size_t)2 * contact_tot + 2))) return rc;
    // the one-CTA-per-world kernels of the step spread over a cooperative grid: worlds beyond 512 bodies, and batches of
    // up to four worlds of more than 256 (C1, 500 boxes in one world: collide + assemble 0.45 -> 0.26 ms); the test switch
    // GZCUDA_LARGE_KERNELS forces them onto small worlds
    b->large_kernels = max_nb > 512 || W <= 4;
    if (const char *el = getenv("GZCUDA_LARGE_KERNELS")) b->large_kernels = atoi(el) != 0;
    if (b->large_kernels) {
      int p1 = 0, p2 = 0;
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&p1, k_contacts_large, 256, 0));
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&p2, k_assemble1_large, 256, 0));
      b->large_grid = n_sm * std::max(1, std::min(2, std::min(p1, p2)));
      if (std::min(p1, p2) < 1) b->large_kernels = false;
      else if ((rc = dev_alloc(b, &b->large_part, (size_t)b->large_grid + 1))) return rc;
    }
    if ((rc = dev_alloc(b, &d.bs_glob, (size_t)W * 5 * std::max(max_nb, 1)))) return rc;
    if (any_tors && (rc = dev_alloc(b, &d.bs_cntt, (size_t)std::max(nb, 1)))) return rc;
  }
  if (d.pgs_impl >= 1) {
    b->bsched_smem = (size_t)5 * std::max(max_nb, 1) * sizeof(int);
    size_t free_b = 0, total_b = 0;
    CK(cudaMemGetInfo(&free_b, &total_b));
    const size_t per_ent = GZ_ENT * sizeof(double) + sizeof(int) + 3 * sizeof(double);
    long want = 2 * row_tot + 1024;
    const long most = (long)(free_b / 4 / per_ent);
    d.ent_cap = std::max(4096L, std::min(want, most));
    if (d.ent_cap > 0x7ffffff0L) d.ent_cap = 0x7ffffff0L;
    if ((rc = dev_alloc(b, &d.ent, (size_t)GZ_ENT * d.ent_cap))) return rc;
    {   // the pool's last entry: the all-zero dummy entry (no partner) of lanes without rows
      double dm[GZ_ENT] = {0};
      const int meta[2] = {0, GZ_ENT_NOPART};
      memcpy(&dm[19], meta, sizeof(double));
      CK(cudaMemcpy(d.ent + (size_t)GZ_ENT * (d.ent_cap - 1), dm, sizeof(dm), cudaMemcpyHostToDevice));
    }
    if ((rc = dev_alloc(b, &d.ent_row, (size_t)d.ent_cap))) return rc;
    if ((rc = dev_alloc(b, &d.ent_ad2, (size_t)d.ent_cap))) return rc;
    if ((rc = dev_alloc(b, &d.ent_dl, (size_t)2 * d.ent_cap))) return rc;
    if (d.pgs_impl == 2 && d.pgs_dataflow) {
      // k_pgsd: partner entries, and per entry 2 solves x 2 sweep parities x 16-byte slots (tag 0 = never written)
      if ((rc = dev_alloc(b, &d.ent_part, (size_t)d.ent_cap))) return rc;
      if ((rc = dev_alloc(b, &b->pgsd_slots, (size_t)8 * d.ent_cap))) return rc;
      CK(cudaMemset(b->pgsd_slots, 0, (size_t)8 * d.ent_cap * sizeof(unsigned long long)));
      if (const char *ew = getenv("GZCUDA_PGSD_WRAP")) b->pgsd_wrap = std::max(1000ull, (unsigned long long)atoll(ew));
    }
    if (d.pgs_impl == 2 && b->pgs_large_mode == 2) {
      // k_pgsc: CTAs of 8 warps, pgsc_bpw bodies per warp - as few as the resident warps allow (two CTAs of 256
      // threads per SM at ~120 registers): a warp is one instruction stream for all its bodies' rows; clusters of 8
      // CTAs (developer switch: 16); inboxes of 64 bytes per entry in shared memory; every cluster must be resident,
      // else the L2 dataflow kernel takes the world
      int n_sm = 148;
      CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, b->device));
      const int warps_res = n_sm * 2 * 8 * 7 / 8;   // (clusters do not pack every SM: leave an eighth)
      int bpw = 1;
      while (bpw < 16 && cdiv_host(std::max(max_nb, 1), bpw) > warps_res) bpw++;
      if (const char *eb = getenv("GZCUDA_PGSC_BPW")) bpw = std::max(1, std::min(16, atoi(eb)));
      d.pgsc_bpw = bpw;
      d.pgsc_bodies = 8 * bpw;
      const int ncta = cdiv_host(std::max(max_nb, 1), d.pgsc_bodies);
      int clmax = 8;
      if (const char *ec = getenv("GZCUDA_PGSC_CLUSTER")) { const int v = atoi(ec); if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16) clmax = v; }
      int cl = 1;
      while (cl < clmax && cl < ncta) cl <<= 1;
      b->pgsc_cluster = cl;
      b->pgsc_grid = cdiv_host(ncta, cl) * cl;
      const size_t smem_max = ncta <= n_sm ? (size_t)192 * 1024 : (size_t)100 * 1024;   // one or two CTAs per SM
      b->pgsc_smem = std::min(smem_max, (((size_t)64 * d.ent_cap + 1023) / 1024) * 1024);
      if (const char *eg = getenv("GZCUDA_PGSC_GLOB_EVERY")) { int v = std::max(1, atoi(eg)), p2 = 1; while (p2 < v) p2 <<= 1; b->pgsc_glob_every = p2; }
      if (const char *es = getenv("GZCUDA_PGSC_SLEEP")) b->pgsc_sleep = std::max(0, atoi(es));
      bool ok = b->pgsc_grid <= GZ_PGSC_MAXCTA + 8;
      for (int v = 0; v < 2 && ok; v++) {
        const void *fn = v ? (const void *)k_pgsc<true> : (const void *)k_pgsc<false>;
        ok = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b->pgsc_smem) == cudaSuccess;
        if (ok && cl > 8) ok = cudaFuncSetAttribute(fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess;
        if (!ok) break;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(b->pgsc_grid); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = b->pgsc_smem;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int n_clusters = 0;
        ok = cudaOccupancyMaxActiveClusters(&n_clusters, fn, &cfg) == cudaSuccess && n_clusters * cl >= b->pgsc_grid;
      }
      cudaGetLastError();
      if (!ok) b->pgs_large_mode = 1;
      else {
        if ((rc = dev_alloc(b, &d.cta_e0, (size_t)W * GZ_PGSC_TAB))) return rc;
        if ((rc = dev_alloc(b, &d.ent_aux, (size_t)d.ent_cap))) return rc;
      }
    }
    if (d.pgs_impl == 2 && b->pgs_large_mode == 1 && max_nb <= 2048 && !(getenv("GZCUDA_PGSM") && atoi(getenv("GZCUDA_PGSM")) == 0)) {
      // k_pgsm: 64 (up to 1024 bodies) or 128 bodies per CTA, one cluster of up to 16 CTAs (more than 8: non-portable
      // size, B200 allows it), the world's entries in shared memory
      int per = max_nb <= 1024 ? 64 : 128;
      if (const char *ep = getenv("GZCUDA_PGSM_PER")) { const int v = atoi(ep); if ((v == 32 || v == 64 || v == 128) && cdiv_host(std::max(max_nb, 1), v) <= 16) per = v; }
      const int ncta = cdiv_host(std::max(max_nb, 1), per);
      int cl = 1;
      while (cl < ncta) cl <<= 1;
      b->pgsm_cluster = cl; b->pgsm_threads = 2 * per;
      b->pgsm_smem = std::min((size_t)224 * 1024, (((size_t)GZ_PGSM_BYTES * d.ent_cap + 1023) / 1024) * 1024);
      bool ok = cl <= 16;
      for (int v = 0; v < 2 && ok; v++) {
        const void *fn = v ? (const void *)k_pgsm<true> : (const void *)k_pgsm<false>;
        ok = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b->pgsm_smem) == cudaSuccess;
        if (ok && cl > 8) ok = cudaFuncSetAttribute(fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess;
        if (!ok) break;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(cl); cfg.blockDim = dim3(b->pgsm_threads); cfg.dynamicSmemBytes = b->pgsm_smem;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int n_clusters = 0;
        ok = cudaOccupancyMaxActiveClusters(&n_clusters, fn, &cfg) == cudaSuccess && n_clusters >= 1;
      }
      cudaGetLastError();
      if (ok) {
        b->pgsm = true;
        d.pgsc_bodies = per; d.pgsc_bpw = 16;
        if ((rc = dev_alloc(b, &d.cta_e0, (size_t)W * GZ_PGSC_TAB))) return rc;
        if ((rc = dev_alloc(b, &b->pgsm_done, (size_t)W))) return rc;
      }
    }
    if ((rc = dev_alloc(b, &d.ent_cursor, (size_t)2))) return rc;
    if ((rc = dev_alloc(b, &d.bsched, (size_t)nb))) return rc;
    if ((rc = dev_alloc(b, &d.wsb, (size_t)W))) return rc;
    if ((rc = dev_alloc(b, &d.rk_count, (size_t)2 * GZ_PGSB_NKEY))) return rc;
    if ((rc = dev_alloc(b, &d.wkey, (size_t)W))) return rc;
    if ((rc = dev_alloc(b, &d.wperm, (size_t)W))) return rc;
    if ((rc = dev_alloc(b, &d.cls_ctr, (size_t)16))) return rc;
  }
  if (d.pgs_impl == 1) {
    int bp = 1;
    while (bp < max_nb) bp <<= 1;
    PgsbCfg &c = b->pgsb;
    c.lw = 2 * bp;
    b->pgsb_multi = c.lw > 32;
    c.wpc = b->pgsb_multi ? 1 : 32 / c.lw;
    b->pgsb_threads = b->pgsb_multi ? c.lw : 32;
    b->pgsb_ctas = cdiv_host(W, c.wpc);
    // Shared memory per CTA for staged entries (176 B each).  One-warp CTAs come in two classes: the small one
    // (7 CTAs per SM) takes the CTAs whose worlds fit it, the large one (3 per SM) the rest; what fits neither
    // reads the pool.  A multi-warp CTA (one world) has one class.
    // Shared memory per CTA for staged entries (176 B each).  One-warp CTAs come in classes of ascending budget,
    // one launch each (a CTA belongs to the smallest class that holds the entries of its worlds; what fits none
    // reads the pool); class q is sized so that 9 - q CTAs share an SM (233 KB, 1 KB reserved per CTA), the last
    // one takes a whole SM.  A multi-warp CTA (one world) has one class.
    const size_t ent_b = GZ_ENTS * sizeof(double);
    const size_t xch_b = b->pgsb_multi ? (size_t)2 * c.lw * sizeof(double) : 0;
    const size_t red_b = b->pgsb_multi ? (size_t)9 * c.lw * sizeof(double) : 0;
    size_t byt[GZ_PGSB_NCLS];
    b->pgsb_ncls = b->pgsb_multi ? 1 : GZ_PGSB_NCLS;
    for (int i = 0; i < GZ_PGSB_NCLS; i++) byt[i] = (i == GZ_PGSB_NCLS - 1) ? (size_t)200 * 1024 : ((size_t)233 * 1024) / (9 - i) - 1024 - 64;
    if (b->pgsb_multi) byt[0] = (size_t)128 * 1024;
    if (const char *e = getenv("GZCUDA_PGSB_KB")) {   // comma-separated, ascending: replaces the classes
      const char *q = e;
      int n = 0;
      while (n < GZ_PGSB_NCLS && *q) {
        byt[n++] = (size_t)std::min(200, std::max(1, atoi(q))) * 1024;
        while (*q && *q != ',') q++;
        if (*q == ',') q++;
      }
      if (!b->pgsb_multi && n > 0) b->pgsb_ncls = n;
    }
    size_t mx_b = 0;
    for (int i = 0; i < b->pgsb_ncls; i++) {
      b->pgsbc[i] = c;
      b->pgsbc[i].nres = (int)(byt[i] / ent_b) - 1;   // (one slot is the dummy entry)
      b->pgsbc[i].cls = b->pgsb_multi ? -1 : i;
      b->pgsbc[i].last = i == b->pgsb_ncls - 1;
      b->pgsbc_smem[i] = std::max((size_t)(b->pgsbc[i].nres + 1) * ent_b + xch_b, red_b);
      mx_b = std::max(mx_b, b->pgsbc_smem[i]);
      {
        int n_sm = 148;
        CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, b->device));
        const int per_sm = (int)std::min<size_t>(16, std::max<size_t>(1, ((size_t)233 * 1024) / (b->pgsbc_smem[i] + 1024)));
        b->pgsbc_grid[i] = std::max(1, std::min(b->pgsb_ctas, n_sm * per_sm));
      }
      if (i > 0) {   // class 0 runs on the batch's stream; the larger the class, the higher the priority of its stream
        int plo = 0, phi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&plo, &phi));
        CK(cudaStreamCreateWithPriority(&b->pgsb_stream[i], cudaStreamNonBlocking, std::min(plo, phi + (b->pgsb_ncls - 1 - i))));
        CK(cudaEventCreateWithFlags(&b->pgsb_join[i], cudaEventDisableTiming));
      }
    }
    b->pgsb = b->pgsbc[0];
    b->pgsb_smem = b->pgsbc_smem[0];
    const int mx_smem = (int)mx_b;
    for (auto f : {k_pgsb<false, false>, k_pgsb<false, true>, k_pgsb<true, false>, k_pgsb<true, true>}) {
      if (mx_smem > 48 * 1024) CK(cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, mx_smem));
      CK(cudaFuncSetAttribute(f, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    }
  }

  // Broadphase (k_collide.cuh): all pairs for small worlds, sort-and-sweep from bp_sap_min geoms on,
  // if the sorted keys and the n x n hit bits of the largest world fit in shared memory
  {
    d.bp_sap_min = 64;
    if (const char *e = getenv("GZCUDA_BP_SAP_MIN")) d.bp_sap_min = std::max(2, atoi(e));
    b->bp_smem = (size_t)(max_ng + 1) * sizeof(int);
    b->bp_threads = b->bp_t_small;
    if (max_ng >= d.bp_sap_min) {
      size_t npad = 32;
      while (npad < (size_t)max_ng) npad <<= 1;
      const size_t need = npad * 20 + ((size_t)max_ng + 1) * 4 + (size_t)max_ng * (((size_t)max_ng + 31) / 32) * 4 + 16;
      bool force_large = d.pgs_impl == 2 && W <= 4;   // a few worlds of hundreds of geoms: one CTA each leaves the GPU idle (C1: 0.26 -> 0.14 ms)
      if (const char *e = getenv("GZCUDA_BP_LARGE")) force_large = atoi(e) != 0;   // test hook
      if (need <= (size_t)220 * 1024 && !force_large) {
        b->bp_smem = std::max(b->bp_smem, need);
        b->bp_threads = 512;
        if (b->bp_smem > 48 * 1024)
          CK(cudaFuncSetAttribute(k_broadphase, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b->bp_smem));
      } else {
        // the largest world does not fit one CTA's shared memory: the multi-CTA broadphase (k_bpl.cuh) for every world
        b->bp_large = true;
        BplBuf &Q = b->bpl;
        const size_t nt = ((size_t)max_ng + 1 + GZ_BPL_TILE - 1) / GZ_BPL_TILE * GZ_BPL_TILE;   // whole tiles, at least one padding element
        const size_t nwords = ((size_t)max_ng + 31) / 32 + 1;
        Q.npad = (int)nt;
        Q.big = GZ_BPL_BIG;
        Q.item_cap = (int)std::min<size_t>(8 * nt, (size_t)1 << 30);
        if (const char *e = getenv("GZCUDA_BPL_ITEM_CAP")) Q.item_cap = std::max(0, std::min(Q.item_cap, atoi(e)));   // developer / test switches (a small table: everything through the queue)
        if (const char *e = getenv("GZCUDA_BPL_BIG")) Q.big = std::max(1, atoi(e));
        if ((rc = dev_alloc(b, &Q.key, nt))) return rc;
        if ((rc = dev_alloc(b, &Q.tkey, nt))) return rc;
        if ((rc = dev_alloc(b, &Q.maxx, (size_t)max_ng))) return rc;
        if ((rc = dev_alloc(b, &Q.idx, nt))) return rc;
        if ((rc = dev_alloc(b, &Q.tidx, nt))) return rc;
        if ((rc = dev_alloc(b, &Q.cnt, (size_t)max_ng + 1))) return rc;
        if ((rc = dev_alloc(b, &Q.fill, (size_t)max_ng))) return rc;
        if ((rc = dev_alloc(b, &Q.tmp, (size_t)std::max(max_pc, 1)))) return rc;
        if ((rc = dev_alloc(b, &Q.queue, (size_t)2 * max_ng + 2))) return rc;
        if ((rc = dev_alloc(b, &Q.ctl, (size_t)4 + GZ_BPL_NBIG))) return rc;
        if ((rc = dev_alloc(b, &Q.item, (size_t)std::max(Q.item_cap, 1)))) return rc;
        if ((rc = dev_alloc(b, &Q.yz, (size_t)4 * nt))) return rc;
        if ((rc = dev_alloc(b, &Q.bits, (size_t)GZ_BPL_NBIG * 2 * nwords))) return rc;
        b->bp_smem = 16;
      }
    }
    if (b->bp_smem > (size_t)220 * 1024) return fail(GZCUDA_E_INVALID, "world has too many geoms for the broadphase's row counts");
    if (b->bp_smem > 48 * 1024)
      CK(cudaFuncSetAttribute(k_broadphase, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b->bp_smem));
  }

  if (d.pgs_impl != 0) return 0;   // what follows configures round 1's row-slot kernels (worlds of 257..512 bodies)
  // PGS launch shape (k_pgs_sweep.cuh): one warp per CTA, a fixed shared-memory budget per CTA
  // that holds the accumulators of its worlds, their lambdas and as many row records as fit.
  d.sch_rcap = std::min(GZ_RMAX, 2 * std::max(max_rc, 1) + 64);
  // 32 candidate periods probed at once need 32 x (2 max_nb ints + sch_rpar shorts) of shared memory
  d.sch_rpar = (max_nb <= 16) ? 96 : 0;
  if (const char *e = getenv("GZCUDA_SCH_RPAR")) d.sch_rpar = std::max(0, std::min(512, atoi(e)));
  d.pgs_rt = 28; d.pgs_kmin = 1; d.pgs_kmax = 16; d.pgs_chain_min = 28;
  if (const char *e = getenv("GZCUDA_PGS_CHAIN_MIN")) d.pgs_chain_min = std::max(2, atoi(e));
  if (const char *e = getenv("GZCUDA_PGS_RT")) d.pgs_rt = std::max(1, atoi(e));
  if (const char *e = getenv("GZCUDA_PGS_KMIN")) d.pgs_kmin = std::max(1, std::min(16, atoi(e)));
  if (const char *e = getenv("GZCUDA_PGS_KMAX")) d.pgs_kmax = std::max(d.pgs_kmin, std::min(16, atoi(e)));
  for (int *kk : {&d.pgs_kmin, &d.pgs_kmax}) { int v = 1; while (2 * v <= *kk) v <<= 1; *kk = v; }
  PgsCfg &c = b->pgs;
  c.part = 0; c.rsplit = 0;
  c.nbp = std::max(max_nb, 1);
  size_t smem = 52 * 1024;
  if (const char *e = getenv("GZCUDA_PGS_SMEM_KB")) smem = (size_t)atoi(e) * 1024;
  const size_t ca_bytes = (size_t)12 * (c.nbp + 1) * sizeof(double);
  const size_t ring_bytes = (size_t)GZ_RING_DOUBLES * sizeof(double);
  smem = std::max(smem, ca_bytes + ring_bytes + 4096);
  if (smem > (size_t)227 * 1024) return fail(GZCUDA_E_INVALID, "PGS accumulators of one world exceed shared memory");
  c.wpw_max = (int)std::min<size_t>(16, std::max<size_t>(1, ((smem - ring_bytes) / 2) / ca_bytes));
  d.pgs_wpw_max = c.wpw_max;
  c.smem_doubles = (int)(smem / sizeof(double));
  b->pgs_smem = smem;
  if (b->pgs_smem > 48 * 1024) {
    for (int v = 0; v < 4; v++)
      CK(cudaFuncSetAttribute(pgs_kernel(v & 1, v & 2), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b->pgs_smem));
  }
  // chain-mode worlds (k_pgs_chain.cuh): one world per CTA, resident in shared memory; one launch per
  // size class (ascending), so that small worlds do not take a large world's share of an SM
  size_t csmem[GZ_CHAIN_NCLS] = {32 * 1024, 56 * 1024, 104 * 1024, 200 * 1024};
  if (const char *e = getenv("GZCUDA_PGS_CHAIN_KB")) {   // comma-separated, ascending
    const char *q = e;
    for (int i = 0; i < GZ_CHAIN_NCLS && *q; i++) {
      csmem[i] = (size_t)atoi(q) * 1024;
      while (*q && *q != ',') q++;
      if (*q == ',') q++;
    }
  }
  for (int q = 0; q < GZ_CHAIN_NCLS; q++) {
    csmem[q] = std::min<size_t>((size_t)227 * 1024, std::max<size_t>(csmem[q], 16 * 1024));
    b->pgs_chain[q] = c;
    b->pgs_chain[q].smem_doubles = (int)(csmem[q] / sizeof(double));
    d.pgs_chain_doubles[q] = b->pgs_chain[q].smem_doubles;
    b->pgs_chain_smem[q] = csmem[q];
  }
  if (!(getenv("GZCUDA_PGS_CARVEOUT") && atoi(getenv("GZCUDA_PGS_CARVEOUT")) == 0)) {
    // k_pgs and the k_pgs_chain launches share the SMs: ask for the largest shared-memory carve-out for all
    // of them, so that an SM configured for one can take CTAs of the other
    for (int v = 0; v < 4; v++) {
      CK(cudaFuncSetAttribute(pgs_kernel(v & 1, v & 2), cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
      CK(cudaFuncSetAttribute(pgs_chain_kernel(v & 1, v & 2), cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    }
  }
  {
    const size_t mx = *std::max_element(csmem, csmem + GZ_CHAIN_NCLS);
    for (int v = 0; v < 4; v++)
      CK(cudaFuncSetAttribute(pgs_chain_kernel(v & 1, v & 2), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mx));
  }
  return 0;
}

gzcuda_batch *gzcuda_batch_create(const gz_scene *scene, const gz_params *params, int device) {
  if (!scene || !params || scene->n_worlds <= 0) { fail(GZCUDA_E_INVALID, "null or empty scene"); return nullptr; }
  if (params->friction_model != 0 && params->friction_model != 2) {
    fail(GZCUDA_E_INVALID, "friction_model: pyramid (0) and box (2) are supported, cone (1) is not");
    return nullptr;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    fail(GZCUDA_E_CUDA, "no CUDA device: libgzcuda has no CPU fallback");
    return nullptr;
  }
  if (device < 0 || device >= ndev) { fail(GZCUDA_E_INVALID, "device index out of range"); return nullptr; }
  if (cudaSetDevice(device) != cudaSuccess) { fail(GZCUDA_E_CUDA, "cudaSetDevice failed"); return nullptr; }
  gzcuda_batch *b = new gzcuda_batch;
  b->device = device;
  int plo0 = 0, phi0 = 0;
  cudaDeviceGetStreamPriorityRange(&plo0, &phi0);
  b->reg_first = getenv("GZCUDA_PGS_REGFIRST") && atoi(getenv("GZCUDA_PGS_REGFIRST")) != 0;
  bool ok = cudaStreamCreateWithPriority(&b->stream, cudaStreamNonBlocking, b->reg_first ? phi0 : plo0) == cudaSuccess &&
            cudaEventCreate(&b->ev0) == cudaSuccess && cudaEventCreate(&b->ev1) == cudaSuccess;
  for (auto &e : b->evs) ok = ok && cudaEventCreate(&e) == cudaSuccess;
  ok = ok && cudaStreamCreateWithFlags(&b->stream_aux, cudaStreamNonBlocking) == cudaSuccess &&
       cudaEventCreateWithFlags(&b->ev_aux_fork, cudaEventDisableTiming) == cudaSuccess &&
       cudaEventCreateWithFlags(&b->ev_aux_join, cudaEventDisableTiming) == cudaSuccess;
  {
    int plo = 0, phi = 0;
    cudaDeviceGetStreamPriorityRange(&plo, &phi);
    ok = ok && cudaEventCreateWithFlags(&b->ev_fork, cudaEventDisableTiming) == cudaSuccess;
    for (int q = 0; q < GZ_CHAIN_NCLS; q++)
      // (greatest priority = lowest number) the larger the class, the longer its worlds run: first on the SMs
      ok = ok && cudaStreamCreateWithPriority(&b->stream_chain[q], cudaStreamNonBlocking,
                                              b->reg_first ? (q >= 2 ? phi : plo) : std::min(plo, phi + (GZ_CHAIN_NCLS - 1 - q))) == cudaSuccess &&
           cudaEventCreateWithFlags(&b->ev_join[q], cudaEventDisableTiming) == cudaSuccess;
  }
  if (const char *e = getenv("GZCUDA_PGS_RSPLIT")) b->pgs_rsplit = std::max(0, atoi(e));
  if (b->pgs_rsplit > 0)
    ok = ok && cudaStreamCreateWithPriority(&b->stream_hi, cudaStreamNonBlocking, phi0) == cudaSuccess &&
         cudaEventCreateWithFlags(&b->ev_join_hi, cudaEventDisableTiming) == cudaSuccess;
  if (!ok) { fail(GZCUDA_E_CUDA, "stream/event creation failed"); gzcuda_batch_destroy(b); return nullptr; }
  if (build(b, scene, params) != 0) { gzcuda_batch_destroy(b); return nullptr; }
  return b;
}

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }
// blocks per world of a (gx, n_worlds) grid whose blocks stride over a world's items: enough to cover the slab's
// capacity when the batch is small, but no more than keeps the whole grid near 8192 blocks (a block that finds
// nothing to do still costs launch time: 57k mostly idle blocks made k_rowfinish the largest kernel outside the sweeps)
static inline int item_blocks(int cap, int threads, int n_worlds) {
  return std::max(1, std::min(cdiv(std::max(cap, 1), threads), 8192 / std::max(n_worlds, 1)));
}

static int launch_collision(gzcuda_batch *b) {
  DevBatch &d = b->d;
  const int W = d.n_worlds, T = 128;
  k_geom<<<cdiv(std::max(d.ng, 1), T), T, 0, b->stream>>>(d);
  if (b->bp_large) {
    for (int w = 0; w < W; w++) {   // large worlds come one or a few per batch
      const int nwarps = std::max(1, b->h_world[w].ng), blocks = std::min(cdiv(nwarps, 8), 148 * 8);
      const int ntile = cdiv(b->h_world[w].ng + 1, GZ_BPL_TILE);
      k_bpl_tilesort<<<ntile, GZ_BPL_TILE, 0, b->stream>>>(d, w, b->bpl);
      k_bpl_merge<<<4 * ntile, 256, 0, b->stream>>>(d, w, b->bpl);
      k_bpl_span<<<cdiv(nwarps, 256), 256, 0, b->stream>>>(d, w, b->bpl);
      k_bpl_sweep<<<std::min(148 * 8, cdiv(4 * nwarps, 8)), 256, 0, b->stream>>>(d, w, b->bpl);
      k_bpl_tail<<<blocks, 256, 0, b->stream>>>(d, w, b->bpl);
      k_bpl_scan<<<1, 1024, 0, b->stream>>>(d, w, b->bpl);
      k_bpl_place<<<148, 256, 0, b->stream>>>(d, w, b->bpl);
      k_bpl_emit<<<blocks, 256, 0, b->stream>>>(d, w, b->bpl);
      k_bpl_emit_big<<<GZ_BPL_NBIG, 1024, 0, b->stream>>>(d, w, b->bpl);
    }
    b->last_launches += 9 * W - 1;
  } else k_broadphase<<<W, b->bp_threads, b->bp_smem, b->stream>>>(d);
  if (b->has_cylinders) k_narrow<true><<<dim3(item_blocks(d.max_pair_cap, 32, W), W), 32, 0, b->stream>>>(d);
  else k_narrow<false><<<dim3(item_blocks(d.max_pair_cap, 32, W), W), 32, 0, b->stream>>>(d);
  if (b->large_kernels) {
    for (int w = 0; w < W; w++) {
      unsigned *bar = reinterpret_cast<unsigned *>(d.xch_g + (size_t)2 * b->pgsl_grid * 256 + (size_t)9 * b->pgsl_grid);
      void *args[] = {(void *)&d, (void *)&w, (void *)&b->large_part, (void *)&bar};
      cudaLaunchCooperativeKernel((void *)k_contacts_large, dim3(b->large_grid), dim3(256), args, 0, b->stream);
    }
    b->last_launches += W - 1;
  } else
    k_contacts<<<W, d.max_nb > 512 ? 1024 : b->small_t, 0, b->stream>>>(d);   // (one CTA per world: a large world gets the largest CTA)
  b->last_launches += 4;
  return 0;
}

// developer switch GZCUDA_DEBUG_SYNC=1: synchronise at a few points of the step and name the phase a fault came from
static void dbg_point(gzcuda_batch *b, const char *what) {
  static const bool on = getenv("GZCUDA_DEBUG_SYNC") && atoi(getenv("GZCUDA_DEBUG_SYNC")) != 0;
  if (!on) return;
  const cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { fprintf(stderr, "gzcuda debug: %s after %s (step %lld)\n", cudaGetErrorString(e), what, (long long)b->d.step_index); fflush(stderr); }
}

static int launch_physics(gzcuda_batch *b, bool stage_events, int prof_slot = -1) {
  DevBatch &d = b->d;
  const int W = d.n_worlds, T = 128;
  if (b->large_kernels) {
    for (int w = 0; w < W; w++) {
      unsigned *bar = reinterpret_cast<unsigned *>(d.xch_g + (size_t)2 * b->pgsl_grid * 256 + (size_t)9 * b->pgsl_grid);
      void *args[] = {(void *)&d, (void *)&w, (void *)&bar};
      cudaLaunchCooperativeKernel((void *)k_assemble1_large, dim3(b->large_grid), dim3(256), args, 0, b->stream);
    }
    b->last_launches += W - 1;
  } else
    k_assemble1<<<W, d.max_nb > 512 ? 1024 : b->small_t, 0, b->stream>>>(d);   // islands, body prologue, row counts
  dbg_point(b, "collision + k_assemble1");
  // the joint rows (one thread per joint: a long dependent chain on few threads) run beside the contact rows and
  // the schedule, which do not depend on them; k_rowfinish needs all three
  const bool fork_joint = d.nj > 0 && d.pgs_impl >= 1;
  if (fork_joint) {
    cudaEventRecord(b->ev_aux_fork, b->stream);
    cudaStreamWaitEvent(b->stream_aux, b->ev_aux_fork, 0);
    k_rows_joint<<<cdiv(d.nj, 32), 32, 0, b->stream_aux>>>(d);
    cudaEventRecord(b->ev_aux_join, b->stream_aux);
    b->last_launches++;
  } else if (d.nj > 0) { k_rows_joint<<<cdiv(d.nj, T), T, 0, b->stream>>>(d); b->last_launches++; }
  k_rows_contact<<<dim3(item_blocks(d.max_contact_cap, 32, W), W), 32, 0, b->stream>>>(d);
  dbg_point(b, "k_rows_joint + k_rows_contact");
  if (d.pgs_impl >= 1) {
    // body-owner sweeps: schedule (integer work on the row counts), row prologue + entries, sweeps
    if (d.pgs_impl == 2) {
      for (int w = 0; w < W; w++) {
        unsigned *bar = reinterpret_cast<unsigned *>(d.xch_g + (size_t)2 * b->pgsl_grid * 256 + (size_t)9 * b->pgsl_grid);
        int relax = d.pgs_dataflow ? 0 : 1;   // round offsets are only needed by the round-synchronous sweeps
        void *args[] = {(void *)&d, (void *)&w, (void *)&bar, (void *)&relax};
        cudaLaunchCooperativeKernel((void *)k_bsched_large, dim3(b->bsl_grid), dim3(256), args, 0, b->stream);
      }
      b->last_launches += W - 1;
    } else
      k_bsched<<<W, b->bsched_t, d.bs_glob ? 0 : b->bsched_smem, b->stream>>>(d);
    k_border<<<cdiv(W, 256), 256, 0, b->stream>>>(d);
    dbg_point(b, "k_bsched + k_border");
    if (fork_joint) cudaStreamWaitEvent(b->stream, b->ev_aux_join, 0);
    if (d.p.dynamic_inertia && d.nj > 0) { k_dyn_inertia<<<cdiv(W, 64), 64, 0, b->stream>>>(d); b->last_launches++; }
    k_rowfinish<<<dim3(item_blocks(d.max_row_cap, 64, W), W), 64, 0, b->stream>>>(d);
    dbg_point(b, "k_rowfinish");
    if (stage_events) cudaEventRecord(b->evs[2], b->stream);
    if (prof_slot >= 0) cudaEventRecord(b->pgs_ev[2 * prof_slot], b->stream);
    const bool box = d.p.friction_model == 2;
    if (d.pgs_impl == 2) {
      // large worlds: one cooperative launch per world (all CTAs resident, grid barrier per round)
      for (int w = 0; w < W && d.pgs_dataflow; w++) {
        double *partial = d.xch_g + (size_t)2 * b->pgsl_grid * 256;
        unsigned *epoch = reinterpret_cast<unsigned *>(d.xch_g + (size_t)2 * b->pgsl_grid * 256 + (size_t)9 * b->pgsl_grid) + 8;
        unsigned *ticket = epoch + 1;
        // 32-bit tags: long before they could repeat, clear the slots and start the epochs again
        if ((++b->pgsd_launches + 2) * (unsigned long long)(d.p.iters + d.p.friction_iters + 1) >= b->pgsd_wrap) {
          cudaMemsetAsync(b->pgsd_slots, 0, (size_t)8 * d.ent_cap * sizeof(unsigned long long), b->stream);
          cudaMemsetAsync(epoch, 0, sizeof(unsigned), b->stream);
          b->pgsd_launches = 1;
        }
        if (b->pgs_large_mode == 2) {
          cudaLaunchConfig_t cfg = {};
          cfg.gridDim = dim3(b->pgsc_grid); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = b->pgsc_smem; cfg.stream = b->stream;
          cudaLaunchAttribute at[1];
          at[0].id = cudaLaunchAttributeClusterDimension;
          at[0].val.clusterDim.x = b->pgsc_cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
          cfg.attrs = at; cfg.numAttrs = 1;
          const int *tab = d.cta_e0 + (size_t)w * GZ_PGSC_TAB;
          const unsigned *ep = epoch;
          const int smem_entries = (int)(b->pgsc_smem / 64), glob_every = b->pgsc_glob_every - 1, sleep_ns = b->pgsc_sleep;
          if (box) cudaLaunchKernelEx(&cfg, k_pgsc<true>, d, w, b->pgsd_slots, tab, ep, smem_entries, glob_every, sleep_ns);
          else cudaLaunchKernelEx(&cfg, k_pgsc<false>, d, w, b->pgsd_slots, tab, ep, smem_entries, glob_every, sleep_ns);
          k_pgs_large_stats<<<b->pgsl_grid, 256, 0, b->stream>>>(d, w, partial, ticket, epoch, (const int *)nullptr);
          b->last_launches++;
          continue;
        }
        const int *skip = nullptr;
        if (b->pgsm) {
          cudaLaunchConfig_t cfg = {};
          cfg.gridDim = dim3(b->pgsm_cluster); cfg.blockDim = dim3(b->pgsm_threads); cfg.dynamicSmemBytes = b->pgsm_smem; cfg.stream = b->stream;
          cudaLaunchAttribute at[1];
          at[0].id = cudaLaunchAttributeClusterDimension;
          at[0].val.clusterDim.x = b->pgsm_cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
          cfg.attrs = at; cfg.numAttrs = 1;
          const int *tab = d.cta_e0 + (size_t)w * GZ_PGSC_TAB;
          const int smem_entries = (int)(b->pgsm_smem / GZ_PGSM_BYTES) & ~1;   // (even: the arrays behind the inboxes stay 16-byte aligned)
          int *done = b->pgsm_done + w;
          if (box) cudaLaunchKernelEx(&cfg, k_pgsm<true>, d, w, tab, smem_entries, done);
          else cudaLaunchKernelEx(&cfg, k_pgsm<false>, d, w, tab, smem_entries, done);
          skip = done;
          b->last_launches += 2;
        }
        void *args[] = {(void *)&d, (void *)&w, (void *)&b->pgsd_slots, (void *)&partial, (void *)&epoch, (void *)&skip};
        cudaLaunchCooperativeKernel(box ? (void *)k_pgsd<true> : (void *)k_pgsd<false>, dim3(b->pgsl_grid), dim3(256), args, 0, b->stream);
        if (b->pgsm) k_pgs_large_stats<<<b->pgsl_grid, 256, 0, b->stream>>>(d, w, partial, ticket, epoch, skip);
      }
      for (int w = 0; w < W && !d.pgs_dataflow; w++) {
        double *xch = d.xch_g, *partial = d.xch_g + (size_t)2 * b->pgsl_grid * 256;
        unsigned *bar = reinterpret_cast<unsigned *>(d.xch_g + (size_t)2 * b->pgsl_grid * 256 + (size_t)9 * b->pgsl_grid);
        void *args[] = {(void *)&d, (void *)&w, (void *)&xch, (void *)&partial, (void *)&bar};
        cudaLaunchCooperativeKernel(box ? (void *)k_pgsl<true> : (void *)k_pgsl<false>, dim3(b->pgsl_grid), dim3(256), args, 0, b->stream);
      }
      b->last_launches += W - 1;
    } else if (b->pgsb_multi) {
      if (box) k_pgsb<true, true><<<b->pgsb_ctas, b->pgsb_threads, b->pgsb_smem, b->stream>>>(d, b->pgsb, 0);
      else k_pgsb<true, false><<<b->pgsb_ctas, b->pgsb_threads, b->pgsb_smem, b->stream>>>(d, b->pgsb, 0);
    } else {
      // one launch per shared-memory class, side by side (fork / join by events), the largest class first and on
      // the stream of highest priority: its worlds run longest
      cudaEventRecord(b->ev_fork, b->stream);
      for (int q = b->pgsb_ncls - 1; q >= 0; q--) {
        cudaStream_t sq = q == 0 ? b->stream : b->pgsb_stream[q];
        if (q > 0) cudaStreamWaitEvent(sq, b->ev_fork, 0);
        const int lower = q > 0 ? b->pgsbc[q - 1].nres : -1;
        if (box) k_pgsb<false, true><<<b->pgsbc_grid[q], 32, b->pgsbc_smem[q], sq>>>(d, b->pgsbc[q], lower);
        else k_pgsb<false, false><<<b->pgsbc_grid[q], 32, b->pgsbc_smem[q], sq>>>(d, b->pgsbc[q], lower);
        { char nm[96]; snprintf(nm, sizeof nm, "k_pgsb class %d (grid %d, smem %zu, nres %d, wpc %d, lw %d, lower %d)", q, b->pgsbc_grid[q], b->pgsbc_smem[q], b->pgsbc[q].nres, b->pgsbc[q].wpc, b->pgsbc[q].lw, lower); dbg_point(b, nm); }
        if (q > 0) cudaEventRecord(b->pgsb_join[q], sq);
      }
      for (int q = 1; q < b->pgsb_ncls; q++) cudaStreamWaitEvent(b->stream, b->pgsb_join[q], 0);
      b->last_launches += b->pgsb_ncls - 1;
    }
    if (prof_slot >= 0) cudaEventRecord(b->pgs_ev[2 * prof_slot + 1], b->stream);
    if (stage_events) cudaEventRecord(b->evs[3], b->stream);
    if (d.nj > 0) {   // lambda write-back to the joints beside the integration
      cudaEventRecord(b->ev_aux_fork, b->stream);
      cudaStreamWaitEvent(b->stream_aux, b->ev_aux_fork, 0);
      k_lambda_save<<<cdiv(d.nj, T), T, 0, b->stream_aux>>>(d);
      cudaEventRecord(b->ev_aux_join, b->stream_aux);
      b->last_launches++;
    }
    dbg_point(b, "the sweeps");
    k_integrate<<<cdiv(std::max(d.nb, 1), 64), 64, 0, b->stream>>>(d);
    dbg_point(b, "k_integrate");
    if (d.nj > 0) cudaStreamWaitEvent(b->stream, b->ev_aux_join, 0);
    b->last_launches += 7;
    return 0;
  }
  const size_t sch_smem = (size_t)3 * std::max(d.max_nb, 1) * sizeof(int) + (size_t)d.sch_rcap * sizeof(unsigned short) + 16 +
                          (d.sch_rpar > 0 ? (size_t)32 * (2 * std::max(d.max_nb, 1) * sizeof(int) + d.sch_rpar * sizeof(unsigned short)) : 0);
  if (d.step_index > 0 && gz_epoch(d.step_index) == 1u) {   // the epochs start over (every 2^24 - 1 steps): no record of the last period may survive
    cudaMemsetAsync(d.prec, 0, (size_t)GZ_REC * d.prec_cap * sizeof(double), b->stream);
    cudaMemsetAsync(d.runrec, 0, (size_t)GZ_REC * std::max<size_t>(b->row_tot, 1) * sizeof(double), b->stream);
  }
  k_schedule<<<W, 32, sch_smem, b->stream>>>(d);
  k_slab<<<1 + cdiv((W + GZ_NBUCKET) * 16, GZ_SLAB_T * 8), GZ_SLAB_T, 0, b->stream>>>(d);
  if (d.p.dynamic_inertia && d.nj > 0) { k_dyn_inertia<<<cdiv(W, 64), 64, 0, b->stream>>>(d); b->last_launches++; }
  k_rowfinish<<<dim3(item_blocks(d.max_row_cap, 64, W), W), 64, 0, b->stream>>>(d);
  if (stage_events) cudaEventRecord(b->evs[2], b->stream);
  if (prof_slot >= 0) cudaEventRecord(b->pgs_ev[2 * prof_slot], b->stream);
  // at most one warp per world plus a partly filled one per bucket; CTAs beyond the step's warp count exit at once
  // chain-mode warps (heavy worlds, first in warp order) are their own instantiation on a second, higher-priority stream
  const bool chain_on = d.pgs_chain_min < (1 << 24);
  const PgsKernel kp = pgs_kernel(d.pgs_prof, d.p.friction_model == 2);
  const PgsChainKernel kc = pgs_chain_kernel(d.pgs_prof, d.p.friction_model == 2);
  if (chain_on && b->reg_first) {
    cudaEventRecord(b->ev_fork, b->stream);
    kp<<<W + GZ_NBUCKET, 32, b->pgs_smem, b->stream>>>(d, b->pgs);
  }
  const bool split = chain_on && !b->reg_first && b->pgs_rsplit > 0;
  PgsCfg cfg_lo = b->pgs;
  if (split) {   // the long-period warps of k_pgs first, above everything else
    PgsCfg cfg_hi = b->pgs;
    cfg_hi.part = 1; cfg_hi.rsplit = b->pgs_rsplit;
    cfg_lo.part = 2; cfg_lo.rsplit = b->pgs_rsplit;
    cudaEventRecord(b->ev_fork, b->stream);
    cudaStreamWaitEvent(b->stream_hi, b->ev_fork, 0);
    kp<<<W + GZ_NBUCKET, 32, b->pgs_smem, b->stream_hi>>>(d, cfg_hi);
    cudaEventRecord(b->ev_join_hi, b->stream_hi);
    b->last_launches++;
  }
  if (chain_on) {
    if (!b->reg_first && !split) cudaEventRecord(b->ev_fork, b->stream);
    for (int q = GZ_CHAIN_NCLS - 1; q >= 0; q--) {   // the large worlds first: they take longest
      cudaStreamWaitEvent(b->stream_chain[q], b->ev_fork, 0);
      kc<<<W, 32, b->pgs_chain_smem[q], b->stream_chain[q]>>>(d, b->pgs_chain[q], q);
      cudaEventRecord(b->ev_join[q], b->stream_chain[q]);
      b->last_launches++;
    }
  }
  if (!(chain_on && b->reg_first)) {
    kp<<<W + GZ_NBUCKET, 32, b->pgs_smem, b->stream>>>(d, cfg_lo);
  }
  if (split) cudaStreamWaitEvent(b->stream, b->ev_join_hi, 0);
  if (chain_on)
    for (int q = 0; q < GZ_CHAIN_NCLS; q++) cudaStreamWaitEvent(b->stream, b->ev_join[q], 0);
  if (prof_slot >= 0) cudaEventRecord(b->pgs_ev[2 * prof_slot + 1], b->stream);
  if (stage_events) cudaEventRecord(b->evs[3], b->stream);
  if (d.nj > 0) { k_lambda_save<<<cdiv(d.nj, T), T, 0, b->stream>>>(d); b->last_launches++; }
  k_integrate<<<cdiv(std::max(d.nb, 1), T), T, 0, b->stream>>>(d);
  b->last_launches += 7;
  return 0;
}

// wrench_in (host, [nb][6], may be null): added to the bodies' accumulators before the first step - uploaded and applied
// on the auxiliary stream beside the first step's collision pipeline, ahead of the controllers as before
static int run_steps(gzcuda_batch *b, int n_steps, const double *wrench_in = nullptr) {
  if (b->profiling) {
    while ((int)b->pgs_ev.size() < 2 * n_steps) {
      cudaEvent_t e;
      CK(cudaEventCreate(&e));
      b->pgs_ev.push_back(e);
    }
    b->pgs_pairs = n_steps;
  } else {
    b->pgs_pairs = 0;
  }
  for (int s = 0; s < n_steps; s++) {
    const bool last = (s == n_steps - 1);
    if (last) cudaEventRecord(b->evs[0], b->stream);
    if (b->have_velcmd) {
      k_velcmd<<<cdiv(b->d.n_worlds, 128), 128, 0, b->stream>>>(b->d);
      b->last_launches++;
    }
    // the plugin stand-in (PD torques into tacc) runs beside the collision pipeline, which does not read the
    // accumulators; k_assemble1 (body prologue) is the first kernel that does
    const bool up = (s == 0 && wrench_in != nullptr);
    const bool fork_ctl = b->have_pd || up;
    if (fork_ctl) {
      cudaEventRecord(b->ev_aux_fork, b->stream);
      cudaStreamWaitEvent(b->stream_aux, b->ev_aux_fork, 0);
      if (up) {
        CK(cudaMemcpyAsync(b->d.io_in, wrench_in, sizeof(double) * 6 * (size_t)b->d.nb, cudaMemcpyHostToDevice, b->stream_aux));
        k_apply_wrench<<<cdiv(std::max(b->d.nb, 1), 128), 128, 0, b->stream_aux>>>(b->d);
        b->last_launches++;
      }
      if (b->have_pd) {
        k_control<<<cdiv(std::max(b->d.nb, 1), 64), 64, 0, b->stream_aux>>>(b->d);
        b->last_launches++;
      }
      cudaEventRecord(b->ev_aux_join, b->stream_aux);
    }
    launch_collision(b);
    if (fork_ctl) cudaStreamWaitEvent(b->stream, b->ev_aux_join, 0);
    if (last) cudaEventRecord(b->evs[1], b->stream);
    launch_physics(b, last, b->profiling ? s : -1);
    if (last) cudaEventRecord(b->evs[4], b->stream);
    b->d.step_index++;
  }
  b->stage_valid = b->stage_valid || n_steps > 0;
  return 0;
}

int gzcuda_batch_step(gzcuda_batch *b, int n_steps) {
  if (!b || n_steps < 0) return fail(GZCUDA_E_INVALID, "bad batch or step count");
  CK(cudaSetDevice(b->device));
  b->last_launches = 0;
  CK(cudaEventRecord(b->ev0, b->stream));
  int rc = run_steps(b, n_steps);
  if (rc) return rc;
  CK(cudaEventRecord(b->ev1, b->stream));
  CK(cudaGetLastError());
  return 0;
}

// Host-buffer variant of gzcuda_batch_step: what the Gazebo glue calls once per World::Update.
int gzcuda_batch_step_io(gzcuda_batch *b, int n_steps, const double *wrench_in, double *state_out) {
  if (!b || n_steps < 0) return fail(GZCUDA_E_INVALID, "bad batch or step count");
  CK(cudaSetDevice(b->device));
  const int T = 128, nb = std::max(b->d.nb, 1);
  b->last_launches = 0;
  CK(cudaEventRecord(b->ev0, b->stream));
  if (wrench_in && n_steps == 0) {   // no step to run beside: apply at once
    CK(cudaMemcpyAsync(b->d.io_in, wrench_in, sizeof(double) * 6 * (size_t)b->d.nb, cudaMemcpyHostToDevice, b->stream));
    k_apply_wrench<<<cdiv(nb, T), T, 0, b->stream>>>(b->d);
    b->last_launches++;
  }
  int rc = run_steps(b, n_steps, wrench_in);
  if (rc) return rc;
  if (state_out) {
    k_gather_state<<<cdiv(nb, T), T, 0, b->stream>>>(b->d);
    b->last_launches++;
    CK(cudaMemcpyAsync(state_out, b->d.io_out, sizeof(double) * 13 * (size_t)b->d.nb, cudaMemcpyDeviceToHost, b->stream));
  }
  CK(cudaEventRecord(b->ev1, b->stream));
  CK(cudaGetLastError());
  return check_sticky(b);   // synchronises; GZCUDA_E_CAPACITY if any of the steps overflowed a slab
}

// State log at a decimated rate (the WorldState stream of gz log / ~/pose/info, gazebo/physics/World.cc:2757-2850,
// gazebo/util/LogRecord.cc:193): n_steps steps in one call, the state of every body after each `every`-th step copied into
// frames[frame][nb][13] (pos 3, quat 4, lvel 3, avel 3) by a copy stream that runs beside the next steps (two device staging
// frames), no host synchronisation until the end.  Returns the number of frames written.
int gzcuda_batch_step_log(gzcuda_batch *b, int n_steps, int every, double *frames, int max_frames) {
  if (!b || n_steps < 0 || every <= 0 || !frames || max_frames < 0) return fail(GZCUDA_E_INVALID, "bad argument");
  CK(cudaSetDevice(b->device));
  const int T = 128, nb = std::max(b->d.nb, 1);
  const size_t frame_doubles = (size_t)13 * b->d.nb;
  if (!b->stream_copy) {
    CK(cudaStreamCreateWithFlags(&b->stream_copy, cudaStreamNonBlocking));
    for (int f = 0; f < 2; f++) {
      CK(cudaEventCreateWithFlags(&b->ev_frame[f], cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&b->ev_copied[f], cudaEventDisableTiming));
      CK(cudaMalloc(&b->log_frame[f], std::max<size_t>(frame_doubles, 1) * sizeof(double)));
    }
  }
  b->last_launches = 0;
  CK(cudaEventRecord(b->ev0, b->stream));
  int done = 0, nframes = 0;
  double *const io_out = b->d.io_out;
  while (done < n_steps) {
    const int n = std::min(every, n_steps - done);
    int rc = run_steps(b, n);
    if (rc) { b->d.io_out = io_out; return rc; }
    done += n;
    if (n == every && nframes < max_frames) {
      const int f = nframes & 1;
      if (nframes >= 2) CK(cudaStreamWaitEvent(b->stream, b->ev_copied[f], 0));   // the frame's previous copy has left
      b->d.io_out = b->log_frame[f];
      k_gather_state<<<cdiv(nb, T), T, 0, b->stream>>>(b->d);
      b->d.io_out = io_out;
      b->last_launches++;
      CK(cudaEventRecord(b->ev_frame[f], b->stream));
      CK(cudaStreamWaitEvent(b->stream_copy, b->ev_frame[f], 0));
      CK(cudaMemcpyAsync(frames + (size_t)nframes * frame_doubles, b->log_frame[f], frame_doubles * sizeof(double),
                         cudaMemcpyDeviceToHost, b->stream_copy));
      CK(cudaEventRecord(b->ev_copied[f], b->stream_copy));
      nframes++;
    }
  }
  CK(cudaEventRecord(b->ev1, b->stream));
  CK(cudaGetLastError());
  const int rc_cap = check_sticky(b);
  CK(cudaStreamSynchronize(b->stream_copy));
  if (rc_cap) return rc_cap;
  return nframes;
}

int gzcuda_batch_set_profiling(gzcuda_batch *b, int on) {
  if (!b) return fail(GZCUDA_E_INVALID, "null batch");
  b->profiling = on != 0;
  return 0;
}

// Sum of the k_pgs launch durations of the last step call (needs profiling on) and their count.
int gzcuda_batch_pgs_ms(gzcuda_batch *b, float *total_ms, int *launches) {
  if (!b) return fail(GZCUDA_E_INVALID, "null batch");
  int rc = gzcuda_batch_sync(b);
  if (rc) return rc;
  float tot = 0;
  for (int i = 0; i < b->pgs_pairs; i++) {
    float t = 0;
    CK(cudaEventElapsedTime(&t, b->pgs_ev[2 * i], b->pgs_ev[2 * i + 1]));
    tot += t;
  }
  if (total_ms) *total_ms = tot;
  if (launches) *launches = b->pgs_pairs;
  return 0;
}

// Running totals since creation (or the last reset): constraint rows, friction rows, contacts,
// world-steps.  Used for the algorithmic-bytes figure of the roofline.
int gzcuda_batch_get_counters(gzcuda_batch *b, unsigned long long out[4], int reset) {
  if (!b || !out) return fail(GZCUDA_E_INVALID, "null argument");
  int rc = gzcuda_batch_sync(b);
  if (rc) return rc;
  CK(cudaMemcpy(out, b->d.acc, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  if (reset) CK(cudaMemset(b->d.acc, 0, 4 * sizeof(unsigned long long)));
  return 0;
}

int gzcuda_batch_update_collision(gzcuda_batch *b) {
  if (!b) return fail(GZCUDA_E_INVALID, "null batch");
  CK(cudaSetDevice(b->device));
  launch_collision(b);
  CK(cudaGetLastError());
  return 0;
}

int gzcuda_batch_update_physics(gzcuda_batch *b) {
  if (!b) return fail(GZCUDA_E_INVALID, "null batch");
  CK(cudaSetDevice(b->device));
  launch_physics(b, false);
  b->d.step_index++;
  CK(cudaGetLastError());
  return 0;
}

int gzcuda_world_step(gzcuda_batch *b, double step_size) {
  if (!b || !(step_size > 0)) return 0;
  b->d.p.step_size = step_size;
  if (gzcuda_batch_update_physics(b) != 0) return 0;
  return gzcuda_batch_sync(b) == 0 ? 1 : 0;
}

int gzcuda_batch_sync(gzcuda_batch *b) {
  if (!b) return fail(GZCUDA_E_INVALID, "null batch");
  CK(cudaSetDevice(b->device));
  return check_sticky(b);
}

// Sticky capacity flag (DevBatch::sticky): a step that dropped pairs, contacts, rows or whole worlds somewhere in a
// multi-step call must not go unnoticed.  Reads it back (stream-ordered, after everything queued so far), reports
// GZCUDA_E_CAPACITY once and clears it.
static int check_sticky(gzcuda_batch *b) {
  CK(cudaMemcpyAsync(b->h_sticky, b->d.sticky, 4 * sizeof(int), cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  if (b->h_sticky[0]) {
    char buf[200];
    snprintf(buf, sizeof(buf), "capacity overflow during a step (mask %d: 1 pairs, 2 contacts, 4 rows, 8 solver limits, 16 record pool, 32 MPR iteration cap; "
             "last world %d): results of that step are incomplete", b->h_sticky[0], b->h_sticky[1] - 1);
    CK(cudaMemsetAsync(b->d.sticky, 0, 4 * sizeof(int), b->stream));
    return fail(GZCUDA_E_CAPACITY, buf);
  }
  return 0;
}

static int check_overflow(gzcuda_batch *b) {
  std::vector<WorldStep> ws(b->d.n_worlds);
  CK(cudaMemcpy(ws.data(), b->d.wstep, ws.size() * sizeof(WorldStep), cudaMemcpyDeviceToHost));
  for (size_t w = 0; w < ws.size(); w++)
    if (ws[w].overflow) {
      char buf[128];
      snprintf(buf, sizeof(buf), "world %zu overflowed a per-world slab (mask %d)", w, ws[w].overflow);
      return fail(GZCUDA_E_CAPACITY, buf);
    }
  return 0;
}

int gzcuda_batch_get_body_state(gzcuda_batch *b, double *pos, double *quat, double *lvel, double *avel) {
  if (!b) return fail(GZCUDA_E_INVALID, "null batch");
  CK(cudaSetDevice(b->device));
  const int nb = b->d.nb;
  k_gather_state<<<cdiv(std::max(nb, 1), 128), 128, 0, b->stream>>>(b->d);
  std::vector<double> h((size_t)13 * nb);
  CK(cudaMemcpyAsync(h.data(), b->d.io_out, h.size() * sizeof(double), cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  for (int i = 0; i < nb; i++) {
    const double *o = h.data() + 13 * (size_t)i;
    for (int k = 0; k < 3; k++) {
      if (pos) pos[3 * i + k] = o[k];
      if (lvel) lvel[3 * i + k] = o[7 + k];
      if (avel) avel[3 * i + k] = o[10 + k];
    }
    if (quat) for (int k = 0; k < 4; k++) quat[4 * i + k] = o[3 + k];
  }
  return check_overflow(b);
}

static int set_body_state_impl(gzcuda_batch *b, int first, int count, const double *pos, const double *quat,
                               const double *lvel, const double *avel, int raw);

int gzcuda_batch_set_body_state(gzcuda_batch *b, int first, int count, const double *pos,
                                const double *quat, const double *lvel, const double *avel) {
  return set_body_state_impl(b, first, count, pos, quat, lvel, avel, 0);
}

int gzcuda_batch_set_body_state_raw(gzcuda_batch *b, int first, int count, const double *pos,
                                    const double *quat, const double *lvel, const double *avel) {
  return set_body_state_impl(b, first, count, pos, quat, lvel, avel, 1);
}

int gzcuda_batch_set_body_enabled(gzcuda_batch *b, int first, int count, const int *enabled) {
  if (!b || !enabled || first < 0 || count < 0 || first + count > b->d.nb) return fail(GZCUDA_E_INVALID, "bad body range");
  if (count == 0) return 0;
  CK(cudaSetDevice(b->device));
  CK(cudaMemcpyAsync(b->d.flag_stage, enabled, (size_t)count * sizeof(int), cudaMemcpyHostToDevice, b->stream));
  k_set_enabled<<<cdiv(count, 128), 128, 0, b->stream>>>(b->d, first, count);
  CK(cudaGetLastError());
  return 0;
}

static int set_body_state_impl(gzcuda_batch *b, int first, int count, const double *pos, const double *quat,
                               const double *lvel, const double *avel, int raw) {
  if (!b || first < 0 || count < 0 || first + count > b->d.nb) return fail(GZCUDA_E_INVALID, "bad body range");
  if (count == 0) return 0;
  CK(cudaSetDevice(b->device));
  // stream-ordered: stage [count][13] (NaN = group not given), one small kernel writes the touched fields
  const double nan = std::nan("");
  std::vector<double> h((size_t)13 * count);
  for (int i = 0; i < count; i++) {
    double *o = h.data() + 13 * (size_t)i;
    for (int k = 0; k < 3; k++) {
      o[k] = pos ? pos[3 * i + k] : nan;
      o[7 + k] = lvel ? lvel[3 * i + k] : nan;
      o[10 + k] = avel ? avel[3 * i + k] : nan;
    }
    for (int k = 0; k < 4; k++) o[3 + k] = quat ? quat[4 * i + k] : nan;
  }
  CK(cudaMemcpyAsync(b->d.set_stage, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice, b->stream));   // pageable source: staged before the call returns
  k_set_state<<<cdiv(count, 128), 128, 0, b->stream>>>(b->d, first, count, raw);
  CK(cudaGetLastError());
  return 0;
}

int gzcuda_batch_add_body_wrench(gzcuda_batch *b, const double *force, const double *torque) {
  if (!b) return fail(GZCUDA_E_INVALID, "null batch");
  CK(cudaSetDevice(b->device));
  const int nb = b->d.nb;
  if (nb == 0) return 0;
  std::vector<double> h((size_t)6 * nb, 0.0);
  for (int i = 0; i < nb; i++)
    for (int k = 0; k < 3; k++) {
      if (force) h[6 * (size_t)i + k] = force[3 * i + k];
      if (torque) h[6 * (size_t)i + 3 + k] = torque[3 * i + k];
    }
  CK(cudaMemcpyAsync(b->d.io_in, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice, b->stream));
  k_apply_wrench<<<cdiv(nb, 128), 128, 0, b->stream>>>(b->d);
  CK(cudaGetLastError());
  return 0;
}

int gzcuda_batch_get_body_flags(gzcuda_batch *b, int *flags) {
  if (!b || !flags) return fail(GZCUDA_E_INVALID, "null argument");
  CK(cudaSetDevice(b->device));
  const int nb = b->d.nb;
  if (nb == 0) return 0;
  k_gather_flags<<<cdiv(nb, 128), 128, 0, b->stream>>>(b->d);
  CK(cudaMemcpyAsync(flags, b->d.flag_stage, (size_t)nb * sizeof(int), cudaMemcpyDeviceToHost, b->stream));
  return check_sticky(b);
}

int gzcuda_batch_set_joint_param(gzcuda_batch *b, int joint, int slot, double value) {
  if (!b || joint < 0 || joint >= b->d.nj || slot < 0 || slot >= GZ_JP_COUNT) return fail(GZCUDA_E_INVALID, "bad joint/slot");
  CK(cudaSetDevice(b->device));
  double *dst = slot < GZ_JP_HOT ? &b->d.joint[joint].param[slot] : b->d.joint_ext + (size_t)(GZ_JP_COUNT - GZ_JP_HOT) * joint + slot - GZ_JP_HOT;
  CK(cudaMemcpyAsync(dst, &value, sizeof(double), cudaMemcpyHostToDevice, b->stream));   // stream-ordered, 8 bytes
  return 0;
}

int gzcuda_batch_get_joint_state(gzcuda_batch *b, double *lambda, double *lambda_erp, double *cum_angle) {
  if (!b) return fail(GZCUDA_E_INVALID, "null batch");
  int rc = gzcuda_batch_sync(b);
  if (rc) return rc;
  std::vector<Joint> hj(b->d.nj);
  if (b->d.nj) CK(cudaMemcpy(hj.data(), b->d.joint, hj.size() * sizeof(Joint), cudaMemcpyDeviceToHost));
  for (int j = 0; j < b->d.nj; j++) {
    for (int k = 0; k < 6; k++) {
      if (lambda) lambda[6 * j + k] = hj[j].lambda[k];
      if (lambda_erp) lambda_erp[6 * j + k] = hj[j].lambda_erp[k];
    }
    if (cum_angle) cum_angle[j] = hj[j].cum_angle;
  }
  return 0;
}

int gzcuda_batch_set_hinge_pd(gzcuda_batch *b, const double *pd) {
  if (!b || !pd) return fail(GZCUDA_E_INVALID, "null argument");
  CK(cudaSetDevice(b->device));
  const int nj = b->d.nj;
  if (nj == 0) return 0;
  bool any = false;
  for (int j = 0; j < nj; j++) any = any || pd[6 * j + 5] > 0;
  CK(cudaMemcpyAsync(b->d.jstage, pd, (size_t)6 * nj * sizeof(double), cudaMemcpyHostToDevice, b->stream));
  k_set_pd<<<cdiv(nj, 128), 128, 0, b->stream>>>(b->d);
  CK(cudaGetLastError());
  b->have_pd = any;
  return 0;
}

int gzcuda_batch_set_joint_velocity_cmd(gzcuda_batch *b, const double *vel) {
  if (!b || !vel) return fail(GZCUDA_E_INVALID, "null argument");
  CK(cudaSetDevice(b->device));
  const int nj = b->d.nj;
  if (nj == 0) return 0;
  bool any = false;
  for (int j = 0; j < nj; j++) any = any || (vel[j] == vel[j] && b->h_joint_type[j] == GZ_JOINT_HINGE);   // NaN = no command
  CK(cudaMemcpyAsync(b->d.jstage, vel, (size_t)nj * sizeof(double), cudaMemcpyHostToDevice, b->stream));
  k_set_velcmd<<<cdiv(nj, 128), 128, 0, b->stream>>>(b->d);
  CK(cudaGetLastError());
  b->have_velcmd = any;
  return 0;
}

int gzcuda_batch_get_stats(gzcuda_batch *b, double *rms_residual, int *num_contacts, int *num_rows) {
  if (!b) return fail(GZCUDA_E_INVALID, "null batch");
  int rc = gzcuda_batch_sync(b);
  if (rc) return rc;
  std::vector<WorldStep> ws(b->d.n_worlds);
  CK(cudaMemcpy(ws.data(), b->d.wstep, ws.size() * sizeof(WorldStep), cudaMemcpyDeviceToHost));
  for (int w = 0; w < b->d.n_worlds; w++) {
    if (rms_residual) for (int k = 0; k < 4; k++) rms_residual[4 * w + k] = ws[w].rms_residual[k];
    if (num_contacts) num_contacts[w] = ws[w].n_contacts;
    if (num_rows) {
      num_rows[w] = ws[w].m;
      const WorldInfo &wi = b->h_world[w];
      if (wi.rpc == 4 && ws[w].m > 0 && ws[w].n_contacts > 0) {   // the placeholder rows of pairs without torsional friction are not rows
        std::vector<Contact> hc(ws[w].n_contacts);
        CK(cudaMemcpy(hc.data(), b->d.contact + wi.contact0, hc.size() * sizeof(Contact), cudaMemcpyDeviceToHost));
        for (const Contact &c : hc) num_rows[w] -= c.tors ? 0 : 1;
      }
    }
  }
  return 0;
}

// Instrumentation of the last step's k_pgs: per world, loop trips and row updates of its longest
// lane, and the cycles its sweeps took (out[nw][3]).
int gzcuda_batch_get_pgs_debug(gzcuda_batch *b, int *out) {
  if (!b || !out) return fail(GZCUDA_E_INVALID, "null argument");
  int rc = gzcuda_batch_sync(b);
  if (rc) return rc;
  std::vector<WorldStep> ws(b->d.n_worlds);
  CK(cudaMemcpy(ws.data(), b->d.wstep, ws.size() * sizeof(WorldStep), cudaMemcpyDeviceToHost));
  for (int w = 0; w < b->d.n_worlds; w++) { out[3 * w] = ws[w].pad[0]; out[3 * w + 1] = ws[w].pad[1]; out[3 * w + 2] = ws[w].n_fric; }
  return 0;
}

// Developer instrumentation: when the sweeps of each world started and
// ended in the last step, ns of the GPU's global timer, out[nw][2].
int gzcuda_batch_get_pgs_timeline(gzcuda_batch *b, unsigned long long *out) {
  if (!b || !out) return fail(GZCUDA_E_INVALID, "null argument");
  int rc = gzcuda_batch_sync(b);
  if (rc) return rc;
  CK(cudaMemcpy(out, b->d.prof + 16, 2 * (size_t)b->d.n_worlds * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  return 0;
}

// Developer instrumentation (GZCUDA_PGS_PROF=1 at creation): cycle sums over all k_pgs warps since
// creation: fetch issue, wait, ring load, row update + sync, active lanes, rounds.
int gzcuda_batch_get_pgs_profile(gzcuda_batch *b, unsigned long long out[16]) {
  if (!b || !out) return fail(GZCUDA_E_INVALID, "null argument");
  int rc = gzcuda_batch_sync(b);
  if (rc) return rc;
  CK(cudaMemcpy(out, b->d.prof, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  CK(cudaMemset(b->d.prof, 0, 16 * sizeof(unsigned long long)));
  return 0;
}

int gzcuda_batch_set_param(gzcuda_batch *b, const char *key, double v) {
  if (!b || !key) return fail(GZCUDA_E_INVALID, "null argument");
  gz_params &p = b->d.p;
  std::string k(key);
  if (k == "iters") p.iters = (int)v;
  else if (k == "sor") p.sor_w = v;
  else if (k == "cfm") p.cfm = v;
  else if (k == "erp") p.erp = v;
  else if (k == "contact_max_correcting_vel") p.contact_max_correcting_vel = v;
  else if (k == "contact_surface_layer") p.contact_surface_layer = v;
  else if (k == "max_contacts") p.max_contacts = (int)v;
  else if (k == "friction_iters") p.friction_iters = (int)v;
  else if (k == "warm_start_factor") p.warm_start = v;
  else if (k == "contact_residual_smoothing") p.smooth_contacts = v;
  else if (k == "contact_sor_scale") p.contact_sor_scale = v;
  else if (k == "max_step_size") p.step_size = v;
  else if (k == "dynamic_moi_rescaling") p.dynamic_inertia = v != 0;   // ODEPhysics::SetParam("use_dynamic_moi_rescaling")
  else if (k == "gravity_x") p.gravity[0] = v;
  else if (k == "gravity_y") p.gravity[1] = v;
  else if (k == "gravity_z") p.gravity[2] = v;
  else if (k == "friction_model") {   // ODEPhysics::SetParam takes "pyramid_model" / "box_model" / "cone_model": 0 / 2 / 1 here
    if ((int)v != 0 && (int)v != 2) return fail(GZCUDA_E_INVALID, "friction_model: pyramid (0) and box (2) are supported");
    p.friction_model = (int)v;
  }
  else return fail(GZCUDA_E_INVALID, "unknown parameter key: " + k);
  return 0;
}

int gzcuda_batch_get_pairs(gzcuda_batch *b, int world, int *pairs, int cap) {
  if (!b || world < 0 || world >= b->d.n_worlds) return fail(GZCUDA_E_INVALID, "bad world");
  int rc = gzcuda_batch_sync(b);
  if (rc) return rc;
  WorldStep ws;
  CK(cudaMemcpy(&ws, b->d.wstep + world, sizeof(ws), cudaMemcpyDeviceToHost));
  const WorldInfo &wi = b->h_world[world];
  std::vector<int> hp(2 * (size_t)ws.n_pairs);
  if (ws.n_pairs) CK(cudaMemcpy(hp.data(), b->d.pair + 2 * (size_t)wi.pair0, hp.size() * sizeof(int), cudaMemcpyDeviceToHost));
  for (int i = 0; i < ws.n_pairs && i < cap; i++) {
    pairs[2 * i] = hp[2 * i] - wi.geom0;
    pairs[2 * i + 1] = hp[2 * i + 1] - wi.geom0;
  }
  return ws.n_pairs;
}

int gzcuda_batch_ray_query(gzcuda_batch *b, int n_rays, const int *ray_world, const double *start, const double *dir,
                           const double *length, double *out_range, int *out_geom) {
  if (!b || n_rays < 0 || (n_rays > 0 && (!ray_world || !start || !dir || !length || !out_range))) return fail(GZCUDA_E_INVALID, "null argument");
  if (n_rays == 0) return 0;
  for (int k = 0; k < n_rays; k++)
    if (ray_world[k] < 0 || ray_world[k] >= b->d.n_worlds) return fail(GZCUDA_E_INVALID, "ray in a world the batch does not have");
  CK(cudaSetDevice(b->device));
  // scan buffers grow with the largest scan asked for and are kept
  if (n_rays > b->ray_cap) {
    if (b->ray_buf) cudaFree(b->ray_buf);
    b->ray_buf = nullptr; b->ray_cap = 0;
    CK(cudaMalloc((void **)&b->ray_buf, (size_t)n_rays * (8 * sizeof(double) + 2 * sizeof(int))));
    b->ray_cap = n_rays;
  }
  const size_t n = (size_t)n_rays;
  double *d_start = reinterpret_cast<double *>(b->ray_buf), *d_dir = d_start + 3 * n, *d_len = d_dir + 3 * n, *d_range = d_len + n;
  int *d_world = reinterpret_cast<int *>(d_range + n), *d_geom = d_world + n;
  CK(cudaMemcpyAsync(d_start, start, 3 * n * sizeof(double), cudaMemcpyHostToDevice, b->stream));
  CK(cudaMemcpyAsync(d_dir, dir, 3 * n * sizeof(double), cudaMemcpyHostToDevice, b->stream));
  CK(cudaMemcpyAsync(d_len, length, n * sizeof(double), cudaMemcpyHostToDevice, b->stream));
  CK(cudaMemcpyAsync(d_world, ray_world, n * sizeof(int), cudaMemcpyHostToDevice, b->stream));
  k_geom<<<cdiv(std::max(b->d.ng, 1), 128), 128, 0, b->stream>>>(b->d);   // poses and AABBs as of now (idempotent)
  RayQuery Q;
  Q.n = n_rays; Q.world = d_world; Q.start = d_start; Q.dir = d_dir; Q.length = d_len; Q.out_range = d_range; Q.out_geom = d_geom;
  k_rays<<<std::min(cdiv(n_rays, 8), 148 * 8), 256, 0, b->stream>>>(b->d, Q);
  CK(cudaMemcpyAsync(out_range, d_range, n * sizeof(double), cudaMemcpyDeviceToHost, b->stream));
  if (out_geom) CK(cudaMemcpyAsync(out_geom, d_geom, n * sizeof(int), cudaMemcpyDeviceToHost, b->stream));
  CK(cudaGetLastError());
  return check_sticky(b);   // synchronises
}

int gzcuda_batch_get_contacts(gzcuda_batch *b, int world, int *geoms, double *data, int cap) {
  if (!b || world < 0 || world >= b->d.n_worlds) return fail(GZCUDA_E_INVALID, "bad world");
  int rc = gzcuda_batch_sync(b);
  if (rc) return rc;
  WorldStep ws;
  CK(cudaMemcpy(&ws, b->d.wstep + world, sizeof(ws), cudaMemcpyDeviceToHost));
  const WorldInfo &wi = b->h_world[world];
  std::vector<Contact> hc(ws.n_contacts);
  if (ws.n_contacts) CK(cudaMemcpy(hc.data(), b->d.contact + wi.contact0, hc.size() * sizeof(Contact), cudaMemcpyDeviceToHost));
  for (int i = 0; i < ws.n_contacts && i < cap; i++) {
    geoms[2 * i] = hc[i].g1 - wi.geom0;
    geoms[2 * i + 1] = hc[i].g2 - wi.geom0;
    for (int k = 0; k < 3; k++) { data[7 * i + k] = hc[i].pos[k]; data[7 * i + 3 + k] = hc[i].normal[k]; }
    data[7 * i + 6] = hc[i].depth;
  }
  return ws.n_contacts;
}

int gzcuda_batch_get_joint_feedback(gzcuda_batch *b, double *fb) {
  if (!b || !fb) return fail(GZCUDA_E_INVALID, "null argument");
  int rc = gzcuda_batch_sync(b);
  if (rc) return rc;
  const int nj = b->d.nj;
  if (nj == 0) return 0;
  double *dev = b->d.fb_stage;   // preallocated at creation
  k_feedback_joint<<<cdiv(nj, 128), 128, 0, b->stream>>>(b->d, dev);
  CK(cudaMemcpyAsync(fb, dev, (size_t)12 * nj * sizeof(double), cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  return nj;
}

// ---- gzcuda_batch_get_joint_feedback without any allocation --------------------------------------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
// forbid it).  The scratch k_feedback_joint writes lives in this file-scope device array: GZCUDA_NOALLOC_MAX_JOINTS
// x 12 doubles = 3 MiB, held by every GPU that loads the library until the process exits.  In host code the
// array's NAME is a host-side stand-in, not a device address; the address is asked for once per device
// (cudaGetSymbolAddress with that device current) and that pointer is what the kernel gets.  All calls share the
// array and it keeps the last call's contents: the kernel overwrites every element of [0, 12 nj) it hands back, so
// nothing has to be cleared; calls that overlap on the GPU (two batches' streams) would race and must be serialised
// by the caller.
static __device__ __align__(256) double g_noalloc_fb_scratch[12 * GZCUDA_NOALLOC_MAX_JOINTS];
static double *g_noalloc_fb_ptr[64] = {};   // device address of the array, per device ordinal

int gzcuda_batch_get_joint_feedback_noalloc(gzcuda_batch *b, double *fb) {
  if (!b || !fb) return fail(GZCUDA_E_INVALID, "null argument");
  const int nj = b->d.nj;
  if (nj > GZCUDA_NOALLOC_MAX_JOINTS) {
    char buf[160];
    snprintf(buf, sizeof(buf), "gzcuda_batch_get_joint_feedback_noalloc: %d joints exceed the static scratch of %d "
             "(GZCUDA_NOALLOC_MAX_JOINTS); use gzcuda_batch_get_joint_feedback", nj, GZCUDA_NOALLOC_MAX_JOINTS);
    return fail(GZCUDA_E_INVALID, buf);
  }
  if (b->device < 0 || b->device >= 64) return fail(GZCUDA_E_INVALID, "device ordinal beyond 63");
  int rc = gzcuda_batch_sync(b);   // (sets the batch's device current)
  if (rc) return rc;
  if (nj == 0) return 0;
  if (!g_noalloc_fb_ptr[b->device]) {   // set-up, once per device, with that device current
    double *ptr = nullptr;
    CK(cudaGetSymbolAddress((void **)&ptr, g_noalloc_fb_scratch));
    if (!ptr) return fail(GZCUDA_E_CUDA, "cudaGetSymbolAddress returned a null address");
    g_noalloc_fb_ptr[b->device] = ptr;
  }
  double *dev = g_noalloc_fb_ptr[b->device];
  k_feedback_joint<<<cdiv(nj, 128), 128, 0, b->stream>>>(b->d, dev);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(fb, dev, (size_t)12 * nj * sizeof(double), cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  return nj;
}

int gzcuda_batch_get_contact_feedback(gzcuda_batch *b, int world, double *fb, int cap) {
  if (!b || !fb || world < 0 || world >= b->d.n_worlds) return fail(GZCUDA_E_INVALID, "bad argument");
  int rc = gzcuda_batch_sync(b);
  if (rc) return rc;
  WorldStep ws;
  CK(cudaMemcpy(&ws, b->d.wstep + world, sizeof(ws), cudaMemcpyDeviceToHost));
  const int n = ws.n_contacts;
  if (n == 0) return 0;
  double *dev = b->d.fb_stage;   // preallocated at creation ([max contact_cap][12])
  k_feedback_contact<<<cdiv(n, 128), 128, 0, b->stream>>>(b->d, world, dev);
  CK(cudaMemcpyAsync(fb, dev, (size_t)12 * std::min(n, cap) * sizeof(double), cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  return n;
}

int gzcuda_batch_last_step_ms(gzcuda_batch *b, float *ms, long *launches) {
  if (!b) return fail(GZCUDA_E_INVALID, "null batch");
  int rc = gzcuda_batch_sync(b);
  if (rc) return rc;
  float t = 0;
  CK(cudaEventElapsedTime(&t, b->ev0, b->ev1));
  b->last_ms = t;
  if (ms) *ms = t;
  if (launches) *launches = b->last_launches;
  return 0;
}

int gzcuda_batch_stage_ms(gzcuda_batch *b, float ms[4]) {
  if (!b || !ms) return fail(GZCUDA_E_INVALID, "null argument");
  int rc = gzcuda_batch_sync(b);
  if (rc) return rc;
  if (!b->stage_valid) return fail(GZCUDA_E_INVALID, "no step has run yet");
  for (int k = 0; k < 4; k++) CK(cudaEventElapsedTime(&ms[k], b->evs[k], b->evs[k + 1]));
  return 0;
}

}  // extern "C"
