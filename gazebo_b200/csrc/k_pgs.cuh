// k_pgs.cuh — row prologue, dataflow schedule, the SOR-PGS sweeps, and integration.
//
//   k_rowfinish  rhs / rhs_erp (quickstep.cpp:403-437), cfm /= h (:443), warm start (:473-495),
//                iMJ = M^-1 J^T (quickstep_util.cpp:79-98), Ad and the Ad scaling
//                (quickstep_pgs_lcp.cpp:1012-1028,1091-1107)
//   k_schedule   order[] of PGS_LCP with row_reorder1 (quickstep_pgs_lcp.cpp:1115-1150) and the
//                dataflow tickets / executor queues described below
//   k_pgs        initial caccel = iMJ * lambda0 (multiply_invM_JT quickstep_util.cpp:101-119),
//                then ComputeRows (quickstep_pgs_lcp.cpp:52-954) on shared-memory-staged rows
//   k_integrate  lambda write-back (quickstep.cpp:523-541), dxUpdateBodies
//                (quickstep_update_bodies.cpp:41-368), dxStepBody (util.cpp:274-379)
//
// Exact-order parallel Gauss-Seidel.  The reference sweeps rows strictly in order[]
// sequence.  Two row updates commute EXACTLY when they share no body, so any schedule
// that keeps, for every body, the order in which rows touch it gives bit-identical
// lambda and caccel.  Each row gets a ticket per body = how many earlier rows (in
// order[] sequence) touch that body; a row may run once both its bodies' completion
// counters equal its tickets.  Sweeps pipeline automatically (sweep s+1 of a row starts
// as soon as its bodies are done with sweep s).  The lambda and lambda_erp solves are
// independent problems (split impulse, quickstep_pgs_lcp.cpp:526-542) and run on
// separate threads with separate counters.
#pragma once
#include "k_rows2.cuh"

namespace gz {

// per-world int scratch (B.scratch): bodies of the rows in order[] sequence
struct WScratch { int *rb1, *rb2; };
__device__ inline WScratch wscratch(const DevBatch &B, int w) {
  WScratch s;
  int *base = B.scratch + (size_t)w * 2 * B.max_row_cap;
  s.rb1 = base; s.rb2 = base + B.max_row_cap;
  return s;
}

// ---- packed row record: what the sweeps stream, 34 doubles = 272 B (16-byte aligned) -------
// [0..11] J (scaled by Ad)   [12..23] iMJ   [24] rhs  [25] rhs_erp  [26] Adcfm  [27] lo  [28] hi
// [29] lambda  [30] lambda_erp  [31] Ad (statistics)  [32..33] RecInts
// Records of a world are stored executor by executor (an executor = the body that owns the row:
// the higher-indexed of its two bodies), each executor's rows in order[] sequence, so a thread
// walks a contiguous range and the friction rows of an executor are the tail of its range.
#define GZ_REC 34
#define GZ_REC_INTS 32
struct RecInts {
  int findex_p;           // -1 bilateral, -2 contact normal, >= 0 PACKED index of the normal row
  unsigned short b1, b2;  // world-local bodies, b2 = 0xFFFF: none
  unsigned short t1, t2;  // tickets: rows touching b1 / b2 that come earlier in order[]
  unsigned short run;     // rows that follow in the same block (same bodies, consecutive tickets)
  unsigned short pad;
};
// per-world schedule table (ints): for body/executor b: nA, nF, qs, qf, qe
#define GZ_SCHED 5

__device__ inline int *wsched(const DevBatch &B, int w) { return B.sched + (size_t)w * GZ_SCHED * B.max_nb; }
__device__ inline double *wprec(const DevBatch &B, const WorldInfo &wi) { return B.prec + (size_t)wi.row0 * GZ_REC; }

// One CTA per world: order[] positions (closed form), per-body tickets, per-executor queues,
// packed position of every row.  Needs only b1/b2 of the rows, which k_rows_* wrote.
// Row::t1a/t2a receive the tickets, Row::next_a the rank inside the owner's queue and
// Row::next_f the packed index.
__global__ void k_schedule(DevBatch B) {
  const int w = blockIdx.x;
  const WorldInfo wi = B.world[w];
  const WorldStep &ws = B.wstep[w];
  const int m = ws.m, nc = ws.n_contacts, mj = m - 3 * nc;
  int *order = B.order + wi.row0;
  Row *rows = B.row + wi.row0;
  WScratch S = wscratch(B, w);
  int *sch = wsched(B, w);
  // order[]: bilateral rows in index order, contact normals in REVERSE index order, friction
  // rows in index order (quickstep_pgs_lcp.cpp:1115-1150).  Persistent-joint rows are all
  // bilateral and every contact has exactly (normal, fric1, fric2): positions are closed-form.
  for (int i = threadIdx.x; i < m; i += blockDim.x) {
    int pos;
    if (i < mj) pos = i;
    else {
      int c = (i - mj) / 3, f = (i - mj) % 3;
      pos = (f == 0) ? mj + (nc - 1 - c) : mj + nc + 2 * c + (f - 1);
    }
    order[pos] = i;
    S.rb1[pos] = rows[i].b1;
    S.rb2[pos] = rows[i].b2;
  }
  __syncthreads();
  const int nfr0 = mj + nc;   // positions >= nfr0 are friction rows
  for (int b = threadIdx.x; b < wi.nb; b += blockDim.x) {
    int cA = 0, cF = 0, nQ = 0, nQF = 0;
    for (int k = 0; k < m; k++) {
      int x1 = S.rb1[k], x2 = S.rb2[k];
      if (x1 != b && x2 != b) continue;
      Row &r = rows[order[k]];
      bool fr = k >= nfr0;
      if (x1 == b) r.t1a = cA; else r.t2a = cA;
      cA++; if (fr) cF++;
      int owner = (x2 >= 0) ? (x1 > x2 ? x1 : x2) : x1;
      if (owner == b) { r.owner = b; r.next_a = nQ; nQ++; if (fr) nQF++; }
    }
    sch[GZ_SCHED * b + 0] = cA; sch[GZ_SCHED * b + 1] = cF;
    sch[GZ_SCHED * b + 2] = nQ; sch[GZ_SCHED * b + 3] = nQF;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int acc = 0;
    for (int b = 0; b < wi.nb; b++) {
      int nQ = sch[GZ_SCHED * b + 2], nQF = sch[GZ_SCHED * b + 3];
      sch[GZ_SCHED * b + 2] = acc; sch[GZ_SCHED * b + 3] = acc + nQ - nQF; sch[GZ_SCHED * b + 4] = acc + nQ;
      acc += nQ;
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < m; i += blockDim.x) {
    Row &r = rows[i];
    r.next_f = sch[GZ_SCHED * r.owner + 2] + r.next_a;
  }
}

// grid (ceil(max_row_cap / blockDim), n_worlds): one thread per row.  Finishes the row and
// writes its packed record.
__global__ void k_rowfinish(DevBatch B) {
  const int w = blockIdx.y;
  const WorldInfo wi = B.world[w];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B.wstep[w].m) return;
  Row &r = B.row[wi.row0 + i];
  const double stepsize1 = 1.0 / B.p.step_size;
  const Body &b1 = B.body[wi.body0 + r.b1];
  const BodyStep &s1 = B.bstep[wi.body0 + r.b1];
  double J[12], iMJ[12];
  for (int j = 0; j < 12; j++) J[j] = r.J[j];
  double t[6];
  for (int j = 0; j < 3; j++) t[j] = b1.facc[j] * b1.invMass + b1.lvel[j] * stepsize1;
  mul0_331(t + 3, s1.invMOI, b1.tacc);
  for (int k = 0; k < 3; k++) t[3 + k] += b1.avel[k] * stepsize1;
  double sum = 0;
  for (int j = 0; j < 6; j++) sum += J[j] * t[j];
  double k1 = b1.invMass;
  for (int j = 0; j < 3; j++) iMJ[j] = k1 * J[j];
  mul0_331(iMJ + 3, s1.invMOI, J + 3);
  if (r.b2 >= 0) {
    const Body &b2 = B.body[wi.body0 + r.b2];
    const BodyStep &s2 = B.bstep[wi.body0 + r.b2];
    for (int j = 0; j < 3; j++) t[j] = b2.facc[j] * b2.invMass + b2.lvel[j] * stepsize1;
    mul0_331(t + 3, s2.invMOI, b2.tacc);
    for (int k = 0; k < 3; k++) t[3 + k] += b2.avel[k] * stepsize1;
    for (int j = 0; j < 6; j++) sum += J[6 + j] * t[j];
    double k2 = b2.invMass;
    for (int j = 0; j < 3; j++) iMJ[6 + j] = k2 * J[6 + j];
    mul0_331(iMJ + 9, s2.invMOI, J + 9);
  } else {
    for (int j = 0; j < 6; j++) iMJ[6 + j] = 0;
  }
  double rhs_erp = r.c * stepsize1 - sum;
  double rhs;
  if (fabs(r.c) > r.c_v_max) rhs = r.c_v_max * stepsize1 - sum;
  else rhs = r.c * stepsize1 - sum;
  double cfm = r.cfm * stepsize1;
  // warm start: lambda of the previous step * warm_start for persistent joints, 0 for contacts
  double lam = 0, lam_erp = 0;
  if (B.p.warm_start > 0 && r.src >= 0) {
    const Joint &j = B.joint[r.src];
    int k = i - j.row0;
    lam = j.lambda[k] * B.p.warm_start;
    lam_erp = j.lambda_erp[k] * B.p.warm_start;
  }
  double d = 0;
  d += iMJ[0] * J[0] + iMJ[1] * J[1] + iMJ[2] * J[2] + iMJ[3] * J[3] + iMJ[4] * J[4] + iMJ[5] * J[5];
  if (r.b2 >= 0)
    d += iMJ[6] * J[6] + iMJ[7] * J[7] + iMJ[8] * J[8] + iMJ[9] * J[9] + iMJ[10] * J[10] + iMJ[11] * J[11];
  double Ad;
  if (r.findex < 0) Ad = B.p.sor_w / (d + cfm);
  else Ad = B.p.contact_sor_scale * B.p.sor_w / (d + cfm);
  r.Ad = Ad;
  r.lambda = lam; r.lambda_erp = lam_erp;
  double *rec = wprec(B, wi) + (size_t)r.next_f * GZ_REC;
  for (int j = 0; j < 12; j++) { rec[j] = J[j] * Ad; rec[12 + j] = iMJ[j]; }
  rec[24] = rhs * Ad;
  rec[25] = rhs_erp * Ad;
  rec[26] = Ad * cfm;
  rec[27] = r.lo; rec[28] = r.hi;
  rec[29] = lam; rec[30] = lam_erp; rec[31] = Ad;
  RecInts ri;
  ri.run = 0; ri.pad = 0;
  B.order[wi.row0 + r.next_f] = i;   // order[] is free after k_schedule: packed index -> row
  ri.findex_p = (r.findex >= 0) ? B.row[wi.row0 + r.findex].next_f : r.findex;
  ri.b1 = (unsigned short)r.b1; ri.b2 = (r.b2 >= 0) ? (unsigned short)r.b2 : (unsigned short)0xFFFF;
  ri.t1 = (unsigned short)r.t1a; ri.t2 = (unsigned short)r.t2a;
  *reinterpret_cast<RecInts *>(rec + GZ_REC_INTS) = ri;
}

}  // namespace gz
#include "k_pgs_sweep.cuh"
namespace gz {

// One thread per joint: save lambda for the next step's warm start.
__global__ void k_lambda_save(DevBatch B) {
  int ji = blockIdx.x * blockDim.x + threadIdx.x;
  if (ji >= B.nj) return;
  Joint &j = B.joint[ji];
  if (j.m == 0 || !(B.p.warm_start > 0)) return;
  const WorldInfo wi = B.world[j.world];
  if (B.wstep[j.world].m == 0) return;
  const Row *rows = B.row + wi.row0 + j.row0;
  for (int k = 0; k < j.m; k++) { j.lambda[k] = rows[k].lambda; j.lambda_erp[k] = rows[k].lambda_erp; }
}

// One thread per body: dxUpdateBodies + dxStepBody.
__global__ void k_integrate(DevBatch B) {
  int bi = blockIdx.x * blockDim.x + threadIdx.x;
  if (bi >= B.nb) return;
  const BodyStep &bs = B.bstep[bi];
  if (!bs.active) return;
  Body &b = B.body[bi];
  const double h = B.p.step_size;
  for (int j = 0; j < 3; j++) {
    b.lvel[j] += h * bs.caccel_erp[j];
    b.avel[j] += h * bs.caccel_erp[3 + j];
  }
  double him = h * b.invMass;
  double tt[3], tmp[3];
  for (int j = 0; j < 3; j++) { b.lvel[j] += him * b.facc[j]; tt[j] = b.tacc[j] * h; }
  mul0_331(tmp, bs.invMOI, tt);
  b.avel[0] = b.avel[0] + tmp[0]; b.avel[1] = b.avel[1] + tmp[1]; b.avel[2] = b.avel[2] + tmp[2];
  // dxStepBody (infinitesimal rotation branch; Gazebo never sets the finite-rotation flags)
  for (int j = 0; j < 3; j++) b.pos[j] += h * b.lvel[j];
  double dq[4];
  dq_from_w(dq, b.avel, b.q);
  for (int j = 0; j < 4; j++) b.q[j] += h * dq[j];
  normalize4(b.q);
  q_to_R(b.q, b.R);
  if (b.lin_damp != 0) {
    double sp = dot3(b.lvel, b.lvel);
    if (sp > B.p.damping_linear_threshold_sq) { double k = 1 - b.lin_damp; b.lvel[0] *= k; b.lvel[1] *= k; b.lvel[2] *= k; }
  }
  if (b.ang_damp != 0) {
    double sp = dot3(b.avel, b.avel);
    if (sp > B.p.damping_angular_threshold_sq) { double k = 1 - b.ang_damp; b.avel[0] *= k; b.avel[1] *= k; b.avel[2] *= k; }
  }
  for (int j = 0; j < 3; j++) {
    double dv = 1.00 * h * (bs.caccel[j] - bs.caccel_erp[j]);
    double da = 1.00 * h * (bs.caccel[3 + j] - bs.caccel_erp[3 + j]);
    b.lvel[j] += dv;
    b.avel[j] += da;
  }
  for (int j = 0; j < 3; j++) { b.facc[j] = 0; b.tacc[j] = 0; }
}

// Host exchange without touching the whole Body record: external wrench in, pose/velocity out.
__global__ void k_apply_wrench(DevBatch B) {
  int bi = blockIdx.x * blockDim.x + threadIdx.x;
  if (bi >= B.nb) return;
  Body &b = B.body[bi];
  const double *in = B.io_in + 6 * (size_t)bi;
  for (int k = 0; k < 3; k++) { b.facc[k] += in[k]; b.tacc[k] += in[3 + k]; }
}
__global__ void k_gather_state(DevBatch B) {
  int bi = blockIdx.x * blockDim.x + threadIdx.x;
  if (bi >= B.nb) return;
  const Body &b = B.body[bi];
  double *o = B.io_out + 13 * (size_t)bi;
  for (int k = 0; k < 3; k++) { o[k] = b.pos[k]; o[7 + k] = b.lvel[k]; o[10 + k] = b.avel[k]; }
  for (int k = 0; k < 4; k++) o[3 + k] = b.q[k];
}

}  // namespace gz
