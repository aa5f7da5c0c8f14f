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

// ---- packed row record: what the sweeps stream, 33 doubles = 264 B --------------------------
// [0..11] J (scaled by Ad)   [12..23] iMJ   [24] rhs  [25] rhs_erp  [26] Adcfm  [27] lo  [28] hi
// [29] lambda  [30] lambda_erp  [31..32] RecInts
// Records of a world are stored executor by executor (an executor = the body that owns the row:
// the higher-indexed of its two bodies), each executor's rows in order[] sequence, so a thread
// walks a contiguous range and the friction rows of an executor are the tail of its range.
#define GZ_REC 33
struct RecInts {
  int orig;               // world-local row index (write-back, statistics)
  int findex_p;           // -1 bilateral, -2 contact normal, >= 0 PACKED index of the normal row
  unsigned short b1, b2;  // world-local bodies, b2 = 0xFFFF: none
  unsigned short t1, t2;  // tickets: rows touching b1 / b2 that come earlier in order[]
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
  rec[29] = lam; rec[30] = lam_erp;
  RecInts ri;
  ri.orig = i;
  ri.findex_p = (r.findex >= 0) ? B.row[wi.row0 + r.findex].next_f : r.findex;
  ri.b1 = (unsigned short)r.b1; ri.b2 = (r.b2 >= 0) ? (unsigned short)r.b2 : (unsigned short)0xFFFF;
  ri.t1 = (unsigned short)r.t1a; ri.t2 = (unsigned short)r.t2a;
  *reinterpret_cast<RecInts *>(rec + 31) = ri;
}

// The SOR-PGS sweeps.  L = 2 * nbp lanes per world (nbp = executors, a power of two >= nb;
// lane = 2 * executor + solve), WPC worlds per CTA.  Dynamic shared memory per CTA:
//   per world slot: ca[2][nbp][6] doubles | stat[nbp][9] doubles | done[2][nbp] ints | nA[nbp], nF[nbp] ints
//   pool: WPC * budget packed records shared by the CTA's worlds (first come first served);
//         records that do not fit are streamed from the packed slab in global memory instead.
struct PgsCfg { int L, nbp, wpc, budget; };

__device__ inline size_t pgs_slot_doubles(int nbp) { return (size_t)12 * nbp + 9 * nbp + 2 * nbp; }

__global__ void k_pgs(DevBatch B, PgsCfg cfg) {
  extern __shared__ double shd[];
  const int L = cfg.L, nbp = cfg.nbp;
  const int slot = threadIdx.x / L, l = threadIdx.x % L;
  const int w = blockIdx.x * cfg.wpc + slot;
  const bool live = w < B.n_worlds;
  const size_t slotd = pgs_slot_doubles(nbp);
  double *ca = shd + slot * slotd;                       // [2][nbp][6]
  double *stat = ca + 12 * nbp;                          // [nbp][9]
  volatile int *done = (volatile int *)(stat + 9 * nbp); // [2][nbp]
  int *nAs = (int *)(stat + 9 * nbp) + 2 * nbp;          // [nbp]
  int *nFs = nAs + nbp;                                  // [nbp]
  double *pool = shd + cfg.wpc * slotd;
  __shared__ int pool_off[33], pool_cap[33];

  WorldInfo wi;
  int nb = 0, m = 0;
  const int *sch = nullptr;
  const double *gslab = nullptr;
  if (live) { wi = B.world[w]; nb = wi.nb; m = B.wstep[w].m; sch = wsched(B, w); gslab = wprec(B, wi); }
  if (threadIdx.x == 0) {
    int used = 0;
    const int total = cfg.wpc * cfg.budget;
    for (int k = 0; k < cfg.wpc; k++) {
      int wk = blockIdx.x * cfg.wpc + k;
      int mk = (wk < B.n_worlds) ? B.wstep[wk].m : 0;
      int c = mk < total - used ? mk : total - used;
      pool_off[k] = used; pool_cap[k] = c; used += c;
    }
  }
  __syncthreads();
  // stage the packed records (coalesced copy), zero the counters, load the per-body tables
  for (int k = 0; k < cfg.wpc; k++) {
    int wk = blockIdx.x * cfg.wpc + k;
    if (wk >= B.n_worlds) break;
    const WorldInfo wk_i = B.world[wk];
    const double *src = wprec(B, wk_i);
    double *dst = pool + (size_t)pool_off[k] * GZ_REC;
    const int n = pool_cap[k] * GZ_REC;
    for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
  }
  const int cap = pool_cap[slot];
  double *mypool = pool + (size_t)pool_off[slot] * GZ_REC;
  for (int i = l; i < nbp; i += L) {
    done[i] = 0; done[nbp + i] = 0;
    nAs[i] = (live && i < nb) ? sch[GZ_SCHED * i + 0] : 0;
    nFs[i] = (live && i < nb) ? sch[GZ_SCHED * i + 1] : 0;
    for (int t = 0; t < 6; t++) { ca[6 * i + t] = 0; ca[6 * (nbp + i) + t] = 0; }
  }
  __syncthreads();
#define GZ_RECPTR(idx) (((idx) < cap) ? (mypool + (size_t)(idx) * GZ_REC) : (const_cast<double *>(gslab) + (size_t)(idx) * GZ_REC))
  // initial caccel = iMJ * lambda0 over the warm-started (persistent-joint) rows, row index order
  if (live && B.p.warm_start > 0 && m > 0) {
    for (int b = l; b < nb; b += L) {
      const Body &bd = B.body[wi.body0 + b];
      if (!B.bstep[wi.body0 + b].active) continue;
      for (int q = bd.joint_begin; q < bd.joint_end; q++) {
        const Joint &j = B.joint[B.body_joint[q]];
        for (int rr = 0; rr < j.m; rr++) {
          const int pidx = B.row[wi.row0 + j.row0 + rr].next_f;
          const double *rec = GZ_RECPTR(pidx);
          const RecInts ri = *reinterpret_cast<const RecInts *>(rec + 31);
          const double *im = (ri.b1 == b) ? rec + 12 : rec + 18;
          for (int t = 0; t < 6; t++) {
            ca[6 * b + t] += im[t] * rec[29];
            ca[6 * (nbp + b) + t] += im[t] * rec[30];
          }
        }
      }
    }
  }
  __syncthreads();
  const int s = l & 1, e = l >> 1;
  const int iters = B.p.iters, total = B.p.iters + B.p.friction_iters;
  const double smooth = B.p.smooth_contacts;
  double *cas = ca + 6 * (size_t)nbp * s;
  volatile int *dn = done + nbp * s;
  double acc[3] = {0, 0, 0}, err[3] = {0, 0, 0};
  int cnt[3] = {0, 0, 0};
  int qs = 0, qf = 0, qe = 0;
  if (live && e < nb && m > 0) { qs = sch[GZ_SCHED * e + 2]; qf = sch[GZ_SCHED * e + 3]; qe = sch[GZ_SCHED * e + 4]; }
  if (qe > qs) {
    int sweep = 0, idx = qs;
    if (iters == 0) { idx = qf; if (qf == qe) sweep = total; }
    // header of the current row, kept in registers while the lane waits for its turn
    double *rec = nullptr;
    int b1 = 0, b2 = -1, fi = -1, tg1 = 0, tg2 = 0;
    bool need_header = true;
    while (sweep < total) {
      if (need_header) {
        rec = GZ_RECPTR(idx);
        const RecInts ri = *reinterpret_cast<const RecInts *>(rec + 31);
        b1 = ri.b1; b2 = (ri.b2 == 0xFFFF) ? -1 : (int)ri.b2; fi = ri.findex_p;
        if (sweep < iters) {
          tg1 = sweep * nAs[b1] + ri.t1;
          if (b2 >= 0) tg2 = sweep * nAs[b2] + ri.t2;
        } else {   // friction-only sweeps: ticket among the friction rows of the body
          tg1 = iters * nAs[b1] + (sweep - iters) * nFs[b1] + (ri.t1 - (nAs[b1] - nFs[b1]));
          if (b2 >= 0) tg2 = iters * nAs[b2] + (sweep - iters) * nFs[b2] + (ri.t2 - (nAs[b2] - nFs[b2]));
        }
        need_header = false;
      }
      const bool ready = (dn[b1] == tg1) && (b2 < 0 || dn[b2] == tg2);
      if (ready) {
        __threadfence_block();
        double *c1 = cas + 6 * b1;
        double *c2 = (b2 >= 0) ? cas + 6 * b2 : nullptr;
        const double old_lambda = rec[29 + s];
        double delta = rec[24 + s] - old_lambda * rec[26];
        delta -= c1[0] * rec[0] + c1[1] * rec[1] + c1[2] * rec[2] + c1[3] * rec[3] + c1[4] * rec[4] + c1[5] * rec[5];
        if (c2) delta -= c2[0] * rec[6] + c2[1] * rec[7] + c2[2] * rec[8] + c2[3] * rec[9] + c2[4] * rec[10] + c2[5] * rec[11];
        double hi_act, lo_act;
        if (fi >= 0) {
          const double *nrec = GZ_RECPTR(fi);
          hi_act = fabs(rec[28] * nrec[29 + s]);   // pyramid friction: |mu * lambda_normal|
          lo_act = -hi_act;
        } else { hi_act = rec[28]; lo_act = rec[27]; }
        double nl = old_lambda + delta;
        if (nl < lo_act) { delta = lo_act - old_lambda; nl = lo_act; }
        else if (nl > hi_act) { delta = hi_act - old_lambda; nl = hi_act; }
        // SMOOTH_LAMBDA on the lambda solve only; delta is not recomputed (quickstep_pgs_lcp.cpp:687-708)
        if (s == 0 && fi != -1) nl = (1.0 - smooth) * nl + smooth * old_lambda;
        rec[29 + s] = nl;
        for (int t = 0; t < 6; t++) c1[t] += delta * rec[12 + t];
        if (c2) for (int t = 0; t < 6; t++) c2[t] += delta * rec[18 + t];
        // Statistics are reset every sweep (friction class) or every non-friction sweep (the
        // other two), so only the last such sweeps survive: accumulate just there.
        if (s == 0 && (sweep == iters - 1 || sweep == total - 1)) {
          const int cls = (fi == -1) ? 0 : (fi == -2 ? 1 : 2);
          if (sweep == iters - 1 || cls == 2) {
            const RecInts ri = *reinterpret_cast<const RecInts *>(rec + 31);
            const double Ad = B.row[wi.row0 + ri.orig].Ad;
            double Ad2 = 0.0;
            if (!dequal(Ad, 0.0)) Ad2 = 1.0 / (Ad * Ad);
            const double d2 = delta * delta;
            acc[cls] += d2; err[cls] += d2 * Ad2; cnt[cls]++;
          }
        }
        __threadfence_block();
        dn[b1] = tg1 + 1;
        if (b2 >= 0) dn[b2] = tg2 + 1;
        idx++;
        need_header = true;
        if (idx == qe) {
          sweep++;
          if (sweep == total - 1 && total - 1 != iters - 1) { acc[2] = 0; err[2] = 0; cnt[2] = 0; }
          if (sweep < iters) idx = qs;
          else { idx = qf; if (qf == qe) sweep = total; }
        }
      }
    }
  }
  if (s == 0) {
    double *st = stat + 9 * e;
    for (int k = 0; k < 3; k++) { st[k] = acc[k]; st[3 + k] = err[k]; st[6 + k] = (double)cnt[k]; }
  }
  __syncthreads();
  if (live) {
    // write lambda back to the rows (k_lambda_save copies them to the joints) and caccel to the bodies
    Row *rows = B.row + wi.row0;
    for (int idx = qs + s; idx < qe; idx += 2) {     // the two solve lanes of an executor split its rows
      const double *rec = GZ_RECPTR(idx);
      const RecInts ri = *reinterpret_cast<const RecInts *>(rec + 31);
      rows[ri.orig].lambda = rec[29];
      rows[ri.orig].lambda_erp = rec[30];
    }
    for (int i = l; i < nb; i += L) {
      BodyStep &bs = B.bstep[wi.body0 + i];
      for (int t = 0; t < 6; t++) { bs.caccel[t] = ca[6 * i + t]; bs.caccel_erp[t] = ca[6 * (nbp + i) + t]; }
    }
    if (l == 0) {
      WorldStep &ws = B.wstep[w];
      double a[3] = {0, 0, 0}, er[3] = {0, 0, 0}, c[3] = {0, 0, 0};
      if (m > 0)
        for (int q = 0; q < nb; q++)
          for (int k = 0; k < 3; k++) { a[k] += stat[9 * q + k]; er[k] += stat[9 * q + 3 + k]; c[k] += stat[9 * q + 6 + k]; }
      // quickstep_pgs_lcp.cpp:814-855
      const double tot = c[0] + c[1] + c[2];
      for (int k = 0; k < 3; k++) {
        ws.rms_dlambda[k] = sqrt(c[k] > 0 ? a[k] / c[k] : 0.0);
        ws.rms_residual[k] = sqrt(c[k] > 0 ? er[k] / c[k] : 0.0);
      }
      ws.rms_dlambda[3] = sqrt((a[0] + a[1] + a[2] > 0) ? (a[0] + a[1] + a[2]) / tot : 0.0);
      ws.rms_residual[3] = sqrt((er[0] + er[1] + er[2] > 0) ? (er[0] + er[1] + er[2]) / tot : 0.0);
      ws.num_contacts_stat = (int)c[1];
    }
  }
#undef GZ_RECPTR
}

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
