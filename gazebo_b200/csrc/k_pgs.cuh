// k_pgs.cuh — row prologue, dataflow schedule, the SOR-PGS sweeps, and integration.
//
//   k_rowfinish  rhs / rhs_erp (quickstep.cpp:403-437), cfm /= h (:443), warm start (:473-495),
//                iMJ = M^-1 J^T (quickstep_util.cpp:79-98), Ad and the Ad scaling
//                (quickstep_pgs_lcp.cpp:1012-1028,1091-1107)
//   k_order      order[] of PGS_LCP with row_reorder1 (quickstep_pgs_lcp.cpp:1115-1150),
//                initial caccel = iMJ * lambda0 (multiply_invM_JT quickstep_util.cpp:101-119),
//                and the dataflow tickets described below
//   k_pgs        ComputeRows (quickstep_pgs_lcp.cpp:52-954)
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

// grid (ceil(max_row_cap / blockDim), n_worlds): one thread per row
__global__ void k_rowfinish(DevBatch B) {
  const int w = blockIdx.y;
  const WorldInfo wi = B.world[w];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B.wstep[w].m) return;
  Row &r = B.row[wi.row0 + i];
  const double stepsize1 = 1.0 / B.p.step_size;
  const Body &b1 = B.body[wi.body0 + r.b1];
  const BodyStep &s1 = B.bstep[wi.body0 + r.b1];
  double t[6];
  for (int j = 0; j < 3; j++) t[j] = b1.facc[j] * b1.invMass + b1.lvel[j] * stepsize1;
  mul0_331(t + 3, s1.invMOI, b1.tacc);
  for (int k = 0; k < 3; k++) t[3 + k] += b1.avel[k] * stepsize1;
  double sum = 0;
  for (int j = 0; j < 6; j++) sum += r.J[j] * t[j];
  double k1 = b1.invMass;
  for (int j = 0; j < 3; j++) r.iMJ[j] = k1 * r.J[j];
  mul0_331(r.iMJ + 3, s1.invMOI, r.J + 3);
  if (r.b2 >= 0) {
    const Body &b2 = B.body[wi.body0 + r.b2];
    const BodyStep &s2 = B.bstep[wi.body0 + r.b2];
    for (int j = 0; j < 3; j++) t[j] = b2.facc[j] * b2.invMass + b2.lvel[j] * stepsize1;
    mul0_331(t + 3, s2.invMOI, b2.tacc);
    for (int k = 0; k < 3; k++) t[3 + k] += b2.avel[k] * stepsize1;
    for (int j = 0; j < 6; j++) sum += r.J[6 + j] * t[j];
    double k2 = b2.invMass;
    for (int j = 0; j < 3; j++) r.iMJ[6 + j] = k2 * r.J[6 + j];
    mul0_331(r.iMJ + 9, s2.invMOI, r.J + 9);
  } else {
    for (int j = 0; j < 6; j++) r.iMJ[6 + j] = 0;
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
  r.lambda = lam; r.lambda_erp = lam_erp;
  // Ad
  double d = 0;
  d += r.iMJ[0] * r.J[0] + r.iMJ[1] * r.J[1] + r.iMJ[2] * r.J[2] + r.iMJ[3] * r.J[3] + r.iMJ[4] * r.J[4] + r.iMJ[5] * r.J[5];
  if (r.b2 >= 0)
    d += r.iMJ[6] * r.J[6] + r.iMJ[7] * r.J[7] + r.iMJ[8] * r.J[8] + r.iMJ[9] * r.J[9] + r.iMJ[10] * r.J[10] + r.iMJ[11] * r.J[11];
  double Ad;
  if (r.findex < 0) Ad = B.p.sor_w / (d + cfm);
  else Ad = B.p.contact_sor_scale * B.p.sor_w / (d + cfm);
  for (int j = 0; j < 12; j++) r.J[j] *= Ad;
  r.rhs = rhs * Ad;
  r.rhs_erp = rhs_erp * Ad;
  r.Adcfm = Ad * cfm;
  r.Ad = Ad;
  r.cfm = cfm;
}

// per-world int scratch layout (B.scratch + w * scratch_stride):
//   [0, cap)        order position -> b1 | (b2 << 16)  ... stored as two ints below
//   rb1[cap], rb2[cap], isfr[cap], nA[max_nb], nF[max_nb], headA[max_nb], headF[max_nb]
struct WScratch {
  int *rb1, *rb2, *nA, *nF, *headA, *headF;
};
__device__ inline WScratch wscratch(const DevBatch &B, int w) {
  WScratch s;
  size_t stride = (size_t)2 * B.max_row_cap + 4 * B.max_nb;
  int *base = B.scratch + (size_t)w * stride;
  s.rb1 = base; s.rb2 = base + B.max_row_cap;
  s.nA = base + 2 * B.max_row_cap; s.nF = s.nA + B.max_nb; s.headA = s.nF + B.max_nb; s.headF = s.headA + B.max_nb;
  return s;
}

// One CTA per world.
__global__ void k_order(DevBatch B) {
  const int w = blockIdx.x;
  const WorldInfo wi = B.world[w];
  const WorldStep &ws = B.wstep[w];
  const int m = ws.m, nc = ws.n_contacts, mj = m - 3 * nc;
  int *order = B.order + wi.row0;
  Row *rows = B.row + wi.row0;
  WScratch S = wscratch(B, w);
  // order[]: bilateral rows in index order, contact normals in REVERSE index order, friction
  // rows in index order.  Persistent-joint rows are all bilateral and every contact has
  // exactly (normal, fric1, fric2), so positions are closed-form.
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
  // tickets and per-executor queues: one thread per body walks order[] once
  for (int b = threadIdx.x; b < wi.nb; b += blockDim.x) {
    int cA = 0, cF = 0, tailA = -1, tailF = -1, hA = -1, hF = -1;
    for (int k = 0; k < m; k++) {
      int x1 = S.rb1[k], x2 = S.rb2[k];
      if (x1 != b && x2 != b) continue;
      int ri = order[k];
      Row &r = rows[ri];
      bool fr = k >= nfr0;
      if (x1 == b) { r.t1a = cA; if (fr) r.t1f = cF; }
      else         { r.t2a = cA; if (fr) r.t2f = cF; }
      cA++; if (fr) cF++;
      int owner = (x2 >= 0) ? (x1 > x2 ? x1 : x2) : x1;
      if (owner == b) {
        r.owner = b; r.next_a = -1; r.next_f = -1;
        if (tailA >= 0) rows[tailA].next_a = ri; else hA = ri;
        tailA = ri;
        if (fr) { if (tailF >= 0) rows[tailF].next_f = ri; else hF = ri; tailF = ri; }
      }
    }
    S.nA[b] = cA; S.nF[b] = cF; S.headA[b] = hA; S.headF[b] = hF;
    // initial caccel = sum over this body's warm-started rows, in row index order
    BodyStep &bs = B.bstep[wi.body0 + b];
    const Body &bd = B.body[wi.body0 + b];
    if (B.p.warm_start > 0 && bs.active) {
      for (int q = bd.joint_begin; q < bd.joint_end; q++) {
        const Joint &j = B.joint[B.body_joint[q]];
        for (int rr = 0; rr < j.m; rr++) {
          const Row &r = rows[j.row0 + rr];
          const double *im = (r.b1 == b) ? r.iMJ : r.iMJ + 6;
          for (int t = 0; t < 6; t++) {
            bs.caccel[t] += im[t] * r.lambda;
            bs.caccel_erp[t] += im[t] * r.lambda_erp;
          }
        }
      }
    }
  }
}

// One CTA per world; thread = (executor e = tid >> 1, solve s = tid & 1).
// dynamic smem: ca[2][nb][6] doubles, done[2][nb] ints, stats
__global__ void k_pgs(DevBatch B) {
  extern __shared__ double shd[];
  const int w = blockIdx.x;
  const WorldInfo wi = B.world[w];
  WorldStep &ws = B.wstep[w];
  const int nb = wi.nb, m = ws.m;
  double *ca = shd;                                   // [2][nb][6]
  volatile int *done = (volatile int *)(shd + 12 * (size_t)B.max_nb);   // [2][nb]
  double *stat = (double *)(shd + 12 * (size_t)B.max_nb) + ((2 * B.max_nb + 1) / 2);   // [blockDim/2][9]
  WScratch S = wscratch(B, w);
  Row *rows = B.row + wi.row0;
  for (int i = threadIdx.x; i < nb; i += blockDim.x) {
    const BodyStep &bs = B.bstep[wi.body0 + i];
    for (int t = 0; t < 6; t++) { ca[6 * i + t] = bs.caccel[t]; ca[6 * (nb + i) + t] = bs.caccel_erp[t]; }
    done[i] = 0; done[nb + i] = 0;
  }
  __syncthreads();
  const int s = threadIdx.x & 1;
  const int iters = B.p.iters, total = B.p.iters + B.p.friction_iters;
  const double smooth = B.p.smooth_contacts;
  double *cas = ca + 6 * (size_t)nb * s;
  volatile int *dn = done + nb * s;
  double acc[3] = {0, 0, 0}, err[3] = {0, 0, 0};
  int cnt[3] = {0, 0, 0};
  // a thread serves executors e, e + E, ... (E = blockDim/2) one after the other would
  // deadlock, so worlds with more bodies than executors fold bodies onto executors by
  // queue concatenation at schedule time; here E >= nb is guaranteed by the launcher.
  const int e = threadIdx.x >> 1;
  if (e < nb && m > 0) {
    int sweep = 0;
    int cur = (iters > 0) ? S.headA[e] : S.headF[e];
    const int nAe_dummy = 0; (void)nAe_dummy;
    // skip to the first sweep that has rows for this executor
    while (sweep < total && cur < 0) { sweep++; cur = (sweep < iters) ? S.headA[e] : S.headF[e]; if (S.headA[e] < 0 && S.headF[e] < 0) sweep = total; }
    while (sweep < total) {
      Row &r = rows[cur];
      const int b1 = r.b1, b2 = r.b2;
      const bool fsweep = sweep >= iters;
      int tg1, tg2 = 0;
      if (!fsweep) { tg1 = sweep * S.nA[b1] + r.t1a; if (b2 >= 0) tg2 = sweep * S.nA[b2] + r.t2a; }
      else {
        tg1 = iters * S.nA[b1] + (sweep - iters) * S.nF[b1] + r.t1f;
        if (b2 >= 0) tg2 = iters * S.nA[b2] + (sweep - iters) * S.nF[b2] + r.t2f;
      }
      bool ready = (dn[b1] == tg1) && (b2 < 0 || dn[b2] == tg2);
      if (ready) {
        __threadfence_block();
        const int fi = r.findex;
        double *c1 = cas + 6 * b1;
        double *c2 = (b2 >= 0) ? cas + 6 * b2 : nullptr;
        const double old_lambda = s ? r.lambda_erp : r.lambda;
        double delta = (s ? r.rhs_erp : r.rhs) - old_lambda * r.Adcfm;
        delta -= c1[0] * r.J[0] + c1[1] * r.J[1] + c1[2] * r.J[2] + c1[3] * r.J[3] + c1[4] * r.J[4] + c1[5] * r.J[5];
        if (c2) delta -= c2[0] * r.J[6] + c2[1] * r.J[7] + c2[2] * r.J[8] + c2[3] * r.J[9] + c2[4] * r.J[10] + c2[5] * r.J[11];
        double hi_act, lo_act;
        if (fi >= 0) {
          const Row &nr = rows[fi];
          hi_act = fabs(r.hi * (s ? nr.lambda_erp : nr.lambda));   // pyramid friction
          lo_act = -hi_act;
        } else { hi_act = r.hi; lo_act = r.lo; }
        double nl = old_lambda + delta;
        if (nl < lo_act) { delta = lo_act - old_lambda; nl = lo_act; }
        else if (nl > hi_act) { delta = hi_act - old_lambda; nl = hi_act; }
        if (s == 0) {
          if (fi != -1) nl = (1.0 - smooth) * nl + smooth * old_lambda;   // SMOOTH_LAMBDA; delta is not recomputed
          r.lambda = nl;
        } else {
          r.lambda_erp = nl;
        }
        for (int t = 0; t < 6; t++) c1[t] += delta * r.iMJ[t];
        if (c2) for (int t = 0; t < 6; t++) c2[t] += delta * r.iMJ[6 + t];
        if (s == 0) {
          int cls = (fi == -1) ? 0 : (fi == -2 ? 1 : 2);
          double Ad2 = 0.0;
          if (!dequal(r.Ad, 0.0)) Ad2 = 1.0 / (r.Ad * r.Ad);
          double d2 = delta * delta;
          acc[cls] += d2; err[cls] += d2 * Ad2; cnt[cls]++;
        }
        __threadfence_block();
        dn[b1] = tg1 + 1;
        if (b2 >= 0) dn[b2] = tg2 + 1;
        // advance
        cur = fsweep ? r.next_f : r.next_a;
        while (cur < 0) {
          sweep++;
          if (sweep >= total) break;
          // statistics are per sweep: class 2 restarts every sweep, classes 0/1 while sweep < iters
          acc[2] = 0; err[2] = 0; cnt[2] = 0;
          if (sweep < iters) { acc[0] = acc[1] = 0; err[0] = err[1] = 0; cnt[0] = cnt[1] = 0; }
          cur = (sweep < iters) ? S.headA[e] : S.headF[e];
        }
      }
    }
  }
  if (s == 0) {
    double *st = stat + 9 * (threadIdx.x >> 1);
    for (int k = 0; k < 3; k++) { st[k] = acc[k]; st[3 + k] = err[k]; st[6 + k] = (double)cnt[k]; }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nb; i += blockDim.x) {
    BodyStep &bs = B.bstep[wi.body0 + i];
    for (int t = 0; t < 6; t++) { bs.caccel[t] = ca[6 * i + t]; bs.caccel_erp[t] = ca[6 * (nb + i) + t]; }
  }
  if (threadIdx.x == 0) {
    double a[3] = {0, 0, 0}, er[3] = {0, 0, 0}, c[3] = {0, 0, 0};
    if (m > 0) {
      int ne = blockDim.x >> 1; if (ne > nb) ne = nb;
      for (int q = 0; q < ne; q++)
        for (int k = 0; k < 3; k++) { a[k] += stat[9 * q + k]; er[k] += stat[9 * q + 3 + k]; c[k] += stat[9 * q + 6 + k]; }
    }
    // quickstep_pgs_lcp.cpp:814-855
    double tot = c[0] + c[1] + c[2];
    for (int k = 0; k < 3; k++) {
      ws.rms_dlambda[k] = sqrt(c[k] > 0 ? a[k] / c[k] : 0.0);
      ws.rms_residual[k] = sqrt(c[k] > 0 ? er[k] / c[k] : 0.0);
    }
    ws.rms_dlambda[3] = sqrt((a[0] + a[1] + a[2] > 0) ? (a[0] + a[1] + a[2]) / tot : 0.0);
    ws.rms_residual[3] = sqrt((er[0] + er[1] + er[2] > 0) ? (er[0] + er[1] + er[2]) / tot : 0.0);
    ws.num_contacts_stat = (int)c[1];
  }
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
