// k_pgs_sweep.cuh — the SOR-PGS sweeps (ComputeRows, quickstep_pgs_lcp.cpp:52-954) as an
// exact-order dataflow over shared-memory-staged packed rows.
//
// Threads: a world gets TW threads (TW = 32 when 2*nbp <= 32, else 2*nbp); lane l < 2*nbp is
// (executor e = l >> 1, solve s = l & 1).  A CTA holds WPC worlds.  When a world fits one warp
// (the batched small-world case) its lanes run in lockstep and need no memory fences; larger
// worlds span several warps and use __threadfence_block around the completion counters.
//
// Shared memory per CTA: per world slot ca[2][nbp+1][6] (slot nbp is a dummy "no body"),
// stat[nbp][9], done[2][nbp+1], nA[nbp], nF[nbp]; then a pool of WPC*budget packed records
// handed out first come first served.  Records that do not fit are read from the packed slab in
// global memory (same layout), so capacity is a performance knob, never a correctness limit.
#pragma once

namespace gz {

struct PgsCfg { int L, nbp, tw, wpc, budget; };

__host__ __device__ inline size_t pgs_slot_doubles(int nbp) {
  // ca 12*(nbp+1) | stat 9*nbp | done 2*(nbp+1) ints + nA,nF 2*nbp ints -> (4*nbp+2)/2 doubles
  size_t n = (size_t)12 * (nbp + 1) + 9 * nbp + 2 * nbp + 2;
  return (n + 1) & ~(size_t)1;   // even: keeps the record pool 16-byte aligned
}

// the per-row constants a lane keeps in registers while it waits for its turn
struct RowRegs {
  double J[12], iM[12], rhs, adcfm, lo, hi, Ad;
  int b1, b2, fi, t1, t2, run;
};

__device__ __forceinline__ void load_row(RowRegs &R, const double *rec, const int S) {
#pragma unroll
  for (int k = 0; k < 12; k++) { R.J[k] = rec[k]; R.iM[k] = rec[12 + k]; }
  R.rhs = rec[24 + S]; R.adcfm = rec[26]; R.lo = rec[27]; R.hi = rec[28]; R.Ad = rec[31];
  const RecInts ri = *reinterpret_cast<const RecInts *>(rec + GZ_REC_INTS);
  R.b1 = ri.b1; R.b2 = (ri.b2 == 0xFFFF) ? -1 : (int)ri.b2; R.fi = ri.findex_p;
  R.t1 = ri.t1; R.t2 = ri.t2; R.run = ri.run;
}

// One solve (S = 0: lambda with smoothing and statistics, S = 1: lambda_erp) of one executor.
// S is a run-time value on purpose: the two solve lanes of an executor sit in the same warp and
// must share one instruction stream.  SMEM = every record of this world is in the shared pool
// (pure LDS path); otherwise records are addressed generically (pool or global slab).
// The next row is fetched into registers right after a row update, inside the same divergent
// region, so a loop trip is "check" or "check + update + fetch" and nothing else.
template <bool MULTIWARP, bool SMEM>
__device__ __forceinline__ void pgs_lane(const DevBatch &B, const int S, const int qs, const int qf, const int qe,
                                         const int cap, double *mypool, double *gslab,
                                         double *cas, volatile int *dn, const int *nAs, const int *nFs,
                                         const int nbp, double *st, int *dbg_iters) {
  const int iters = B.p.iters, total = B.p.iters + B.p.friction_iters;
  const double smooth = B.p.smooth_contacts;
  double acc[3] = {0, 0, 0}, err[3] = {0, 0, 0};
  int cnt[3] = {0, 0, 0};
  int loops = 0, execs = 0;
  if (qe > qs) {
    int sweep = 0, idx = qs;
    if (iters == 0) { idx = qf; if (qf == qe) sweep = total; }
    RowRegs R;
    double *lam = nullptr;      // &lambda (S=0) or &lambda_erp (S=1) of the current row
    int tg1 = 0, tg2 = 0, d2i = nbp;
#define GZ_FETCH()                                                                               \
    {                                                                                            \
      double *rec = (SMEM || idx < cap) ? mypool + (size_t)idx * GZ_REC : gslab + (size_t)idx * GZ_REC; \
      load_row(R, rec, S);                                                                       \
      lam = rec + 29 + S;                                                                        \
      d2i = (R.b2 >= 0) ? R.b2 : nbp;   /* dummy counter slot nbp stays 0 */                     \
      if (sweep < iters) {                                                                       \
        tg1 = sweep * nAs[R.b1] + R.t1;                                                          \
        tg2 = (R.b2 >= 0) ? sweep * nAs[R.b2] + R.t2 : 0;                                        \
      } else { /* friction-only sweeps: ticket among the friction rows of the body */            \
        tg1 = iters * nAs[R.b1] + (sweep - iters) * nFs[R.b1] + (R.t1 - (nAs[R.b1] - nFs[R.b1])); \
        tg2 = (R.b2 >= 0) ? iters * nAs[R.b2] + (sweep - iters) * nFs[R.b2] + (R.t2 - (nAs[R.b2] - nFs[R.b2])) : 0; \
      }                                                                                          \
    }
    if (sweep < total) GZ_FETCH();
    while (sweep < total) {
      loops++;
      const bool ready = (dn[R.b1] == tg1) && (dn[d2i] == tg2);
      if (ready) {
        if (MULTIWARP) __threadfence_block(); else asm volatile("" ::: "memory");
        // A block = R.run + 1 consecutive rows of this queue on the same bodies with consecutive
        // tickets: the bodies are acquired once, their accumulators live in registers across the
        // block, and the counters advance by the block length at the end.
        const int b1 = R.b1, b2 = R.b2, nblk = R.run + 1;
        double *c1 = cas + 6 * b1, *c2 = cas + 6 * b2;
        double a1[6], a2[6];
#pragma unroll
        for (int t = 0; t < 6; t++) { a1[t] = c1[t]; a2[t] = (b2 >= 0) ? c2[t] : 0.0; }
        // Software pipeline without register copies: every field of R is reloaded from the NEXT
        // row of the block right after its last use, so the loads overlap the dependent FP chain
        // of the current row.  (When there is no next row the current record is re-read, which
        // is harmless.)  Safe: a row's lambda is written only by its own update, and a friction
        // row's normal lives in an earlier block of this same lane.
        double old_lambda = *lam;
        double nlam = 0.0;
        if (R.fi >= 0) nlam = ((SMEM || R.fi < cap) ? mypool + (size_t)R.fi * GZ_REC : gslab + (size_t)R.fi * GZ_REC)[29 + S];
        for (int k = 0; k < nblk; k++) {
          const int nidx = (k + 1 < nblk) ? idx + 1 : idx;
          double *rn = (SMEM || nidx < cap) ? mypool + (size_t)nidx * GZ_REC : gslab + (size_t)nidx * GZ_REC;
          double delta = R.rhs - old_lambda * R.adcfm;
          delta -= a1[0] * R.J[0] + a1[1] * R.J[1] + a1[2] * R.J[2] + a1[3] * R.J[3] + a1[4] * R.J[4] + a1[5] * R.J[5];
          if (b2 >= 0)
            delta -= a2[0] * R.J[6] + a2[1] * R.J[7] + a2[2] * R.J[8] + a2[3] * R.J[9] + a2[4] * R.J[10] + a2[5] * R.J[11];
          const double rhs_n = rn[24 + S], adcfm_n = rn[26], old_n = rn[29 + S];
#pragma unroll
          for (int t = 0; t < 12; t++) R.J[t] = rn[t];
          double hi_act, lo_act;
          const int fi = R.fi;
          if (fi >= 0) {
            hi_act = fabs(R.hi * nlam);   // pyramid friction: |mu * lambda_normal|
            lo_act = -hi_act;
          } else { hi_act = R.hi; lo_act = R.lo; }
          double nl = old_lambda + delta;
          if (nl < lo_act) { delta = lo_act - old_lambda; nl = lo_act; }
          else if (nl > hi_act) { delta = hi_act - old_lambda; nl = hi_act; }
          // SMOOTH_LAMBDA on the lambda solve only; delta is not recomputed (quickstep_pgs_lcp.cpp:687-708)
          if (S == 0 && fi != -1) nl = (1.0 - smooth) * nl + smooth * old_lambda;
          *lam = nl;
#pragma unroll
          for (int t = 0; t < 6; t++) a1[t] += delta * R.iM[t];
          if (b2 >= 0) {
#pragma unroll
            for (int t = 0; t < 6; t++) a2[t] += delta * R.iM[6 + t];
          }
#pragma unroll
          for (int t = 0; t < 12; t++) R.iM[t] = rn[12 + t];
          // Statistics restart every sweep (friction class) or every non-friction sweep (the
          // other two), so only the last such sweeps survive: accumulate just there.
          if (S == 0 && (sweep == iters - 1 || sweep == total - 1)) {
            const int cls = (fi == -1) ? 0 : (fi == -2 ? 1 : 2);
            if (sweep == iters - 1 || cls == 2) {
              double Ad2 = 0.0;
              if (!dequal(R.Ad, 0.0)) Ad2 = 1.0 / (R.Ad * R.Ad);
              const double d2 = delta * delta;
              acc[cls] += d2; err[cls] += d2 * Ad2; cnt[cls]++;
            }
          }
          // remaining fields of the next row
          R.rhs = rhs_n; R.adcfm = adcfm_n; R.lo = rn[27]; R.hi = rn[28]; R.Ad = rn[31];
          R.fi = reinterpret_cast<const RecInts *>(rn + GZ_REC_INTS)->findex_p;
          old_lambda = old_n;
          lam = rn + 29 + S;
          if (R.fi >= 0) nlam = ((SMEM || R.fi < cap) ? mypool + (size_t)R.fi * GZ_REC : gslab + (size_t)R.fi * GZ_REC)[29 + S];
          idx++;
          execs++;
        }
#pragma unroll
        for (int t = 0; t < 6; t++) c1[t] = a1[t];
        if (b2 >= 0) {
#pragma unroll
          for (int t = 0; t < 6; t++) c2[t] = a2[t];
        }
        if (MULTIWARP) __threadfence_block(); else asm volatile("" ::: "memory");
        dn[b1] = tg1 + nblk;
        if (b2 >= 0) dn[b2] = tg2 + nblk;
        if (idx == qe) {
          sweep++;
          if (S == 0 && sweep == total - 1 && total - 1 != iters - 1) { acc[2] = 0; err[2] = 0; cnt[2] = 0; }
          if (sweep < iters) idx = qs;
          else { idx = qf; if (qf == qe) sweep = total; }
        }
        if (sweep < total) GZ_FETCH();
      }
    }
#undef GZ_FETCH
  }
  dbg_iters[0] = loops; dbg_iters[1] = execs;
  if (S == 0) {
    for (int k = 0; k < 3; k++) { st[k] = acc[k]; st[3 + k] = err[k]; st[6 + k] = (double)cnt[k]; }
  }
}

}  // namespace gz

namespace gz {

template <bool MULTIWARP>
__global__ void k_pgs(DevBatch B, PgsCfg cfg) {
  extern __shared__ __align__(16) double shd[];
  const int L = cfg.L, nbp = cfg.nbp, TW = cfg.tw;
  const int slot = threadIdx.x / TW, l = threadIdx.x % TW;
  const int w = blockIdx.x * cfg.wpc + slot;
  const bool live = w < B.n_worlds;
  const size_t slotd = pgs_slot_doubles(nbp);
  double *ca = shd + slot * slotd;                              // [2][nbp+1][6]
  double *stat = ca + 12 * (nbp + 1);                           // [nbp][9]
  volatile int *done = (volatile int *)(stat + 9 * nbp);        // [2][nbp+1]
  int *nAs = (int *)(stat + 9 * nbp) + 2 * (nbp + 1);           // [nbp]
  int *nFs = nAs + nbp;                                         // [nbp]
  double *pool = shd + cfg.wpc * slotd;
  __shared__ int pool_off[33], pool_cap[33];

  WorldInfo wi;
  int nb = 0, m = 0;
  const int *sch = nullptr;
  double *gslab = nullptr;
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
  // stage the packed records with 16-byte copies (records are 272 B, slabs 16-byte aligned)
  for (int k = 0; k < cfg.wpc; k++) {
    int wk = blockIdx.x * cfg.wpc + k;
    if (wk >= B.n_worlds) break;
    const WorldInfo wk_i = B.world[wk];
    const double2 *src = reinterpret_cast<const double2 *>(wprec(B, wk_i));
    double2 *dst = reinterpret_cast<double2 *>(pool + (size_t)pool_off[k] * GZ_REC);
    const int n = pool_cap[k] * (GZ_REC / 2);
    for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
  }
  const int cap = pool_cap[slot];
  double *mypool = pool + (size_t)pool_off[slot] * GZ_REC;
  for (int i = l; i <= nbp; i += TW) {
    done[i] = 0; done[nbp + 1 + i] = 0;
    for (int t = 0; t < 6; t++) { ca[6 * i + t] = 0; ca[6 * (nbp + 1 + i) + t] = 0; }
    if (i < nbp) {
      nAs[i] = (live && i < nb) ? sch[GZ_SCHED * i + 0] : 0;
      nFs[i] = (live && i < nb) ? sch[GZ_SCHED * i + 1] : 0;
    }
  }
  __syncthreads();
  // initial caccel = iMJ * lambda0 over the warm-started (persistent-joint) rows, in row index
  // order (multiply_invM_JT, quickstep_util.cpp:101-119)
  if (live && B.p.warm_start > 0 && m > 0) {
    for (int b = l; b < nb; b += TW) {
      const Body &bd = B.body[wi.body0 + b];
      if (!B.bstep[wi.body0 + b].active) continue;
      for (int q = bd.joint_begin; q < bd.joint_end; q++) {
        const Joint &j = B.joint[B.body_joint[q]];
        for (int rr = 0; rr < j.m; rr++) {
          const int pidx = B.row[wi.row0 + j.row0 + rr].next_f;
          const double *rec = (pidx < cap) ? mypool + (size_t)pidx * GZ_REC : gslab + (size_t)pidx * GZ_REC;
          const RecInts ri = *reinterpret_cast<const RecInts *>(rec + GZ_REC_INTS);
          const double *im = (ri.b1 == b) ? rec + 12 : rec + 18;
          for (int t = 0; t < 6; t++) {
            ca[6 * b + t] += im[t] * rec[29];
            ca[6 * (nbp + 1 + b) + t] += im[t] * rec[30];
          }
        }
      }
    }
  }
  __syncthreads();
  const int s = l & 1, e = l >> 1;
  int qs = 0, qf = 0, qe = 0;
  if (live && l < L && e < nb && m > 0) { qs = sch[GZ_SCHED * e + 2]; qf = sch[GZ_SCHED * e + 3]; qe = sch[GZ_SCHED * e + 4]; }
  // block lengths: walk the executor's queue backwards; a row joins the block of its successor
  // when both touch the same bodies with consecutive tickets (and the successor does not start
  // the friction-only part of the queue)
  if (s == 0) {
    int run_next = 0, nb1 = -1, nb2 = -1, nt1 = 0, nt2 = 0;
    for (int idx = qe - 1; idx >= qs; idx--) {
      double *rec = (idx < cap) ? mypool + (size_t)idx * GZ_REC : gslab + (size_t)idx * GZ_REC;
      RecInts ri = *reinterpret_cast<const RecInts *>(rec + GZ_REC_INTS);
      int run = 0;
      if (idx + 1 < qe && idx + 1 != qf && ri.b1 == nb1 && ri.b2 == nb2 && ri.t1 + 1 == nt1 &&
          (ri.b2 == 0xFFFF || ri.t2 + 1 == nt2) && run_next < 60000)
        run = run_next + 1;
      ri.run = (unsigned short)run;
      *reinterpret_cast<RecInts *>(rec + GZ_REC_INTS) = ri;
      run_next = run; nb1 = ri.b1; nb2 = ri.b2; nt1 = ri.t1; nt2 = ri.t2;
    }
  }
  if (MULTIWARP) __syncthreads(); else __syncwarp();
  int dbg[2] = {0, 0};
  const long long t0 = clock64();
  if (m <= cap)   // uniform per world (per warp in the single-warp case): all rows are in the pool
    pgs_lane<MULTIWARP, true>(B, s, qs, qf, qe, cap, mypool, gslab, ca + 6 * (size_t)(nbp + 1) * s,
                              done + (nbp + 1) * s, nAs, nFs, nbp, stat + 9 * (e < nbp ? e : 0), dbg);
  else
    pgs_lane<MULTIWARP, false>(B, s, qs, qf, qe, cap, mypool, gslab, ca + 6 * (size_t)(nbp + 1) * s,
                               done + (nbp + 1) * s, nAs, nFs, nbp, stat + 9 * (e < nbp ? e : 0), dbg);
  const long long t1 = clock64();
  if (live) {   // instrumentation: longest lane of the world (loop trips, row updates) and its cycles
    WorldStep &wsd = B.wstep[w];
    atomicMax(&wsd.pad[0], dbg[0]);
    atomicMax(&wsd.pad[1], dbg[1]);
    atomicMax(&wsd.n_fric, (int)(t1 - t0));
  }
  if (MULTIWARP) __syncthreads(); else __syncwarp();
  if (live) {
    // lambda back to the rows (k_lambda_save copies them to the joints), caccel to the bodies
    Row *rows = B.row + wi.row0;
    for (int idx = qs + s; idx < qe; idx += 2) {   // the two solve lanes of an executor split its rows
      const double *rec = (idx < cap) ? mypool + (size_t)idx * GZ_REC : gslab + (size_t)idx * GZ_REC;
      const int orig = B.order[wi.row0 + idx];
      rows[orig].lambda = rec[29];
      rows[orig].lambda_erp = rec[30];
    }
    for (int i = l; i < nb; i += TW) {
      BodyStep &bs = B.bstep[wi.body0 + i];
      for (int t = 0; t < 6; t++) { bs.caccel[t] = ca[6 * i + t]; bs.caccel_erp[t] = ca[6 * (nbp + 1 + i) + t]; }
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
}

}  // namespace gz
