// k_pgs_sweep.cuh — the SOR-PGS sweeps (ComputeRows, quickstep_pgs_lcp.cpp:52-954) run from the
// static modulo schedule k_schedule built (k_pgs.cuh).
//
// One warp per CTA.  A warp holds 16/k worlds of the same class k (lanes per solve); lane l is
// (world slot l / 2k, schedule lane j = (l % 2k) / 2, solve S = l & 1: lambda or lambda_erp).
// In round T a lane takes the record at (slot T mod R, lane j) of its world and, if that row
// is due (sweep = T / R - stage inside the sweep range), updates it.  All lanes of the warp run
// the same instruction stream with no waiting and no inter-lane traffic except through the
// bodies' accumulators; rows that run in the same round share no body by construction, so one
// __syncwarp per round orders everything and the result is bit-identical to the sequential sweep.
//
// With 4096 small worlds there is about one such warp per SM sub-partition, so the kernel runs at
// the LATENCY of one round; everything below is arranged to keep that short:
//  * row records stream from the packed slab (L2-resident: k_rowfinish wrote it just before)
//    into a shared-memory ring with bulk async copies (cp.async.bulk + mbarrier, one copy of
//    k * 272 bytes per world and round, issued by the world's first lane) several rounds ahead;
//  * the record of round T+1 is read from the ring into registers while round T computes;
//  * the two 6-term dot products are issued interleaved and without branches.
// Shared memory per CTA: accumulators ca[world][solve][nbp+1][6] (slot nbp = "no body"), the
// ring, the mbarriers, then lambda[world][slot][2] (spilled to global memory if it does not fit).
#pragma once

namespace gz {

#define GZ_RING_D 2                 // ring stages
#ifndef GZ_RING_G
#define GZ_RING_G 4                 // rounds per stage (2, 3 or 4): one fetch (and one mbarrier wait) per G rounds
#endif
#define GZ_STAGE_DOUBLES (GZ_RING_G * GZ_ROUND_DOUBLES)
#define GZ_RING_DOUBLES (GZ_RING_D * GZ_STAGE_DOUBLES + 2 * GZ_RING_D)   // records + one mbarrier per stage (padded)

// part / rsplit: k_pgs can be launched in two parts (GZCUDA_PGS_RSPLIT): part 1 takes the warps whose period is
// >= rsplit (they run longest) and goes first at high priority, part 2 the others; part 0 = all warps
struct PgsCfg { int smem_doubles, nbp, wpw_max, part, rsplit; };

// the constants of one row, in registers
struct RowRegs {
  double J[12], iM[12], rhs, adcfm, lo, hi;
  int fi, b1, b2, stage, orig;
};

// Reads one record; a record whose epoch is not this step's is an empty slot (or a stale one in a
// part of the slab no world of this warp owns) and is turned into a harmless no-op row.
// epoch_sh = epoch << 8; dummy_bb = (nbp << 16): body 1 = 0, body 2 = the all-zero dummy.
__device__ __forceinline__ void load_row(RowRegs &R, const double *rec, const int S, const unsigned epoch_sh,
                                         const unsigned dummy_bb) {
  const double2 *r2 = reinterpret_cast<const double2 *>(rec);
#pragma unroll
  for (int k = 0; k < 6; k++) {
    const double2 a = r2[k], b = r2[6 + k];
    R.J[2 * k] = a.x; R.J[2 * k + 1] = a.y;
    R.iM[2 * k] = b.x; R.iM[2 * k + 1] = b.y;
  }
  const double2 d = r2[13];
  R.rhs = rec[24 + S]; R.adcfm = d.x; R.lo = d.y; R.hi = rec[28];
  const int4 iv = *reinterpret_cast<const int4 *>(rec + GZ_REC_INTS);   // RecInts: fi | b1,b2 | epoch,stage | orig
  const bool valid = ((unsigned)iv.z ^ epoch_sh) < 256u;
  const unsigned bb = valid ? (unsigned)iv.y : dummy_bb;
  R.fi = valid ? iv.x : GZ_FI_EMPTY;
  R.b1 = bb & 0xFFFFu;
  R.b2 = bb >> 16;
  R.stage = iv.z & 255;
  R.orig = iv.w;
}

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ unsigned smid() {
  unsigned s;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(s));
  return s & 511u;
}

// ---- mbarrier / bulk-copy wrappers (PTX ISA: mbarrier, cp.async.bulk) ---------------------------
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void *src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// One row update of one solve.  Mirrors quickstep_pgs_lcp.cpp:488-808 for the pyramid friction
// model Gazebo uses by default: delta, limits from lo/hi or |hi * lambda[findex]|, clamp,
// SMOOTH_LAMBDA on the lambda solve only (delta is not recomputed after it), caccel update.
// Branch-free, stores under the predicate `act`, so that the compiler can schedule a whole
// round as one basic block.  A row without a second body has J2 = iMJ2 = 0 and reads the
// all-zero dummy accumulator; the reference skips that term, and so does the select below.
struct PgsStat { double acc[3], err[3]; int cnt[3]; };

// what a row update reads before it can start: the accumulators of its bodies and two lambdas
struct RowIn { double a1[6], a2[6], old_lambda, nlam; };

__device__ __forceinline__ void pgs_row_in(RowIn &I, const RowRegs &R, const int S, const int idx, const double *cas,
                                           const double *lam, const bool box) {
  const double2 *p1 = reinterpret_cast<const double2 *>(cas + 6 * R.b1);
  const double2 *p2 = reinterpret_cast<const double2 *>(cas + 6 * R.b2);
#pragma unroll
  for (int t = 0; t < 3; t++) {
    const double2 x = p1[t], y = p2[t];
    I.a1[2 * t] = x.x; I.a1[2 * t + 1] = x.y; I.a2[2 * t] = y.x; I.a2[2 * t + 1] = y.y;
  }
  I.old_lambda = lam[2 * idx + S];
  I.nlam = lam[2 * (R.fi >= 0 ? R.fi : idx) + S];
  if (box) I.nlam = 1.0;   // box friction (quickstep_pgs_lcp.cpp:588-594): limits +-hi, i.e. |hi * 1|
}

// The arithmetic of one row update given the accumulators of its bodies: returns delta, sets nl.
template <bool TWO>
__device__ __forceinline__ double pgs_delta(const RowRegs &R, const double *a1, const double *a2, const bool two,
                                            const double old_lambda, const double nlam, const int S, const double smooth,
                                            double &nl_out) {
  const int fi = R.fi;
  const double hf = fabs(R.hi * nlam);   // pyramid friction: |mu * lambda_normal|
  const double hi_act = fi >= 0 ? hf : R.hi, lo_act = fi >= 0 ? -hf : R.lo;
  // the two dot products, term by term in the reference's left-to-right order, interleaved
  double d1 = a1[0] * R.J[0], d2 = 0.0;
  if (TWO) d2 = a2[0] * R.J[6];
#pragma unroll
  for (int t = 1; t < 6; t++) { d1 += a1[t] * R.J[t]; if (TWO) d2 += a2[t] * R.J[6 + t]; }
  double delta = R.rhs - old_lambda * R.adcfm;
  delta -= d1;
  if (TWO) { const double delta2 = delta - d2; delta = two ? delta2 : delta; }
  double nl = old_lambda + delta;
  const double dlo = lo_act - old_lambda, dhi = hi_act - old_lambda;
  const bool below = nl < lo_act, above = !below && nl > hi_act;
  delta = below ? dlo : (above ? dhi : delta);
  nl = below ? lo_act : (above ? hi_act : nl);
  const double sm = (1.0 - smooth) * nl + smooth * old_lambda;
  nl_out = (S == 0 && fi != -1) ? sm : nl;
  return delta;
}

// rms_dlambda / rms_constraint_residual (quickstep_pgs_lcp.cpp:700-760,814-855): the sums
// restart every sweep, so only the last sweep that touches a class survives.
__device__ __forceinline__ void pgs_stat(PgsStat &st, const int fi, const bool on01, const bool on2, const double delta,
                                         const double Ad2) {
  const int cls = (fi == -1) ? 0 : (fi == -2 ? 1 : 2);
  const bool on = (cls == 2) ? on2 : on01;
  const double dd = delta * delta, e2 = dd * Ad2;
  if (on && cls == 0) { st.acc[0] += dd; st.err[0] += e2; st.cnt[0]++; }
  if (on && cls == 1) { st.acc[1] += dd; st.err[1] += e2; st.cnt[1]++; }
  if (on && cls == 2) { st.acc[2] += dd; st.err[2] += e2; st.cnt[2]++; }
}

template <bool STATS, bool TWO>
__device__ __forceinline__ void pgs_row(const RowRegs &R, const RowIn &I, const bool act, const int S, const int idx,
                                        double *cas, double *lam, const int nbp, const double smooth, const double Ad2,
                                        const bool stat01, const bool stat2, PgsStat &st) {
  // TWO = false: no due row of this warp has a second body in this round (warp-uniform), so the
  // second body's half of the work is not even issued
  double *c1 = cas + 6 * R.b1, *c2 = cas + 6 * R.b2;
  const bool two = TWO && R.b2 < nbp;
  const double *a1 = I.a1, *a2 = I.a2;
  double nl;
  const double delta = pgs_delta<TWO>(R, a1, a2, two, I.old_lambda, I.nlam, S, smooth, nl);
  if (act) lam[2 * idx + S] = nl;
  {
    double2 *p1 = reinterpret_cast<double2 *>(c1);
    double2 *p2 = reinterpret_cast<double2 *>(c2);
    double2 o1[3], o2[3];
#pragma unroll
    for (int t = 0; t < 3; t++) {
      o1[t] = make_double2(a1[2 * t] + delta * R.iM[2 * t], a1[2 * t + 1] + delta * R.iM[2 * t + 1]);
      if (TWO) o2[t] = make_double2(a2[2 * t] + delta * R.iM[6 + 2 * t], a2[2 * t + 1] + delta * R.iM[6 + 2 * t + 1]);
    }
    if (act) {
#pragma unroll
      for (int t = 0; t < 3; t++) p1[t] = o1[t];
    }
    if (TWO && act && two) {
#pragma unroll
      for (int t = 0; t < 3; t++) p2[t] = o2[t];
    }
  }
  if (STATS) pgs_stat(st, R.fi, act && S == 0 && stat01, act && S == 0 && stat2, delta, Ad2);
}

template <bool PROF, bool BOX>
__global__ void __launch_bounds__(32) k_pgs(DevBatch B, PgsCfg cfg) {
  extern __shared__ __align__(16) double shd[];
  const int lane = threadIdx.x;
  const int nbp = cfg.nbp;
  // ---- which bucket, which warp of it (buckets are laid out heaviest first by k_slab)
  const int cta = blockIdx.x;
  const int *swarp0 = bk_row(B, GZ_BK_SWARP0);
  if (cta >= swarp0[GZ_NBUCKET]) return;
  int lo_i = 0, hi_i = GZ_NBUCKET - 1;
  while (lo_i < hi_i) {   // largest i with swarp0[i] <= cta
    const int mid = (lo_i + hi_i + 1) >> 1;
    if (swarp0[mid] <= cta) lo_i = mid; else hi_i = mid - 1;
  }
  const int bucket = bk_row(B, GZ_BK_SBUCKET)[lo_i];
  const int k = 1 << ((bucket / GZ_NRUNG) % 5), LW = 2 * k, R = rung_value(bucket % GZ_NRUNG);
  if (bucket >= 5 * GZ_NRUNG) return;          // chain-mode worlds: k_pgs_chain (k_pgs_chain.cuh)
  if (cfg.part == 1 && R < cfg.rsplit) return;
  if (cfg.part == 2 && R >= cfg.rsplit) return;
  const int wpw = bucket_wpw(B, bucket);
  const int wb = cta - swarp0[lo_i];
  const size_t slab_rec0 = (size_t)bk_row(B, GZ_BK_REC0)[bucket] + (size_t)wb * R * 16;
  const double *slab = B.prec + slab_rec0 * GZ_REC;
  const int wslot = lane / LW, jl = (lane % LW) >> 1, S = lane & 1, lw = lane % LW;
  const int w = (wslot < wpw) ? B.warp_world[cta * 16 + wslot] : -1;
  const bool live = w >= 0;
  WorldInfo wi;
  wi.nb = 0; wi.row0 = 0; wi.body0 = 0;
  int nstage = 0;
  if (live) {
    wi = B.world[w];
    nstage = reinterpret_cast<const WSched *>(B.wsched)[w].nstage;
  }
  const int nslots = R * k;
  // ---- carve shared memory: ca | ring + mbarriers | lambda (first come first served over world slots)
  const int ca_doubles = 12 * (nbp + 1);
  double *ca = shd + (size_t)(wslot < wpw ? wslot : 0) * ca_doubles;
  double *ring = shd + (size_t)wpw * ca_doubles;
  // lambda[nslots][2] per world; a world whose lambdas do not fit spills to its global region
  const int lam_need = 2 * nslots;
  int lam_off = wpw * ca_doubles + GZ_RING_DOUBLES, my_lam_off = -1;
  for (int s = 0; s < wpw; s++) {
    const int need = __shfl_sync(0xffffffffu, lam_need, s * LW);
    const bool fits = lam_off + need <= cfg.smem_doubles;
    if (s == wslot && fits) my_lam_off = lam_off;
    if (fits) lam_off += need;
  }
  double *lam = (my_lam_off >= 0) ? shd + my_lam_off : B.lamg + 2 * (3 * (size_t)wi.row0 + 128 * (size_t)(w < 0 ? 0 : w));
  bool live_ok = live;
  if (live && my_lam_off < 0 && nslots > 3 * wi.row_cap + 128) {   // cannot happen with sane inputs
    if (lw == 0) B.wstep[w].overflow |= 8;
    live_ok = false;
  }
  if (!live) lam = shd;
  const unsigned ring_s = (unsigned)__cvta_generic_to_shared(ring);
  const unsigned bar_s = ring_s + GZ_RING_D * GZ_STAGE_DOUBLES * 8;
  if (lane < GZ_RING_D) mbar_init(bar_s + 8 * lane, 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  const unsigned epoch = gz_epoch(B.step_index);
  WSlab sl;
  sl.base = const_cast<double *>(slab); sl.k = k; sl.wslot = wslot; sl.warp = cta; sl.width = 16;
  // ---- lambda0 and the accumulators
  if (wslot < wpw)
    for (int i = lw; i < ca_doubles; i += LW) ca[i] = 0.0;
  __syncwarp();
  if (live_ok) {
    const Row *rows = B.row + wi.row0;
    for (int i = lw; i < nslots; i += LW) {
      const RecInts *ri = reinterpret_cast<const RecInts *>(slab_rec(sl, i) + GZ_REC_INTS);
      double l0 = 0.0, l1 = 0.0;
      if (((unsigned)ri->stage_epoch >> 8) == epoch && ri->fi >= GZ_FI_EMPTY + 1) { const Row &r = rows[ri->orig]; l0 = r.lambda; l1 = r.lambda_erp; }
      lam[2 * i] = l0; lam[2 * i + 1] = l1;
    }
    // initial caccel = iMJ * lambda0 over the warm-started (persistent-joint) rows, in row index
    // order per body (multiply_invM_JT, quickstep_util.cpp:101-119)
    if (B.p.warm_start > 0) {
      for (int b = lw; b < wi.nb; b += LW) {
        const Body &bd = B.body[wi.body0 + b];
        if (!B.bstep[wi.body0 + b].active) continue;
        double s0[6] = {0, 0, 0, 0, 0, 0}, s1[6] = {0, 0, 0, 0, 0, 0};
        for (int q = bd.joint_begin; q < bd.joint_end; q++) {
          const Joint &j = B.joint[B.body_joint[q]];
          for (int rr = 0; rr < j.m; rr++) {
            const Row &r = rows[j.row0 + rr];
            const double *rec = slab_rec(sl, r.next_f);
            const double *im = (r.b1 == b) ? rec + 12 : rec + 18;
            for (int t = 0; t < 6; t++) { s0[t] += im[t] * r.lambda; s1[t] += im[t] * r.lambda_erp; }
          }
        }
        for (int t = 0; t < 6; t++) { ca[6 * b + t] = s0[t]; ca[6 * (nbp + 1 + b) + t] = s1[t]; }
      }
    }
  }
  __syncwarp();
  // ---- the sweeps
  const int iters = B.p.iters, total = B.p.iters + B.p.friction_iters;
  const double smooth = B.p.smooth_contacts;
  constexpr bool box = BOX;   // box friction model (friction_model == 2), chosen at launch
  const bool on = live_ok && total > 0;
  const int max_nstage = __reduce_max_sync(0xffffffffu, nstage);
  const int rounds = (total > 0 && max_nstage > 0) ? (total + max_nstage - 1) * R : 0;
  double *cas = ca + (size_t)6 * (nbp + 1) * S;
  PgsStat st;
  for (int t = 0; t < 3; t++) { st.acc[t] = 0; st.err[t] = 0; st.cnt[t] = 0; }
  const int pair = lane >> 1;
  const unsigned epoch_sh = epoch << 8, dummy_bb = (unsigned)nbp << 16;
  const int ngroups = (rounds + GZ_RING_G - 1) / GZ_RING_G;
  int fs = 0;                                   // producer side (lane 0): first slot of the next group to fetch
  // one group = GZ_RING_G consecutive rounds = consecutive slots of the slab (wrapping at R)
#define GZ_FETCH(g)                                                                               \
  if (lane == 0 && (g) < ngroups) {                                                               \
    const unsigned bar = bar_s + 8 * ((g) & (GZ_RING_D - 1));                                     \
    unsigned dst = ring_s + (unsigned)(((g) & (GZ_RING_D - 1)) * GZ_STAGE_DOUBLES * 8);           \
    mbar_expect_tx(bar, GZ_STAGE_DOUBLES * 8);                                                    \
    int left = GZ_RING_G;                                                                         \
    while (left > 0) {                                                                            \
      const int n = min(left, R - fs);                                                            \
      bulk_g2s(dst, slab + (size_t)fs * GZ_ROUND_DOUBLES, (unsigned)(n * GZ_ROUND_DOUBLES * 8), bar); \
      dst += (unsigned)(n * GZ_ROUND_DOUBLES * 8);                                                \
      left -= n; fs += n;                                                                         \
      if (fs == R) fs = 0;                                                                        \
    }                                                                                             \
  }
#define GZ_RECP(g, i) (ring + (size_t)((g) & (GZ_RING_D - 1)) * GZ_STAGE_DOUBLES + (size_t)(i) * GZ_ROUND_DOUBLES + pair * GZ_REC)
#define GZ_ROUND(CUR, NXT, g, i, NG, NI)        /* CUR = record of round (g, i); reads NXT = (NG, NI) */ \
  {                                                                                               \
    long long c0 = 0, c3 = 0;                                                                     \
    if (prof) c0 = clock64();                                                                     \
    const int sweep = period - CUR.stage;                                                         \
    const bool act = on && (unsigned)sweep < (unsigned)(CUR.fi >= 0 ? total : (CUR.fi >= -2 ? iters : 0)); \
    const bool st_round = (period >= iters - 1 && period < iters - 1 + max_nstage) ||             \
                          (period >= total - 1 && period < total - 1 + max_nstage);               \
    const int idx = on ? slot * k + jl : 0;                                                       \
    RowIn in;                                                                                     \
    pgs_row_in(in, CUR, S, idx, cas, lam, box); /* issue these loads first, then the next record's */ \
    load_row(NXT, GZ_RECP(NG, NI), S, epoch_sh, dummy_bb);                                        \
    const unsigned due = __ballot_sync(0xffffffffu, act);                                         \
    const unsigned due2 = __ballot_sync(0xffffffffu, act && CUR.b2 < nbp);                        \
    if (!due) { /* nothing due in this round (friction-only sweeps, pipeline ends) */ }           \
    else if (st_round)                                                                            \
      pgs_row<true, true>(CUR, in, act, S, idx, cas, lam, nbp, smooth, GZ_RECP(g, i)[29], sweep == iters - 1, sweep == total - 1, st); \
    else if (due2)                                                                                \
      pgs_row<false, true>(CUR, in, act, S, idx, cas, lam, nbp, smooth, 0.0, false, false, st);       \
    else                                                                                          \
      pgs_row<false, false>(CUR, in, act, S, idx, cas, lam, nbp, smooth, 0.0, false, false, st);      \
    __syncwarp();                                                                                 \
    if (++slot == R) { slot = 0; period++; }                                                      \
    if (prof) {                                                                                   \
      c3 = clock64();                                                                             \
      pc[3] += c3 - c0;                                                                           \
      pc[4] += __popc(__ballot_sync(0xffffffffu, act));                                           \
    }                                                                                             \
  }
  constexpr bool prof = PROF;
  long long pc[5] = {0, 0, 0, 0, 0};
  int slot = 0, period = 0;
  const unsigned long long g0 = gtimer();
  const long long t0 = clock64();
  RowRegs A, Bn;
  if (ngroups > 0) {
    GZ_FETCH(0);
    mbar_wait(bar_s, 0);
    load_row(A, GZ_RECP(0, 0), S, epoch_sh, dummy_bb);
#if GZ_RING_G == 3
    for (int g = 0; g < ngroups; g += 2) {      // two groups per trip: the A / Bn roles swap after an odd group
      GZ_FETCH(g + 1);                          // into the stage group g - 1 used
      GZ_ROUND(A, Bn, g, 0, g, 1);
      GZ_ROUND(Bn, A, g, 1, g, 2);
      if (g + 1 < ngroups) mbar_wait(bar_s + 8 * ((g + 1) & (GZ_RING_D - 1)), ((g + 1) / GZ_RING_D) & 1);
      GZ_ROUND(A, Bn, g, 2, g + 1, 0);          // past the last group this reads a stale stage: never used
      if (g + 1 < ngroups) {
        GZ_FETCH(g + 2);
        GZ_ROUND(Bn, A, g + 1, 0, g + 1, 1);
        GZ_ROUND(A, Bn, g + 1, 1, g + 1, 2);
        if (g + 2 < ngroups) mbar_wait(bar_s + 8 * ((g + 2) & (GZ_RING_D - 1)), ((g + 2) / GZ_RING_D) & 1);
        GZ_ROUND(Bn, A, g + 1, 2, g + 2, 0);
      }
    }
#else
    for (int g = 0; g < ngroups; g++) {
      GZ_FETCH(g + 1);                          // into the stage group g - 1 used
      GZ_ROUND(A, Bn, g, 0, g, 1);
#if GZ_RING_G == 4
      GZ_ROUND(Bn, A, g, 1, g, 2);
      GZ_ROUND(A, Bn, g, 2, g, 3);
#endif
      if (g + 1 < ngroups) mbar_wait(bar_s + 8 * ((g + 1) & (GZ_RING_D - 1)), ((g + 1) / GZ_RING_D) & 1);
      GZ_ROUND(Bn, A, g, GZ_RING_G - 1, g + 1, 0);   // past the last group this reads a stale stage: never used
    }
#endif
  }
#undef GZ_ROUND
#undef GZ_RECP
#undef GZ_FETCH
  const long long t1 = clock64();
  __syncwarp();
  // ---- statistics: reduce over the 2k lanes of the world (lambda-solve lanes carry the sums)
  for (int d = 1; d < LW; d <<= 1) {
#pragma unroll
    for (int t = 0; t < 3; t++) {
      st.acc[t] += __shfl_xor_sync(0xffffffffu, st.acc[t], d);
      st.err[t] += __shfl_xor_sync(0xffffffffu, st.err[t], d);
      st.cnt[t] += __shfl_xor_sync(0xffffffffu, st.cnt[t], d);
    }
  }
  if (live_ok) {
    // lambda back to the rows (k_lambda_save copies them to the joints), caccel to the bodies
    Row *rw = B.row + wi.row0;
    for (int i = lw; i < nslots; i += LW) {
      const RecInts *ri = reinterpret_cast<const RecInts *>(slab_rec(sl, i) + GZ_REC_INTS);
      if (((unsigned)ri->stage_epoch >> 8) == epoch && ri->fi >= GZ_FI_EMPTY + 1) {
        rw[ri->orig].lambda = lam[2 * i]; rw[ri->orig].lambda_erp = lam[2 * i + 1];
      }
    }
    for (int i = lw; i < wi.nb; i += LW) {
      BodyStep &bs = B.bstep[wi.body0 + i];
      for (int t = 0; t < 6; t++) { bs.caccel[t] = ca[6 * i + t]; bs.caccel_erp[t] = ca[6 * (nbp + 1 + i) + t]; }
    }
    if (lw == 0) {
      WorldStep &ws = B.wstep[w];
      const double *a = st.acc, *er = st.err;
      const double c[3] = {(double)st.cnt[0], (double)st.cnt[1], (double)st.cnt[2]};
      // quickstep_pgs_lcp.cpp:814-855
      const double tot = c[0] + c[1] + c[2];
      for (int q = 0; q < 3; q++) {
        ws.rms_dlambda[q] = sqrt(c[q] > 0 ? a[q] / c[q] : 0.0);
        ws.rms_residual[q] = sqrt(c[q] > 0 ? er[q] / c[q] : 0.0);
      }
      ws.rms_dlambda[3] = sqrt((a[0] + a[1] + a[2] > 0) ? (a[0] + a[1] + a[2]) / tot : 0.0);
      ws.rms_residual[3] = sqrt((er[0] + er[1] + er[2] > 0) ? (er[0] + er[1] + er[2]) / tot : 0.0);
      ws.num_contacts_stat = st.cnt[1];
      // instrumentation: rounds the warp ran, this world's period * 256 + k, SM cycles of the sweeps
      ws.pad[0] = rounds; ws.pad[1] = R * 256 + k; ws.n_fric = (int)(t1 - t0);
    }
  }
  if (live && lw == 0) { B.prof[16 + 2 * w] = g0; B.prof[16 + 2 * w + 1] = (gtimer() & ~511ull) | smid(); }   // low 9 bits: the SM it ran on
  if (prof && lane == 0) {
    for (int q = 0; q < 5; q++) atomicAdd(&B.prof[q], (unsigned long long)pc[q]);
    atomicAdd(&B.prof[5], (unsigned long long)rounds);
  }
}

}  // namespace gz
