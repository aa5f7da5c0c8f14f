// k_pgs_chain.cuh — the SOR-PGS sweeps (ComputeRows, quickstep_pgs_lcp.cpp:52-954) of CHAIN-MODE
// worlds: worlds in which one body carries so many rows (a robot lying on its core: 60-90 contact
// rows on one body) that the sweep is one long dependent chain and the round-per-row schedule of
// k_pgs (k_pgs_sweep.cuh) would run for R = 60-90 rounds per sweep.
//
// k_schedule (k_pgs.cuh) turns every maximal run of consecutive rows of order[] that touch the same
// bodies into ONE schedule item.  Here a warp owns one world; lane l is (schedule lane j = l / 2,
// solve S = l & 1).  The world's row records (the run array, in order[] sequence) are copied into
// shared memory once (bulk async copy) and stay there for the 60 sweeps, with the lambdas, the
// accumulators and the 16-byte heads of the items [R][16].  In a round a lane reads the head at
// (slot, j) and, if the item is due, runs its rows back to back: the accumulators of the item's
// bodies stay in registers from the first row to the last, a row costs one shared-memory record
// read (issued one row ahead) and the dependent arithmetic.  Rows execute in exactly the reference's
// order per body, so the result is bit-identical to the sequential sweep.
//
// Shared memory per CTA: ca[solve][nbp+1][6] | run statistics [32][10] | mbarrier | lambda[m][2] |
// records [m + 1][34] | heads [R][16] x 16 bytes.  k_schedule only puts a world in chain mode if
// this fits one of the two launch sizes (B.pgs_chain_doubles).
#pragma once

namespace gz {

// ---- the run loop ------------------------------------------------------------------------------------
// Everything it touches (accumulators, lambdas, records) is in shared memory, addressed from the
// dynamic shared array so that the loads are LDS.  "Some due item of this round has a second body"
// and "this sweep feeds the statistics" are warp-uniform and compile-time here; the row class
// (bilateral fi = -1 / contact normal -2 / friction >= 0) is per-lane data handled with selects, so
// the lanes of a round run items of different classes in lockstep.
struct RunRec { double J[12], iM[12], rhs, adcfm, lo, hi, ad2; int fi; };

template <bool TWO, bool STATS>
__device__ __forceinline__ void run_load(RunRec &X, const double *rec, const int S) {
  const double2 *r2 = reinterpret_cast<const double2 *>(rec);
#pragma unroll
  for (int k = 0; k < (TWO ? 6 : 3); k++) {
    const double2 a = r2[k], b = r2[6 + k];
    X.J[2 * k] = a.x; X.J[2 * k + 1] = a.y;
    X.iM[2 * k] = b.x; X.iM[2 * k + 1] = b.y;
  }
  const double2 d = r2[13];
  X.rhs = rec[24 + S]; X.adcfm = d.x; X.lo = d.y;
  if (STATS) { const double2 e = r2[14]; X.hi = e.x; X.ad2 = e.y; }
  else X.hi = rec[28];
  X.fi = *reinterpret_cast<const int *>(rec + GZ_REC_INTS);
}

// One row update of one solve (quickstep_pgs_lcp.cpp:488-808, pyramid friction): delta, limits from
// lo/hi or |hi * lambda[findex]|, clamp, SMOOTH_LAMBDA on the lambda solve only (delta is not
// recomputed after it), caccel update in registers.
template <bool TWO, bool STATS, bool BOX>
__device__ __forceinline__ void run_row(const RunRec &X, double *a1, double *a2, double *lp, const double *lamS,
                                        const int S, const double smooth, const double smooth1, const bool stat01,
                                        const bool stat2, PgsStat &st) {
  const double old_lambda = *lp;
  const bool fric = X.fi >= 0;
  const double nlam = BOX ? 1.0 : lamS[2 * (fric ? X.fi : 0)];   // box friction: limits +-hi
  // pyramid friction: |mu * lambda_normal| and its negative, by the sign bit (integer pipe, not FP64)
  const double hl = X.hi * nlam;
  const double hf = __hiloint2double(__double2hiint(hl) & 0x7fffffff, __double2loint(hl));
  const double nhf = __hiloint2double(__double2hiint(hl) | 0x80000000, __double2loint(hl));
  const double hi_act = fric ? hf : X.hi, lo_act = fric ? nhf : X.lo;
  double d1 = a1[0] * X.J[0], d2 = 0.0;
  if (TWO) d2 = a2[0] * X.J[6];
#pragma unroll
  for (int t = 1; t < 6; t++) { d1 += a1[t] * X.J[t]; if (TWO) d2 += a2[t] * X.J[6 + t]; }
  double delta = X.rhs - old_lambda * X.adcfm;
  delta -= d1;
  if (TWO) delta -= d2;   // a row without a second body: J2 = 0 and the all-zero dummy accumulator, d2 = +0.0, delta unchanged
  double nl = old_lambda + delta;
  const double dlo = lo_act - old_lambda, dhi = hi_act - old_lambda;
  const bool below = nl < lo_act, above = !below && nl > hi_act;
  delta = below ? dlo : (above ? dhi : delta);
  nl = below ? lo_act : (above ? hi_act : nl);
  const double sm = smooth1 * nl + smooth * old_lambda;   // smooth1 = 1.0 - smooth
  *lp = (S == 0 && X.fi != -1) ? sm : nl;
#pragma unroll
  for (int t = 0; t < 6; t++) a1[t] += delta * X.iM[t];
  if (TWO) {
#pragma unroll
    for (int t = 0; t < 6; t++) a2[t] += delta * X.iM[6 + t];
  }
  if (STATS) pgs_stat(st, X.fi, S == 0 && stat01, S == 0 && stat2, delta, X.ad2);
}

// offsets are in doubles from the start of the dynamic shared array; flags: 1 = sweep iters - 1
// (bilateral and normal statistics), 2 = sweep total - 1 (friction statistics); BOX = box friction model
template <bool TWO, bool STATS, bool BOX>
__device__ __noinline__ void pgs_run_sm(const int fi, const int orig, const int b1, const int b2, const int S,
                                        const int cas_off, const int lam_off, const int run_off, const int nbp,
                                        const double smooth, const int flags, const int stat_off) {
  extern __shared__ __align__(16) double shd[];
  const int len = GZ_FI_RUN - fi, roff = orig & 0x3FFFFFFF;
  const double *rp = shd + run_off + roff * GZ_REC;
  double *lp = shd + lam_off + 2 * roff + S;
  const double *lamS = shd + lam_off + S;
  double *c1 = shd + cas_off + 6 * b1, *c2 = shd + cas_off + 6 * b2;
  const bool stat01 = flags & 1, stat2 = flags & 2;
  const double smooth1 = 1.0 - smooth;
  PgsStat st;
  for (int t = 0; t < 3; t++) { st.acc[t] = 0; st.err[t] = 0; st.cnt[t] = 0; }
  double a1[6], a2[6];
  {
    const double2 *p1 = reinterpret_cast<const double2 *>(c1), *p2 = reinterpret_cast<const double2 *>(c2);
#pragma unroll
    for (int t = 0; t < 3; t++) {
      const double2 x = p1[t];
      a1[2 * t] = x.x; a1[2 * t + 1] = x.y;
      if (TWO) { const double2 y = p2[t]; a2[2 * t] = y.x; a2[2 * t + 1] = y.y; }
    }
  }
  RunRec X, Y;
  run_load<TWO, STATS>(X, rp, S);
  int i = 0;
  for (; i + 2 <= len; i += 2) {
    run_load<TWO, STATS>(Y, rp + GZ_REC, S);
    run_row<TWO, STATS, BOX>(X, a1, a2, lp, lamS, S, smooth, smooth1, stat01, stat2, st);
    run_load<TWO, STATS>(X, rp + 2 * GZ_REC, S);   // at the end one record past the item: still inside the array (m + 1 records)
    run_row<TWO, STATS, BOX>(Y, a1, a2, lp + 2, lamS, S, smooth, smooth1, stat01, stat2, st);
    rp += 2 * GZ_REC; lp += 4;
  }
  if (i < len) run_row<TWO, STATS, BOX>(X, a1, a2, lp, lamS, S, smooth, smooth1, stat01, stat2, st);
  double2 *p1 = reinterpret_cast<double2 *>(c1);
#pragma unroll
  for (int t = 0; t < 3; t++) p1[t] = make_double2(a1[2 * t], a1[2 * t + 1]);
  if (TWO && b2 < nbp) {
    double2 *p2 = reinterpret_cast<double2 *>(c2);
#pragma unroll
    for (int t = 0; t < 3; t++) p2[t] = make_double2(a2[2 * t], a2[2 * t + 1]);
  }
  if (STATS) {
    double *stat_sm = shd + stat_off;
#pragma unroll
    for (int t = 0; t < 3; t++) { stat_sm[t] += st.acc[t]; stat_sm[3 + t] += st.err[t]; stat_sm[6 + t] += (double)st.cnt[t]; }
  }
}

// cls: shared-memory size class of this launch (0 small, 1 large); worlds of the other class exit
template <bool PROF, bool BOX>
__global__ void __launch_bounds__(32) k_pgs_chain(DevBatch B, PgsCfg cfg, int cls) {
  extern __shared__ __align__(16) double shd[];
  const int lane = threadIdx.x;
  const int nbp = cfg.nbp;
  // ---- which bucket, which warp of it: chain-mode buckets come first in k_slab's order
  const int cta = blockIdx.x;
  const int *swarp0 = bk_row(B, GZ_BK_SWARP0);
  if (cta >= swarp0[GZ_NBUCKET]) return;
  int lo_i = 0, hi_i = GZ_NBUCKET - 1;
  while (lo_i < hi_i) {   // largest i with swarp0[i] <= cta
    const int mid = (lo_i + hi_i + 1) >> 1;
    if (swarp0[mid] <= cta) lo_i = mid; else hi_i = mid - 1;
  }
  const int bucket = bk_row(B, GZ_BK_SBUCKET)[lo_i];
  if (bucket / GZ_NRUNG != 5 + cls) return;
  const int wb = cta - swarp0[lo_i];
  const RecInts *heads_g = reinterpret_cast<const RecInts *>(
      B.prec + ((size_t)bk_row(B, GZ_BK_REC0)[bucket] + (size_t)wb * rung_value(bucket % GZ_NRUNG)) * GZ_REC);
  const int jl = lane >> 1, S = lane & 1;
  const int w = B.warp_world[cta * 16];
  if (w < 0) return;
  const WorldInfo wi = B.world[w];
  const int m = B.wstep[w].m;
  const WSched sc = reinterpret_cast<const WSched *>(B.wsched)[w];
  const int nstage = sc.nstage, R = sc.R;   // (the bucket's rung is that of m, not R)
  const int nheads = R * GZ_CHAIN_K;
  const double *runrec_g = B.runrec + (size_t)wi.row0 * GZ_REC;
  // ---- carve shared memory (k_schedule checked that it fits)
  const int ca_doubles = 12 * (nbp + 1);
  double *ca = shd;
  const int stat_off = ca_doubles;
  const unsigned bar_s = (unsigned)__cvta_generic_to_shared(shd + stat_off + 320);
  const int lam_off = ca_doubles + GZ_CHAIN_FIXED_DOUBLES;
  const int run_off = lam_off + 2 * m;
  const int head_off = run_off + (m + 1) * GZ_REC;
  double *lam = shd + lam_off;
  int4 *heads = reinterpret_cast<int4 *>(shd + head_off);
  if (head_off + 2 * nheads > cfg.smem_doubles) {   // cannot happen: k_schedule's test
    if (lane == 0) B.wstep[w].overflow |= 8;
    return;
  }
  if (lane == 0) mbar_init(bar_s, 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncwarp();
  if (lane == 0) {   // the copy runs while lambda0, the heads and the accumulators are set up
    const unsigned bytes = (unsigned)m * GZ_REC * 8u, chunk = 32768u;
    const unsigned dst = (unsigned)__cvta_generic_to_shared(shd + run_off);
    mbar_expect_tx(bar_s, bytes);
    for (unsigned o = 0; o < bytes; o += chunk)
      bulk_g2s(dst + o, reinterpret_cast<const char *>(runrec_g) + o, min(chunk, bytes - o), bar_s);
  }
  const unsigned epoch = gz_epoch(B.step_index);
  for (int i = lane; i < ca_doubles + 320; i += 32) shd[i] = 0.0;
  for (int i = lane; i < nheads; i += 32) {   // a head whose epoch is not this step's is an empty slot
    int4 h = *reinterpret_cast<const int4 *>(heads_g + i);
    if (((unsigned)h.z >> 8) != epoch) h.x = GZ_FI_EMPTY;
    heads[i] = h;
  }
  const Row *rows = B.row + wi.row0;
  const int *order = B.order + wi.row0;
  for (int i = lane; i < m; i += 32) {
    const Row &r = rows[order[i]];
    lam[2 * i] = r.lambda; lam[2 * i + 1] = r.lambda_erp;
  }
  __syncwarp();
  // initial caccel = iMJ * lambda0 over the warm-started (persistent-joint) rows, in row index
  // order per body (multiply_invM_JT, quickstep_util.cpp:101-119)
  if (B.p.warm_start > 0) {
    for (int b = lane; b < wi.nb; b += 32) {
      const Body &bd = B.body[wi.body0 + b];
      if (!B.bstep[wi.body0 + b].active) continue;
      double s0[6] = {0, 0, 0, 0, 0, 0}, s1[6] = {0, 0, 0, 0, 0, 0};
      for (int q = bd.joint_begin; q < bd.joint_end; q++) {
        const Joint &j = B.joint[B.body_joint[q]];
        for (int rr = 0; rr < j.m; rr++) {
          const Row &r = rows[j.row0 + rr];
          const double *rec = runrec_g + (size_t)r.t2a * GZ_REC;
          const double *im = (r.b1 == b) ? rec + 12 : rec + 18;
          for (int t = 0; t < 6; t++) { s0[t] += im[t] * r.lambda; s1[t] += im[t] * r.lambda_erp; }
        }
      }
      for (int t = 0; t < 6; t++) { ca[6 * b + t] = s0[t]; ca[6 * (nbp + 1 + b) + t] = s1[t]; }
    }
  }
  __syncwarp();
  mbar_wait(bar_s, 0);
  // ---- the sweeps
  const int iters = B.p.iters, total = B.p.iters + B.p.friction_iters;
  const double smooth = B.p.smooth_contacts;
  const int rounds = (total > 0 && nstage > 0) ? (total + nstage - 1) * R : 0;
  const int cas_off = 6 * (nbp + 1) * S;
  int slot = 0, period = 0;
  long long pc[5] = {0, 0, 0, 0, 0};
  const unsigned long long g0 = gtimer();
  const long long t0 = clock64();
  for (int T = 0; T < rounds; T++) {
    long long c0 = 0;
    if (PROF) c0 = clock64();
    const int4 h = heads[slot * GZ_CHAIN_K + jl];   // RecInts: fi | b1, b2 | epoch, stage | orig
    const int sweep = period - (h.z & 255);
    const bool st_round = (period >= iters - 1 && period < iters - 1 + nstage) ||
                          (period >= total - 1 && period < total - 1 + nstage);
    const bool act = h.x <= GZ_FI_RUN - 1 && (unsigned)sweep < (unsigned)(((h.w >> 30) & 1) ? total : iters);
    const int b1 = h.y & 0xFFFF, b2 = (unsigned)h.y >> 16;
    // one code path for all lanes of the round: the two-body loop if any due item has a second body
    // (running the single-body items first and the others after them was measured: slower)
    const unsigned due2 = __ballot_sync(0xffffffffu, act && b2 < nbp);
    if (act) {
      if (st_round) {
        const int flags = (sweep == iters - 1 ? 1 : 0) | (sweep == total - 1 ? 2 : 0);
        if (due2) pgs_run_sm<true, true, BOX>(h.x, h.w, b1, b2, S, cas_off, lam_off, run_off, nbp, smooth, flags, stat_off + 10 * lane);
        else pgs_run_sm<false, true, BOX>(h.x, h.w, b1, b2, S, cas_off, lam_off, run_off, nbp, smooth, flags, stat_off + 10 * lane);
      } else {
        if (due2) pgs_run_sm<true, false, BOX>(h.x, h.w, b1, b2, S, cas_off, lam_off, run_off, nbp, smooth, 0, 0);
        else pgs_run_sm<false, false, BOX>(h.x, h.w, b1, b2, S, cas_off, lam_off, run_off, nbp, smooth, 0, 0);
      }
    }
    __syncwarp();
    if (++slot == R) { slot = 0; period++; }
    if (PROF) {
      const int mx = __reduce_max_sync(0xffffffffu, act ? GZ_FI_RUN - h.x : 0);
      const long long c1 = clock64();
      if (mx > 0) { pc[0] += c1 - c0; pc[1]++; pc[2] += mx; } else { pc[3] += c1 - c0; pc[4]++; }
    }
  }
  const long long t1 = clock64();
  __syncwarp();
  // ---- statistics: reduce over the lanes (lambda-solve lanes carry the sums)
  PgsStat st;
  {
    const double *stat_sm = shd + stat_off + 10 * lane;
#pragma unroll
    for (int t = 0; t < 3; t++) { st.acc[t] = stat_sm[t]; st.err[t] = stat_sm[3 + t]; st.cnt[t] = (int)stat_sm[6 + t]; }
  }
  for (int d = 1; d < 32; d <<= 1) {
#pragma unroll
    for (int t = 0; t < 3; t++) {
      st.acc[t] += __shfl_xor_sync(0xffffffffu, st.acc[t], d);
      st.err[t] += __shfl_xor_sync(0xffffffffu, st.err[t], d);
      st.cnt[t] += __shfl_xor_sync(0xffffffffu, st.cnt[t], d);
    }
  }
  // lambda back to the rows (k_lambda_save copies them to the joints), caccel to the bodies
  Row *rw = B.row + wi.row0;
  for (int i = lane; i < m; i += 32) {
    Row &r = rw[order[i]];
    r.lambda = lam[2 * i]; r.lambda_erp = lam[2 * i + 1];
  }
  for (int i = lane; i < wi.nb; i += 32) {
    BodyStep &bs = B.bstep[wi.body0 + i];
    for (int t = 0; t < 6; t++) { bs.caccel[t] = ca[6 * i + t]; bs.caccel_erp[t] = ca[6 * (nbp + 1 + i) + t]; }
  }
  if (lane == 0) {
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
    ws.pad[0] = rounds; ws.pad[1] = R * 256 + GZ_CHAIN_K; ws.n_fric = (int)(t1 - t0);
  }
  if (lane == 0) { B.prof[16 + 2 * w] = g0; B.prof[16 + 2 * w + 1] = (gtimer() & ~511ull) | smid(); }
  if (PROF && lane == 0) {
    for (int q = 0; q < 5; q++) atomicAdd(&B.prof[8 + q], (unsigned long long)pc[q]);
    atomicAdd(&B.prof[13], (unsigned long long)(t1 - t0));
    atomicAdd(&B.prof[14], 1ull);
  }
}

}  // namespace gz
