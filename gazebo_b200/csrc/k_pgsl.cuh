// k_pgsl.cuh — the SOR-PGS sweeps (ComputeRows, quickstep_pgs_lcp.cpp:52-954) of a LARGE world: more bodies than one
// CTA can own (C4: a 20 000-body rubble pile in one island).  Same data structure and the same arithmetic as k_pgsb
// (k_pgsb.cuh): per-body entry lists in order[] sequence, lane (b, S) owns body b and keeps its accumulator in
// registers, a two-body row is run by its two owner lanes in the same round.  What changes is the scale: the lanes
// of the world are spread over a cooperative grid (one launch per world, all CTAs resident), the partial dot
// products are exchanged through a double-buffered array in global memory (L2) and one GRID barrier per round
// orders them; entries are read from the pool (L2: the pile's entries, ~100 MB at C4, stay L2-resident over the
// 60 sweeps).  Per body the rows still execute in exactly the reference's order: bit-identical.
#pragma once
#include <cooperative_groups.h>

namespace gz {

template <bool BOX>
__global__ void __launch_bounds__(256) k_pgsl(DevBatch B, const int w, double *xch, double *partial) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  const int gl = blockIdx.x * blockDim.x + threadIdx.x, L = gridDim.x * blockDim.x;
  const int bl = gl >> 1, S = gl & 1;
  const WorldInfo wi = B.world[w];
  const int4 sc = B.wsb[w];
  const int R = sc.x, nstage = sc.y, nent = sc.w;
  int cnt = 0, e0 = sc.z, fric0 = 0;
  if (bl < wi.nb && nent > 0) { const int4 bs = B.bsched[wi.body0 + bl]; e0 = bs.x; cnt = bs.y; fric0 = bs.z; }
  double *const lp = B.ent + (size_t)e0 * GZ_ENT;
  double *const dl = B.ent_dl + 2 * (size_t)e0 + S;
  // initial caccel = iMJ * lambda0 over the warm-started (persistent-joint) rows (multiply_invM_JT, quickstep_util.cpp:101-119)
  double a[6] = {0, 0, 0, 0, 0, 0};
  if (B.p.warm_start > 0) {
    for (int q = 0; q < cnt; q++) {
      const double *p = lp + (size_t)q * GZ_ENT;
      if (((__double2loint(p[19]) >> 27) & 3) != 0) break;
      const double l0 = p[13 + 2 * S];
#pragma unroll
      for (int t = 0; t < 6; t++) a[t] += p[6 + t] * l0;
    }
  }
  const int iters = B.p.iters, total = B.p.iters + B.p.friction_iters;
  const double smooth = B.p.smooth_contacts, smooth1 = 1.0 - smooth;
  int i = (iters > 0) ? 0 : fric0, sweep = 0;
  int T_due = 0x7fffffff;
  if (i < cnt && total > 0) T_due = __double2loint(lp[(size_t)i * GZ_ENT + 19]) & GZ_ENT_OFFMASK;
  const int rounds = (total > 0 && nstage > 0 && nent > 0) ? (total + nstage - 1) * R : 0;
  for (int T = 0; T < rounds; T++) {
    const bool due = (T_due == T);
    double *p = lp + (size_t)i * GZ_ENT;
    double J[6] = {0, 0, 0, 0, 0, 0}, d = 0.0;
    if (due) {
      const double2 *p2 = reinterpret_cast<const double2 *>(p);
#pragma unroll
      for (int k = 0; k < 3; k++) { const double2 x = p2[k]; J[2 * k] = x.x; J[2 * k + 1] = x.y; }
      // this body's half of the dot product, term by term in the reference's left-to-right order
      d = a[0] * J[0];
#pragma unroll
      for (int t = 1; t < 6; t++) d += a[t] * J[t];
    }
    double *x = xch + (size_t)(T & 1) * L;
    x[gl] = d;
    grid.sync();
    if (due) {
      const double2 *p2 = reinterpret_cast<const double2 *>(p);
      double iM[6];
#pragma unroll
      for (int k = 0; k < 3; k++) { const double2 y = p2[3 + k]; iM[2 * k] = y.x; iM[2 * k + 1] = y.y; }
      const double2 rl = p2[6 + S], cl = p2[8], hm = p2[9];
      const double rhs = rl.x, old_lambda = rl.y, adcfm = cl.x, lo = cl.y, hi = hm.x;
      const int mx = __double2loint(hm.y), my = __double2hiint(hm.y);
      const int cls = (mx >> 27) & 3;
      const bool fric = cls == 2, isb2 = (mx >> 29) & 1;
      const int part = my & 0xFFFF;
      const bool two = part != GZ_ENT_NOPART;
      // pyramid friction: limits +-|mu * lambda_normal| (quickstep_pgs_lcp.cpp:572-582); box: +-mu (:588-594)
      double nlam = 1.0;
      if (!BOX && fric) nlam = (p - (size_t)((unsigned)my >> 16) * GZ_ENT)[13 + 2 * S];
      const double pd = two ? x[2 * part + S] : 0.0;
      double delta = rhs - old_lambda * adcfm;
      const double d1 = isb2 ? pd : d, d2 = isb2 ? d : pd;
      delta -= d1;
      const double delta2 = delta - d2;
      delta = two ? delta2 : delta;
      const double hl = hi * nlam;
      const double hf = fabs(hl);
      const double hi_act = fric ? hf : hi, lo_act = fric ? -hf : lo;
      double nl = old_lambda + delta;
      const double dlo = lo_act - old_lambda, dhi = hi_act - old_lambda;
      const bool below = nl < lo_act, above = !below && nl > hi_act;
      delta = below ? dlo : (above ? dhi : delta);
      nl = below ? lo_act : (above ? hi_act : nl);
      // SMOOTH_LAMBDA on the lambda solve of non-bilateral rows; delta is not recomputed after it (:687-691 vs :706)
      const double sm = smooth1 * nl + smooth * old_lambda;
      p[13 + 2 * S] = (S == 0 && cls != 0) ? sm : nl;
      dl[2 * (size_t)i] = delta;
#pragma unroll
      for (int t = 0; t < 6; t++) a[t] += delta * iM[t];
      // move on
      int inx = i + 1, nsw = sweep;
      if (inx == cnt) { nsw = sweep + 1; inx = nsw < iters ? 0 : fric0; }
      if (nsw < total && inx < cnt) {
        i = inx; sweep = nsw;
        T_due = sweep * R + (__double2loint(lp[(size_t)i * GZ_ENT + 19]) & GZ_ENT_OFFMASK);
      } else {
        T_due = 0x7fffffff;
      }
    }
  }
  // ---- results: caccel to the bodies; then over all entries of the world with all lanes: lambda back to the rows
  // and the statistics of the last update of every row (quickstep_pgs_lcp.cpp:700-760, 814-855)
  if (bl < wi.nb) {
    BodyStep &bs = B.bstep[wi.body0 + bl];
    double *o = S ? bs.caccel_erp : bs.caccel;
#pragma unroll
    for (int t = 0; t < 6; t++) o[t] = a[t];
  }
  grid.sync();
  double st[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};   // acc[3], err[3], cnt[3]
  {
    const double *wp = B.ent + (size_t)sc.z * GZ_ENT;
    const double *wdl = B.ent_dl + 2 * (size_t)sc.z;
    Row *rw = B.row + wi.row0;
    for (int q = gl; q < nent; q += L) {
      const int r = B.ent_row[sc.z + q];
      if (r < 0) continue;   // the row's body-1 entry carries it
      const double *p = wp + (size_t)q * GZ_ENT;
      rw[r].lambda = p[13]; rw[r].lambda_erp = p[15];
      const int cls = (__double2loint(p[19]) >> 27) & 3;
      if ((cls == 2 ? total : iters) <= 0) continue;
      const double de = wdl[2 * (size_t)q], dd = de * de, e2 = dd * B.ent_ad2[sc.z + q];
      st[cls] += dd; st[3 + cls] += e2; st[6 + cls] += 1.0;
    }
  }
  // deterministic reduction: lanes of a warp, warps of a CTA in order, CTAs in order
  __shared__ double red[8][9];
  for (int dd = 1; dd < 32; dd <<= 1)
#pragma unroll
    for (int q = 0; q < 9; q++) st[q] += __shfl_xor_sync(0xffffffffu, st[q], dd);
  if ((threadIdx.x & 31) == 0)
    for (int q = 0; q < 9; q++) red[threadIdx.x >> 5][q] = st[q];
  __syncthreads();
  if (threadIdx.x < 9) {
    double s = 0;
    for (int k = 0; k < (int)(blockDim.x >> 5); k++) s += red[k][threadIdx.x];
    partial[9 * (size_t)blockIdx.x + threadIdx.x] = s;
  }
  grid.sync();
  if (gl == 0) {
    double t[9];
    for (int q = 0; q < 9; q++) { double s = 0; for (int k = 0; k < (int)gridDim.x; k++) s += partial[9 * (size_t)k + q]; t[q] = s; }
    WorldStep &ws = B.wstep[w];
    const double tot = t[6] + t[7] + t[8];
    for (int q = 0; q < 3; q++) {
      ws.rms_dlambda[q] = sqrt(t[6 + q] > 0 ? t[q] / t[6 + q] : 0.0);
      ws.rms_residual[q] = sqrt(t[6 + q] > 0 ? t[3 + q] / t[6 + q] : 0.0);
    }
    const double sa = t[0] + t[1] + t[2], se = t[3] + t[4] + t[5];
    ws.rms_dlambda[3] = sqrt(sa > 0 ? sa / tot : 0.0);
    ws.rms_residual[3] = sqrt(se > 0 ? se / tot : 0.0);
    ws.num_contacts_stat = (int)t[7];
    ws.pad[0] = rounds; ws.pad[1] = R * 256;
  }
}

}  // namespace gz
