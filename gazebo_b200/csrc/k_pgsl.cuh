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

// Grid barrier of the rounds (the grid is launched cooperatively, so all its CTAs are resident): one arrival counter in
// L2 that only grows; a CTA's thread 0 publishes the CTA's writes (fence), arrives, and spins on an acquire load
// until everyone of this round has arrived.  About 1 us against ~2.5 us for cooperative_groups' grid.sync().
__device__ __forceinline__ void pgsl_barrier(unsigned *ctr, unsigned &target) {
  __syncthreads();
  target += gridDim.x;
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(ctr, 1u);
    unsigned v;
    do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory"); } while ((int)(v - target) < 0);
  }
  __syncthreads();
}

template <bool BOX>
__global__ void __launch_bounds__(256) k_pgsl(DevBatch B, const int w, double *xch, double *partial, unsigned *bar) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  unsigned bar_target = *bar;   // (the counter is left where the last launch ended: read before anyone arrives)
  grid.sync();
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
    // where this lane goes next (known before the row is computed)
    int inx = i + 1, nsw = sweep;
    if (inx == cnt) { nsw = sweep + 1; inx = nsw < iters ? 0 : fric0; }
    const bool more = due && nsw < total && inx < cnt;
    double J[6] = {0, 0, 0, 0, 0, 0}, iM[6] = {0, 0, 0, 0, 0, 0}, d = 0.0;
    double rhs = 0, old_lambda = 0, adcfm = 0, lo = 0, hi = 0, nlam = 1.0;
    int mx = 0, my = GZ_ENT_NOPART, off_next = 0;
    if (due) {
      // ONE trip to L2 for everything the row needs (and the round of the lane's next row): all loads are issued here
      const double2 *p2 = reinterpret_cast<const double2 *>(p);
      double2 v[10];
#pragma unroll
      for (int k = 0; k < 10; k++) v[k] = (k == 6 || k == 7) ? (k == 6 + S ? p2[k] : make_double2(0.0, 0.0)) : p2[k];
      if (more) off_next = __double2loint(lp[(size_t)inx * GZ_ENT + 19]) & GZ_ENT_OFFMASK;
#pragma unroll
      for (int k = 0; k < 3; k++) { J[2 * k] = v[k].x; J[2 * k + 1] = v[k].y; iM[2 * k] = v[3 + k].x; iM[2 * k + 1] = v[3 + k].y; }
      rhs = v[6 + S].x; old_lambda = v[6 + S].y; adcfm = v[8].x; lo = v[8].y; hi = v[9].x;
      mx = __double2loint(v[9].y); my = __double2hiint(v[9].y);
      // pyramid friction: limits +-|mu * lambda_normal| (quickstep_pgs_lcp.cpp:572-582); box: +-mu (:588-594)
      if (((mx >> 27) & 3) == 2 && (!BOX || ((mx >> 30) & 1))) nlam = (p - (size_t)((unsigned)my >> 16) * GZ_ENT)[13 + 2 * S];
      // this body's half of the dot product, term by term in the reference's left-to-right order
      d = a[0] * J[0];
#pragma unroll
      for (int t = 1; t < 6; t++) d += a[t] * J[t];
    }
    double *x = xch + (size_t)(T & 1) * L;
    x[gl] = d;
    pgsl_barrier(bar, bar_target);
    if (due) {
      const int cls = (mx >> 27) & 3;
      const bool fric = cls == 2, isb2 = (mx >> 29) & 1;
      const int part = my & 0xFFFF;
      const bool two = part != GZ_ENT_NOPART;
      const double pd = two ? __ldcg(&x[2 * part + S]) : 0.0;   // (L2: another SM wrote it)
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
      if (more) { i = inx; sweep = nsw; T_due = sweep * R + off_next; }
      else T_due = 0x7fffffff;
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

namespace gz {

// k_bsched for a LARGE world (k_pgsb.cuh describes what it computes: every row's round offset and its place in the
// entry lists of its bodies, the world's period and its slice of the entry pool).  The small-world kernel does it in
// one serial pass; at C4 (80 000 rows) that pass was the largest kernel of the step (27 ms).  Here, cooperative grid:
//   1  contacts per body (atomics), joint rows per body
//   2  exclusive scans over the bodies (one CTA): list bases, contact-list bases
//   3  contact lists per body (unordered fill), 4  sorted per body (short lists), from which every row's list index
//      follows in closed form: [joint rows in index order | normals in DEscending contact order | friction rows in
//      ascending contact order] is the order[] sequence restricted to the body (quickstep_pgs_lcp.cpp:1115-1150)
//   5  round offsets off(r) = 1 + max over its bodies of off(previous row in the body's list), by chaotic relaxation
//      over all rows until nothing changes (monotone, so in-place updates converge to the same least fixpoint the
//      serial pass computes); one grid barrier per sweep over the rows
//   6  period R, stages, pool allocation, k_border's sort key
// tab = B.bs_glob + w * 5 * max_nb: [cntc | njr | base | cbase | fill]; clist after all worlds' tables; lp = the two
// spare arrays of the world's row scratch (entry -> order position).
__global__ void __launch_bounds__(256) k_bsched_large(DevBatch B, const int w, unsigned *bar, const int relax) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  unsigned bar_target = *bar;
  grid.sync();
  const int gt = blockIdx.x * blockDim.x + threadIdx.x, GT = gridDim.x * blockDim.x;
  const WorldInfo wi = B.world[w];
  WorldStep &ws = B.wstep[w];
  const int nb = wi.nb, m = ws.m, nc = ws.n_contacts, rpc = wi.rpc, mj = m - rpc * nc, pfric = mj + nc;
  int *tab = B.bs_glob + (size_t)w * 5 * B.max_nb;
  int *cntt = rpc == 4 ? B.bs_cntt + wi.body0 : nullptr;   // torsional rows per body (worlds of four-row contacts)
  int *cntc = tab, *njr = tab + B.max_nb, *base = tab + 2 * B.max_nb, *cbase = tab + 3 * B.max_nb, *fill = tab + 4 * B.max_nb;
  const WScratchB S = wscratchb(B, w);
  int *lp = S.idx2 + B.max_row_cap;   // [2 * max_row_cap]
  int *clist = B.bs_clist + 2 * (size_t)wi.contact0;
  int *flag = reinterpret_cast<int *>(bar) + 1, *acc = reinterpret_cast<int *>(bar) + 4;   // three rotating sweep flags; period and last offset
  __shared__ int s_scan[256], s_carry[2], s_flag;
  const int par = (int)(B.step_index & 1);
  if (w == 0 && gt == 0) B.ent_cursor[par ^ 1] = 0ull;
  if (w == 0) for (int q = gt; q < GZ_PGSB_NKEY; q += GT) B.rk_count[(par ^ 1) * GZ_PGSB_NKEY + q] = 0;
  // ---- 1
  for (int b = gt; b < nb; b += GT) {
    cntc[b] = 0; fill[b] = 0;
    if (cntt) cntt[b] = 0;
    const Body &bd = B.body[wi.body0 + b];
    int r = 0;
    if (m > 0) for (int q = bd.joint_begin; q < bd.joint_end; q++) r += B.joint[B.body_joint[q]].m;
    njr[b] = r;
  }
  if (gt == 0) { flag[0] = flag[1] = flag[2] = 0; acc[0] = 1; acc[1] = -1; }
  pgsl_barrier(bar, bar_target);
  for (int k = gt; k < wi.nj; k += GT) {
    const Joint &j = B.joint[wi.joint0 + k];
    const int b1 = j.b1 - wi.body0, b2 = j.b2 >= 0 ? j.b2 - wi.body0 : -1;
    if (m > 0) for (int r = 0; r < j.m; r++) { S.rb1[j.row0 + r] = b1; S.rb2[j.row0 + r] = b2; }
  }
  for (int c = gt; c < nc && m > 0; c += GT) {
    const Contact &C = B.contact[wi.contact0 + c];
    const bool live = B.bstep[C.b1].active != 0;   // a contact of an island that is not stepped: rows without entries
    const int b1 = live ? C.b1 - wi.body0 : -2, b2 = C.b2 >= 0 ? C.b2 - wi.body0 : -1;
    const int pn = mj + (nc - 1 - c), pf = pfric + (rpc - 1) * c;
    S.rb1[pn] = b1; S.rb2[pn] = b2; S.rb1[pf] = b1; S.rb2[pf] = b2; S.rb1[pf + 1] = b1; S.rb2[pf + 1] = b2;
    if (rpc == 4) { S.rb1[pf + 2] = C.tors ? b1 : -2; S.rb2[pf + 2] = b2; }   // (no torsional friction: a placeholder, never swept)
    if (live) { atomicAdd(&cntc[b1], 1); if (b2 >= 0) atomicAdd(&cntc[b2], 1); }
    if (live && cntt && C.tors) { atomicAdd(&cntt[b1], 1); if (b2 >= 0) atomicAdd(&cntt[b2], 1); }
  }
  pgsl_barrier(bar, bar_target);
  // ---- 2 (CTA 0): base = exclusive scan of njr + 3 cntc (+ cntt), cbase = exclusive scan of cntc; both scans in one
  // pass, eight bodies per thread and trip (one body per thread and scan: 632 block barriers at C4, 50 us)
  if (blockIdx.x == 0) {
    const int t = threadIdx.x, lane = t & 31, wp = t >> 5;
    int carry0 = 0, carry1 = 0;   // (the same in every thread)
    for (int b0 = 0; b0 < nb; b0 += 256 * 8) {
      const int bb = b0 + t * 8;
      int e[8], c[8];
#pragma unroll
      for (int x = 0; x < 8; x++) {
        const int b = bb + x;
        c[x] = b < nb ? cntc[b] : 0;
        e[x] = b < nb ? njr[b] + 3 * c[x] + (cntt ? cntt[b] : 0) : 0;
      }
      int se = 0, sc_ = 0;
#pragma unroll
      for (int x = 0; x < 8; x++) { const int ee = e[x], cc = c[x]; e[x] = se; c[x] = sc_; se += ee; sc_ += cc; }   // exclusive, local
      int v0 = se, v1 = sc_;
      for (int d = 1; d < 32; d <<= 1) {
        const int u0 = __shfl_up_sync(0xffffffffu, v0, d), u1 = __shfl_up_sync(0xffffffffu, v1, d);
        if (lane >= d) { v0 += u0; v1 += u1; }
      }
      if (lane == 31) { s_scan[wp] = v0; s_scan[8 + wp] = v1; }
      __syncthreads();
      if (wp == 0) {
        int s0 = lane < 8 ? s_scan[lane] : 0, s1 = lane < 8 ? s_scan[8 + lane] : 0;
        for (int d = 1; d < 8; d <<= 1) {
          const int u0 = __shfl_up_sync(0xffffffffu, s0, d), u1 = __shfl_up_sync(0xffffffffu, s1, d);
          if (lane >= d) { s0 += u0; s1 += u1; }
        }
        if (lane < 8) { s_scan[lane] = s0; s_scan[8 + lane] = s1; }
      }
      __syncthreads();
      const int before0 = carry0 + (wp > 0 ? s_scan[wp - 1] : 0) + v0 - se;
      const int before1 = carry1 + (wp > 0 ? s_scan[8 + wp - 1] : 0) + v1 - sc_;
#pragma unroll
      for (int x = 0; x < 8; x++) { const int b = bb + x; if (b < nb) { base[b] = before0 + e[x]; cbase[b] = before1 + c[x]; } }
      carry0 += s_scan[7]; carry1 += s_scan[15];
      __syncthreads();   // (s_scan is rewritten by the next trip)
    }
    if (t == 0) s_carry[0] = carry0;
    if (t == 0) {
      int nent = s_carry[0];
      long long wbase = 0;
      if (nent > 0) {
        wbase = (long long)atomicAdd(&B.ent_cursor[par], (unsigned long long)nent);
        if (wbase + nent > B.ent_cap - 1) { ws.overflow |= 16; nent = 0; }   // pool exhausted (its last entry is the dummy entry)
      }
      B.wsb[w] = make_int4(1, 0, (int)wbase, nent);   // period and stages follow in phase 6
    }
  }
  pgsl_barrier(bar, bar_target);
  const int4 sc = B.wsb[w];
  const int nent = sc.w, wbase = sc.z;
  // ---- 3
  for (int c = gt; c < nc && nent > 0; c += GT) {
    const Contact &C = B.contact[wi.contact0 + c];
    if (!B.bstep[C.b1].active) continue;
    const int b1 = C.b1 - wi.body0, b2 = C.b2 >= 0 ? C.b2 - wi.body0 : -1;
    clist[cbase[b1] + atomicAdd(&fill[b1], 1)] = c;
    if (b2 >= 0) clist[cbase[b2] + atomicAdd(&fill[b2], 1)] = c;
  }
  pgsl_barrier(bar, bar_target);
  // ---- 4
  for (int b = gt; b < nb; b += GT) {
    const int ncb = nent > 0 ? cntc[b] : 0, nj_rows = nent > 0 ? njr[b] : 0;
    int *cl = clist + cbase[b];
    for (int x = 1; x < ncb; x++) {   // insertion sort, ascending contact index
      const int v = cl[x];
      int y = x - 1;
      while (y >= 0 && cl[y] > v) { cl[y + 1] = cl[y]; y--; }
      cl[y + 1] = v;
    }
    int *lpb = lp + base[b];
    int at = 0;
    const Body &bd = B.body[wi.body0 + b];
    if (nent > 0)
      for (int q = bd.joint_begin; q < bd.joint_end; q++) {
        const Joint &j = B.joint[B.body_joint[q]];
        const bool is1 = (j.b1 - wi.body0) == b;
        for (int r = 0; r < j.m; r++) { (is1 ? S.idx1 : S.idx2)[j.row0 + r] = at; lpb[at] = j.row0 + r; at++; }
      }
    int tb = 0;   // torsional rows of the body's earlier contacts
    for (int k = 0; k < ncb; k++) {
      const int c = cl[k];
      const Contact &C = B.contact[wi.contact0 + c];
      const bool is1 = (C.b1 - wi.body0) == b;
      int *idx = is1 ? S.idx1 : S.idx2;
      const int pn = mj + (nc - 1 - c), pf = pfric + (rpc - 1) * c;
      const int in = nj_rows + (ncb - 1 - k), if0 = nj_rows + ncb + 2 * k + tb;
      idx[pn] = in; lpb[in] = pn;
      idx[pf] = if0; lpb[if0] = pf;
      idx[pf + 1] = if0 + 1; lpb[if0 + 1] = pf + 1;
      if (rpc == 4 && C.tors) { idx[pf + 2] = if0 + 2; lpb[if0 + 2] = pf + 2; tb++; }
    }
    B.bsched[wi.body0 + b] = nent > 0 ? make_int4(wbase + base[b], nj_rows + 3 * ncb + tb, nj_rows + ncb, 0) : make_int4(0, 0, 0, 0);
  }
  for (int pos = gt; pos < m; pos += GT) S.off[pos] = (nent > 0 && S.rb1[pos] >= 0) ? 0 : -1;
  if (B.cta_e0)   // k_pgsc: where the entries of every 256 bodies begin
    for (int c = gt; c < GZ_PGSC_TAB; c += GT) {
      const int b = c * B.pgsc_bodies;
      B.cta_e0[(size_t)w * GZ_PGSC_TAB + c] = wbase + ((b < nb && nent > 0) ? base[b] : nent);
    }
  pgsl_barrier(bar, bar_target);
  // ---- 5
  for (int it = 0; it < (1 << 20) && nent > 0 && relax; it++) {
    int changed = 0;
    for (int pos = gt; pos < m; pos += GT) {
      const int x1 = S.rb1[pos];
      if (x1 < 0) continue;
      const int x2 = S.rb2[pos], i1 = S.idx1[pos];
      int o = 0;
      if (i1 > 0) o = __ldcg(&S.off[lp[base[x1] + i1 - 1]]) + 1;
      if (x2 >= 0) { const int i2 = S.idx2[pos]; if (i2 > 0) o = max(o, __ldcg(&S.off[lp[base[x2] + i2 - 1]]) + 1); }
      if (o != S.off[pos]) { S.off[pos] = o; changed = 1; }
    }
    if (threadIdx.x == 0) s_flag = 0;
    __syncthreads();
    if (changed) s_flag = 1;
    __syncthreads();
    if (threadIdx.x == 0 && s_flag) atomicAdd(&flag[it % 3], 1);
    pgsl_barrier(bar, bar_target);
    if (gt == 0) flag[(it + 2) % 3] = 0;   // (last read after the previous barrier, next raised after the next one)
    if (__ldcg(&flag[it % 3]) == 0) break;
  }
  // ---- 6
  {
    int Rl = 1, mo = -1;
    for (int b = gt; b < nb && nent > 0 && relax; b += GT) {
      const int4 bs = B.bsched[wi.body0 + b];
      if (bs.y > 0) {
        const int first = S.off[lp[base[b]]], last = S.off[lp[base[b] + bs.y - 1]];
        Rl = max(Rl, last - first + 1); mo = max(mo, last);
      }
    }
    Rl = __reduce_max_sync(0xffffffffu, Rl); mo = __reduce_max_sync(0xffffffffu, mo);
    if ((threadIdx.x & 31) == 0) { atomicMax(&acc[0], Rl); atomicMax(&acc[1], mo); }
  }
  pgsl_barrier(bar, bar_target);
  if (gt == 0) {
    const int R = __ldcg(&acc[0]), maxoff = __ldcg(&acc[1]);
    int ne = nent;
    if (maxoff > GZ_ENT_OFFMASK) { ws.overflow |= 16; ne = 0; acc[0] = -1; }
    B.wsb[w] = make_int4(R, maxoff < 0 ? 0 : maxoff / R + 1, wbase, ne);
    const int key = min((ne + 3) >> 2, GZ_PGSB_NKEY - 1);
    B.wkey[w] = key | (atomicAdd(&B.rk_count[par * GZ_PGSB_NKEY + key], 1) << 10);
  }
  pgsl_barrier(bar, bar_target);
  if (__ldcg(&acc[0]) < 0) {   // offsets beyond the entry format: the world is not solved, the step is flagged
    for (int pos = gt; pos < m; pos += GT) S.off[pos] = -1;
    for (int b = gt; b < nb; b += GT) B.bsched[wi.body0 + b] = make_int4(0, 0, 0, 0);
  }
}

}  // namespace gz

namespace gz {

// ---- k_pgsd: the sweeps of a large world as DATAFLOW.  Same lanes, entries and arithmetic as k_pgsl, but no rounds and
// no grid barrier: a lane walks its body's entry list at its own pace and meets another lane only at a two-body row,
// where each publishes its half of the dot product in the row's slot and polls the partner's.  A slot carries value and
// tag in two 8-byte words {32 bits of the double, tag} written by one 16-byte store (relaxed, GPU scope): each word is atomic, so a
// reader that finds the tag in both has the value (no fence, no second trip).  tag = epoch * (sweeps + 1) + sweep + 1,
// the epoch counted on the device per launch; slots are double-buffered by the sweep's parity: a lane can only publish
// sweep s + 2 of a row after its partner has published s + 1, i.e. has consumed s.
// Progress: the grid is launched cooperatively (all lanes resident), a warp only spins while NONE of its lanes can
// advance (so no lane waits on a lane of its own warp that is held at a reconvergence point), and the rows a lane
// waits on precede its own in order[] (a total order), so some lane of the grid can always advance.  The result does not depend on timing:
// every row still sees exactly the accumulators the serial sweep would give it.
__device__ __forceinline__ void slot_put(unsigned long long *s, double v, unsigned tag) {
  const unsigned long long t = (unsigned long long)tag << 32;
  const unsigned long long w0 = t | (unsigned)__double2loint(v), w1 = t | (unsigned)__double2hiint(v);
  asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(s), "l"(w0), "l"(w1) : "memory");
}
__device__ __forceinline__ bool slot_get(const unsigned long long *s, unsigned tag, double &v) {
  unsigned long long w0, w1;
  asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(s) : "memory");
  v = __hiloint2double((int)(unsigned)w1, (int)(unsigned)w0);
  return (unsigned)(w0 >> 32) == tag && (unsigned)(w1 >> 32) == tag;
}

template <bool BOX>
__global__ void __launch_bounds__(256) k_pgsd(DevBatch B, const int w, unsigned long long *slots, double *partial, unsigned *epoch_ctr, const int *skip) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  if (skip && *skip) return;   // k_pgsm (one cluster, on chip) has taken the world this step
  const unsigned epoch = *epoch_ctr;   // (advanced by the last lane standing, after everybody has read it)
  const int gl = blockIdx.x * blockDim.x + threadIdx.x, L = gridDim.x * blockDim.x;
  const int bl = gl >> 1, S = gl & 1;
  const WorldInfo wi = B.world[w];
  const int4 sc = B.wsb[w];
  const int nent = sc.w;
  int cnt = 0, e0 = sc.z, fric0 = 0;
  if (bl < wi.nb && nent > 0) { const int4 bs = B.bsched[wi.body0 + bl]; e0 = bs.x; cnt = bs.y; fric0 = bs.z; }
  double *const lp = B.ent + (size_t)e0 * GZ_ENT;
  double *const dl = B.ent_dl + 2 * (size_t)e0 + S;
  const int *const part = B.ent_part + e0;
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
  const unsigned tag0 = epoch * (unsigned)(total + 1) + 1u;
  int i = (iters > 0) ? 0 : fric0, sweep = 0;
  bool alive = i < cnt && total > 0;
  // The row in flight: what finishing it needs (everything that does not wait for the partner is computed when the
  // row is loaded), and the NEXT row's Jacobian half and partner, loaded ahead: from the partner's value to this lane's
  // next publication there are only the clamp, the accumulator update and the next dot product.
  double iM[6] = {0, 0, 0, 0, 0, 0}, Jn[6] = {0, 0, 0, 0, 0, 0};
  double d = 0.0, delta0 = 0, old_lambda = 0, lo_act = 0, hi_act = 0;
  int cls = 0, pe = -1, pen = -1, ni = 0, nsweep = 0;
  bool isb2 = false, has_next = false;
  auto load_J = [&](int q) {   // Jacobian half (x Ad) and partner entry of list place q
    const double2 *p2 = reinterpret_cast<const double2 *>(lp + (size_t)q * GZ_ENT);
    const double2 v0 = p2[0], v1 = p2[1], v2 = p2[2];
    pen = part[q];
    Jn[0] = v0.x; Jn[1] = v0.y; Jn[2] = v1.x; Jn[3] = v1.y; Jn[4] = v2.x; Jn[5] = v2.y;
  };
  auto load_rest = [&](int q) {   // the rest of the row at list place q, and where the lane goes after it
    const double *p = lp + (size_t)q * GZ_ENT;
    const double2 *p2 = reinterpret_cast<const double2 *>(p);
    const double2 v3 = p2[3], v4 = p2[4], v5 = p2[5], vs = p2[6 + S], v8 = p2[8], v9 = p2[9];
    iM[0] = v3.x; iM[1] = v3.y; iM[2] = v4.x; iM[3] = v4.y; iM[4] = v5.x; iM[5] = v5.y;
    const int mx = __double2loint(v9.y), my = __double2hiint(v9.y);
    cls = (mx >> 27) & 3; isb2 = (mx >> 29) & 1;
    // pyramid friction: limits +-|mu * lambda_normal| (quickstep_pgs_lcp.cpp:572-582); box: +-mu (:588-594)
    double nlam = 1.0;
    if (cls == 2 && (!BOX || ((mx >> 30) & 1))) nlam = (p - (size_t)((unsigned)my >> 16) * GZ_ENT)[13 + 2 * S];   // (bit 30: a torsional row)
    old_lambda = vs.y;
    delta0 = vs.x - old_lambda * v8.x;
    const double hl = v9.x * nlam;
    const double hf = fabs(hl);
    hi_act = cls == 2 ? hf : v9.x; lo_act = cls == 2 ? -hf : v8.y;
    ni = q + 1; nsweep = sweep;
    if (ni == cnt) { nsweep = sweep + 1; ni = nsweep < iters ? 0 : fric0; }
    has_next = nsweep < total && ni < cnt;
    if (has_next) load_J(ni);
  };
  if (alive) {
    load_J(i);
    d = a[0] * Jn[0];
#pragma unroll
    for (int t = 1; t < 6; t++) d += a[t] * Jn[t];
    pe = pen;
    if (pe >= 0) slot_put(slots + 2 * (((size_t)(e0 + i) * 2 + S) * 2 + (sweep & 1)), d, tag0 + (unsigned)sweep);
    load_rest(i);
  }
  // where and for what the lane polls: kept in registers across the poll loop (left to the compiler, the loop
  // recomputed S from the thread index, the slot address and the tag on every trip: 26 instructions a poll)
  const unsigned long long *pslot = slots + 2 * (((size_t)max(pe, 0) * 2 + S) * 2 + (sweep & 1));
  unsigned ptag = tag0 + (unsigned)sweep;
  // The warp stays together: ALL its lanes poll back to back until some lane can finish its row (a waiting lane's
  // poll interval is one L2 trip, not one pass of its neighbours through the row code), then those lanes finish.
  int n_outer = 0, n_spin = 0;   // developer counters: passes of this warp through the row code, polls
  const long long t_begin = clock64();
  while (__any_sync(0xffffffffu, alive)) {
    double pd = 0.0;
    bool ready;
    n_outer++;
    do {
      n_spin++;
      ready = alive;
      if (alive && pe >= 0) ready = slot_get(pslot, ptag, pd);
    } while (!__any_sync(0xffffffffu, ready));
    if (ready) {
      // this body's half and the partner's, subtracted in the row's body order (body 1 first)
      double delta = delta0;
      const double d1 = isb2 ? pd : d, d2 = isb2 ? d : pd;
      delta -= d1;
      const double delta2 = delta - d2;
      delta = pe >= 0 ? delta2 : delta;
      double nl = old_lambda + delta;
      const double dlo = lo_act - old_lambda, dhi = hi_act - old_lambda;
      const bool below = nl < lo_act, above = !below && nl > hi_act;
      delta = below ? dlo : (above ? dhi : delta);
      nl = below ? lo_act : (above ? hi_act : nl);
#pragma unroll
      for (int t = 0; t < 6; t++) a[t] += delta * iM[t];
      double dn = 0.0;
      if (has_next) {   // the next row's half goes out before anything else
        dn = a[0] * Jn[0];
#pragma unroll
        for (int t = 1; t < 6; t++) dn += a[t] * Jn[t];
        if (pen >= 0) slot_put(slots + 2 * (((size_t)(e0 + ni) * 2 + S) * 2 + (nsweep & 1)), dn, tag0 + (unsigned)nsweep);
      }
      // SMOOTH_LAMBDA on the lambda solve of non-bilateral rows; delta is not recomputed after it (:687-691 vs :706)
      const double sm = smooth1 * nl + smooth * old_lambda;
      lp[(size_t)i * GZ_ENT + 13 + 2 * S] = (S == 0 && cls != 0) ? sm : nl;
      dl[2 * (size_t)i] = delta;
      i = ni; sweep = nsweep; alive = has_next; d = dn; pe = pen;
      pslot = slots + 2 * (((size_t)max(pe, 0) * 2 + S) * 2 + (sweep & 1)); ptag = tag0 + (unsigned)sweep;
      if (alive) load_rest(i);
    }
  }
  if ((threadIdx.x & 31) == 0) {   // gzcuda_batch_get_pgs_debug: the longest warp's passes, polls, cycles / 16
    WorldStep &wsd = B.wstep[w];
    atomicMax(&wsd.pad[0], n_outer); atomicMax(&wsd.pad[1], n_spin); atomicMax(&wsd.n_fric, (int)((clock64() - t_begin) >> 4));
  }
  if (bl < wi.nb) {
    BodyStep &bs = B.bstep[wi.body0 + bl];
    double *o = S ? bs.caccel_erp : bs.caccel;
#pragma unroll
    for (int t = 0; t < 6; t++) o[t] = a[t];
  }
  grid.sync();
  if (gl == 0) *epoch_ctr = epoch + 1u;
  double st[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};   // acc[3], err[3], cnt[3]
  {
    const double *wp = B.ent + (size_t)sc.z * GZ_ENT;
    const double *wdl = B.ent_dl + 2 * (size_t)sc.z;
    Row *rw = B.row + wi.row0;
    for (int q = gl; q < nent; q += L) {
      const int r = B.ent_row[sc.z + q];
      if (r < 0) continue;   // the row's body-1 entry carries it
      const double *pq = wp + (size_t)q * GZ_ENT;   // (L2 loads: a line this SM cached may hold a neighbour's stale half)
      rw[r].lambda = __ldcg(&pq[13]); rw[r].lambda_erp = __ldcg(&pq[15]);
      const int cls = (__double2loint(__ldcg(&pq[19])) >> 27) & 3;
      if ((cls == 2 ? total : iters) <= 0) continue;
      const double de = __ldcg(&wdl[2 * (size_t)q]), dd = de * de, e2 = dd * B.ent_ad2[sc.z + q];
      st[cls] += dd; st[3 + cls] += e2; st[6 + cls] += 1.0;
    }
  }
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
  }
}

}  // namespace gz
