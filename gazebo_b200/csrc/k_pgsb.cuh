// k_pgsb.cuh — body-owner SOR-PGS: ComputeRows (quickstep_pgs_lcp.cpp:52-954) with one lane per
// (body, solve) instead of one lane per row.
//
// The reference sweeps rows strictly in order[] sequence.  Two row updates commute EXACTLY when they
// share no body, so any execution that keeps, for every body, the order in which rows touch it is
// bit-identical (same argument as round 1's row-slot kernels).  Here the per-body order is not a
// constraint a scheduler has to satisfy but the data structure itself:
//
//   * every body has an ENTRY LIST: its half of every row that touches it, in order[] sequence
//     (bilateral rows, contact normals in reverse index order, friction rows; :1115-1150);
//   * lane (b, S) owns body b for the whole kernel and keeps caccel[b] (S = 0) or caccel_erp[b]
//     (S = 1) in SIX REGISTERS; the lambda / lambda_erp solves are independent problems (split
//     impulse, :526-542) on adjacent lanes reading the same entry;
//   * a two-body row is run by its two owner lanes in the same round: each computes the 6-term dot
//     product of its own half (the reference subtracts the two sums separately, :526-531), the
//     partial sums are exchanged with ONE shuffle, both lanes then compute the same delta and each
//     updates its own accumulator and its own copy of lambda.  Nothing else is shared, so there is
//     no shared-memory traffic on the dependent chain and no barrier in the one-warp case;
//   * the round of a row is off(r) = 1 + max over its bodies of the round of their previous row
//     (one serial pass, k_bsched: no greedy, no period probing), the sweep period of a world is
//     R = the longest span first..last of any body, and row r of sweep s runs in round s*R + off(r)
//     (software pipelining of the sweep loop: sweeps overlap wherever bodies allow);
//   * entries (160 B) are staged in shared memory once per step and stay there for the 60 sweeps;
//     what does not fit the CTA's budget is read from the pool (L2) with the same code path.
//
// Worlds of up to 16 bodies share a warp (32 / (2 * padded bodies) worlds per warp, event-driven
// rounds: the warp jumps to the next round in which any lane is due).  Larger worlds take a CTA of
// their own (up to 256 bodies): the exchange goes through shared memory and one barrier per round.
#pragma once

namespace gz {

// ---- entry: what a lane streams per row of its body, 20 doubles = 160 B ------------------------------
// [0..5] J half (scaled by Ad)  [6..11] iMJ half  [12] rhs*Ad  [13] lambda  [14] rhs_erp*Ad  [15] lambda_erp
// [16] Ad*cfm  [17] lo  [18] hi  [19] meta {int x, y}:
//   x = off (bits 0..26) | class << 27 (0 bilateral, 1 contact normal, 2 friction) | is-body-2 << 29
//   y = partner body (world-local, 0xFFFF none) | back << 16 (friction: distance, in list entries, to the normal row)
#define GZ_ENT_OFFMASK 0x7FFFFFF
#define GZ_ENT_NOPART 0xFFFF

__device__ __forceinline__ unsigned long long pgsb_gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned pgsb_smid() {
  unsigned s;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(s));
  return s & 511u;
}

#define GZ_BS_SM 384   // rows up to which k_bsched works in shared memory
struct WScratchB { int *rb1, *rb2, *off, *idx1, *idx2; };
__device__ inline WScratchB wscratchb(const DevBatch &B, int w) {
  WScratchB s;
  const size_t n = B.max_row_cap;
  int *base = B.scratch + (size_t)w * 7 * n;
  s.rb1 = base; s.rb2 = base + n; s.off = base + 2 * n; s.idx1 = base + 3 * n; s.idx2 = base + 4 * n;
  return s;
}

// order[] position of row i (quickstep_pgs_lcp.cpp:1115-1150 with row_reorder1): bilateral rows in index
// order, contact normals in REVERSE index order, friction rows in index order.  mj = persistent-joint rows.
// rpc = rows per contact (3; 4 in a world with torsional friction: the torsional row follows the contact's two
// friction rows).
__device__ __forceinline__ int order_pos(int i, int mj, int nc, int rpc) {
  if (i < mj) return i;
  const int c = (i - mj) / rpc, f = (i - mj) % rpc;
  return f == 0 ? mj + (nc - 1 - c) : mj + nc + (rpc - 1) * c + (f - 1);
}

// One CTA per world: bodies of the rows by order[] position (parallel), then ONE serial pass that gives
// every row its round offset and its place in the entry lists of its bodies, then the world's period and
// its slice of the entry pool.  Integer work only; needs the row counts (k_assemble1), not the rows.
// Dynamic shared memory: 5 * max_nb ints.
__global__ void __launch_bounds__(128) k_bsched(DevBatch B) {
  extern __shared__ int bs_sh[];
  const int w = blockIdx.x, tid = threadIdx.x;
  const WorldInfo wi = B.world[w];
  WorldStep &ws = B.wstep[w];
  const int nb = wi.nb, m = ws.m, nc = ws.n_contacts, rpc = wi.rpc, mj = m - rpc * nc;
  int *const tab = B.bs_glob ? B.bs_glob + (size_t)w * 5 * B.max_nb : bs_sh;   // (large worlds: the tables live in global memory)
  int *last = tab, *first = tab + B.max_nb, *cnt = tab + 2 * B.max_nb, *fr0 = tab + 3 * B.max_nb, *base = tab + 4 * B.max_nb;
  __shared__ int s_nent, s_wbase;
  const WScratchB G = wscratchb(B, w);
  // the serial pass below walks these arrays with dependent accesses: worlds of up to GZ_BS_SM rows keep them in
  // shared memory (the row offsets and list places go out to global memory in parallel afterwards)
  __shared__ int sm_arr[5 * GZ_BS_SM];
  const bool in_sm = m <= GZ_BS_SM;
  WScratchB S = G;
  if (in_sm) { S.rb1 = sm_arr; S.rb2 = sm_arr + GZ_BS_SM; S.off = sm_arr + 2 * GZ_BS_SM; S.idx1 = sm_arr + 3 * GZ_BS_SM; S.idx2 = sm_arr + 4 * GZ_BS_SM; }
  const int par = (int)(B.step_index & 1);
  if (w == 0) {   // the other parity's allocator and period histogram: idle this step
    if (tid == 0) B.ent_cursor[par ^ 1] = 0ull;
    for (int q = tid; q < GZ_PGSB_NKEY; q += blockDim.x) B.rk_count[(par ^ 1) * GZ_PGSB_NKEY + q] = 0;
  }
  for (int b = tid; b < nb; b += blockDim.x) { last[b] = -1; first[b] = -1; cnt[b] = 0; fr0[b] = -1; }
  if (m > 0) {
    for (int k = tid; k < wi.nj; k += blockDim.x) {
      const Joint &j = B.joint[wi.joint0 + k];
      const int b1 = j.b1 - wi.body0, b2 = j.b2 >= 0 ? j.b2 - wi.body0 : -1;
      for (int r = 0; r < j.m; r++) { S.rb1[j.row0 + r] = b1; S.rb2[j.row0 + r] = b2; }
    }
    for (int c = tid; c < nc; c += blockDim.x) {
      const Contact &C = B.contact[wi.contact0 + c];
      // a contact whose island is not stepped this step has rows (getInfo1 counted them) but they are never swept
      const bool live = B.bstep[C.b1].active != 0;
      const int b1 = live ? C.b1 - wi.body0 : -2, b2 = C.b2 >= 0 ? C.b2 - wi.body0 : -1;
      const int pn = mj + (nc - 1 - c), pf = mj + nc + (rpc - 1) * c;
      S.rb1[pn] = b1; S.rb2[pn] = b2;
      S.rb1[pf] = b1; S.rb2[pf] = b2;
      S.rb1[pf + 1] = b1; S.rb2[pf + 1] = b2;
      if (rpc == 4) { S.rb1[pf + 2] = C.tors ? b1 : -2; S.rb2[pf + 2] = b2; }   // (a pair without torsional friction: a placeholder, never swept)
    }
  }
  __syncthreads();
  if (B.max_nb <= 32) {
    // ---- worlds of up to 32 bodies: the serial pass is run by ONE WARP with lane b holding body b's counters in
    // registers: a row costs two shuffles and a handful of integer instructions instead of a chain of dependent
    // shared-memory reads and writes on one thread
    if (tid < 32) {
      const int lane = tid, pfric = mj + nc;
      int lst = -1, fst = -1, c = 0, f0 = -1, maxoff = -1;
      for (int pos = 0; pos < m; pos++) {
        const int x1 = S.rb1[pos], x2 = S.rb2[pos];          // same address for all lanes: broadcast
        if (x1 < 0) { if (lane == 0) S.off[pos] = -1; continue; }
        const int y2 = x2 >= 0 ? x2 : x1;
        const int l1 = __shfl_sync(0xffffffffu, lst, x1), l2 = __shfl_sync(0xffffffffu, lst, y2);
        const int c1 = __shfl_sync(0xffffffffu, c, x1), c2 = __shfl_sync(0xffffffffu, c, y2);
        const int o = max(l1, l2) + 1;
        if (lane == x1 || lane == x2) {
          if (pos >= pfric && f0 < 0) f0 = c;
          c++;
          lst = o;
          if (fst < 0) fst = o;
        }
        if (lane == 0) { S.off[pos] = o; S.idx1[pos] = c1; if (x2 >= 0) S.idx2[pos] = c2; }
        maxoff = max(maxoff, o);
      }
      const int span = fst >= 0 ? lst - fst + 1 : 1;
      const int R = __reduce_max_sync(0xffffffffu, span);
      int incl = c;                                           // inclusive scan of the list lengths
      for (int dd = 1; dd < 32; dd <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, dd); if (lane >= dd) incl += t; }
      int nent = __shfl_sync(0xffffffffu, incl, 31);
      if (lane < nb) { cnt[lane] = c; fr0[lane] = f0; base[lane] = incl - c; }
      if (lane == 0) {
        long long wbase = 0;
        if (nent > 0) {
          wbase = (long long)atomicAdd(&B.ent_cursor[par], (unsigned long long)nent);
          if (wbase + nent > B.ent_cap - 1 || maxoff > GZ_ENT_OFFMASK) { ws.overflow |= 16; nent = 0; }   // pool exhausted (its last entry is the dummy entry)
        }
        s_nent = nent; s_wbase = (int)wbase;
        B.wsb[w] = make_int4(R, maxoff < 0 ? 0 : maxoff / R + 1, (int)wbase, nent);
        const int key = min((nent + 3) >> 2, GZ_PGSB_NKEY - 1);   // k_border's sort key, see below
        B.wkey[w] = key | (atomicAdd(&B.rk_count[par * GZ_PGSB_NKEY + key], 1) << 10);
      }
      nent = __shfl_sync(0xffffffffu, nent, 0);
      if (nent == 0)
        for (int pos = lane; pos < m; pos += 32) S.off[pos] = -1;   // rows without entries: k_rowfinish skips them
    }
  } else
  if (tid == 0) {
    int maxoff = -1;
    const int pfric = mj + nc;
    for (int pos = 0; pos < m; pos++) {
      const int x1 = S.rb1[pos], x2 = S.rb2[pos];
      if (x1 < 0) { S.off[pos] = -1; continue; }
      int o = last[x1] + 1;
      if (x2 >= 0) o = max(o, last[x2] + 1);
      const int i1 = cnt[x1]++;
      last[x1] = o;
      if (first[x1] < 0) first[x1] = o;
      if (pos >= pfric && fr0[x1] < 0) fr0[x1] = i1;
      S.idx1[pos] = i1;
      if (x2 >= 0) {
        const int i2 = cnt[x2]++;
        last[x2] = o;
        if (first[x2] < 0) first[x2] = o;
        if (pos >= pfric && fr0[x2] < 0) fr0[x2] = i2;
        S.idx2[pos] = i2;
      }
      S.off[pos] = o;
      maxoff = max(maxoff, o);
    }
    int R = 1, nent = 0;
    for (int b = 0; b < nb; b++) {
      if (first[b] >= 0) R = max(R, last[b] - first[b] + 1);
      base[b] = nent;
      nent += cnt[b];
    }
    long long wbase = 0;
    if (nent > 0) {
      wbase = (long long)atomicAdd(&B.ent_cursor[par], (unsigned long long)nent);
      if (wbase + nent > B.ent_cap - 1 || maxoff > GZ_ENT_OFFMASK) {   // (the pool's last entry is the dummy entry)   // pool exhausted: the world is not solved, the step is flagged
        ws.overflow |= 16;
        nent = 0;
      }
    }
    s_nent = nent; s_wbase = (int)wbase;
    B.wsb[w] = make_int4(R, maxoff < 0 ? 0 : maxoff / R + 1, (int)wbase, nent);
    // k_border sorts the worlds by their number of entries (in steps of 4), largest first: that is their shared-
    // memory need, so the CTAs of a class are filled tightly, and it follows the duration of the sweeps closely,
    // so the worlds of a warp run for about the same number of rounds and the longest ones start first
    const int key = min((nent + 3) >> 2, GZ_PGSB_NKEY - 1);
    B.wkey[w] = key | (atomicAdd(&B.rk_count[par * GZ_PGSB_NKEY + key], 1) << 10);
    if (nent == 0)
      for (int pos = 0; pos < m; pos++) S.off[pos] = -1;   // rows without entries: k_rowfinish skips them
  }
  __syncthreads();
  const int nent = s_nent, wbase = s_wbase;
  if (in_sm)
    for (int pos = tid; pos < m; pos += blockDim.x) { G.off[pos] = S.off[pos]; G.idx1[pos] = S.idx1[pos]; G.idx2[pos] = S.idx2[pos]; }
  for (int b = tid; b < nb; b += blockDim.x)
    B.bsched[wi.body0 + b] = nent > 0 ? make_int4(wbase + base[b], cnt[b], fr0[b] < 0 ? cnt[b] : fr0[b], 0) : make_int4(0, 0, 0, 0);
}

// Places every world in the sweep order: by period, longest first (counting sort: k_bsched counted the periods
// and gave every world its rank among those of equal period).  Each block rebuilds the prefix over the keys.
__global__ void __launch_bounds__(256) k_border(DevBatch B) {
  __shared__ int pre[GZ_PGSB_NKEY];
  const int par = (int)(B.step_index & 1);
  if (blockIdx.x == 0 && threadIdx.x < 16) B.cls_ctr[threadIdx.x] = 0;   // work counters of k_pgsb's class launches
  const int *cnt = B.rk_count + par * GZ_PGSB_NKEY;
  for (int q = threadIdx.x; q < GZ_PGSB_NKEY; q += blockDim.x) pre[q] = cnt[GZ_PGSB_NKEY - 1 - q];   // descending key
  __syncthreads();
  if (threadIdx.x < 32) {   // exclusive scan of GZ_PGSB_NKEY values by one warp
    int carry = 0;
    for (int b0 = 0; b0 < GZ_PGSB_NKEY; b0 += 32) {
      const int v = pre[b0 + threadIdx.x];
      int x = v;
      for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, x, d); if ((int)threadIdx.x >= d) x += t; }
      pre[b0 + threadIdx.x] = carry + x - v;
      carry += __shfl_sync(0xffffffffu, x, 31);
    }
  }
  __syncthreads();
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= B.n_worlds) return;
  const int k = B.wkey[w];
  B.wperm[pre[GZ_PGSB_NKEY - 1 - (k & 1023)] + (k >> 10)] = w;
}

// ---- the sweeps ---------------------------------------------------------------------------------------
template <bool V> struct PgsbTag { static constexpr bool value = V; };
struct PgsbCfg {
  int lw;        // lanes per world = 2 * padded body count (power of two, >= 2)
  int wpc;       // worlds per CTA (32 / lw for one-warp CTAs, 1 for multi-warp CTAs)
  int nres;      // entries the CTA can stage in shared memory (all its worlds together)
  int cls;       // shared-memory class of this launch (ascending budgets), -1: the only launch (multi-warp CTAs)
  int last;      // the last class: it also takes the CTAs that fit no class (they read the pool)
};

// Entries staged in shared memory are padded to 22 doubles (176 B = 11 x 16 B, an odd number of 16-byte bank
// groups) so that the 16 distinct entries a warp reads per 128-bit load spread over all banks.
#define GZ_ENTS 22

// The sweeps of one lane = (body, solve).  RES: the lane's list is in shared memory (stride GZ_ENTS, lofs doubles
// into the dynamic shared array), else in the pool (lp_g, stride GZ_ENT).  The current entry lives in registers;
// its successor is loaded IN PLACE, field by field right after the last use of each field in this pass (J after
// the dot product, iMJ and the scalars after the accumulator update), so the loads run beside the arithmetic and
// no register is ever moved.  A pass is ONE basic block (no branch, so the compiler interleaves the address and
// round arithmetic with the dependent FP64 chain): a lane that is not due, or has finished, "moves on" to the
// entry it already holds and keeps its state through selects and predicated updates; a lane without rows holds
// an all-zero dummy entry.  The last delta of every entry is left beside it (shared memory: the spare doubles of
// the padded entry; pool: ent_dl) for the statistics.
template <bool RES, bool MULTI, bool BOX>
__device__ __forceinline__ int pgsb_sweeps(const DevBatch &B, double *lp_g, const int lofs, const int cnt, const int fric0,
                                           const int R, const int nstage, const int S, const int bl, const int lane_base,
                                           const int LW, const int xch_ofs, const int e0, double (&a)[6]) {
  extern __shared__ __align__(16) double shd[];
  constexpr int STR = RES ? GZ_ENTS : GZ_ENT;
  double *const lp = RES ? shd + lofs : lp_g;
  const int iters = B.p.iters, total = B.p.iters + B.p.friction_iters;
  const double smooth = B.p.smooth_contacts, smooth1 = 1.0 - smooth;
  int sweep = 0;
  double *pc = lp + (size_t)((iters > 0) ? 0 : fric0) * STR;   // current entry
  double *const p0 = lp, *const pfr = lp + (size_t)fric0 * STR, *const pend = lp + (size_t)cnt * STR;
  double *dlc = B.ent_dl + 2 * ((size_t)e0 + ((iters > 0) ? 0 : fric0)) + S;   // pool: where the current entry's delta goes
  double *const dl0 = B.ent_dl + 2 * (size_t)e0 + S, *const dlfr = dl0 + 2 * fric0;
  const bool alive = pc < pend && total > 0;
  if (!alive) pc = RES ? shd + xch_ofs - GZ_ENTS : B.ent + (size_t)(B.ent_cap - 1) * GZ_ENT;   // the all-zero dummy entry
  // the entry a lane holds: two register sets used alternately (a pass reads one and loads the other, so the
  // loads of a pass are all issued at its start and have a whole pass to complete)
  struct EntR { double J[6], iM[6], rhs, lam, adcfm, lo, hi; int mx, my; };
  auto load_ent = [&](EntR &E, const double *p) {
    const double2 *p2 = reinterpret_cast<const double2 *>(p);
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const double2 x = p2[k], y = p2[3 + k];
      E.J[2 * k] = x.x; E.J[2 * k + 1] = x.y; E.iM[2 * k] = y.x; E.iM[2 * k + 1] = y.y;
    }
    const double2 rl = p2[6 + S], cl = p2[8], hm = p2[9];
    E.rhs = rl.x; E.lam = rl.y; E.adcfm = cl.x; E.lo = cl.y; E.hi = hm.x;
    E.mx = __double2loint(hm.y); E.my = __double2hiint(hm.y);
  };
  EntR EA, EB;
  load_ent(EA, pc);
  int T_due = alive ? (EA.mx & GZ_ENT_OFFMASK) : 0x7fffffff;
  const int rounds_multi = (total > 0 && nstage > 0) ? (total + nstage - 1) * R : 0;
  int npass = 0;
  // event-driven rounds: T = the next round in which any lane of the warp is due
  int T = MULTI ? 0 : __reduce_min_sync(0xffffffffu, T_due);
  // One pass: the lanes that are due in round T update the row they hold (C) and move on; every lane loads the
  // entry it holds next (N): its next entry, or the one it holds already if it is not due or has finished.
  auto pass = [&](EntR &C, EntR &N) {
    npass++;
    const bool due = (T_due == T);
    double *pn = pc + STR, *dln = dlc + 2;
    int nsw = sweep;
    if (pn == pend) { nsw = sweep + 1; pn = nsw < iters ? p0 : pfr; dln = nsw < iters ? dl0 : dlfr; }
    const bool more = due && nsw < total && pn != pend;
    pn = more ? pn : pc;
    dln = more ? dln : dlc;
    nsw = more ? nsw : sweep;
    load_ent(N, pn);
    const int cls = (C.mx >> 27) & 3;
    const bool fric = cls == 2;
    // pyramid friction: limits +-|mu * lambda_normal| (quickstep_pgs_lcp.cpp:572-582); box: +-mu (:588-594);
    // (the distance to the normal row is 0 for the other classes: they read their own lambda and ignore it)
    double nlam = 1.0;
    // a torsional row (bit 30) follows the normal force under either model (:554-566)
    if (!BOX || ((C.mx >> 30) & 1)) nlam = (pc - (size_t)((unsigned)C.my >> 16) * STR)[13 + 2 * S];
    // this body's half of the dot product, term by term in the reference's left-to-right order
    double d = a[0] * C.J[0];
#pragma unroll
    for (int t = 1; t < 6; t++) d += a[t] * C.J[t];
    const double old_lambda = C.lam;
    double delta = C.rhs - old_lambda * C.adcfm;
    {
      double pd = 0.0;
      const int part = C.my & 0xFFFF;
      const bool two = part != GZ_ENT_NOPART;
      if (MULTI) {
        double *x = shd + xch_ofs + (size_t)(T & 1) * LW;
        x[threadIdx.x] = d;
        __syncthreads();
        pd = x[two ? 2 * part + S : threadIdx.x];
      } else {
        pd = __shfl_sync(0xffffffffu, d, lane_base + 2 * (two ? part : bl) + S);
      }
      const bool isb2 = (C.mx >> 29) & 1;
      const double d1 = isb2 ? pd : d, d2 = isb2 ? d : pd;
      delta -= d1;
      const double delta2 = delta - d2;
      delta = two ? delta2 : delta;
    }
    const double hl = C.hi * nlam;
    const double hf = __hiloint2double(__double2hiint(hl) & 0x7fffffff, __double2loint(hl));
    const double nhf = __hiloint2double(__double2hiint(hl) | 0x80000000, __double2loint(hl));
    const double hi_act = fric ? hf : C.hi, lo_act = fric ? nhf : C.lo;
    double nl = old_lambda + delta;
    const double dlo = lo_act - old_lambda, dhi = hi_act - old_lambda;
    const bool below = nl < lo_act, over = nl > hi_act, above = !below && over;   // (two independent compares)
    delta = below ? dlo : (above ? dhi : delta);
    nl = below ? lo_act : (above ? hi_act : nl);
    // SMOOTH_LAMBDA on the lambda solve of non-bilateral rows; delta is not recomputed after it (:687-691 vs :706)
    const double sm = smooth1 * nl + smooth * old_lambda;
    const double nlw = (S == 0 && cls != 0) ? sm : nl;
    if (due) pc[13 + 2 * S] = nlw;
    if (due) { if (RES) pc[20 + S] = delta; else *dlc = delta; }
#pragma unroll
    for (int t = 0; t < 6; t++) { const double u = a[t] + delta * C.iM[t]; a[t] = due ? u : a[t]; }
    N.lam = (due && pn == pc) ? nlw : N.lam;   // a one-entry list holds on to the entry it has just updated
    pc = pn; dlc = dln; sweep = nsw;
    T_due = more ? sweep * R + (N.mx & GZ_ENT_OFFMASK) : (due ? 0x7fffffff : T_due);
  };
#define GZ_PGSB_NEXT() T = MULTI ? T + 1 : __reduce_min_sync(0xffffffffu, T_due)
  while (MULTI ? T < rounds_multi : T != 0x7fffffff) {
    pass(EA, EB);
    GZ_PGSB_NEXT();
    if (!(MULTI ? T < rounds_multi : T != 0x7fffffff)) break;
    pass(EB, EA);
    GZ_PGSB_NEXT();
  }
#undef GZ_PGSB_NEXT
  return npass;
}

// Shared memory a work item (the wpc worlds at places item * wpc .. of the sweep order) needs, in entries, rounded
// up the way k_border's sort key is (steps of 4): NON-INCREASING in item, so the items of a shared-memory class
// are a contiguous range.  A world beyond the key's range counts as "does not fit anywhere" (it reads the pool).
__device__ __forceinline__ int pgsb_need(const DevBatch &B, const PgsbCfg &cfg, int item) {
  int n = 0;
  for (int s = 0; s < cfg.wpc; s++) {
    const int q = item * cfg.wpc + s;
    if (q < B.n_worlds) {
      const int key = B.wkey[B.wperm[q]] & 1023;
      n += key >= GZ_PGSB_NKEY - 1 ? (1 << 24) : 4 * key;
    }
  }
  return n;
}
// first item whose need is <= limit (items are sorted by need, largest first)
__device__ __forceinline__ int pgsb_first_fitting(const DevBatch &B, const PgsbCfg &cfg, int n_items, int limit) {
  int lo = 0, hi = n_items;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (pgsb_need(B, cfg, mid) <= limit) hi = mid; else lo = mid + 1;
  }
  return lo;
}

// One work item: the sweeps of the wpc worlds of a warp (or of the one world of a multi-warp CTA).
template <bool MULTI, bool BOX>
__device__ __forceinline__ void pgsb_item(const DevBatch &B, const PgsbCfg &cfg, const int item, const bool res) {
  extern __shared__ __align__(16) double shd[];
  const int tid = threadIdx.x;
  const int LW = cfg.lw;
  const int wslot = MULTI ? 0 : tid / LW, lw = MULTI ? tid : tid % LW;
  const int bl = lw >> 1, S = lw & 1;
  const unsigned long long g0 = pgsb_gtimer();
  const int wq = item * cfg.wpc + wslot;   // place in the sweep order (k_border)
  const bool live_w = wq < B.n_worlds;
  const int w = live_w ? B.wperm[wq] : 0;
  WorldInfo wi;
  wi.nb = 0; wi.body0 = 0; wi.row0 = 0;
  int4 sc = make_int4(1, 0, 0, 0);
  if (live_w) { wi = B.world[w]; sc = B.wsb[w]; }
  const int R = sc.x, nent = sc.w;
  // ---- stage the entries of the CTA's worlds in shared memory (world after world, 22-double stride)
  int soff = 0;   // first staged entry of this world
  if (!MULTI)
    for (int s = 0; s < wslot && item * cfg.wpc + s < B.n_worlds; s++) soff += B.wsb[B.wperm[item * cfg.wpc + s]].w;   // (a partly filled last item)
  if (res) {
    const int n2 = nent * (GZ_ENT / 2);
    const double2 *src = reinterpret_cast<const double2 *>(B.ent + (size_t)sc.z * GZ_ENT);
    double2 *dst = reinterpret_cast<double2 *>(shd + (size_t)soff * GZ_ENTS);
    // 16-byte asynchronous copies: all of them are in flight at once (one wait at the end)
    for (int q = lw; q < n2; q += LW) {
      const unsigned da = (unsigned)__cvta_generic_to_shared(dst + (q / (GZ_ENT / 2)) * (GZ_ENTS / 2) + q % (GZ_ENT / 2));
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(da), "l"(src + q) : "memory");
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
  }
  const int xch_ofs = (cfg.nres + 1) * GZ_ENTS;   // after the staged entries and the dummy entry; MULTI: [2][LW] partial dot products
  if (res) {   // the all-zero dummy entry (no partner, never due) of the lanes without rows
    double *dm = shd + (size_t)cfg.nres * GZ_ENTS;
    for (int q = lw; q < GZ_ENTS; q += LW) dm[q] = (q == 19) ? __hiloint2double(GZ_ENT_NOPART, 0) : 0.0;
  }
  if (MULTI) __syncthreads(); else __syncwarp();
  // ---- this lane's list
  int cnt = 0, e0 = sc.z, fric0 = 0;
  if (live_w && bl < wi.nb) { const int4 bs = B.bsched[wi.body0 + bl]; e0 = bs.x; cnt = bs.y; fric0 = bs.z; }
  const int lofs = (soff + (e0 - sc.z)) * GZ_ENTS;
  double *lp_g = B.ent + (size_t)e0 * GZ_ENT;
  double *lp = res ? shd + lofs : lp_g;
  const int lstr = res ? GZ_ENTS : GZ_ENT;
  // ---- initial caccel = iMJ * lambda0 over the warm-started (persistent-joint) rows, in row index order per
  // body (multiply_invM_JT, quickstep_util.cpp:101-119): the bilateral entries lead every list
  double a[6] = {0, 0, 0, 0, 0, 0};
  if (B.p.warm_start > 0) {
    for (int q = 0; q < cnt; q++) {
      const double *p = lp + (size_t)q * lstr;
      const int mx = __double2loint(p[19]);
      if (((mx >> 27) & 3) != 0) break;
      const double l0 = p[13 + 2 * S];
#pragma unroll
      for (int t = 0; t < 6; t++) a[t] += p[6 + t] * l0;
    }
  }
  const int lane_base = MULTI ? 0 : (tid & 31) - lw;   // first lane of this world in the warp
  int npass;
  if (res) npass = pgsb_sweeps<true, MULTI, BOX>(B, lp_g, lofs, cnt, fric0, R, sc.y, S, bl, lane_base, LW, xch_ofs, e0, a);
  else npass = pgsb_sweeps<false, MULTI, BOX>(B, lp_g, lofs, cnt, fric0, R, sc.y, S, bl, lane_base, LW, xch_ofs, e0, a);
  // ---- results.  caccel to the bodies; then, over ALL entries of the world with all its lanes (coalesced reads
  // of the side arrays): lambda back to the rows (k_lambda_save copies them to the joints), and the statistics.
  // rms_dlambda / rms_constraint_residual (quickstep_pgs_lcp.cpp:700-760, 814-855): the sums restart every sweep,
  // so what survives is the last update of every row (sweep iters - 1 for bilateral and normal rows, total - 1
  // for friction rows): the deltas the sweeps left beside the entries.
  if (live_w && bl < wi.nb) {
    BodyStep &bs = B.bstep[wi.body0 + bl];
    double *o = S ? bs.caccel_erp : bs.caccel;
#pragma unroll
    for (int t = 0; t < 6; t++) o[t] = a[t];
  }
  if (MULTI) __syncthreads(); else __syncwarp();
  double st_acc[3] = {0, 0, 0}, st_err[3] = {0, 0, 0};
  int st_cnt[3] = {0, 0, 0};
  if (live_w) {
    const double *wp = res ? shd + (size_t)soff * GZ_ENTS : B.ent + (size_t)sc.z * GZ_ENT;   // the world's entries
    const double *wdl = res ? wp + 20 : B.ent_dl + 2 * (size_t)sc.z;
    const int dstr = res ? GZ_ENTS : 2;
    const int iters = B.p.iters, total = B.p.iters + B.p.friction_iters;
    Row *rw = B.row + wi.row0;
    for (int q = lw; q < nent; q += LW) {
      const int r = B.ent_row[sc.z + q];
      if (r < 0) continue;   // the row's body-1 entry carries it
      const double ad2 = B.ent_ad2[sc.z + q];
      const double *p = wp + (size_t)q * lstr;
      rw[r].lambda = p[13]; rw[r].lambda_erp = p[15];
      const int cls = (__double2loint(p[19]) >> 27) & 3;
      if ((cls == 2 ? total : iters) <= 0) continue;
      const double de = wdl[(size_t)q * dstr], dd = de * de, e2 = dd * ad2;
      if (cls == 0) { st_acc[0] += dd; st_err[0] += e2; st_cnt[0]++; }
      else if (cls == 1) { st_acc[1] += dd; st_err[1] += e2; st_cnt[1]++; }
      else { st_acc[2] += dd; st_err[2] += e2; st_cnt[2]++; }
    }
  }
  // ---- statistics: reduce over the lanes of the world
  if (MULTI) {
    __syncthreads();
    double *red = shd;   // the staged entries are dead now
    for (int q = 0; q < 3; q++) {
      red[(size_t)(3 * q + 0) * LW + tid] = st_acc[q];
      red[(size_t)(3 * q + 1) * LW + tid] = st_err[q];
      red[(size_t)(3 * q + 2) * LW + tid] = (double)st_cnt[q];
    }
    __syncthreads();
    if (tid < 9) {
      double s = 0;
      for (int k = 0; k < LW; k++) s += red[(size_t)tid * LW + k];
      red[(size_t)tid * LW] = s;
    }
    __syncthreads();
    for (int q = 0; q < 3; q++) {
      st_acc[q] = red[(size_t)(3 * q + 0) * LW]; st_err[q] = red[(size_t)(3 * q + 1) * LW];
      st_cnt[q] = (int)red[(size_t)(3 * q + 2) * LW];
    }
  } else {
    for (int dd = 1; dd < LW; dd <<= 1) {
#pragma unroll
      for (int q = 0; q < 3; q++) {
        st_acc[q] += __shfl_xor_sync(0xffffffffu, st_acc[q], dd);
        st_err[q] += __shfl_xor_sync(0xffffffffu, st_err[q], dd);
        st_cnt[q] += __shfl_xor_sync(0xffffffffu, st_cnt[q], dd);
      }
    }
  }
  if (live_w && lw == 0) {
    WorldStep &ws = B.wstep[w];
    const double c[3] = {(double)st_cnt[0], (double)st_cnt[1], (double)st_cnt[2]};
    const double tot = c[0] + c[1] + c[2];   // quickstep_pgs_lcp.cpp:814-855
    for (int q = 0; q < 3; q++) {
      ws.rms_dlambda[q] = sqrt(c[q] > 0 ? st_acc[q] / c[q] : 0.0);
      ws.rms_residual[q] = sqrt(c[q] > 0 ? st_err[q] / c[q] : 0.0);
    }
    const double sa = st_acc[0] + st_acc[1] + st_acc[2], se = st_err[0] + st_err[1] + st_err[2];
    ws.rms_dlambda[3] = sqrt(sa > 0 ? sa / tot : 0.0);
    ws.rms_residual[3] = sqrt(se > 0 ? se / tot : 0.0);
    ws.num_contacts_stat = st_cnt[1];
    ws.pad[0] = npass; ws.pad[1] = R * 256 + (res ? 1 : 0);   // instrumentation: rounds, period, staged or not
    B.prof[16 + 2 * w] = g0; B.prof[16 + 2 * w + 1] = (pgsb_gtimer() & ~511ull) | pgsb_smid();   // timeline, low 9 bits: the SM
  }
}


// The launch.  Multi-warp CTAs (worlds of more than 16 bodies): one CTA per world, entries staged if they fit.
// One-warp CTAs: ONE LAUNCH PER SHARED-MEMORY CLASS (ascending budgets cfg.nres; nres_below = the budget of the
// class below, -1 for the first), each a set of PERSISTENT warps that take the class's work items - a contiguous
// range of the sweep order, heaviest first - from a counter until none is left: no CTA is launched only to find
// that it belongs to another class, and the tail of a class is shared out dynamically.
template <bool MULTI, bool BOX>
__global__ void __launch_bounds__(MULTI ? 512 : 32) k_pgsb(DevBatch B, PgsbCfg cfg, int nres_below) {
  if (MULTI) {
    const int need = blockIdx.x < B.n_worlds ? B.wsb[B.wperm[blockIdx.x]].w : 0;
    pgsb_item<true, BOX>(B, cfg, blockIdx.x, need <= cfg.nres);
    return;
  }
  const int n_items = (B.n_worlds + cfg.wpc - 1) / cfg.wpc;
  // class range: items that fit this class and not the one below; the last class also takes what fits nowhere
  int lo = 0, hi = 0;
  if (threadIdx.x == 0) {
    lo = cfg.last ? 0 : pgsb_first_fitting(B, cfg, n_items, cfg.nres);
    hi = nres_below >= 0 ? pgsb_first_fitting(B, cfg, n_items, nres_below) : n_items;
  }
  lo = __shfl_sync(0xffffffffu, lo, 0);
  hi = __shfl_sync(0xffffffffu, hi, 0);
  for (;;) {
    int item = 0;
    if (threadIdx.x == 0) item = lo + atomicAdd(&B.cls_ctr[cfg.cls], 1);
    item = __shfl_sync(0xffffffffu, item, 0);
    if (item >= hi) break;
    pgsb_item<false, BOX>(B, cfg, item, pgsb_need(B, cfg, item) <= cfg.nres);
    __syncwarp();   // the next item overwrites the staged entries
  }
}

}  // namespace gz
