// k_pgsm.cuh — the sweeps of ONE world of a few hundred bodies (C1: 500 boxes, one settled pile = one island of ~340
// bodies) on ONE thread-block cluster, everything on chip.
//
// Such a world is a dependency chain (C1: ~3 100 hops per step) with next to no parallel work; what counts is the cost
// of a hop.  Over the whole GPU (k_pgsd) a hop is ~1 460 cycles (store -> L2 -> poll -> row code with its entry coming
// from L1 / L2).  Here: lane = (body, solve) as everywhere; the entries of a CTA's bodies are staged in its shared memory
// once per step and stay for the 60 sweeps; the sweeps are k_pgsd's dataflow (no rounds, no schedule), but the partial
// dot product of a two-body row is PUSHED into an inbox in the partner CTA's shared memory (st.shared::cluster on a
// mapa address, tagged {32 bits of the double, sweep + 1} words as in k_pgsd) and every lane polls its own CTA only.
// First version, measured at C1: rounds (k_pgsl's modulo schedule) with the hardware cluster barrier: 1 070 cycles a
// round - a cluster barrier costs 400 cycles, 680 behind a DSMEM store (scripts/micro/clbar.cu) - 2.0 ms for the sweeps
// plus 0.2 ms for the round offsets, against 2.6 ms for k_pgsd.
// Worlds whose entries do not fit the cluster's shared memory leave the flag down and k_pgsd runs instead.
#pragma once
#include <cooperative_groups.h>

namespace gz {

#define GZ_PGSM_STRIDE 22   // doubles per staged entry: 20 + the last delta of the two solves (statistics)
#define GZ_PGSM_BYTES (GZ_PGSM_STRIDE * 8 + 64 + 4)   // shared memory per entry: the entry, its inbox (2 solves x 2 sweep parities x 16 B), its partner entry

template <bool BOX>
__global__ void __launch_bounds__(256, 1) k_pgsm(DevBatch B, const int w, const int *cta_e0, const int smem_entries, int *done_flag) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  extern __shared__ __align__(16) unsigned long long sm_raw[];
  // [inbox: smem_entries x 8 words (the same place in every CTA)][entries: x GZ_PGSM_STRIDE doubles][partner entry: x int]
  unsigned long long *const inbox = sm_raw;
  double *const sm_ent = reinterpret_cast<double *>(sm_raw + (size_t)smem_entries * 8);
  int *const sm_part = reinterpret_cast<int *>(sm_ent + (size_t)smem_entries * GZ_PGSM_STRIDE);
  __shared__ int s_e0[20];
  const int tid = threadIdx.x, cta = blockIdx.x, ncta = gridDim.x, nthr = blockDim.x;
  if (tid <= ncta) s_e0[tid] = cta_e0[tid];
  __syncthreads();
  const int4 sc = B.wsb[w];
  const int nent = sc.w;
  bool fits = nent > 0;
  for (int c = 0; c < ncta; c++) fits = fits && (s_e0[c + 1] - s_e0[c]) <= smem_entries;
  if (!fits) {   // (the same answer in every thread of the cluster: nobody waits at a barrier)
    if (cta == 0 && tid == 0) *done_flag = 0;
    return;
  }
  const int my_lo = s_e0[cta], my_n = s_e0[cta + 1] - my_lo;
  for (int q = tid; q < my_n * (GZ_ENT / 2); q += nthr) {   // stage: 10 double2 per entry
    const int e = q / (GZ_ENT / 2), k = q % (GZ_ENT / 2);
    reinterpret_cast<double2 *>(sm_ent + (size_t)e * GZ_PGSM_STRIDE)[k] = reinterpret_cast<const double2 *>(B.ent + (size_t)(my_lo + e) * GZ_ENT)[k];
  }
  for (int q = tid; q < my_n; q += nthr) {
    sm_ent[(size_t)q * GZ_PGSM_STRIDE + 20] = 0.0; sm_ent[(size_t)q * GZ_PGSM_STRIDE + 21] = 0.0;
    sm_part[q] = B.ent_part[my_lo + q];
  }
  for (int q = tid; q < my_n * 8; q += nthr) inbox[q] = 0ull;   // tag 0 = never written
  const int gl = cta * nthr + tid;
  const int bl = gl >> 1, S = gl & 1;
  const WorldInfo wi = B.world[w];
  int cnt = 0, e0 = my_lo, fric0 = 0;
  if (bl < wi.nb) { const int4 bs = B.bsched[wi.body0 + bl]; e0 = bs.x; cnt = bs.y; fric0 = bs.z; }
  const int rel0 = e0 - my_lo;
  double *const lp = sm_ent + (size_t)rel0 * GZ_PGSM_STRIDE;
  const unsigned inbox_base = (unsigned)__cvta_generic_to_shared(inbox);
  const int nthr_shift = 31 - __clz(nthr);   // (CTAs of 64, 128 or 256 threads)
  cluster.sync();   // entries staged, inboxes cleared; everybody's shared memory is live
  // initial caccel = iMJ * lambda0 over the warm-started (persistent-joint) rows (multiply_invM_JT, quickstep_util.cpp:101-119)
  double a[6] = {0, 0, 0, 0, 0, 0};
  if (B.p.warm_start > 0) {
    for (int q = 0; q < cnt; q++) {
      const double *p = lp + (size_t)q * GZ_PGSM_STRIDE;
      if (((__double2loint(p[19]) >> 27) & 3) != 0) break;
      const double l0 = p[13 + 2 * S];
#pragma unroll
      for (int t = 0; t < 6; t++) a[t] += p[6 + t] * l0;
    }
  }
  const int iters = B.p.iters, total = B.p.iters + B.p.friction_iters;
  const double smooth = B.p.smooth_contacts, smooth1 = 1.0 - smooth;
  int i = (iters > 0) ? 0 : fric0, sweep = 0;
  bool alive = i < cnt && total > 0;
  // the row in flight (everything that does not wait for the partner is computed when the row is taken up)
  double iM[6] = {0, 0, 0, 0, 0, 0};
  double d = 0.0, delta0 = 0, old_lambda = 0, lo_act = 0, hi_act = 0;
  int cls = 0;
  unsigned raddr = 0;   // this row's inbox (shared::cta address, parity 0); 0 = a one-body row
  bool isb2 = false;
  auto take_up = [&]() {   // row i of this sweep: load, this body's half of the dot product out to the partner, limits
    const double *p = lp + (size_t)i * GZ_PGSM_STRIDE;
    const double2 *p2 = reinterpret_cast<const double2 *>(p);
    const double2 v0 = p2[0], v1 = p2[1], v2 = p2[2], v3 = p2[3], v4 = p2[4], v5 = p2[5], vs = p2[6 + S], v8 = p2[8], v9 = p2[9];
    const int pen = sm_part[rel0 + i];
    const int mx = __double2loint(v9.y), my = __double2hiint(v9.y);
    // this body's half, term by term in the reference's left-to-right order
    d = a[0] * v0.x;
    d += a[1] * v0.y; d += a[2] * v1.x; d += a[3] * v1.y; d += a[4] * v2.x; d += a[5] * v2.y;
    raddr = 0;
    if (pen >= 0) {
      const int pcta = (2 * (my & 0xFFFF) + S) >> nthr_shift;
      unsigned remote;
      const unsigned local = inbox_base + inbox_smem_off(pen - s_e0[pcta], S) + 16u * (sweep & 1);
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(pcta));
      inbox_put_smem(remote, d, (unsigned)sweep + 1u);
      raddr = inbox_base + inbox_smem_off(rel0 + i, S);
    }
    iM[0] = v3.x; iM[1] = v3.y; iM[2] = v4.x; iM[3] = v4.y; iM[4] = v5.x; iM[5] = v5.y;
    cls = (mx >> 27) & 3; isb2 = (mx >> 29) & 1;
    // pyramid friction: limits +-|mu * lambda_normal| (quickstep_pgs_lcp.cpp:572-582); box: +-mu (:588-594)
    double nlam = 1.0;
    if (cls == 2 && (!BOX || ((mx >> 30) & 1))) nlam = (p - (size_t)((unsigned)my >> 16) * GZ_PGSM_STRIDE)[13 + 2 * S];
    old_lambda = vs.y;
    delta0 = vs.x - old_lambda * v8.x;
    const double hl = v9.x * nlam;
    const double hf = fabs(hl);
    hi_act = cls == 2 ? hf : v9.x; lo_act = cls == 2 ? -hf : v8.y;
  };
  if (alive) take_up();
  int n_outer = 0, n_spin = 0;   // developer counters: passes of this warp through the row code, polls
  const long long t_begin = clock64();
  while (__any_sync(0xffffffffu, alive)) {
    double pd = 0.0;
    bool ready;
    n_outer++;
    do {
      n_spin++;
      ready = alive;
      if (alive && raddr != 0) ready = inbox_get_smem(raddr + 16u * (sweep & 1), (unsigned)sweep + 1u, pd);
    } while (!__any_sync(0xffffffffu, ready));
    if (ready) {
      double delta = delta0;
      const double d1 = isb2 ? pd : d, d2 = isb2 ? d : pd;
      delta -= d1;
      const double delta2 = delta - d2;
      delta = raddr != 0 ? delta2 : delta;
      double nl = old_lambda + delta;
      const double dlo = lo_act - old_lambda, dhi = hi_act - old_lambda;
      const bool below = nl < lo_act, above = !below && nl > hi_act;
      delta = below ? dlo : (above ? dhi : delta);
      nl = below ? lo_act : (above ? hi_act : nl);
#pragma unroll
      for (int t = 0; t < 6; t++) a[t] += delta * iM[t];
      // SMOOTH_LAMBDA on the lambda solve of non-bilateral rows; delta is not recomputed after it (:687-691 vs :706)
      const double sm = smooth1 * nl + smooth * old_lambda;
      double *p = lp + (size_t)i * GZ_PGSM_STRIDE;
      p[13 + 2 * S] = (S == 0 && cls != 0) ? sm : nl;
      p[20 + S] = delta;
      if (++i == cnt) { sweep++; i = sweep < iters ? 0 : fric0; }
      alive = sweep < total && i < cnt;
      if (alive) take_up();   // (shared memory: the next row's half goes out a few dozen cycles after the update)
    }
  }
  if ((tid & 31) == 0) {   // gzcuda_batch_get_pgs_debug: the longest warp's passes, polls, cycles / 16
    WorldStep &wsd = B.wstep[w];
    atomicMax(&wsd.pad[0], n_outer); atomicMax(&wsd.pad[1], n_spin); atomicMax(&wsd.n_fric, (int)((clock64() - t_begin) >> 4));
  }
  if (bl < wi.nb) {
    BodyStep &bs = B.bstep[wi.body0 + bl];
    double *o = S ? bs.caccel_erp : bs.caccel;
#pragma unroll
    for (int t = 0; t < 6; t++) o[t] = a[t];
  }
  __syncthreads();
  // lambda and the last deltas back to the pool: k_pgs_large_stats reads them there
  for (int q = tid; q < my_n; q += nthr) {
    const double *p = sm_ent + (size_t)q * GZ_PGSM_STRIDE;
    double *g = B.ent + (size_t)(my_lo + q) * GZ_ENT;
    g[13] = p[13]; g[15] = p[15];
    B.ent_dl[2 * (size_t)(my_lo + q)] = p[20]; B.ent_dl[2 * (size_t)(my_lo + q) + 1] = p[21];
  }
  if (cta == 0 && tid == 0) *done_flag = 1;
  cluster.sync();   // peers may store into this CTA's shared memory until they are done
}

}  // namespace gz
