// k_pgsc.cuh — the dataflow sweeps of a large world (k_pgsd, k_pgsl.cuh) with the exchange of two-body rows PUSHED into
// the receiver's shared memory wherever the two bodies sit in the same thread-block cluster.
//
// k_pgsd measured: a lane that waits for its partner's half of a dot product polls a slot in L2, ~540 cycles a poll,
// ~1460 cycles from one publication on the critical chain to the next; the chain (the rows of a pile that share bodies
// one after the other, 60 sweeps) is what the kernel's duration is.  Here every entry (row, body) has an INBOX: the
// partner's lane stores its half there.  The inbox of an entry lives in the shared memory of the CTA that owns the
// entry's body when the partner's CTA is in the same cluster (st.shared::cluster through DSMEM, then a ~30-cycle local
// poll); otherwise in global memory as in k_pgsd.  Sender and receiver evaluate the same predicate, so they agree on
// the place.  CTA = 64 or 128 bodies x 2 solves (few warps per SM: a warp that polls shared memory takes issue slots
// from the warps that have rows to run), clusters of 8 CTAs (16 where the device allows it).
//
// Slot format, tags, double buffering by sweep parity and the progress argument are k_pgsd's (k_pgsl.cuh); shared
// memory is zeroed (tag 0 = never written) before the cluster's first barrier, and a CTA leaves only after the
// cluster's last barrier (its peers may store into its shared memory until then).
// All CTAs must be resident (a lane polls for values other CTAs produce): the host launches at most
// cudaOccupancyMaxActiveClusters clusters and nothing else runs beside the kernel on the batch's streams.
#pragma once
#include <cooperative_groups.h>

namespace gz {

__device__ __forceinline__ void inbox_put_smem(unsigned addr, double v, unsigned tag) {
  const unsigned long long t = (unsigned long long)tag << 32;
  const unsigned long long w0 = t | (unsigned)__double2loint(v), w1 = t | (unsigned)__double2hiint(v);
  asm volatile("st.relaxed.cluster.shared::cluster.v2.u64 [%0], {%1, %2};" ::"r"(addr), "l"(w0), "l"(w1) : "memory");
}
__device__ __forceinline__ bool inbox_get_smem(unsigned addr, unsigned tag, double &v) {
  unsigned long long w0, w1;
  asm volatile("ld.relaxed.cluster.shared::cta.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "r"(addr) : "memory");
  v = __hiloint2double((int)(unsigned)w1, (int)(unsigned)w0);
  return (unsigned)(w0 >> 32) == tag && (unsigned)(w1 >> 32) == tag;
}

// inbox of entry e for solve S, sweep parity 0: byte offset in the owner CTA's shared memory / word index in global memory
__device__ __forceinline__ unsigned inbox_smem_off(int rel, int S) { return (unsigned)(((rel * 2 + S) * 2) * 16); }
__device__ __forceinline__ size_t inbox_glob_idx(int e, int S) { return 2 * (((size_t)e * 2 + S) * 2); }

template <bool BOX>
__global__ void __launch_bounds__(256, 2) k_pgsc(DevBatch B, const int w, unsigned long long *gslots, const int *cta_e0,
                                                 const unsigned *epoch_ctr, const int smem_entries, const int glob_mask, const int sleep_ns) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  extern __shared__ __align__(16) unsigned long long inbox[];   // [entries of this CTA][solve][parity][2 words]
  __shared__ int s_e0[GZ_PGSC_TAB];
  const unsigned epoch = *epoch_ctr;   // (advanced by k_pgs_large_stats, the next kernel on the stream)
  const int tid = threadIdx.x, cta = blockIdx.x, ncta = gridDim.x;
  const int CL = (int)cluster.dim_blocks().x;   // a power of two
  const int cl_shift = 31 - __clz(CL);
  for (int q = tid; q <= ncta; q += blockDim.x) s_e0[q] = cta_e0[q];
  __syncthreads();
  const int my_lo = s_e0[cta], my_n = s_e0[cta + 1] - my_lo;
  const bool my_smem = my_n <= smem_entries;
  if (my_smem)
    for (int q = tid; q < my_n * 8; q += blockDim.x) inbox[q] = 0ull;
  cluster.sync();
  const unsigned inbox_base = (unsigned)__cvta_generic_to_shared(inbox);
  // lane -> (body, solve): warp k of the CTA owns bodies [k * bpw, (k + 1) * bpw) of the CTA's, two lanes each
  const int slot = (tid & 31) >> 1, S = tid & 1;
  const int bl = (slot < B.pgsc_bpw) ? cta * B.pgsc_bodies + (tid >> 5) * B.pgsc_bpw + slot : 0x7fffffff;
  const WorldInfo wi = B.world[w];
  const int4 sc = B.wsb[w];
  const int nent = sc.w;
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
  const unsigned tag0 = epoch * (unsigned)(total + 1) + 1u;
  int i = (iters > 0) ? 0 : fric0, sweep = 0;
  bool alive = i < cnt && total > 0;
  // Three rows per lane are in registers: the row in flight (what finishing it needs, loaded one pass ago), the NEXT row
  // (its Jacobian half and the places of its inboxes, decoded: its dot product goes out right after the accumulator
  // update) and the row AFTER that (Jacobian half and partner record, still arriving).  Nothing a pass loads is used
  // in the same pass, so a pass never waits for memory it has just asked for - the whole warp would wait with it, the
  // lanes whose partner's value has just arrived included.
  double iM[6] = {0, 0, 0, 0, 0, 0}, Jn[6] = {0, 0, 0, 0, 0, 0}, J2[6] = {0, 0, 0, 0, 0, 0};
  double d = 0.0, r_rhs = 0, r_lam = 0, r_adcfm = 0, r_lo = 0, r_hi = 0, r_nlam = 1.0;
  int r_mx = 0;
  int ni = 0, nsweep = 0, n2i = 0, n2sweep = 0, back_n = 0;
  // where the partner's half arrives (recv) and where this lane's half goes (send): mode 0 none (one-body row),
  // 1 shared memory (recv: byte address in this CTA; send: shared::cluster address), 2 global memory (word index)
  int rmode = 0, rmode_n = 0, smode_n = 0;
  unsigned long long raddr = 0, raddr_n = 0, saddr_n = 0;
  int4 aux2 = make_int4(-1, 0, 0, 0);
  bool has_next = false, has2 = false;
  const int4 *const aux = B.ent_aux + e0;
  auto step_to = [&](int q, int sw, int &nq, int &nsw) {   // the lane's row after (q, sw); false when it was the last
    nq = q + 1; nsw = sw;
    if (nq == cnt) { nsw = sw + 1; nq = nsw < iters ? 0 : fric0; }
    return nsw < total && nq < cnt;
  };
  auto issue_J = [&](int q) {   // loads only
    const double2 *p2 = reinterpret_cast<const double2 *>(lp + (size_t)q * GZ_ENT);
    const double2 v0 = p2[0], v1 = p2[1], v2 = p2[2];
    aux2 = aux[q];
    J2[0] = v0.x; J2[1] = v0.y; J2[2] = v1.x; J2[3] = v1.y; J2[4] = v2.x; J2[5] = v2.y;
  };
  auto decode_J = [&](int q) {   // (J2, aux2) of list place q become the next row
#pragma unroll
    for (int t = 0; t < 6; t++) Jn[t] = J2[t];
    const int pen = aux2.x, pcta = aux2.y;
    back_n = aux2.z;
    rmode_n = 0; smode_n = 0;
    if (pen >= 0) {
      const bool same = (pcta >> cl_shift) == (cta >> cl_shift);
      const int p_lo = s_e0[pcta], p_n = s_e0[pcta + 1] - p_lo;
      if (same && p_n <= smem_entries) {
        unsigned remote;
        const unsigned local = inbox_base + inbox_smem_off(pen - p_lo, S);
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(pcta & (CL - 1)));
        smode_n = 1; saddr_n = remote;
      } else { smode_n = 2; saddr_n = inbox_glob_idx(pen, S); }
      if (same && my_smem) { rmode_n = 1; raddr_n = inbox_base + inbox_smem_off(e0 + q - my_lo, S); }
      else { rmode_n = 2; raddr_n = inbox_glob_idx(e0 + q, S); }
    }
  };
  auto publish_next = [&](double v, int sw) {
    if (smode_n == 1) inbox_put_smem((unsigned)saddr_n + 16u * (sw & 1), v, tag0 + (unsigned)sw);
    else if (smode_n == 2) slot_put(gslots + saddr_n + 2 * (sw & 1), v, tag0 + (unsigned)sw);
  };
  auto issue_rest = [&](int q, int back) {   // loads only; the normal row's lambda sits `back` entries before (0: none)
    const double *p = lp + (size_t)q * GZ_ENT;
    const double2 *p2 = reinterpret_cast<const double2 *>(p);
    const double2 v3 = p2[3], v4 = p2[4], v5 = p2[5], vs = p2[6 + S], v8 = p2[8], v9 = p2[9];
    r_nlam = (p - (size_t)back * GZ_ENT)[13 + 2 * S];
    iM[0] = v3.x; iM[1] = v3.y; iM[2] = v4.x; iM[3] = v4.y; iM[4] = v5.x; iM[5] = v5.y;
    r_rhs = vs.x; r_lam = vs.y; r_adcfm = v8.x; r_lo = v8.y; r_hi = v9.x; r_mx = __double2loint(v9.y);
  };
  if (alive) {
    issue_J(i); decode_J(i);
    d = a[0] * Jn[0];
#pragma unroll
    for (int t = 1; t < 6; t++) d += a[t] * Jn[t];
    publish_next(d, sweep);
    rmode = rmode_n; raddr = raddr_n;
    issue_rest(i, back_n);
    has_next = step_to(i, sweep, ni, nsweep);
    if (has_next) {
      issue_J(ni); decode_J(ni);
      has2 = step_to(ni, nsweep, n2i, n2sweep);
      if (has2) issue_J(n2i);
    }
  }
  int n_outer = 0, n_spin = 0;   // developer counters: passes of this warp through the row code, polls
  const long long t_begin = clock64();
  while (__any_sync(0xffffffffu, alive)) {
    double pd = 0.0;
    bool ready;
    n_outer++;
    int spin = 0;
    do {
      n_spin++;
      if (sleep_ns > 0 && spin > 0) __nanosleep(sleep_ns);   // a spinning warp takes issue slots from the warps that have rows to run
      ready = alive && rmode == 0;
      if (alive && rmode == 1) ready = inbox_get_smem((unsigned)raddr + 16u * (sweep & 1), tag0 + (unsigned)sweep, pd);
      // lanes that wait on global memory look there every (glob_mask + 1)-th poll only: a trip to L2 holds the whole
      // warp, the lanes polling shared memory included
      if (alive && rmode == 2 && (spin & glob_mask) == 0) ready = slot_get(gslots + raddr + 2 * (sweep & 1), tag0 + (unsigned)sweep, pd);
      spin++;
    } while (!__any_sync(0xffffffffu, ready));
    if (ready) {
      const int cls = (r_mx >> 27) & 3;
      const bool isb2 = (r_mx >> 29) & 1;
      // pyramid friction: limits +-|mu * lambda_normal| (quickstep_pgs_lcp.cpp:572-582); box: +-mu (:588-594)
      const double nlam = (cls == 2 && (!BOX || ((r_mx >> 30) & 1))) ? r_nlam : 1.0;   // (bit 30: a torsional row)
      const double old_lambda = r_lam;
      double delta = r_rhs - old_lambda * r_adcfm;
      const double hl = r_hi * nlam;
      const double hf = fabs(hl);
      const double hi_act = cls == 2 ? hf : r_hi, lo_act = cls == 2 ? -hf : r_lo;
      const double d1 = isb2 ? pd : d, d2 = isb2 ? d : pd;
      delta -= d1;
      const double delta2 = delta - d2;
      delta = rmode != 0 ? delta2 : delta;
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
        publish_next(dn, nsweep);
      }
      // SMOOTH_LAMBDA on the lambda solve of non-bilateral rows; delta is not recomputed after it (:687-691 vs :706)
      const double sm = smooth1 * nl + smooth * old_lambda;
      lp[(size_t)i * GZ_ENT + 13 + 2 * S] = (S == 0 && cls != 0) ? sm : nl;
      dl[2 * (size_t)i] = delta;
      i = ni; sweep = nsweep; alive = has_next; d = dn; rmode = rmode_n; raddr = raddr_n;
      if (alive) {
        issue_rest(i, back_n);
        has_next = has2; ni = n2i; nsweep = n2sweep;
        if (has_next) {
          decode_J(ni);   // (loaded a pass ago)
          has2 = step_to(ni, nsweep, n2i, n2sweep);
          if (has2) issue_J(n2i);
        }
      }
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
  cluster.sync();   // peers may store into this CTA's shared memory until they are done
}

// After k_pgsc (stream order): lambda back to the rows and the statistics of the last update of every row
// (quickstep_pgs_lcp.cpp:700-760, 814-855), over all entries of the world with all lanes.  Deterministic: lanes of a
// warp, warps of a CTA in order, and the CTA that arrives last adds the CTAs' partial sums in index order.
__global__ void __launch_bounds__(256) k_pgs_large_stats(DevBatch B, const int w, double *partial, unsigned *ticket, unsigned *epoch_ctr, const int *only_if) {
  if (only_if && !*only_if) return;   // (k_pgsd has run instead of k_pgsm and has done this itself)
  const int gl = blockIdx.x * blockDim.x + threadIdx.x, L = gridDim.x * blockDim.x;
  const WorldInfo wi = B.world[w];
  const int4 sc = B.wsb[w];
  const int nent = sc.w, iters = B.p.iters, total = B.p.iters + B.p.friction_iters;
  double st[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};   // acc[3], err[3], cnt[3]
  {
    const double *wp = B.ent + (size_t)sc.z * GZ_ENT;
    const double *wdl = B.ent_dl + 2 * (size_t)sc.z;
    Row *rw = B.row + wi.row0;
    for (int q = gl; q < nent; q += L) {
      const int r = B.ent_row[sc.z + q];
      if (r < 0) continue;   // the row's body-1 entry carries it
      const double *pq = wp + (size_t)q * GZ_ENT;
      rw[r].lambda = pq[13]; rw[r].lambda_erp = pq[15];
      const int cls = (__double2loint(pq[19]) >> 27) & 3;
      if ((cls == 2 ? total : iters) <= 0) continue;
      const double de = wdl[2 * (size_t)q], dd = de * de, e2 = dd * B.ent_ad2[sc.z + q];
      if (cls == 0) { st[0] += dd; st[3] += e2; st[6] += 1.0; }
      else if (cls == 1) { st[1] += dd; st[4] += e2; st[7] += 1.0; }
      else { st[2] += dd; st[5] += e2; st[8] += 1.0; }
    }
  }
  __shared__ double red[8][9];
  __shared__ bool s_last;
  for (int dd = 1; dd < 32; dd <<= 1)
#pragma unroll
    for (int q = 0; q < 9; q++) st[q] += __shfl_xor_sync(0xffffffffu, st[q], dd);
  if ((threadIdx.x & 31) == 0)
#pragma unroll
    for (int q = 0; q < 9; q++) red[threadIdx.x >> 5][q] = st[q];
  __syncthreads();
  if (threadIdx.x < 9) {
    double s = 0;
    for (int k = 0; k < (int)(blockDim.x >> 5); k++) s += red[k][threadIdx.x];
    partial[9 * (size_t)blockIdx.x + threadIdx.x] = s;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    __threadfence();
    double t[9];
#pragma unroll
    for (int q = 0; q < 9; q++) { double s = 0; for (int k = 0; k < (int)gridDim.x; k++) s += __ldcg(&partial[9 * (size_t)k + q]); t[q] = s; }
    WorldStep &ws = B.wstep[w];
    const double tot = t[6] + t[7] + t[8];
#pragma unroll
    for (int q = 0; q < 3; q++) {
      ws.rms_dlambda[q] = sqrt(t[6 + q] > 0 ? t[q] / t[6 + q] : 0.0);
      ws.rms_residual[q] = sqrt(t[6 + q] > 0 ? t[3 + q] / t[6 + q] : 0.0);
    }
    const double sa = t[0] + t[1] + t[2], se = t[3] + t[4] + t[5];
    ws.rms_dlambda[3] = sqrt(sa > 0 ? sa / tot : 0.0);
    ws.rms_residual[3] = sqrt(se > 0 ? se / tot : 0.0);
    ws.num_contacts_stat = (int)t[7];
    *ticket = 0u;
    *epoch_ctr += 1u;
  }
}

}  // namespace gz
