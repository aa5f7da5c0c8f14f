// k_pgs.cuh — row prologue, static modulo schedule, the SOR-PGS sweeps, and integration.
//
//   k_schedule   order[] of PGS_LCP with row_reorder1 (quickstep_pgs_lcp.cpp:1115-1150) and the
//                static modulo schedule described below
//   k_slab       lays the step's PGS warps and their record slabs out (prefix over the buckets)
//   k_rowfinish  rhs / rhs_erp (quickstep.cpp:403-437), cfm /= h (:443), warm start (:473-495),
//                iMJ = M^-1 J^T (quickstep_util.cpp:79-98), Ad and the Ad scaling
//                (quickstep_pgs_lcp.cpp:1012-1028,1091-1107); writes the packed records
//   k_pgs        initial caccel = iMJ * lambda0 (multiply_invM_JT quickstep_util.cpp:101-119),
//                then ComputeRows (quickstep_pgs_lcp.cpp:52-954), k_pgs_sweep.cuh
//   k_integrate  lambda write-back (quickstep.cpp:523-541), dxUpdateBodies
//                (quickstep_update_bodies.cpp:41-368), dxStepBody (util.cpp:274-379)
//
// Exact-order parallel Gauss-Seidel.  The reference sweeps rows strictly in order[]
// sequence.  Two row updates commute EXACTLY when they share no body, so any schedule
// that keeps, for every body, the order in which rows touch it gives bit-identical
// lambda and caccel.  Every sweep has the same dependency structure, so the schedule is
// built once per step as a MODULO schedule (software pipelining of the sweep loop): row r
// gets an offset off(r); in sweep s it runs in round s*R + off(r), on lane (off(r) mod R, j).
// Constraints: consecutive rows on a body have increasing offsets, the first row of a body in
// sweep s+1 comes after its last row of sweep s (last - first + 1 <= R), and at most k rows
// share a round residue.  k (lanes per solve) is picked per world from its row count, so a
// warp holds 16/k worlds and every lane has a row in (almost) every round.  The lambda and
// lambda_erp solves are independent problems (split impulse, quickstep_pgs_lcp.cpp:526-542)
// and run on adjacent lanes that share the row record.
//
// Worlds are grouped into BUCKETS of equal (k, R) — R is drawn from a geometric ladder — and
// a PGS warp takes 16/k worlds of one bucket.  Because they share R, the records the warp needs
// in one round are stored contiguously (slot-major, then lane pair): one 4352-byte bulk copy
// per round feeds the whole warp.
#pragma once
#include "k_rows2.cuh"
#include "k_pgsb.cuh"

namespace gz {

// per-world int scratch (B.scratch): bodies of the rows in order[] sequence, then the slot and
// stage the scheduler gave each position
struct WScratch { int *rb1, *rb2, *slot, *stage, *it_pos, *it_len, *runoff; };
__device__ inline WScratch wscratch(const DevBatch &B, int w) {
  WScratch s;
  const size_t n = B.max_row_cap;
  int *base = B.scratch + (size_t)w * 7 * n;
  s.rb1 = base; s.rb2 = base + n; s.slot = base + 2 * n; s.stage = base + 3 * n;
  s.it_pos = base + 4 * n; s.it_len = base + 5 * n; s.runoff = base + 6 * n;
  return s;
}

// ---- packed row record: what the sweeps stream, 34 doubles = 272 B ---------------------------
// [0..11] J (scaled by Ad)   [12..23] iMJ   [24] rhs  [25] rhs_erp  [26] Adcfm  [27] lo  [28] hi
// [29] 1/Ad^2 (statistics)  [30..31] RecInts (16 bytes, 16-byte aligned)  [32..33] padding: the
// 272-byte stride puts the 16 records a warp reads per round in distinct shared-memory banks.
// A warp's slab holds R rounds of 16 records: record of (slot s, lane pair p) at (s * 16 + p).
// A record is live only if its epoch equals the current step's.  Epochs run 1 .. 2^24 - 1 (0 is what a never-written slot
// holds); when they start over the host clears the record pool (launch_physics), so a record of 2^24 - 1 steps ago cannot
// pass for a live one.
#define GZ_REC 34
#define GZ_REC_INTS 30
#define GZ_ROUND_DOUBLES (16 * GZ_REC)
#define GZ_FI_EMPTY (-3)
#define GZ_FI_RUN (-4)     // fi <= GZ_FI_RUN - 1: head of a run of (GZ_FI_RUN - fi) rows (chain mode)
#define GZ_CHAIN_K 16      // lane pairs of a chain-mode world: it has a warp to itself
#define GZ_CHAIN_FIXED_DOUBLES (32 * 10 + 2)   // k_pgs_chain shared memory: run statistics, mbarrier (padded)
struct RecInts {
  int fi;                 // -1 bilateral, -2 contact normal, >= 0 world-local slot index (s * k + j) of the normal row
  unsigned short b1, b2;  // world-local bodies, b2 = max_nb (the dummy accumulator slot): none
  int stage_epoch;        // (step epoch << 8) | off(r) / R
  int orig;               // world-local row index (write-back)
};
struct WSched { int R, k, nstage, bucket, idx, nslots, nrun, pad1; };

// ---- buckets: (chain mode, c = log2(k) in 0..4, rb = rung of the R ladder); id = (chain * 5 + c) * GZ_NRUNG + rb;
// for chain-mode worlds k = GZ_CHAIN_K and c is the shared-memory size class (0 .. GZ_CHAIN_NCLS - 1, ascending) ----
#define GZ_RMAX 9984   // rung_value(GZ_NRUNG - 1); GZ_NRUNG / GZ_NBUCKET live in gz_batch.h
__host__ __device__ inline int rung_value(int rb) {
  if (rb <= 32) return rb;
  if (rb <= 48) return 32 + 2 * (rb - 32);
  if (rb <= 64) return 64 + 4 * (rb - 48);
  if (rb <= 80) return 128 + 8 * (rb - 64);
  if (rb <= 96) return 256 + 16 * (rb - 80);
  if (rb <= 112) return 512 + 32 * (rb - 96);
  if (rb <= 128) return 1024 + 64 * (rb - 112);
  return 2048 + 256 * (rb - 128);
}
__host__ __device__ inline int rung_ceil(int R) {   // smallest rung whose value is >= R (R <= GZ_RMAX)
  if (R <= 32) return R < 1 ? 1 : R;
  if (R <= 64) return 32 + (R - 32 + 1) / 2;
  if (R <= 128) return 48 + (R - 64 + 3) / 4;
  if (R <= 256) return 64 + (R - 128 + 7) / 8;
  if (R <= 512) return 80 + (R - 256 + 15) / 16;
  if (R <= 1024) return 96 + (R - 512 + 31) / 32;
  if (R <= 2048) return 112 + (R - 1024 + 63) / 64;
  return 128 + (R - 2048 + 255) / 256;
}
// bk_tab rows
#define GZ_BK_WARP0 0     // [bucket] first PGS warp (CTA) of the bucket
#define GZ_BK_REC0 1      // [bucket] first record of the bucket's slabs
#define GZ_BK_SWARP0 2    // [i] first warp of the i-th bucket in heavy-first order; [GZ_NBUCKET] = total warps
#define GZ_BK_SBUCKET 3   // [i] bucket id of the i-th bucket in heavy-first order
__device__ inline int *bk_row(const DevBatch &B, int row) { return B.bk_tab + row * (GZ_NBUCKET + 1); }
// worlds per PGS warp.  A chain-mode world has a warp to itself: a round lasts as long as its longest
// run, so the runs of several worlds in one warp would add up.
__device__ inline int bucket_wpw(const DevBatch &B, int bucket) {
  return bucket >= 5 * GZ_NRUNG ? 1 : min(16 >> ((bucket / GZ_NRUNG) % 5), B.pgs_wpw_max);
}

// where a world's records live once k_slab has run.  A chain-mode world has a warp to itself; its
// rows live in the world's run array (B.runrec, order[] sequence) and its slab holds only the 16-byte
// heads of its schedule items, [R][GZ_CHAIN_K] RecInts: R * 256 bytes, i.e. "width" 1 record per round
// (its bucket's rung is that of its row count m >= R, see k_schedule).
struct WSlab { double *base; int k, wslot, warp, width; };
__device__ inline int bucket_width(int bucket) { return bucket >= 5 * GZ_NRUNG ? 1 : 16; }
__device__ inline RecInts *chain_head(const WSlab &s, int local) { return reinterpret_cast<RecInts *>(s.base) + local; }
__device__ inline WSlab wslab(const DevBatch &B, const WSched &sc) {
  const int per = bucket_wpw(B, sc.bucket);
  WSlab s;
  s.k = sc.k; s.wslot = sc.idx % per; s.width = bucket_width(sc.bucket);
  const int wb = sc.idx / per;
  const int w0 = bk_row(B, GZ_BK_WARP0)[sc.bucket];
  s.warp = w0 < 0 ? -1 : w0 + wb;   // -1: the record pool was exhausted (k_slab flagged it)
  s.base = B.prec + ((size_t)bk_row(B, GZ_BK_REC0)[sc.bucket] + (size_t)wb * rung_value(sc.bucket % GZ_NRUNG) * s.width) * GZ_REC;
  return s;
}
__device__ inline double *slab_rec(const WSlab &s, int local) {   // local = slot * k + j
  return s.base + ((size_t)(local / s.k) * s.width + s.wslot * s.k + (local % s.k)) * GZ_REC;
}

// One warp per world: order[] positions (closed form), schedule items, k, then the greedy modulo
// scheduler on lane 0 (integer work only), then where every row's record and lambda live.
//
// A schedule ITEM is one row, or - in CHAIN mode - a maximal run of consecutive rows (in order[]
// sequence, inside one of the three segments bilateral / normals / frictions) that touch the same
// bodies.  Chain mode is for worlds whose busiest body carries a long serial chain (a robot lying
// on all its boxes: 80 rows on the core): k_pgs_chain (k_pgs_chain.cuh) runs an item in ONE round,
// its rows back to back with the bodies' accumulators in registers, instead of one round per row.
// It needs the whole world in shared memory; a world that does not fit is scheduled row by row.
// Row::next_f = lambda index, Row::t1a = stage, Row::t2a = index in the world's run array (-1: the
// row has its own slab slot, which is next_f), Row::t1f = slab slot | len << 20 for the first row of
// an item of a chain-mode world (-1 otherwise).
__global__ void __launch_bounds__(32) k_schedule(DevBatch B) {
  // dynamic shared memory: cnt, last, first [max_nb] ints, then occ [sch_rcap] shorts
  extern __shared__ int sch_sh[];
  int *cnt = sch_sh, *last = sch_sh + B.max_nb, *first = sch_sh + 2 * B.max_nb;
  unsigned short *occ = reinterpret_cast<unsigned short *>(sch_sh + 3 * B.max_nb);
  const int nb_cap = max(B.max_nb, 1);   // the probe tables follow occ: [nb_cap][32] ints x 2, [sch_rpar][32] shorts
  const int w = blockIdx.x, lane = threadIdx.x;
  const WorldInfo wi = B.world[w];
  WorldStep &ws = B.wstep[w];
  const int m = ws.m, nc = ws.n_contacts, mj = m - 3 * nc, nb = wi.nb;
  WSched *wsch = reinterpret_cast<WSched *>(B.wsched) + w;
  bool refuse = (m == 0 || m > 16 * GZ_RMAX);
  long long tc0 = clock64();   // GZCUDA_PGS_PROF: cycles of the phases, summed over worlds into prof[6], [7], [15]
  int *order = B.order + wi.row0;
  Row *rows = B.row + wi.row0;
  WScratch S = wscratch(B, w);
  int k = 1, rb = 1, R = 1, nstage = 1, nitems = 0, nrun = 0, chain = 0, cls = 0;
  if (!refuse) {
    // order[]: bilateral rows in index order, contact normals in REVERSE index order, friction
    // rows in index order (quickstep_pgs_lcp.cpp:1115-1150).  Persistent-joint rows are all
    // bilateral and every contact has exactly (normal, fric1, fric2): positions are closed-form.
    for (int i = lane; i < m; i += 32) {
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
    __syncwarp();
    if (B.pgs_prof && lane == 0) { const long long c = clock64(); atomicAdd(&B.prof[6], (unsigned long long)(c - tc0)); tc0 = c; }
    // chain mode needs the whole world resident in k_pgs_chain's shared memory: accumulators, fixed
    // part, lambda[m][2], m + 1 records, and (checked once R is known) the heads [R][GZ_CHAIN_K]
    const int res_doubles = 12 * (max(B.max_nb, 1) + 1) + GZ_CHAIN_FIXED_DOUBLES + 2 * m + (m + 1) * GZ_REC;
    for (int pass = 0; pass < 2; pass++) {   // pass 0 may choose chain mode; pass 1 is the fallback if it does not fit
      for (int b = lane; b < nb; b += 32) cnt[b] = 0;
      __syncwarp();
      for (int pos = lane; pos < m; pos += 32) {
        atomicAdd(&cnt[S.rb1[pos]], 1);
        if (S.rb2[pos] >= 0) atomicAdd(&cnt[S.rb2[pos]], 1);
      }
      __syncwarp();
      int maxcnt = 0;
      for (int b = lane; b < nb; b += 32) maxcnt = max(maxcnt, cnt[b]);
      maxcnt = __reduce_max_sync(0xffffffffu, maxcnt);
      chain = (pass == 0 && maxcnt >= B.pgs_chain_min && res_doubles + 2 * GZ_CHAIN_K <= B.pgs_chain_doubles[GZ_CHAIN_NCLS - 1]) ? 1 : 0;
      // ---- items
      if (chain) {
        __syncwarp();
        for (int b = lane; b < nb; b += 32) cnt[b] = 0;
        __syncwarp();
        if (lane == 0) {
          int pos = 0;
          nitems = 0;
          while (pos < m) {
            const int x1 = S.rb1[pos], x2 = S.rb2[pos];
            const int seg_end = pos < mj ? mj : (pos < mj + nc ? mj + nc : m);
            int L = 1;
            while (pos + L < seg_end && L < 2000 && S.rb1[pos + L] == x1 && S.rb2[pos + L] == x2) L++;
            S.it_pos[nitems] = pos; S.it_len[nitems] = L;
            cnt[x1]++;
            if (x2 >= 0) cnt[x2]++;
            nitems++; pos += L;
          }
        }
        nitems = __shfl_sync(0xffffffffu, nitems, 0);
        __syncwarp();
        maxcnt = 0;
        for (int b = lane; b < nb; b += 32) maxcnt = max(maxcnt, cnt[b]);
        maxcnt = __reduce_max_sync(0xffffffffu, maxcnt);
      } else {
        nitems = m;
      }
      // lanes per solve: the smallest power of two that brings items/k down to the period the busiest
      // body (or the batch-wide target period) imposes anyway; a chain-mode world has the whole warp
      if (chain) k = GZ_CHAIN_K;
      else {
        const int target = max(maxcnt, B.pgs_rt);
        k = B.pgs_kmin;
        while (k < B.pgs_kmax && (nitems + k - 1) / k > target) k <<= 1;
      }
      rb = rung_ceil(max((nitems + k - 1) / k, maxcnt));
      // probe: lane l runs the greedy for rung rb + l (feasibility only, tables per lane in shared
      // memory); the serial ladder below then starts at the first rung that works
      if (B.sch_rpar > 0 && rb < GZ_NRUNG && rung_value(rb) <= B.sch_rpar) {
        int *plast = reinterpret_cast<int *>(occ + ((B.sch_rcap + 1) & ~1)) , *pfirst = plast + 32 * nb_cap;
        unsigned short *pocc = reinterpret_cast<unsigned short *>(pfirst + 32 * nb_cap);
        const int rbl = rb + lane;
        const bool valid = rbl < GZ_NRUNG && rung_value(min(rbl, GZ_NRUNG - 1)) <= B.sch_rpar;
        const int Rl = valid ? rung_value(rbl) : 1;
        __syncwarp();   // it_pos / it_len written by lane 0
        int pf = valid ? 0 : 1, pneed = Rl, pmax = 0;
        for (int s = 0; s < Rl; s++) pocc[s * 32 + lane] = 0;
        for (int b = 0; b < nb; b++) { plast[b * 32 + lane] = -1; pfirst[b * 32 + lane] = -1; }
        for (int it = 0; it < nitems && !pf; it++) {
          const int pos = chain ? S.it_pos[it] : it;
          const int x1 = S.rb1[pos], x2 = S.rb2[pos];
          int o = plast[x1 * 32 + lane] + 1;
          if (x2 >= 0) o = max(o, plast[x2 * 32 + lane] + 1);
          int s = o % Rl, n = 0;
          while (pocc[s * 32 + lane] == k) {
            o++; s++;
            if (s == Rl) s = 0;
            if (++n > Rl) { pf = 1; break; }
          }
          if (pf) break;
          pocc[s * 32 + lane]++;
          plast[x1 * 32 + lane] = o;
          if (pfirst[x1 * 32 + lane] < 0) pfirst[x1 * 32 + lane] = o;
          if (x2 >= 0) { plast[x2 * 32 + lane] = o; if (pfirst[x2 * 32 + lane] < 0) pfirst[x2 * 32 + lane] = o; }
          pmax = max(pmax, o);
        }
        if (!pf)
          for (int b = 0; b < nb; b++)
            if (pfirst[b * 32 + lane] >= 0) pneed = max(pneed, plast[b * 32 + lane] - pfirst[b * 32 + lane] + 1);
        const unsigned okm = __ballot_sync(0xffffffffu, !pf && pneed == Rl && pmax / Rl < 255);
        const unsigned vm = __ballot_sync(0xffffffffu, valid);
        rb += okm ? __ffs(okm) - 1 : __popc(vm);   // none works: go on after the probed rungs (rb stays warp-uniform)
        __syncwarp();
      }
      if (B.pgs_prof && lane == 0) { const long long c = clock64(); atomicAdd(&B.prof[7], (unsigned long long)(c - tc0)); tc0 = c; }
      R = rung_value(min(rb, GZ_NRUNG - 1));
      bool done = false;
      for (int attempt = 0; !done; attempt++) {
        const bool give_up = attempt >= 24 || rb >= GZ_NRUNG || R > B.sch_rcap;
        if (give_up && chain) break;                                                   // pass 1 will do it
        if (give_up && nitems > min(GZ_RMAX, B.sch_rcap)) { refuse = true; break; }   // cannot happen with sane inputs
        if (give_up) {   // always valid: the sequential sweep itself (one lane per solve)
          k = 1; rb = rung_ceil(nitems); R = rung_value(rb); nstage = 1;
          for (int it = lane; it < nitems; it += 32) { S.slot[it] = it; S.stage[it] = 0; }
          done = true;
          break;
        }
        for (int s = lane; s < R; s += 32) occ[s] = 0;
        for (int b = lane; b < nb; b += 32) { last[b] = -1; first[b] = -1; }
        __syncwarp();
        int fail = 0, need = R, maxoff = 0;
        if (lane == 0) {
          for (int it = 0; it < nitems; it++) {
            const int pos = chain ? S.it_pos[it] : it;
            const int x1 = S.rb1[pos], x2 = S.rb2[pos];
            int o = last[x1] + 1;
            if (x2 >= 0) o = max(o, last[x2] + 1);
            int s = o % R, n = 0;
            while (occ[s] == k) {
              o++; s++;
              if (s == R) s = 0;
              if (++n > R) { fail = 1; break; }
            }
            if (fail) break;
            const int ln = occ[s]++;
            S.slot[it] = s * k + ln;      // per ITEM here; expanded to rows below
            S.stage[it] = o / R;
            last[x1] = o;
            if (first[x1] < 0) first[x1] = o;
            if (x2 >= 0) { last[x2] = o; if (first[x2] < 0) first[x2] = o; }
            maxoff = max(maxoff, o);
          }
          if (!fail)
            for (int b = 0; b < nb; b++)
              if (first[b] >= 0) need = max(need, last[b] - first[b] + 1);
        }
        fail = __shfl_sync(0xffffffffu, fail, 0);
        need = __shfl_sync(0xffffffffu, need, 0);
        maxoff = __shfl_sync(0xffffffffu, maxoff, 0);
        if (!fail && need == R && maxoff / R < 255) { done = true; nstage = maxoff / R + 1; }
        else { rb++; R = rung_value(rb); }   // next rung (3-6 % longer): jumping to `need` overshoots badly
        __syncwarp();
      }
      if (refuse) break;
      if (chain && done) {   // does it fit, and in which shared-memory class
        const int need_doubles = res_doubles + 2 * R * GZ_CHAIN_K;
        cls = 0;
        while (cls < GZ_CHAIN_NCLS && need_doubles > B.pgs_chain_doubles[cls]) cls++;
        if (cls < GZ_CHAIN_NCLS) break;
      } else if (done) break;
      chain = 0;
    }
  }
  if (refuse) {
    // nothing to solve (or a step this solver refuses): constraint accelerations are zero
    for (int b = lane; b < nb; b += 32) {
      BodyStep &bs = B.bstep[wi.body0 + b];
      for (int t = 0; t < 6; t++) { bs.caccel[t] = 0; bs.caccel_erp[t] = 0; }
    }
    if (lane == 0) {
      WSched z; z.R = 1; z.k = 1; z.nstage = 0; z.bucket = -1; z.idx = 0; z.nslots = 0; z.nrun = 0; z.pad1 = 0; *wsch = z;
      for (int q = 0; q < 4; q++) { ws.rms_dlambda[q] = 0; ws.rms_residual[q] = 0; }
      ws.num_contacts_stat = 0;
      if (m > 0) ws.overflow |= 8;   // needs the large-world solver
    }
    return;
  }
  const int nslots = R * k;
  __syncwarp();
  if (B.pgs_prof && lane == 0) { const long long c = clock64(); atomicAdd(&B.prof[15], (unsigned long long)(c - tc0)); tc0 = c; }
  // ---- rows: lambda index, stage, run membership
  if (!chain) {
    for (int pos = lane; pos < m; pos += 32) {
      Row &r = rows[order[pos]];
      r.next_f = S.slot[pos]; r.t1a = S.stage[pos]; r.t2a = -1; r.t1f = -1;
    }
  } else {
    // every row of a chain-mode world lives in the run array at its order[] position, which is also
    // its lambda index; the first row of an item carries the item's head
    nrun = m;
    for (int it = lane; it < nitems; it += 32) {
      const int pos = S.it_pos[it], L = S.it_len[it], slot = S.slot[it], stage = S.stage[it];
      for (int i = 0; i < L; i++) {
        Row &r = rows[order[pos + i]];
        r.next_f = pos + i; r.t1a = stage; r.t2a = pos + i;
        r.t1f = (i == 0) ? (slot | (L << 20)) : -1;
      }
    }
  }
  if (lane == 0) {
    int c = 0;
    while ((1 << c) < k) c++;
    // a chain-mode world is filed under the rung of its ROW COUNT: that, not its period, is how long it
    // runs (k_slab orders the warps longest first), and m records hold its R <= m rounds of heads
    WSched z; z.R = R; z.k = k; z.nstage = nstage; z.bucket = chain ? (5 + cls) * GZ_NRUNG + rung_ceil(m) : c * GZ_NRUNG + rb;
    z.nslots = nslots; z.nrun = nrun; z.pad1 = 0;
    z.idx = atomicAdd(&B.bk_count[z.bucket], 1);
    *wsch = z;
  }
}

// Warps and record slabs of every bucket, heaviest first so that the longest warps are the first to
// be scheduled on the SMs: chain-mode buckets by row count, then the others by period.  Block 0 does
// the prefix over the buckets (256 threads, GZ_SLAB_IT consecutive buckets each); the other blocks
// clear the warp -> world table.
#define GZ_SLAB_T 256
#define GZ_SLAB_IT ((GZ_NBUCKET + GZ_SLAB_T - 1) / GZ_SLAB_T)
__device__ inline int slab_bucket(int i) {   // i-th bucket in heavy-first order
  const int chain = i < 5 * GZ_NRUNG ? 1 : 0, r = i - (1 - chain) * 5 * GZ_NRUNG;
  const int rb = GZ_NRUNG - 1 - r / 5, c = 4 - r % 5;
  return (chain * 5 + c) * GZ_NRUNG + rb;
}
__global__ void __launch_bounds__(GZ_SLAB_T) k_slab(DevBatch B) {
  const int t = threadIdx.x;
  if (blockIdx.x > 0) {
    const int n = (B.n_worlds + GZ_NBUCKET) * 16;
    for (int i = (blockIdx.x - 1) * GZ_SLAB_T + t; i < n; i += (gridDim.x - 1) * GZ_SLAB_T) B.warp_world[i] = -1;
    return;
  }
  __shared__ int sw[GZ_SLAB_T];
  __shared__ long long sr[GZ_SLAB_T];
  int *warp0 = bk_row(B, GZ_BK_WARP0), *rec0 = bk_row(B, GZ_BK_REC0);
  int *swarp0 = bk_row(B, GZ_BK_SWARP0), *sbucket = bk_row(B, GZ_BK_SBUCKET);
  int nw[GZ_SLAB_IT];
  long long nr[GZ_SLAB_IT];
  int lw = 0;
  long long lr = 0;
#pragma unroll
  for (int q = 0; q < GZ_SLAB_IT; q++) {
    const int i = t * GZ_SLAB_IT + q;
    nw[q] = 0; nr[q] = 0;
    if (i < GZ_NBUCKET) {
      const int b = slab_bucket(i), per = bucket_wpw(B, b);
      nw[q] = (B.bk_count[b] + per - 1) / per;
      nr[q] = (long long)nw[q] * rung_value(b % GZ_NRUNG) * bucket_width(b);
    }
    lw += nw[q]; lr += nr[q];
  }
  sw[t] = lw; sr[t] = lr;
  __syncthreads();
  for (int d = 1; d < GZ_SLAB_T; d <<= 1) {   // inclusive scan over the threads
    const int aw = t >= d ? sw[t - d] : 0;
    const long long ar = t >= d ? sr[t - d] : 0;
    __syncthreads();
    sw[t] += aw; sr[t] += ar;
    __syncthreads();
  }
  const long long total = sr[GZ_SLAB_T - 1];
  if (total <= B.prec_cap) {
    int warps = sw[t] - lw;
    long long recs = sr[t] - lr;
#pragma unroll
    for (int q = 0; q < GZ_SLAB_IT; q++) {
      const int i = t * GZ_SLAB_IT + q;
      if (i < GZ_NBUCKET) {
        const int b = slab_bucket(i);
        warp0[b] = warps; rec0[b] = (int)recs;
        swarp0[i] = warps; sbucket[i] = b;
        warps += nw[q]; recs += nr[q];
      }
    }
    if (t == GZ_SLAB_T - 1) swarp0[GZ_NBUCKET] = sw[t];
  } else if (t == 0) {   // pool exhausted: buckets that do not fit are dropped and the step flagged
    int warps = 0;
    long long recs = 0;
    for (int i = 0; i < GZ_NBUCKET; i++) {
      const int b = slab_bucket(i), per = bucket_wpw(B, b);
      int n = (B.bk_count[b] + per - 1) / per;
      long long nrec = (long long)n * rung_value(b % GZ_NRUNG) * bucket_width(b);
      warp0[b] = warps; rec0[b] = (int)recs;
      swarp0[i] = warps; sbucket[i] = b;
      if (recs + nrec > B.prec_cap) {
        if (n > 0) B.wstep[0].overflow |= 16;
        n = 0; nrec = 0; warp0[b] = -1;
      }
      warps += n; recs += nrec;
    }
    swarp0[GZ_NBUCKET] = warps;
  }
}

// grid (ceil(max_row_cap / blockDim), n_worlds): one thread per row.  Finishes the row and
// writes its packed record.
__device__ __forceinline__ void rowfinish_one(DevBatch &B, const int w, const WorldInfo &wi, const int i) {
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
  if (B.pgs_impl >= 1) {
    // body-owner sweeps (k_pgsb.cuh, k_pgsl.cuh): one entry per body of the row, at the place k_bsched gave it
    const int nc = B.wstep[w].n_contacts, mj = B.wstep[w].m - wi.rpc * nc;
    const WScratchB S = wscratchb(B, w);
    const int pos = order_pos(i, mj, nc, wi.rpc);
    const int off = S.off[pos];
    if (off < 0) return;   // a row of an island that is not stepped, or a world the entry pool could not take
    const int cls = r.findex == -1 ? 0 : (r.findex == -2 ? 1 : 2);
    int back1 = 0, back2 = 0;
    if (cls == 2) {
      const int pn = order_pos(r.findex, mj, nc, wi.rpc);
      back1 = S.idx1[pos] - S.idx1[pn];
      if (r.b2 >= 0) back2 = S.idx2[pos] - S.idx2[pn];
    }
    double Ad2 = 0.0;   // quickstep_pgs_lcp.cpp:745-752
    if (!dequal(Ad, 0.0)) Ad2 = 1.0 / (Ad * Ad);
    const double rhsA = rhs * Ad, rhsA_erp = rhs_erp * Ad, adcfm = Ad * cfm;
    const int ent1 = B.bsched[wi.body0 + r.b1].x + S.idx1[pos], ent2 = r.b2 >= 0 ? B.bsched[wi.body0 + r.b2].x + S.idx2[pos] : -1;
    for (int h = 0; h < (r.b2 >= 0 ? 2 : 1); h++) {
      const int e = h ? ent2 : ent1;
      if (B.ent_part) B.ent_part[e] = h ? ent1 : ent2;
      if (B.ent_aux) {   // k_pgsc: partner entry, the CTA that owns the partner's body, distance to the normal row, class
        const int pbody = h ? r.b1 : r.b2;
        B.ent_aux[e] = make_int4(h ? ent1 : ent2, pbody >= 0 ? pbody / B.pgsc_bodies : 0, h ? back2 : back1, cls);
      }
      double2 *e2 = reinterpret_cast<double2 *>(B.ent + (size_t)e * GZ_ENT);
      for (int j = 0; j < 3; j++) {
        e2[j] = make_double2(J[6 * h + 2 * j] * Ad, J[6 * h + 2 * j + 1] * Ad);
        e2[3 + j] = make_double2(iMJ[6 * h + 2 * j], iMJ[6 * h + 2 * j + 1]);
      }
      e2[6] = make_double2(rhsA, lam);
      e2[7] = make_double2(rhsA_erp, lam_erp);
      e2[8] = make_double2(adcfm, r.lo);
      // bit 30: a torsional row ("index - constraint_index >= 3", quickstep_pgs_lcp.cpp:554-566): its limits follow the
      // normal force under the box model too
      const int mx = off | (cls << 27) | (h << 29) | ((cls == 2 && i - r.findex >= 3) ? (1 << 30) : 0);
      const int partner = h ? r.b1 : (r.b2 >= 0 ? r.b2 : GZ_ENT_NOPART);
      const int my = partner | ((h ? back2 : back1) << 16);
      e2[9] = make_double2(r.hi, __hiloint2double(my, mx));
      B.ent_row[e] = h ? -1 : i;
      B.ent_ad2[e] = Ad2;
    }
    return;
  }
  const WSched sc = reinterpret_cast<const WSched *>(B.wsched)[w];
  const WSlab sl = wslab(B, sc);
  if (sl.warp < 0) return;
  if (i == 0) B.warp_world[sl.warp * 16 + sl.wslot] = w;
  const unsigned epoch_sh = gz_epoch(B.step_index) << 8;
  // the row's record: its own slab slot, or its place in the world's run array (chain mode)
  double *rec = (r.t2a < 0) ? slab_rec(sl, r.next_f) : B.runrec + ((size_t)wi.row0 + r.t2a) * GZ_REC;
  double2 *rec2 = reinterpret_cast<double2 *>(rec);
  for (int j = 0; j < 6; j++) {
    rec2[j] = make_double2(J[2 * j] * Ad, J[2 * j + 1] * Ad);
    rec2[6 + j] = make_double2(iMJ[2 * j], iMJ[2 * j + 1]);
  }
  rec2[12] = make_double2(rhs * Ad, rhs_erp * Ad);
  rec2[13] = make_double2(Ad * cfm, r.lo);
  double Ad2 = 0.0;   // quickstep_pgs_lcp.cpp:745-752
  if (!dequal(Ad, 0.0)) Ad2 = 1.0 / (Ad * Ad);
  rec2[14] = make_double2(r.hi, Ad2);
  RecInts ri;
  ri.fi = (r.findex >= 0) ? B.row[wi.row0 + r.findex].next_f : r.findex;   // lambda index of the normal row
  ri.b1 = (unsigned short)r.b1; ri.b2 = (r.b2 >= 0) ? (unsigned short)r.b2 : (unsigned short)B.max_nb;   // none = the all-zero dummy slot
  ri.stage_epoch = (int)(epoch_sh | (unsigned)r.t1a);
  ri.orig = i;
  *reinterpret_cast<RecInts *>(rec + GZ_REC_INTS) = ri;
  if (r.t1f >= 0) {
    // first row of a schedule item of a chain-mode world: the item's head sits in the slot the
    // scheduler gave it.  fi = GZ_FI_RUN - len, orig = first index in the run array | friction flag << 30
    RecInts hi;
    hi.fi = GZ_FI_RUN - (r.t1f >> 20);
    hi.b1 = ri.b1; hi.b2 = ri.b2;
    hi.stage_epoch = ri.stage_epoch;
    hi.orig = r.t2a | ((r.findex >= 0) ? (1 << 30) : 0);
    *chain_head(sl, r.t1f & 0xFFFFF) = hi;
  }
}

// grid (gx, n_worlds): one thread per row, gx blocks striding over the world's rows
__global__ void k_rowfinish(DevBatch B) {
  const int w = blockIdx.y;
  const WorldInfo wi = B.world[w];
  const int m = B.wstep[w].m;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) rowfinish_one(B, w, wi, i);
}

}  // namespace gz
#include "k_pgs_sweep.cuh"
#include "k_pgs_chain.cuh"
#include "k_pgsl.cuh"
#include "k_pgsc.cuh"
#include "k_large.cuh"
#include "k_pgsm.cuh"
namespace gz {

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

// Joint feedback (dJointFeedback {f1, t1, f2, t2}: quickstep_update_bodies.cpp:79-119, Multiply1_12q1
// quickstep_util.cpp:42-73): the constraint wrench on each body = sum over the joint's rows of
// (unscaled J row) * lambda, rows in order.  Computed on demand from the row array of the last
// step; out[.][12].  One thread per persistent joint / per contact of one world.
__device__ inline void feedback_rows(const Row *rows, int m, bool two, double *out) {
  double f[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  for (int i = 0; i < m; i++) {
    const double s = rows[i].lambda;
    for (int t = 0; t < 6; t++) f[t] += rows[i].J[t] * s;
    if (two) for (int t = 0; t < 6; t++) f[6 + t] += rows[i].J[6 + t] * s;
  }
  for (int t = 0; t < 12; t++) out[t] = f[t];
}

__global__ void k_feedback_joint(DevBatch B, double *out) {
  int ji = blockIdx.x * blockDim.x + threadIdx.x;
  if (ji >= B.nj) return;
  const Joint &j = B.joint[ji];
  const WorldInfo wi = B.world[j.world];
  const bool live = j.m > 0 && B.wstep[j.world].m > 0;
  feedback_rows(B.row + wi.row0 + j.row0, live ? j.m : 0, j.b2 >= 0, out + 12 * (size_t)ji);
}

__global__ void k_feedback_contact(DevBatch B, int world, double *out) {
  const WorldInfo wi = B.world[world];
  int ci = blockIdx.x * blockDim.x + threadIdx.x;
  if (ci >= B.wstep[world].n_contacts) return;
  const Contact &C = B.contact[wi.contact0 + ci];
  const bool live = B.wstep[world].m > 0 && B.bstep[C.b1].active;
  feedback_rows(B.row + wi.row0 + C.row0, live ? (C.tors ? 4 : 3) : 0, C.b2 >= 0, out + 12 * (size_t)ci);
}

// One thread per body: dxUpdateBodies + dxStepBody.
__global__ void k_integrate(DevBatch B) {
  int bi = blockIdx.x * blockDim.x + threadIdx.x;
  if (bi >= B.nb) return;
  const BodyStep &bs = B.bstep[bi];
  {
    // last kernel of the step: a capacity overflow of this step (pairs 1, contacts 2, rows 4, solver 8, record /
    // entry pool 16) is ORed into the batch's STICKY word, which the host checks at every synchronising call
    const int w = B.body[bi].world;
    if (bi == B.world[w].body0) {
      const int ov = B.wstep[w].overflow;
      if (ov) { atomicOr(&B.sticky[0], ov); atomicMax(&B.sticky[1], w + 1); }
    }
  }
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
// Setters that a per-step plugin loop calls: stream-ordered, only the touched fields, no read-modify-write of the
// whole body array on the host.  in[count][13] = pos 3, quat 4, lvel 3, avel 3; NaN in the first element of a
// group = leave that group as it is.  A new quaternion is normalised and R rebuilt (dBodySetQuaternion, ode.cpp:414-427).
__global__ void k_set_state(DevBatch B, int first, int count, int raw) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  Body &b = B.body[first + i];
  const double *in = B.set_stage + 13 * (size_t)i;
  if (in[0] == in[0]) for (int k = 0; k < 3; k++) b.pos[k] = in[k];
  if (in[3] == in[3]) {
    for (int k = 0; k < 4; k++) b.q[k] = in[3 + k];
    if (!raw) normalize4(b.q);   // raw: the host's dBodySetQuaternion normalised it already; taken bit for bit
    q_to_R(b.q, b.R);
  }
  if (in[7] == in[7]) for (int k = 0; k < 3; k++) b.lvel[k] = in[7 + k];
  if (in[10] == in[10]) for (int k = 0; k < 3; k++) b.avel[k] = in[10 + k];
}
// dBodyEnable / dBodyDisable (ode.cpp:889-905): enabled[i] != 0 clears the disabled flag and restarts the auto-disable
// counters, 0 sets the flag; a body whose flag already agrees is left alone.
__global__ void k_set_enabled(DevBatch B, int first, int count) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  Body &b = B.body[first + i];
  const bool want = B.flag_stage[i] != 0, is = !(b.flags & GZ_BODY_DISABLED);
  if (want == is) return;
  if (want) { b.flags &= ~GZ_BODY_DISABLED; b.adis_stepsleft = B.p.adis_idle_steps; b.adis_timeleft = B.p.adis_idle_time; }
  else b.flags |= GZ_BODY_DISABLED;
}
__global__ void k_set_pd(DevBatch B) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= B.nj) return;
  for (int k = 0; k < 6; k++) B.joint[j].pd[k] = B.jstage[6 * (size_t)j + k];
}
__global__ void k_set_velcmd(DevBatch B) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= B.nj) return;
  const double v = B.jstage[j];
  const bool on = v == v && B.joint[j].type == GZ_JOINT_HINGE;   // NaN = no command
  B.joint[j].has_vel_cmd = on ? 1 : 0;
  B.joint[j].vel_cmd = on ? v : 0.0;
}
__global__ void k_gather_flags(DevBatch B) {
  int bi = blockIdx.x * blockDim.x + threadIdx.x;
  if (bi >= B.nb) return;
  B.flag_stage[bi] = B.body[bi].flags & GZ_BODY_DISABLED;
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
