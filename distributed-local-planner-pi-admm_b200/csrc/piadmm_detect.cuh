// piadmm_detect.cuh -- collision-edge detection of casadi/main.py:110-118 for generic V, as device functions shared by
// the phase kernels (piadmm.cu) and the persistent cooperative kernel (piadmm_coop.cuh).
//
// The reference recomputes the dense V x V edge_mat in every outer iteration.  Here the O(V^2) work is done ONCE per
// MPC step: every trajectory of a step starts at the same plant state and a roll-out of N steps cannot leave a disc
// around it (reach_radius), so only pairs whose discs come within the collision distance can ever collide in this step.
// Those pairs form the candidate bit matrix `cand` (k_cand / coop phase 0).  Per outer iteration a warp per vehicle walks
// the vehicle's candidate row in ascending j, runs the EXACT point-by-point test of the reference on each candidate
// (lane k = trajectory point k, eight candidates in flight) and allocates neighbour slots for the hits -- ascending
// partner order per vehicle, which is what the reference's row-major walk over np.where(edge_mat == 1) gives every
// vehicle.  The result is exact (the candidate set is a superset by construction), and an outer iteration costs
// O(candidates), not O(V^2).  The round-1 path (bounding boxes + full pair test every iteration, `hits` matrix) is kept
// behind PIADMM_DETECT=matrix_warp|matrix_thread as the A/B baseline; both feed the same slot / count / emit functions.
#pragma once
#include "piadmm_dev.cuh"

// |u| bound assumed by the reach radius of the LINEARISED roll-out (steering is boxed to pi/6 = 0.52 up to the inner
// solver's tolerance and the 4-decimal rounding; 2 rad is far outside anything a converged QP returns)
#define PIADMM_U_BOUND 2.0

// every point of every roll-out of a vehicle with speed v lies within this distance of its start
__host__ __device__ inline double reach_radius(const piadmm_params& p, int N, double v) {
  const double s = v * N * p.dt;              // non-linear roll-out (casadi/PI_ADMM_class.py:95-101): N steps of length v*dt
  if (p.pos_nonlinear) return s * (1.0 + 1e-9) + 1e-9;
  // linearised roll-out (:61-65): k*dt*v along the heading, g*(M u)_k across it, |(M u)_k| <= U k(k-1)/2
  const double g = v * v * p.dt * p.dt / p.L, lat = g * PIADMM_U_BOUND * 0.5 * N * (N - 1);
  return sqrt(s * s + lat * lat) * (1.0 + 1e-9) + 1e-9;
}

// casadi/main.py:56-63 for the per-slot arrays (hat, dual, integ, diff, last_hat = 0) and the slot tables, one WARP per
// slot.  Only slots that were allocated in the previous step (nbr >= 0) are written: an unallocated slot is never
// touched by any phase, so it still holds the zeros of the step before (invariant: nbr < 0  =>  the slot is zero;
// slots_dirty forces the full pass after piadmm_bind).  Config 5 has 32768 slots of 2.5 KB and a few hundred in use:
// 81 MB of stores per MPC step become < 2 MB.
__device__ __forceinline__ void dev_begin_slots(const DevCtx& c, size_t warp_id, size_t n_warps, int lane) {
  const int per = 2 * (c.N + 1);
  const size_t n_slots = (size_t)c.S * c.V * c.D;
  if (n_slots <= 4 * n_warps) {
    // few slots per warp (dense small scenes, config 3: most slots are in use): a WARP per slot
    for (size_t t = warp_id; t < n_slots; t += n_warps) {
      const bool used = c.slots_dirty || c.b.nbr[t] >= 0;
      __syncwarp();
      if (used) {
        for (int el = lane; el < per; el += 32) {
          const size_t o = t * per + el;
          c.b.hat[o] = 0.0; c.b.dual[o] = 0.0; c.b.integ[o] = 0.0; c.b.diff[o] = 0.0; c.b.last_hat[o] = 0.0;
        }
      }
      if (lane == 0) {
        const int d = (int)(t % c.D), i = (int)((t / c.D) % c.V);
        int j = -1;
        if (c.p.al_all && d < c.V - 1) j = d < i ? d : d + 1;      // AL over every j != i: slots pre-assigned, ascending
        c.b.nbr[t] = j;
        c.slot_hit[t] = 0;
      }
    }
    return;
  }
  // large scenes (config 5: 32768 slots, a few hundred in use): a LANE per slot for the table entries and the "was it in
  // use" test -- one coalesced read of nbr per 32 slots instead of 28 dependent rounds per warp (~25 us of the step) --
  // then the warp zeroes the slots that were in use one after the other
  for (size_t base = warp_id * 32; base < n_slots; base += n_warps * 32) {
    const size_t t = base + lane;
    bool used = false;
    if (t < n_slots) {
      used = c.slots_dirty || c.b.nbr[t] >= 0;
      const int d = (int)(t % c.D), i = (int)((t / c.D) % c.V);
      int j = -1;
      if (c.p.al_all && d < c.V - 1) j = d < i ? d : d + 1;
      c.b.nbr[t] = j;
      c.slot_hit[t] = 0;
    }
    unsigned m = __ballot_sync(PIADMM_FULL, used);
    while (m) {
      const size_t tt = base + (__ffs(m) - 1);
      m &= m - 1;
      for (int el = lane; el < per; el += 32) {
        const size_t o = tt * per + el;
        c.b.hat[o] = 0.0; c.b.dual[o] = 0.0; c.b.integ[o] = 0.0; c.b.diff[o] = 0.0; c.b.last_hat[o] = 0.0;
      }
    }
  }
}

// x_memo (include/piadmm.h): solve vehicle a in this outer iteration?
__device__ __forceinline__ bool dev_memo_on(const DevCtx& c) { return c.p.x_memo && c.p.dual_mode != PIADMM_DUAL_PI_OLD; }

// disc[s,i] = (x0, y0, reach, 0) of vehicle (s,i); stored in DevCtx::bbox (unused by the candidate path)
__device__ __forceinline__ void dev_disc_vehicle(const DevCtx& c, long a) {
  const double* xt = c.b.xt + a * 3;
  reinterpret_cast<double4*>(c.bbox)[a] = make_double4(xt[0], xt[1], reach_radius(c.p, c.N, c.b.spd[a]), 0.0);
}

// blockbox[s,jb] = (xmin, xmax, ymin, ymax) over the reach discs of the vehicles 32*jb .. 32*jb+31 of scenario s: one
// warp per block, computed from xt / spd directly (no dependence on dev_disc_vehicle having run)
__device__ __forceinline__ void dev_block_box(const DevCtx& c, int s, int jb, int lane) {
  const int j = jb * 32 + lane;
  double x0 = 1e300, x1 = -1e300, y0 = 1e300, y1 = -1e300;
  if (j < c.V) {
    const size_t a = (size_t)s * c.V + j;
    const double* xt = c.b.xt + a * 3;
    const double r = reach_radius(c.p, c.N, c.b.spd[a]);
    x0 = xt[0] - r; x1 = xt[0] + r; y0 = xt[1] - r; y1 = xt[1] + r;
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    x0 = fmin(x0, __shfl_xor_sync(PIADMM_FULL, x0, o)); x1 = fmax(x1, __shfl_xor_sync(PIADMM_FULL, x1, o));
    y0 = fmin(y0, __shfl_xor_sync(PIADMM_FULL, y0, o)); y1 = fmax(y1, __shfl_xor_sync(PIADMM_FULL, y1, o));
  }
  if (lane == 0) reinterpret_cast<double4*>(c.blockbox)[(size_t)s * ((c.V + 31) / 32) + jb] = make_double4(x0, x1, y0, y1);
}

// word jb of row i of the (symmetric) candidate matrix of scenario s.  Two levels: the disc of i against the box of the
// whole block first (one test rejects 32 pairs; with vehicles in a spatial order almost every word goes this way).
__device__ __forceinline__ uint32_t dev_cand_word(const DevCtx& c, int s, int i, int jb) {
  const double4* disc = reinterpret_cast<const double4*>(c.bbox) + (size_t)s * c.V;
  const double4 di = disc[i];
  const double thr_d = c.p.thr_squared ? c.p.dis_thres : sqrt(c.p.dis_thres);   // the distance whose square is the threshold
  const int j0 = jb * 32, j1 = min(j0 + 32, c.V);
  uint32_t m = 0;
  if (!c.p.always_edge) {
    const double4 bb = reinterpret_cast<const double4*>(c.blockbox)[(size_t)s * ((c.V + 31) / 32) + jb];
    const double gx = fmax(0.0, fmax(bb.x - di.x, di.x - bb.y)), gy = fmax(0.0, fmax(bb.z - di.y, di.y - bb.w));
    const double rr = (di.z + thr_d) * (1.0 + 1e-9);      // (the discs' own radii are inside the box)
    if (gx * gx + gy * gy >= rr * rr) return 0u;
  }
  for (int j = j0; j < j1; j++) {
    if (j == i) continue;
    bool near = c.p.always_edge != 0;
    if (!near) {
      const double4 dj = disc[j];
      const double dx = di.x - dj.x, dy = di.y - dj.y, rr = (di.z + dj.z + thr_d) * (1.0 + 1e-9);
      near = dx * dx + dy * dy < rr * rr;
    }
    if (near) m |= 1u << (j - j0);
  }
  return m;
}

// One warp, one vehicle: walk row i of `cand` (EXACT: exact test of every candidate) or of `hits` (already exact) in
// ascending j and find / allocate the neighbour slot of every colliding partner; slot_hit[s,i,d] = 1 for the slots whose
// partner collides in THIS outer iteration.
template <bool EXACT>
__device__ __forceinline__ void dev_slots_vehicle(const DevCtx& c, int s, int i, int lane) {
  const int N = c.N, np1 = N + 1, words = (c.V + 31) / 32;
  const size_t a = (size_t)s * c.V + i;
  int32_t* nb = c.b.nbr + a * c.D;
  uint8_t* sh = c.slot_hit + a * c.D;
  for (int d = lane; d < c.D; d += 32) sh[d] = 0;
  __syncwarp();
  const uint32_t* row = (EXACT ? c.cand : c.hits) + a * words;
  const bool test = EXACT && !c.p.always_edge;
  const double thr = c.p.thr_squared ? c.p.dis_thres * c.p.dis_thres : c.p.dis_thres;
  const int k = lane <= N ? lane : 0;                        // padded lanes repeat point 0
  double xi = 0.0, yi = 0.0;
  if (test) { const double* pi = c.b.pos + a * 2 * np1; xi = pi[k]; yi = pi[np1 + k]; }
  bool overflow = false;
  for (int w0 = 0; w0 < words; w0 += 32) {
    const uint32_t mine = w0 + lane < words ? row[w0 + lane] : 0u;
    unsigned nz = __ballot_sync(PIADMM_FULL, mine != 0u);
    while (nz) {
      const int src = __ffs(nz) - 1;
      nz &= nz - 1;
      uint32_t m = __shfl_sync(PIADMM_FULL, mine, src);
      const int jbase = (w0 + src) * 32;
      if (test) {                       // exact tests, eight candidates in flight (16 independent loads per lane)
        uint32_t rem = m, hits = 0;
        while (rem) {
          int jj[8];
          double xs[8], ys[8];
#pragma unroll
          for (int q = 0; q < 8; q++) {
            jj[q] = -1;
            if (rem) { jj[q] = __ffs(rem) - 1; rem &= rem - 1; }
          }
#pragma unroll
          for (int q = 0; q < 8; q++) {
            xs[q] = 0.0; ys[q] = 0.0;
            if (jj[q] >= 0) {
              const double* pj = c.b.pos + ((size_t)s * c.V + jbase + jj[q]) * 2 * np1;
              xs[q] = pj[k]; ys[q] = pj[np1 + k];
            }
          }
#pragma unroll
          for (int q = 0; q < 8; q++) {
            if (jj[q] >= 0) {           // warp uniform
              const double dx = xi - xs[q], dy = yi - ys[q];
              if (__any_sync(PIADMM_FULL, dx * dx + dy * dy < thr)) hits |= 1u << jj[q];
            }
          }
        }
        m = hits;
      }
      if (lane == 0) {
        while (m) {
          const int j = jbase + __ffs(m) - 1;
          m &= m - 1;
          int slot = -1;
          for (int d = 0; d < c.D; d++) if (nb[d] == j) { slot = d; break; }
          if (slot < 0) for (int d = 0; d < c.D; d++) if (nb[d] < 0) { nb[d] = j; slot = d; break; }
          if (slot < 0) overflow = true; else sh[slot] = 1;
        }
      }
    }
  }
  if (lane == 0 && overflow) { atomicOr(c.b.status + s, PIADMM_ST_SLOT_OVERFLOW); if (c.any_overflow) *c.any_overflow = 1; }
}

__device__ __forceinline__ int dev_far_slot(const DevCtx& c, int s, int j, int i);

// candidate list of vehicle (s,i) out of its bit row: one warp, ascending j (once per MPC step)
__device__ __forceinline__ void dev_cand_list_vehicle(const DevCtx& c, int s, int i, int lane) {
  const int words = (c.V + 31) / 32;
  const size_t a = (size_t)s * c.V + i;
  const uint32_t* row = c.cand + a * words;
  int32_t* lst = c.cand_list + a * PIADMM_CAND_CAP;
  int n = 0;
  for (int w0 = 0; w0 < words; w0 += 32) {
    const uint32_t mine = w0 + lane < words ? row[w0 + lane] : 0u;
    // exclusive prefix of the popcounts: where this lane's candidates go
    const int pc = __popc(mine);
    int inc = pc;
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(PIADMM_FULL, inc, o); if (lane >= o) inc += t; }
    int at = n + inc - pc;
    uint32_t m = mine;
    while (m) {
      const int j = (w0 + lane) * 32 + __ffs(m) - 1;
      m &= m - 1;
      if (at < PIADMM_CAND_CAP) lst[at] = j;
      at++;
    }
    n += __shfl_sync(PIADMM_FULL, inc, 31);
  }
  if (lane == 0) {
    c.cand_cnt[a] = n <= PIADMM_CAND_CAP ? n : -1;
    if (n > PIADMM_CAND_CAP && c.cand_over) *c.cand_over = 1;
  }
}

// Candidate list of vehicle (s,i) STRAIGHT from the reach discs, one warp, ascending j -- the same tests as dev_cand_word
// (block box first, then the 32 discs of a block that passes, one per lane), so the list is the one dev_cand_list_vehicle
// reads off the bit matrix, without writing and re-reading the V x V/32 words of it (2 MB at config 5, 134 MB at 32768
// vehicles) and one launch less per MPC step.  A vehicle with more than PIADMM_CAND_CAP candidates gets cand_cnt = -1 AND
// its bit row (second pass), which is what the slot phase walks for it; the rows of the other vehicles are not written.
// Only for D <= 32 (with more slots the slot phase walks the bit rows of every vehicle: k_cand + k_cand_list).
__device__ __forceinline__ void dev_cand_direct_vehicle(const DevCtx& c, int s, int i, int lane) {
  const int words = (c.V + 31) / 32;
  const size_t a = (size_t)s * c.V + i;
  const double4* disc = reinterpret_cast<const double4*>(c.bbox) + (size_t)s * c.V;
  const double4* box = reinterpret_cast<const double4*>(c.blockbox) + (size_t)s * words;
  const double4 di = disc[i];
  const double thr_d = c.p.thr_squared ? c.p.dis_thres : sqrt(c.p.dis_thres);
  const double rb = (di.z + thr_d) * (1.0 + 1e-9);
  const bool all = c.p.always_edge != 0;
  int32_t* lst = c.cand_list + a * PIADMM_CAND_CAP;
  uint32_t* row = c.cand + a * words;
  for (int pass = 0; pass < 2; pass++) {            // 0: the list; 1 (only after an overflow): the bit row
    int n = 0;
    if (pass == 1) {
      for (int w = lane; w < words; w += 32) row[w] = 0u;
      __syncwarp();
    }
    for (int w0 = 0; w0 < words; w0 += 32) {
      const int jb = w0 + lane;
      bool pass_box = false;
      if (jb < words) {
        pass_box = all;
        if (!all) {
          const double4 bb = box[jb];
          const double gx = fmax(0.0, fmax(bb.x - di.x, di.x - bb.y)), gy = fmax(0.0, fmax(bb.z - di.y, di.y - bb.w));
          pass_box = !(gx * gx + gy * gy >= rb * rb);
        }
      }
      unsigned pm = __ballot_sync(PIADMM_FULL, pass_box);
      while (pm) {
        const int jbb = w0 + __ffs(pm) - 1;
        pm &= pm - 1;
        const int j = jbb * 32 + lane;
        bool near = false;
        if (j < c.V && j != i) {
          near = all;
          if (!all) {
            const double4 dj = disc[j];
            const double dx = di.x - dj.x, dy = di.y - dj.y, rr = (di.z + dj.z + thr_d) * (1.0 + 1e-9);
            near = dx * dx + dy * dy < rr * rr;
          }
        }
        const unsigned m = __ballot_sync(PIADMM_FULL, near);
        if (pass == 0) {
          const int at = n + __popc(m & ((1u << lane) - 1u));
          if (near && at < PIADMM_CAND_CAP) lst[at] = j;
          n += __popc(m);
        } else if (lane == 0) {
          row[jbb] = m;
        }
      }
    }
    if (pass == 0) {
      if (lane == 0) {
        c.cand_cnt[a] = n <= PIADMM_CAND_CAP ? n : -1;
        if (n > PIADMM_CAND_CAP && c.cand_over) *c.cand_over = 1;
      }
      if (n <= PIADMM_CAND_CAP) break;
    }
  }
}

// Slot phase over the candidate LIST (the default): lane q holds candidate q.  Only pairs with a recomputed trajectory
// on either side are tested again (veh_stamp == it: the x phase of this outer iteration solved the vehicle; vehicles of
// other ranks count as recomputed) -- the state of every other pair (slot_hit of its slot, or "no slot") is the one of
// the previous outer iteration and still exact.  With x_memo most vehicles of a large scene rest after the first
// iteration, and so do their pairs.
__device__ __forceinline__ void dev_slots_vehicle_list(const DevCtx& c, int s, int i, int lane, int it) {
  const int N = c.N, np1 = N + 1;
  const size_t a = (size_t)s * c.V + i, sv = (size_t)s * c.V;
  int32_t* nb = c.b.nbr + a * c.D;
  uint8_t* sh = c.slot_hit + a * c.D;
  const int cnt = c.cand_cnt[a];
  const bool test = !c.p.always_edge;
  const double thr = c.p.thr_squared ? c.p.dis_thres * c.p.dis_thres : c.p.dis_thres;
  const int k = lane <= N ? lane : 0;                        // padded lanes repeat point 0
  const bool sg = c.stamps_global != 0;
  const bool moved_i = c.veh_stamp[a] == it || (!sg && (i < c.own_lo || i >= c.own_hi));
  int nbv = lane < c.D ? nb[lane] : -2;                      // the slot table in registers, lane = slot (D <= 32)
  // a point pair closer than sqrt(thr) needs both coordinate gaps of the boxes below it (margin: one rounding of the gap)
  const double box_d = sqrt(thr) * (1.0 + 1e-9) + 1e-12;
  double4 bi = make_double4(0.0, 0.0, 0.0, 0.0);
  if (test && c.use_tbox) bi = reinterpret_cast<const double4*>(c.tbox)[a];
  double xi = 0.0, yi = 0.0;
  bool have_pos = false, overflow = false;
  for (int base = 0; base < cnt; base += 32) {
    const int q = base + lane;
    const int j = q < cnt ? c.cand_list[a * PIADMM_CAND_CAP + q] : -1;
    bool need = false;
    if (j >= 0) need = moved_i || c.veh_stamp[sv + j] == it || (!sg && (j < c.own_lo || j >= c.own_hi));
    const unsigned mask = __ballot_sync(PIADMM_FULL, need);
    if (!mask) continue;
    unsigned hits = test ? 0u : mask;
    unsigned tmask = mask;                                    // pairs that get the exact test
    if (test && c.use_tbox) {           // the two trajectory boxes first: a lane per candidate, 32 B instead of 2(N+1) points
      bool close = false;
      if (need) {
        const double4 bj = reinterpret_cast<const double4*>(c.tbox)[sv + j];
        const double gx = fmax(bj.x - bi.y, bi.x - bj.y), gy = fmax(bj.z - bi.w, bi.z - bj.w);   // gaps (negative: overlap)
        close = gx < box_d && gy < box_d;
      }
      tmask = __ballot_sync(PIADMM_FULL, close);
    }
    if (test && tmask) {                // exact tests, eight pairs in flight (16 independent loads per lane)
      if (!have_pos) { const double* pi = c.b.pos + a * 2 * np1; xi = pi[k]; yi = pi[np1 + k]; have_pos = true; }
      unsigned rem = tmask;
      while (rem) {
        int src[8], jq[8];
        double xs[8], ys[8];
#pragma unroll
        for (int u = 0; u < 8; u++) {
          src[u] = -1;
          if (rem) { src[u] = __ffs(rem) - 1; rem &= rem - 1; }
          jq[u] = __shfl_sync(PIADMM_FULL, j, src[u] < 0 ? 0 : src[u]);
        }
#pragma unroll
        for (int u = 0; u < 8; u++) {
          xs[u] = 0.0; ys[u] = 0.0;
          if (src[u] >= 0) { const double* pj = c.b.pos + (sv + jq[u]) * 2 * np1; xs[u] = pj[k]; ys[u] = pj[np1 + k]; }
        }
#pragma unroll
        for (int u = 0; u < 8; u++) {
          if (src[u] >= 0) {            // warp uniform
            const double dx = xi - xs[u], dy = yi - ys[u];
            if (__any_sync(PIADMM_FULL, dx * dx + dy * dy < thr)) hits |= 1u << src[u];
          }
        }
      }
    }
    // slot of every tested candidate: all lanes search the register copy of the table at once
    int slot = -1;
    for (int d = 0; d < c.D; d++) { const int v = __shfl_sync(PIADMM_FULL, nbv, d); if (need && v == j) slot = d; }
    const bool hit = (hits >> lane) & 1u;
    if (need && slot >= 0) sh[slot] = hit ? 1 : 0;
    unsigned fresh = __ballot_sync(PIADMM_FULL, need && hit && slot < 0);   // new partners: slots in ascending j
    while (fresh) {
      const int b = __ffs(fresh) - 1;
      fresh &= fresh - 1;
      const int jb = __shfl_sync(PIADMM_FULL, j, b);
      const unsigned fr = __ballot_sync(PIADMM_FULL, lane < c.D && nbv == -1);
      if (!fr) { overflow = true; continue; }
      const int f = __ffs(fr) - 1;
      if (lane == f) { nbv = jb; nb[f] = jb; sh[f] = 1; }
    }
  }
  if (lane == 0 && overflow) { atomicOr(c.b.status + s, PIADMM_ST_SLOT_OVERFLOW); if (c.any_overflow) *c.any_overflow = 1; }
}

// up_cnt[a] = cnt (one thread).  In the cooperative kernel the per-chunk sums follow by the difference, so its edge phase
// adds up V / CTA_THREADS chunk totals instead of up to V counts per chunk (and a vehicle whose slot phase is skipped keeps
// both its count and its share of the total)
#define PIADMM_CHUNK 64                    // == CTA_THREADS (piadmm.cu; static_assert there)
__device__ __forceinline__ void dev_set_up_cnt(const DevCtx& c, int s, int i, int cnt, int old) {     // old = up_cnt[a], read before
  const size_t a = (size_t)s * c.V + i;
  if (!c.use_chunk_tot) { c.up_cnt[a] = cnt; return; }
  if (old != cnt) {
    c.up_cnt[a] = cnt;
    atomicAdd(c.chunk_tot + (size_t)s * ((c.V + PIADMM_CHUNK - 1) / PIADMM_CHUNK) + i / PIADMM_CHUNK, cnt - old);
  }
}
__device__ __forceinline__ void dev_set_up_cnt(const DevCtx& c, int s, int i, int cnt) {
  dev_set_up_cnt(c, s, i, cnt, c.use_chunk_tot ? c.up_cnt[(size_t)s * c.V + i] : 0);
}

// slot phase of one vehicle: candidate list, bit row of `cand` when the list overflowed, or the `hits` matrix
// ... and, from the vehicle's OWN table only: the number of its up-edges (i, j > i) of this outer iteration and its MATLAB
// :202 distance flag.  Without a table overflow anywhere in the step this IS the number of valid edges (a partner
// that collides holds the mirror slot); with one, the count phase (dev_count_vehicle_warp) redoes it with the far side.
template <bool EXACT>
__device__ __forceinline__ void dev_slots(const DevCtx& c, int s, int i, int lane, int it) {
  const int old_cnt = c.use_chunk_tot ? c.up_cnt[(size_t)s * c.V + i] : 0;      // (early: its latency hides behind the tests)
  if (EXACT && c.D <= 32 && c.cand_cnt[(size_t)s * c.V + i] >= 0) dev_slots_vehicle_list(c, s, i, lane, it);
  else dev_slots_vehicle<EXACT>(c, s, i, lane);
  __syncwarp();
  const size_t a = (size_t)s * c.V + i;
  const int32_t* nb = c.b.nbr + a * c.D;
  const uint8_t* sh = c.slot_hit + a * c.D;
  const piadmm_params& p = c.p;
  int cnt = 0;
  bool bad = false;
  for (int d0 = 0; d0 < c.D; d0 += 32) {
    const int d = d0 + lane;
    bool up = false;
    if (d < c.D) {
      const int j = nb[d];
      if (j > i) {
        up = sh[d] != 0;
        if (p.stop_dist_test) {
          const double* xi = c.b.xt + a * 3;
          const double* xj = c.b.xt + ((size_t)s * c.V + j) * 3;
          const double vi = c.b.spd[a], vj = c.b.spd[(size_t)s * c.V + j];
          const double dx = (xi[0] + p.dt * vi * cos(xi[2])) - (xj[0] + p.dt * vj * cos(xj[2]));
          const double dy = (xi[1] + p.dt * vi * sin(xi[2])) - (xj[1] + p.dt * vj * sin(xj[2]));
          if (!(sqrt(dx * dx + dy * dy) > p.dis_thres)) bad = true;
        }
      }
    }
    cnt += __popc(__ballot_sync(PIADMM_FULL, up));
  }
  bad = __any_sync(PIADMM_FULL, bad);
  if (lane == 0) {
    dev_set_up_cnt(c, s, i, cnt, old_cnt);
    if (bad) c.dist_bad[s] = 1;
  }
}

// valid up-edges of vehicle i (dev_count_vehicle) and its MATLAB :202 distance flag, one WARP per vehicle: lane = slot
__device__ __forceinline__ void dev_count_vehicle_warp(const DevCtx& c, int s, int i, int lane) {
  const size_t a = (size_t)s * c.V + i;
  const int32_t* nb = c.b.nbr + a * c.D;
  const uint8_t* sh = c.slot_hit + a * c.D;
  const piadmm_params& p = c.p;
  int cnt = 0;
  bool bad = false;
  for (int d0 = 0; d0 < c.D; d0 += 32) {
    const int d = d0 + lane;
    bool ok = false;
    if (d < c.D) {
      const int j = nb[d];
      if (j > i) {
        if (sh[d]) ok = dev_far_slot(c, s, j, i) >= 0;
        if (p.stop_dist_test) {
          const double* xi = c.b.xt + a * 3;
          const double* xj = c.b.xt + ((size_t)s * c.V + j) * 3;
          const double vi = c.b.spd[a], vj = c.b.spd[(size_t)s * c.V + j];
          const double dx = (xi[0] + p.dt * vi * cos(xi[2])) - (xj[0] + p.dt * vj * cos(xj[2]));
          const double dy = (xi[1] + p.dt * vi * sin(xi[2])) - (xj[1] + p.dt * vj * sin(xj[2]));
          if (!(sqrt(dx * dx + dy * dy) > p.dis_thres)) bad = true;
        }
      }
    }
    cnt += __popc(__ballot_sync(PIADMM_FULL, ok));
  }
  bad = __any_sync(PIADMM_FULL, bad);
  if (lane == 0) {
    dev_set_up_cnt(c, s, i, cnt);
    if (bad) c.dist_bad[s] = 1;
  }
}

// slot of vehicle i in the table of vehicle j, or -1
__device__ __forceinline__ int dev_far_slot(const DevCtx& c, int s, int j, int i) {
  const int32_t* nbj = c.b.nbr + ((size_t)s * c.V + j) * c.D;
  int dj = -1;
  for (int d = 0; d < c.D; d++) if (nbj[d] == i) dj = d;
  return dj;
}

// MATLAB ...PI_antiwindup.m:202 for the allocated slots (i, j > i) of vehicle i: is the one-step-ahead distance
// dis_vec(2) of any pair NOT larger than dis_thres?  (depends on the plant state only)
__device__ __forceinline__ bool dev_dist_bad_vehicle(const DevCtx& c, int s, int i) {
  const piadmm_params& p = c.p;
  const int32_t* nb = c.b.nbr + ((size_t)s * c.V + i) * c.D;
  bool bad = false;
  for (int d = 0; d < c.D; d++) {
    const int j = nb[d];
    if (j <= i) continue;
    const double* xi = c.b.xt + ((size_t)s * c.V + i) * 3;
    const double* xj = c.b.xt + ((size_t)s * c.V + j) * 3;
    const double vi = c.b.spd[(size_t)s * c.V + i], vj = c.b.spd[(size_t)s * c.V + j];
    const double dx = (xi[0] + p.dt * vi * cos(xi[2])) - (xj[0] + p.dt * vj * cos(xj[2]));
    const double dy = (xi[1] + p.dt * vi * sin(xi[2])) - (xj[1] + p.dt * vj * sin(xj[2]));
    if (!(sqrt(dx * dx + dy * dy) > p.dis_thres)) bad = true;
  }
  return bad;
}

// number of edges (i, j > i) of this outer iteration: the partner collides now and BOTH tables hold a slot (a pair
// with a full table on either side is dropped, status already set -- as find_or_alloc_slot in the oracle)
__device__ __forceinline__ int dev_count_vehicle(const DevCtx& c, int s, int i) {
  const size_t a = (size_t)s * c.V + i;
  const int32_t* nb = c.b.nbr + a * c.D;
  const uint8_t* sh = c.slot_hit + a * c.D;
  int cnt = 0;
  for (int d = 0; d < c.D; d++) {
    const int j = nb[d];
    if (j > i && sh[d] && dev_far_slot(c, s, j, i) >= 0) cnt++;
  }
  return cnt;
}

// emit the `cnt` edges of vehicle i at edge_ij / edge_slot [n0, n0 + cnt), ascending j (the order of
// np.where(edge_mat == 1), casadi/main.py:121)
__device__ __forceinline__ void dev_emit_vehicle(const DevCtx& c, int s, int i, int n0, int cnt) {
  const size_t a = (size_t)s * c.V + i;
  const int32_t* nb = c.b.nbr + a * c.D;
  const uint8_t* sh = c.slot_hit + a * c.D;
  int n = n0, lastj = i, emitted = 0;
  while (emitted < cnt) {
    int j = 0x7fffffff, di = -1;
    for (int d = 0; d < c.D; d++) {
      const int cand = nb[d];
      if (sh[d] && cand > lastj && cand < j) { j = cand; di = d; }
    }
    if (di < 0) break;
    lastj = j;
    const int dj = dev_far_slot(c, s, j, i);
    if (dj < 0) continue;
    int32_t* eij = c.b.edge_ij + ((size_t)s * c.E + n) * 2;
    int32_t* esl = c.b.edge_slot + ((size_t)s * c.E + n) * 2;
    eij[0] = i; eij[1] = j; esl[0] = di; esl[1] = dj;
    n++; emitted++;
  }
}

// end of the detection of scenario s with n edges: casadi/main.py:115-118 ("no collision if no edge links": break)
__device__ __forceinline__ void dev_edges_done(const DevCtx& c, int s, int n) {
  c.b.n_outer[s] += 1;
  c.b.edge_n[s] = n;
  if (n == 0 && c.b.flag[s] == 0) c.b.done[s] = 1;
  else c.b.flag[s] = 1;
}

// exclusive scan over the threads of a CTA (blockDim.x a multiple of 32, <= 1024); wsum: 32 ints of shared memory.
// Returns the exclusive prefix of v; *total = sum over the CTA.
__device__ __forceinline__ int cta_excl_scan(int v, int* wsum, int* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  int inc = v;
  for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(PIADMM_FULL, inc, o); if (lane >= o) inc += t; }
  __syncthreads();                       // wsum may still be read from a previous call
  if (lane == 31) wsum[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int w = lane < nwarps ? wsum[lane] : 0;
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(PIADMM_FULL, w, o); if (lane >= o) w += t; }
    wsum[lane] = w;
  }
  __syncthreads();
  *total = wsum[nwarps - 1];
  return (warp ? wsum[warp - 1] : 0) + inc - v;
}
