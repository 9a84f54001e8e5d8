// piadmm_coop.cuh -- the whole outer PI-ADMM loop of casadi/main.py:52-181 for ANY number of vehicles as ONE persistent
// cooperative kernel (north_star (c): "one persistent kernel runs all outer iterations without host round-trips").
//
//   phase 0   re-initialisation (:52-72), reach discs, candidate matrix            -> grid barrier
//   per outer iteration
//     x       vehicle QPs of the owned vehicles (:81-106)  [+ peer stores]          -> grid barrier (cross-GPU when sharded)
//     slots   exact collision test of the candidates, neighbour slots (:110-113)    -> grid barrier
//     count   valid edges per vehicle                                               -> grid barrier
//     edges   sorted edge list, flag / early exit (:115-118)                        -> grid barrier
//     z       edge problems + dual / PI update (:121-162)  [+ peer stores]          -> grid barrier (cross-GPU when sharded)
//     decide  residuals, stop test, histories, last_hat copy (:165-181)             -> grid barrier
//   loop exit on the device when no scenario is active any more.
//
// Vehicle-sharded multi-GPU form (config 5): the SAME kernel runs on every rank with an ownership range; the exchange is
// fused into the compute phases -- the x phase stores every owned pos_old / primal_u row straight into every peer's
// exchange block over NVLink (peer-mapped memory), the z phase does the same with the per-edge residual terms of the
// edges whose veh1 it owns -- and the two grid barriers that follow them are extended to all ranks by flags in peer
// memory.  No NCCL call, no host round trip inside the step.  Every rank sums the same per-edge terms in the same
// order, so stop decisions and histories are bit-identical on all ranks and for every world size.
#pragma once
#include <cooperative_groups.h>
#include "piadmm_detect.cuh"

namespace cg = cooperative_groups;

#define PIADMM_MAX_PEERS 8

struct PeerCtx {
  int rank, world;                               // world <= 1: single GPU, nothing below is used
  double* pos[PIADMM_MAX_PEERS];                 // [V,2,N+1] exchange copy of pos_old on rank r (as mapped in this process)
  double* u[PIADMM_MAX_PEERS];                   // [V,N]
  double* edge_res[PIADMM_MAX_PEERS];            // [E,4]
  int32_t* stamp[PIADMM_MAX_PEERS];              // [V] veh_stamp
  int32_t* need[PIADMM_MAX_PEERS];               // [V] veh_need
  double* tbox[PIADMM_MAX_PEERS];                // [V,4] trajectory boxes
  unsigned long long* flags[PIADMM_MAX_PEERS];   // [PIADMM_MAX_PEERS] barrier epochs signalled TO rank r, indexed by sender
  unsigned long long* epoch;                     // local: barrier epoch reached by the previous launch
  int* aborted;                                  // local: a barrier timed out (later barriers do not spin)
  unsigned long long timeout_ns;
  double *user_pos, *user_u;                     // the caller's bound buffers: filled from the exchange block at the end
};

#define PIADMM_COOP_TRACE_MAX 1024
struct CoopArgs {
  long long* trace;                              // debug (piadmm_debug_coop_trace): globaltimer at every phase boundary, or NULL
  int* n_active;                                 // [2] scenarios still active after the decide phase (double buffered)
  int slot_gate;                                 // skip the slot phase of vehicles without a recomputed pair (default 1)
  PeerCtx peer;
};

__device__ __forceinline__ unsigned long long coop_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Grid barrier extended to all ranks.  Every thread that stored into peer memory fences system-wide, the grid meets,
// CTA 0 publishes the new epoch in every peer's flag row, and then EVERY CTA waits on the local flag row for the peers'
// epochs (one thread per peer, local L2 polling) -- no second grid barrier.  A peer that does not arrive within
// timeout_ns sets PIADMM_ST_PEER_TIMEOUT and stops all further waiting (the kernel then runs to its end with
// meaningless results instead of hanging the GPU).
__device__ __forceinline__ void coop_cross_barrier(cg::grid_group& grid, const DevCtx& c, const PeerCtx& pc,
                                                   unsigned long long& epoch, bool stored) {
  if (pc.world <= 1) { grid.sync(); return; }
  if (stored) __threadfence_system();
  grid.sync();
  epoch++;
  if ((int)threadIdx.x < pc.world && (int)threadIdx.x != pc.rank) {
    const int r = threadIdx.x;
    if (blockIdx.x == 0) {
      unsigned long long* dst = pc.flags[r] + pc.rank;
      asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(dst), "l"(epoch) : "memory");
    }
    const unsigned long long* src = pc.flags[pc.rank] + r;
    const unsigned long long t0 = coop_now_ns();
    unsigned long long seen = 0;
    int spins = 0;
    for (;;) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(src) : "memory");
      if (seen >= epoch) break;
      if ((++spins & 63) == 0) {
        if (*reinterpret_cast<volatile int*>(pc.aborted)) break;
        if (coop_now_ns() - t0 > pc.timeout_ns) {
          *reinterpret_cast<volatile int*>(pc.aborted) = 1;
          atomicOr(c.b.status, PIADMM_ST_PEER_TIMEOUT);
          break;
        }
      }
    }
  }
  __syncthreads();
}

// ---- decide phase of scenario s by one CTA (any blockDim >= 64): residual sums (casadi/main.py:165-173), MATLAB :202
// distance test, stop decision, histories, alias latch, last_iter_hat_pos copy (:174-181).  The sums are formed as 128
// strided partial sums and one binary tree whatever the CTA size, so k_res_all (128 / 1024 threads) and the cooperative
// kernel (64 threads) produce bit-identical histories.  red: 2 x 128 doubles, ish: 2 ints of shared memory.
// COPY = false: the last_hat copy is left to the caller (the cooperative kernel spreads it over the whole grid:
// ~45 000 elements behind two dependent index loads are 0.7 ms for the 64 threads of one CTA at config 5).
// DIST = false: the MATLAB :202 distance test was evaluated by the caller into c.dist_bad[s] (the cooperative kernel
// does it per vehicle in its count phase: V*D slots are 0.25 ms for one 64-thread CTA at config 5).
template <bool COPY = true, bool DIST = true>
__device__ __forceinline__ void dev_res_all(const DevCtx& c, int s, int i_iter, int* cnt, double (*red)[128], int* ish) {
  const piadmm_params& p = c.p;
  const int N = c.N, T = blockDim.x, t = threadIdx.x;
  const int n = c.b.edge_n[s];
  const double rho = (p.dual_mode == PIADMM_DUAL_PI_OLD && n > 0) ? c.rho_new[s] : c.b.rho_s[s];
  if (t == 0) ish[0] = 0;
  for (int q = t; q < 128; q += T) {
    double sk = 0.0, rk = 0.0;
    for (int e = q; e < n; e += 128) {
      const double* er = c.edge_res + ((size_t)s * c.E + e) * 4;
      if (p.residual_global) { sk += sqrt(rho * rho * (er[0] + er[1])); rk += sqrt(er[2] + er[3]); }
      else { sk += 2.0 * sqrt(rho * rho * er[0]); rk += 2.0 * sqrt(er[2]); }
    }
    red[0][q] = sk; red[1][q] = rk;
  }
  __syncthreads();
  for (int o = 64; o >= 1; o >>= 1) {
    for (int q = t; q < o; q += T) { red[0][q] += red[0][q + o]; red[1][q] += red[1][q + o]; }
    __syncthreads();
  }
  if (!DIST) { if (t == 0 && c.dist_bad[s]) ish[0] = 1; __syncthreads(); }
  if (DIST && p.stop_dist_test) {              // MATLAB :202 over the allocated neighbour slots of this scenario
    for (int q = t; q < c.V * c.D; q += T) {
      const int i = q / c.D, j = c.b.nbr[(size_t)s * c.V * c.D + q];
      if (j <= i) continue;
      const double* xi = c.b.xt + ((size_t)s * c.V + i) * 3;
      const double* xj = c.b.xt + ((size_t)s * c.V + j) * 3;
      const double vi = c.b.spd[(size_t)s * c.V + i], vj = c.b.spd[(size_t)s * c.V + j];
      const double dx = (xi[0] + p.dt * vi * cos(xi[2])) - (xj[0] + p.dt * vj * cos(xj[2]));
      const double dy = (xi[1] + p.dt * vi * sin(xi[2])) - (xj[1] + p.dt * vj * sin(xj[2]));
      if (!(sqrt(dx * dx + dy * dy) > p.dis_thres)) ish[0] = 1;
    }
    __syncthreads();
  }
  if (t == 0) {
    const double rk = red[1][0], sk = red[0][0];
    c.b.res_part[(size_t)s * 2] = rk; c.b.res_part[(size_t)s * 2 + 1] = sk;
    c.b.err_rk[(size_t)s * c.I + i_iter] = rk;
    c.b.err_sk[(size_t)s * c.I + i_iter] = sk;
    c.b.rho_s[s] = rho;
    const bool conv = rk <= p.eps_pri && sk <= p.eps_dual && ish[0] == 0;
    bool fin = conv;
    if (conv) { c.b.curr_iter[s] = i_iter + 1; c.b.done[s] = 1; }
    else {
      if (p.dual_res_alias) c.b.aliased[s] = 1;
      if (p.curr_iter_latest) c.b.curr_iter[s] = i_iter + 1;
      if (i_iter + 1 >= c.I) { c.b.done[s] = 1; fin = true; c.b.status[s] |= PIADMM_ST_OUTER_MAXITER; }
    }
    if (cnt && !fin) atomicAdd(cnt, 1);
    ish[1] = conv ? 1 : 0;
  }
  __syncthreads();
  if (COPY && !ish[1]) {       // last_iter_hat_pos = hat_pos_old: only the slots of this iteration's edges changed
    const int per = 2 * (N + 1);
    for (int q = t; q < n * 2 * per; q += T) {
      const int e = q / (2 * per), side = (q / per) & 1, el = q % per;
      const int v = c.b.edge_ij[((size_t)s * c.E + e) * 2 + side];
      if (v < 0) continue;
      const size_t o = ((((size_t)s * c.V + v) * c.D) + c.b.edge_slot[((size_t)s * c.E + e) * 2 + side]) * per + el;
      c.b.last_hat[o] = c.b.hat[o];
    }
  }
  __syncthreads();     // red / ish are reused by the next scenario of this CTA
}

// ------------------------------------------------------------------ the persistent cooperative kernel
// (Who, load_tabs, groups_per_cta, pairs_per_cta, CTA_THREADS: piadmm.cu, defined before this file is included)
// Two builds of the kernel.  COOP_MIN_CTAS (4 CTAs per SM, ~250 registers, no spills): registers over occupancy, as the
// generic-V phase kernels -- a scene of a few hundred vehicles is ONE dependent chain per phase and every spilled value is
// on it (config 3: 0.413 vs 0.465 ms).  COOP_MANY_CTAS (6 CTAs per SM, 168 registers, 300 B of spill stores): half as many
// more resident QP groups -- a scene whose x phase is several waves of the machine gains more from the extra slots than the
// spills cost (4096 / 16384 / 32768 vehicles: 0.497 / 1.094 / 4.120 -> 0.479 / 0.994 / 3.866 ms).  launch_coop picks by the
// number of vehicle QPs of this rank (and the inner tolerance); PIADMM_COOP_OCC=lo|hi forces one (A/B).  Same arithmetic, identical results.
#ifndef COOP_MIN_CTAS
#define COOP_MIN_CTAS 4
#endif
#ifndef COOP_MANY_CTAS
#define COOP_MANY_CTAS 6
#endif

#define COOP_TRACE()                                                                  \
  do {                                                                                \
    if (ca.trace && tid == 0) {                                                       \
      const int n_ = (int)ca.trace[0];                                                \
      if (n_ < PIADMM_COOP_TRACE_MAX - 2) { ca.trace[1 + n_] = (long long)coop_now_ns(); ca.trace[0] = n_ + 1; } \
    }                                                                                 \
  } while (0)

template <int LPQ, int MIN_CTAS>
__global__ void __launch_bounds__(CTA_THREADS, MIN_CTAS) k_coop(const __grid_constant__ DevCtx c,
                                                                  const __grid_constant__ CoopArgs ca) {
  cg::grid_group grid = cg::this_grid();
  __shared__ __align__(16) double Wsm[PIADMM_NTAB * LPQ * LPQ];
  __shared__ __align__(16) double gbuf[groups_per_cta<LPQ>()][2 * LPQ];
  __shared__ __align__(16) double xch[128];
  __shared__ double red[2][128];
  __shared__ int ish[34];
  Who<LPQ> w;
  const PeerCtx& pc = ca.peer;
  const bool sharded = pc.world > 1;
  const int N = c.N, np1 = N + 1;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (size_t)gridDim.x * blockDim.x;
  const long agents = (long)c.S * c.V;
  unsigned long long epoch = sharded ? *pc.epoch : 0ull;
  const bool memo = dev_memo_on(c);
  // slot phase only for the vehicles with something to re-test (veh_need): needs every vehicle's candidates as a LIST
  // (the candidate kernels ran before this launch; PIADMM_SLOT_GATE=0 switches it off, A/B)
  const bool gate = ca.slot_gate && c.D <= 32 && *reinterpret_cast<volatile int*>(c.cand_over) == 0;
  load_tabs<LPQ>(c, Wsm, 0, PIADMM_NTAB);                // (ends with a CTA barrier)
  COOP_TRACE();                                          // [0] start

  // ---- phase 0: casadi/main.py:52-72.  pos / u are only zeroed for the owned vehicles: the rows of the others are
  // written by their owners' x phase, which may already be running on a faster rank.
  {
    dev_begin_slots(c, tid >> 5, nth >> 5, w.lane);
    const size_t own_n = (size_t)(c.own_hi - c.own_lo);
    for (size_t t = tid; t < (size_t)c.S * own_n * 2 * np1; t += nth) {
      const size_t s = t / (own_n * 2 * np1), r = t % (own_n * 2 * np1);
      c.b.pos[(s * c.V + c.own_lo) * 2 * np1 + r] = 0.0;
    }
    for (size_t t = tid; t < (size_t)c.S * own_n * N; t += nth) {
      const size_t s = t / (own_n * N), r = t % (own_n * N);
      c.b.u[(s * c.V + c.own_lo) * N + r] = 0.0;
    }
    for (size_t t = tid; t < (size_t)c.S * c.I; t += nth) { c.b.err_rk[t] = 0.0; c.b.err_sk[t] = 0.0; }
    for (size_t t = tid; t < (size_t)c.S; t += nth) {
      c.b.flag[t] = 0; c.b.done[t] = 0; c.b.aliased[t] = 0; c.b.curr_iter[t] = 1; c.b.n_outer[t] = 0;
      c.b.status[t] = 0; c.b.veh_inner[t] = 0; c.b.edge_inner[t] = 0; c.b.n_veh_solves[t] = 0;
      c.b.n_edge_solves[t] = 0; c.b.edge_n[t] = 0;
      c.rho_new[t] = c.b.rho_s[t];
      c.dist_bad[t] = 0;
    }
    for (size_t t = tid; t < (size_t)agents; t += nth) {
      c.veh_dirty[t] = 1;
      c.up_cnt[t] = 0;
      const int i = (int)(t % c.V);         // stamps of other ranks' vehicles are theirs to write (iteration 0 solves everybody)
      if (i >= c.own_lo && i < c.own_hi) c.veh_stamp[t] = -1;
    }
    for (size_t t = tid; t < (size_t)c.S * ((c.V + CTA_THREADS - 1) / CTA_THREADS); t += nth) c.chunk_tot[t] = 0;
    if (tid == 0) *c.any_overflow = 0;
  }
  // (the candidate matrix of this MPC step was built by k_disc + k_cand right before this launch: an O(V^2/32) pass that
  //  wants the whole machine at full occupancy, not the 8 warps per SM of this kernel)
  grid.sync();
  COOP_TRACE();                                          // [1] begin done; then 7 stamps per outer iteration:
                                                         //     x, slots, count, edges, z, (cross barrier), decide

#pragma unroll 1
  for (int it = 0; it < c.I; it++) {
    if (tid == 0) ca.n_active[it & 1] = 0;
    // (dist_bad[s] is NOT cleared here: the MATLAB :202 flag of a vehicle is over its ALLOCATED slots, which only grow
    //  during a step, so the flag of the scenario is monotone within the step -- and a vehicle whose slot phase is
    //  skipped below would not set it again)
    if (it > 0) {
      // last_iter_hat_pos = hat_pos_old (casadi/main.py:180) for the scenarios that go on, spread over the grid; the
      // edge list of the previous iteration is intact until the edges phase, last_hat is next read in the z phase
      const int per = 2 * np1;
      for (int s = 0; s < c.S; s++) {
        if (c.b.done[s]) continue;
        const size_t tot = (size_t)c.b.edge_n[s] * 2 * per;
        for (size_t q = tid; q < tot; q += nth) {
          const int e = (int)(q / (2 * per)), side = (int)((q / per) & 1), el = (int)(q % per);
          const int v = c.b.edge_ij[((size_t)s * c.E + e) * 2 + side];
          const size_t o = ((((size_t)s * c.V + v) * c.D) + c.b.edge_slot[((size_t)s * c.E + e) * 2 + side]) * per + el;
          c.b.last_hat[o] = c.b.hat[o];
        }
      }
    }
    // ---- x-update of the owned vehicles, casadi/main.py:81-106
    bool stored = false;
    {
      const long items = (agents + groups_per_cta<LPQ>() - 1) / groups_per_cta<LPQ>();
      for (long item = blockIdx.x; item < items; item += gridDim.x) {
        const long a = item * groups_per_cta<LPQ>() + w.group_in_cta;
        const int s = (int)(a / c.V), i = (int)(a % c.V);
        bool valid = a < agents && !c.b.done[s] && i >= c.own_lo && i < c.own_hi;
        if (valid && memo) valid = c.veh_dirty[a] != 0;             // unchanged QP: u / pos_old (here and on the peers) stay
        if (!__syncthreads_or(valid)) continue;
        if (!__any_sync(PIADMM_FULL, valid)) continue;
        double px, py, uk;
        dev_x_update<LPQ, false>(c, s, i, valid, w.k, w.lane, w.half, Wsm, gbuf[w.group_in_cta], xch, nullptr, &px, &py, &uk);
        if (valid && w.k == 0) { c.veh_dirty[a] = 0; c.veh_stamp[a] = it; }
        {   // bounding box of the new pos_old row (all lanes shuffle; only a solved vehicle stores)
          const bool pt = valid && w.k <= N;
          double x0 = pt ? px : 1e300, x1 = pt ? px : -1e300, y0 = pt ? py : 1e300, y1 = pt ? py : -1e300;
#pragma unroll
          for (int o = LPQ / 2; o >= 1; o >>= 1) {
            x0 = fmin(x0, __shfl_xor_sync(PIADMM_FULL, x0, o)); x1 = fmax(x1, __shfl_xor_sync(PIADMM_FULL, x1, o));
            y0 = fmin(y0, __shfl_xor_sync(PIADMM_FULL, y0, o)); y1 = fmax(y1, __shfl_xor_sync(PIADMM_FULL, y1, o));
          }
          if (valid && w.k == 0) {
            const double4 bx = make_double4(x0, x1, y0, y1);
            reinterpret_cast<double4*>(c.tbox)[a] = bx;
            if (sharded)
              for (int r = 0; r < pc.world; r++) if (r != pc.rank) reinterpret_cast<double4*>(pc.tbox[r])[a] = bx;
          }
        }
        if (gate && valid && it > 0) {      // this vehicle and its candidates have a pair to re-test (iteration 0 tests all)
          const int cnt = c.cand_cnt[a];
          const size_t sv = (size_t)s * c.V;
          for (int q = w.k; q <= cnt; q += LPQ) {
            const int j = q < cnt ? c.cand_list[a * PIADMM_CAND_CAP + q] : i;
            c.veh_need[sv + j] = it;
            if (sharded)
              for (int r = 0; r < pc.world; r++) if (r != pc.rank) pc.need[r][sv + j] = it;
          }
        }
        if (sharded && valid) {       // fused exchange: this lane's point / control of the row, to every peer
          const size_t po = ((size_t)s * c.V + i) * 2 * np1, uo = ((size_t)s * c.V + i) * N;
          const int k = w.k;
          for (int r = 0; r < pc.world; r++) {
            if (r == pc.rank) continue;
            if (k <= N) { pc.pos[r][po + k] = px; pc.pos[r][po + np1 + k] = py; }
            if (k < N) pc.u[r][uo + k] = uk;
            if (k == 0) pc.stamp[r][a] = it;
          }
          stored = true;
        }
      }
    }
    coop_cross_barrier(grid, c, pc, epoch, stored);
    COOP_TRACE();
    // ---- collision edges, casadi/main.py:110-118
    if (!gate || it == 0) {
      for (long a = (long)(tid >> 5); a < agents; a += (long)(nth >> 5)) {
        const int s = (int)(a / c.V), i = (int)(a % c.V);
        if (!c.b.done[s]) dev_slots<true>(c, s, i, w.lane, it);
        else if (w.lane == 0) dev_set_up_cnt(c, s, i, 0);
      }
    } else {
      // only the vehicles marked by this iteration's x phase (here or on a peer) have a pair with a recomputed
      // trajectory; the slot tables, hit flags and edge counts of all the others are those of the previous iteration.
      // Lane l looks at vehicle  warp + l * n_warps  (neighbours in the spatial order go to different warps), then the
      // warp serves the marked ones one after the other.
      const long n_warps = (long)(nth >> 5), wid = (long)(tid >> 5);
      for (long a0 = wid; a0 < agents; a0 += n_warps * 32) {
        const long al = a0 + (long)w.lane * n_warps;
        bool todo = false;
        if (al < agents) {
          const int s = (int)(al / c.V);
          if (c.b.done[s]) dev_set_up_cnt(c, s, (int)(al % c.V), 0);
          else todo = c.veh_need[al] == it;
        }
        unsigned m = __ballot_sync(PIADMM_FULL, todo);
        while (m) {
          const long a = a0 + (long)(__ffs(m) - 1) * n_warps;
          m &= m - 1;
          dev_slots<true>(c, (int)(a / c.V), (int)(a % c.V), w.lane, it);
        }
      }
    }
    grid.sync();
    COOP_TRACE();
    // (the slots phase left the per-vehicle edge counts; only a neighbour-table overflow somewhere in this step makes
    //  them depend on the far side's table: then, and only then, a count phase redoes them -- grid uniform)
    if (*reinterpret_cast<volatile int*>(c.any_overflow)) {
      for (long a = (long)(tid >> 5); a < agents; a += (long)(nth >> 5)) {         // a warp per vehicle, lane = slot
        const int s = (int)(a / c.V), i = (int)(a % c.V);
        if (c.b.done[s]) { if (w.lane == 0) c.up_cnt[a] = 0; continue; }
        dev_count_vehicle_warp(c, s, i, w.lane);
      }
      grid.sync();
    }
    COOP_TRACE();
    {
      const int chunks = (c.V + CTA_THREADS - 1) / CTA_THREADS;
      for (long item = blockIdx.x; item < (long)c.S * chunks; item += gridDim.x) {
        const int s = (int)(item / chunks), ch = (int)(item % chunks);
        if (c.b.done[s]) continue;                                    // CTA uniform
        const int32_t* uc = c.up_cnt + (size_t)s * c.V;
        const int32_t* ct = c.chunk_tot + (size_t)s * chunks;
        int part = 0;
        for (int q = threadIdx.x; q < ch; q += CTA_THREADS) part += ct[q];          // chunk totals kept by dev_set_up_cnt
        int base, total;
        cta_excl_scan(part, ish, &base);                              // base = edges of all lower vehicles
        const int i = ch * CTA_THREADS + threadIdx.x;
        const int cnt = i < c.V ? uc[i] : 0;
        const int off = cta_excl_scan(cnt, ish, &total);
        if (cnt > 0) dev_emit_vehicle(c, s, i, base + off, cnt);
        if (ch == chunks - 1 && threadIdx.x == 0) dev_edges_done(c, s, base + total);
      }
    }
    grid.sync();
    COOP_TRACE();
    // ---- edge problems + dual / PI update, casadi/main.py:121-162 (every edge with an owned endpoint)
    stored = false;
    {
      const long pairs = (long)c.S * c.E, items = (pairs + pairs_per_cta<LPQ>() - 1) / pairs_per_cta<LPQ>();
      for (long item = blockIdx.x; item < items; item += gridDim.x) {
        const long pidx = item * pairs_per_cta<LPQ>() + w.pair_in_cta;
        const int s = (int)(pidx / c.E), e = (int)(pidx % c.E);
        // all pairs of this CTA lie in the same scenario or beyond its edge list: cheap CTA-uniform skip of the tail
        if (pairs_per_cta<LPQ>() == 1 && (s >= c.S || c.b.done[s] || e >= c.b.edge_n[s])) continue;
        bool valid = s < c.S && !c.b.done[s] && e < c.b.edge_n[s];
        const int32_t* eij = c.b.edge_ij + ((size_t)s * c.E + e) * 2;
        const int32_t* esl = c.b.edge_slot + ((size_t)s * c.E + e) * 2;
        int v0 = -1, v1 = -1;
        if (valid) {
          v0 = eij[0]; v1 = eij[1];
          valid = v0 >= 0 && ((v0 >= c.own_lo && v0 < c.own_hi) || (v1 >= c.own_lo && v1 < c.own_hi));
        }
        if (!__syncthreads_or(valid)) continue;
        if (!valid) continue;                     // LPQ==16 only: the other pair of the CTA works alone
        dev_z_update<LPQ, false, true>(c, s, e, v0, v1, esl[0], esl[1], e == c.b.edge_n[s] - 1, true, w.k, w.lane,
                                       w.half, Wsm, gbuf[w.group_in_cta], xch, nullptr, true);
        if (w.k == 0) c.veh_dirty[(size_t)s * c.V + (w.half ? v1 : v0)] = 1;      // this endpoint's QP changed
        if (sharded && w.k == 0 && v0 >= c.own_lo && v0 < c.own_hi) {   // residual terms of this half, to every peer
          const size_t eo = ((size_t)s * c.E + e) * 4;
          const double ss = c.edge_res[eo + w.half], rr = c.edge_res[eo + 2 + w.half];
          for (int r = 0; r < pc.world; r++) {
            if (r == pc.rank) continue;
            pc.edge_res[r][eo + w.half] = ss; pc.edge_res[r][eo + 2 + w.half] = rr;
          }
          stored = true;
        }
      }
    }
    COOP_TRACE();                                        // z phase done on this rank (before the barrier)
    coop_cross_barrier(grid, c, pc, epoch, stored);
    COOP_TRACE();
    // ---- residuals + stop test, casadi/main.py:165-181
    for (int s = blockIdx.x; s < c.S; s += gridDim.x) {
      if (c.b.done[s]) continue;
      dev_res_all<false, false>(c, s, it, ca.n_active + (it & 1), red, ish);
    }
    grid.sync();
    COOP_TRACE();
    if (*reinterpret_cast<volatile int*>(ca.n_active + (it & 1)) == 0) break;     // grid uniform
  }
  if (sharded) {
    // results into the caller's buffers (the exchange block is overwritten by the peers' next step), then one more
    // barrier over all ranks: nobody starts the next step before everybody has copied out
    const size_t n_pos = (size_t)agents * 2 * np1, n_u = (size_t)agents * N;
    if (pc.user_pos) for (size_t t = tid; t < n_pos; t += nth) pc.user_pos[t] = c.b.pos[t];
    if (pc.user_u) for (size_t t = tid; t < n_u; t += nth) pc.user_u[t] = c.b.u[t];
    coop_cross_barrier(grid, c, pc, epoch, false);
    if (tid == 0) *pc.epoch = epoch;
  }
}
