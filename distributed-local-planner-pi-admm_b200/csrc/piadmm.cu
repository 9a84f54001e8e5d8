// piadmm.cu -- kernels + C ABI (include/piadmm.h) of the B200-native PI-ADMM solver.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared -Xcompiler -fPIC
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <string>
#include <algorithm>
#include "piadmm_phases.cuh"
#include "piadmm_detect.cuh"

#define CTA_THREADS 64
#ifndef Z32_MIN_CTAS
#define Z32_MIN_CTAS 4    // the LPQ==32 edge kernel is latency critical (few concurrent edges): registers over occupancy
#endif
#ifndef X32_MIN_CTAS
#define X32_MIN_CTAS 4    // generic-V x-update, LPQ==32: 250 registers, no spills (6 CTAs = 168 registers spill: config 3 1.32 vs 1.20 ms)
#endif
#ifndef Z16_MIN_CTAS
#define Z16_MIN_CTAS 4    // generic-V edge kernel, LPQ==16: latency critical like Z32 (229 registers, no spills; N=15 grid 9.34 -> 7.75 ms)
#endif
#ifndef X16_MIN_CTAS
#define X16_MIN_CTAS 6    // generic-V x-update, LPQ==16: 166 registers, no spills (8 CTAs = 128 registers spill; N=15 scenes 1.24 -> 1.16 / 7.76 -> 7.61 ms)
#endif
#ifndef FUSED32_MIN_CTAS
#define FUSED32_MIN_CTAS 6   // resident CTAs per SM the LPQ==32 fused kernel is compiled for
#endif
#ifndef FUSED16_MIN_CTAS
#define FUSED16_MIN_CTAS 7   // resident CTAs per SM the LPQ==16 fused kernel is compiled for (register budget)
#endif
#define PIADMM_CHUNKS 8   // max sub-batches of piadmm_step_host (copies of one overlap the kernels of the others)

// ------------------------------------------------------------------ thread -> (pair, half, k)
template <int LPQ>
struct Who {
  int warp, lane, k, half, pair_in_cta, group_in_cta;
  __device__ __forceinline__ Who() {
    warp = threadIdx.x >> 5;
    lane = threadIdx.x & 31;
    k = lane & (LPQ - 1);
    if (LPQ == 16) { half = lane >> 4; pair_in_cta = warp; group_in_cta = warp * 2 + half; }
    else { half = warp; pair_in_cta = 0; group_in_cta = warp; }
  }
};
template <int LPQ> __host__ __device__ constexpr int pairs_per_cta() { return LPQ == 16 ? 2 : 1; }
template <int LPQ> __host__ __device__ constexpr int groups_per_cta() { return CTA_THREADS / LPQ; }

template <int LPQ>
__device__ __forceinline__ void pair_sync() {
  if (LPQ == 16) __syncwarp(); else __syncthreads();
}

// copy tables [first, first+count) (see PIADMM_NTAB in piadmm_dev.cuh) to shared memory: 16-byte loads,
// eight in flight per thread before the first store
template <int LPQ>
__device__ __forceinline__ void load_tabs(const DevCtx& c, double* Wsm, int first, int count) {
  const double2* src = reinterpret_cast<const double2*>(c.W + first * LPQ * LPQ);
  double2* dst = reinterpret_cast<double2*>(Wsm + first * LPQ * LPQ);
  const int n2 = count * LPQ * LPQ / 2;
  for (int base = threadIdx.x; base < n2; base += blockDim.x * 8) {
    double2 v[8];
#pragma unroll
    for (int u = 0; u < 8; u++) { const int t = base + u * blockDim.x; if (t < n2) v[u] = __ldg(src + t); }
#pragma unroll
    for (int u = 0; u < 8; u++) { const int t = base + u * blockDim.x; if (t < n2) dst[t] = v[u]; }
  }
  __syncthreads();
}
template <int LPQ>
__device__ __forceinline__ void load_W(const DevCtx& c, double* Wsm) { load_tabs<LPQ>(c, Wsm, 0, PIADMM_NTAB); }

#include "piadmm_coop.cuh"
#include "piadmm_joint.cuh"
static_assert(PIADMM_CHUNK == CTA_THREADS, "chunk totals are per CTA_THREADS vehicles");

// ------------------------------------------------------------------ phase kernels (generic V)
__global__ void k_begin_step(const __grid_constant__ DevCtx c) {
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (size_t)gridDim.x * blockDim.x;
  const int N = c.N;
  dev_begin_slots(c, tid >> 5, nth >> 5, threadIdx.x & 31);
  for (size_t t = tid; t < (size_t)c.S * c.V; t += nth) { c.veh_dirty[t] = 1; c.veh_stamp[t] = -1; }
  const size_t n_pos = (size_t)c.S * c.V * 2 * (N + 1);
  for (size_t t = tid; t < n_pos; t += nth) c.b.pos[t] = 0.0;
  const size_t n_u = (size_t)c.S * c.V * N;
  for (size_t t = tid; t < n_u; t += nth) c.b.u[t] = 0.0;
  const size_t n_h = (size_t)c.S * c.I;
  for (size_t t = tid; t < n_h; t += nth) { c.b.err_rk[t] = 0.0; c.b.err_sk[t] = 0.0; }
  for (size_t t = tid; t < (size_t)c.S; t += nth) {
    c.b.flag[t] = 0; c.b.done[t] = 0; c.b.aliased[t] = 0; c.b.curr_iter[t] = 1; c.b.n_outer[t] = 0;
    c.b.status[t] = 0; c.b.veh_inner[t] = 0; c.b.edge_inner[t] = 0; c.b.n_veh_solves[t] = 0;
    c.b.n_edge_solves[t] = 0; c.b.edge_n[t] = 0;
    c.rho_new[t] = c.b.rho_s[t];
    c.dist_bad[t] = 0;
  }
}

template <int LPQ>
__global__ void __launch_bounds__(CTA_THREADS, LPQ == 16 ? X16_MIN_CTAS : X32_MIN_CTAS) k_x_update(const __grid_constant__ DevCtx c) {
  __shared__ __align__(16) double Wsm[PIADMM_NTAB * LPQ * LPQ];
  __shared__ __align__(16) double gbuf[groups_per_cta<LPQ>()][2 * LPQ];
  __shared__ __align__(16) double xch[128];
  Who<LPQ> w;
  const long agents = (long)c.S * c.V, items = (agents + groups_per_cta<LPQ>() - 1) / groups_per_cta<LPQ>();
  bool loaded = false;
  for (long item = blockIdx.x; item < items; item += gridDim.x) {       // persistent CTAs, grid-stride
    const long a = item * groups_per_cta<LPQ>() + w.group_in_cta;
    const int s = (int)(a / c.V), i = (int)(a % c.V);
    bool valid = a < agents && !c.b.done[s] && i >= c.own_lo && i < c.own_hi;
    if (valid && dev_memo_on(c)) valid = c.veh_dirty[a] != 0;       // unchanged QP: u / pos_old stay as they are
    if (!__syncthreads_or(valid)) continue;  // nothing to do for the whole CTA
    if (!loaded) { load_tabs<LPQ>(c, Wsm, 0, 2); loaded = true; }
    if (!__any_sync(PIADMM_FULL, valid)) continue;
    dev_x_update<LPQ, false>(c, s, i, valid, w.k, w.lane, w.half, Wsm, gbuf[w.group_in_cta], xch, nullptr);
    if (valid && w.k == 0) { c.veh_dirty[a] = 0; c.veh_stamp[a] = c.i_iter; }
  }
}

// trajectory bounding boxes [S,V,4] = (xmin, xmax, ymin, ymax): cheap reject for the O(V^2) pair test.
// One warp per vehicle, lane k = trajectory point k (N+1 <= 32): one coalesced load per coordinate row and a shuffle
// butterfly instead of a serial 31-point loop of dependent loads per thread (9 us -> ~2 us per launch).
__global__ void k_bbox(const __grid_constant__ DevCtx c) {
  const long a = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (a >= (long)c.S * c.V) return;
  const int N = c.N, k = lane <= N ? lane : 0;             // padded lanes repeat point 0
  const double* p = c.b.pos + (size_t)a * 2 * (N + 1);
  const double x = p[k], y = p[N + 1 + k];
  double x0 = x, x1 = x, y0 = y, y1 = y;
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    const double a0 = __shfl_xor_sync(PIADMM_FULL, x0, o), a1 = __shfl_xor_sync(PIADMM_FULL, x1, o);
    const double b0 = __shfl_xor_sync(PIADMM_FULL, y0, o), b1 = __shfl_xor_sync(PIADMM_FULL, y1, o);
    x0 = a0 < x0 ? a0 : x0; x1 = a1 > x1 ? a1 : x1; y0 = b0 < y0 ? b0 : y0; y1 = b1 > y1 ? b1 : y1;
  }
  if (lane == 0) reinterpret_cast<double4*>(c.bbox)[a] = make_double4(x0, x1, y0, y1);
}

// hits[s,i,jb] bit jj = 1 iff vehicle j = 32*jb+jj collides with i (casadi/main.py:110-113); the matrix is stored
// symmetric (both (i,j) and (j,i)) so that rows can be scanned per vehicle.  One thread per (i, word of 32 j's) of
// the UPPER triangle (the test is symmetric): bounding boxes (32 B each, the whole table stays in L1) reject almost
// every pair, the few hits are OR-ed into both rows of the zeroed matrix.
__global__ void k_hits(const __grid_constant__ DevCtx c) {
  const int N = c.N, words = (c.V + 31) / 32;
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long total = (long)c.S * c.V * words;
  if (t >= total) return;
  const int jb = (int)(t % words);
  const long si = t / words;
  const int i = (int)(si % c.V), s = (int)(si / c.V);
  if (jb * 32 + 31 <= i || c.b.done[s]) return;          // no j > i in this word
  const double thr = c.p.thr_squared ? c.p.dis_thres * c.p.dis_thres : c.p.dis_thres;
  const double4* bb = reinterpret_cast<const double4*>(c.bbox) + (size_t)s * c.V;
  const double4 bi = bb[i];
  const double* pi = c.b.pos + (((size_t)s * c.V + i) * 2) * (N + 1);
  uint32_t m = 0;
  const int j0 = jb * 32, j1 = min(j0 + 32, c.V);
  for (int j = max(j0, i + 1); j < j1; j++) {
    bool hit = c.p.always_edge != 0;
    if (!hit) {
      const double4 bj = bb[j];
      const double gx = fmax(0.0, fmax(bi.x - bj.y, bj.x - bi.y)), gy = fmax(0.0, fmax(bi.z - bj.w, bj.z - bi.w));
      if (gx * gx + gy * gy < thr) {           // boxes closer than the threshold: exact test
        const double* pj = c.b.pos + (((size_t)s * c.V + j) * 2) * (N + 1);
        for (int k = 0; k <= N && !hit; k++) {
          const double dx = pi[k] - pj[k], dy = pi[N + 1 + k] - pj[N + 1 + k];
          hit = dx * dx + dy * dy < thr;
        }
      }
    }
    if (hit) {
      m |= 1u << (j - j0);
      atomicOr(c.hits + ((size_t)s * c.V + j) * words + (i >> 5), 1u << (i & 31));
    }
  }
  if (m) atomicOr(c.hits + (size_t)si * words + jb, m);
}

// The same test with one WARP per (i, word): lane = one j.  For small scenes (config 3: 64 vehicles = 128 words) the
// thread-per-word form leaves 128 threads with 32 serial pair tests each (28 us per launch); here every pair has its
// own lane.  Large scenes (config 5: 8.4 M pairs, almost all rejected by the boxes) keep the thread-per-word form.
__global__ void k_hits_warp(const __grid_constant__ DevCtx c) {
  const int N = c.N, words = (c.V + 31) / 32, lane = threadIdx.x & 31;
  const long t = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (t >= (long)c.S * c.V * words) return;
  const int jb = (int)(t % words);
  const long si = t / words;
  const int i = (int)(si % c.V), s = (int)(si / c.V);
  if (jb * 32 + 31 <= i || c.b.done[s]) return;          // warp uniform
  const double thr = c.p.thr_squared ? c.p.dis_thres * c.p.dis_thres : c.p.dis_thres;
  const int j = jb * 32 + lane;
  bool hit = false;
  if (j > i && j < c.V) {
    hit = c.p.always_edge != 0;
    if (!hit) {
      const double4* bb = reinterpret_cast<const double4*>(c.bbox) + (size_t)s * c.V;
      const double4 bi = bb[i], bj = bb[j];
      const double gx = fmax(0.0, fmax(bi.x - bj.y, bj.x - bi.y)), gy = fmax(0.0, fmax(bi.z - bj.w, bj.z - bi.w));
      if (gx * gx + gy * gy < thr) {
        const double* pi = c.b.pos + (((size_t)s * c.V + i) * 2) * (N + 1);
        const double* pj = c.b.pos + (((size_t)s * c.V + j) * 2) * (N + 1);
        for (int k = 0; k <= N && !hit; k++) {
          const double dx = pi[k] - pj[k], dy = pi[N + 1 + k] - pj[N + 1 + k];
          hit = dx * dx + dy * dy < thr;
        }
      }
    }
  }
  const unsigned m = __ballot_sync(PIADMM_FULL, hit);
  if (hit) atomicOr(c.hits + ((size_t)s * c.V + j) * words + (i >> 5), 1u << (i & 31));
  if (lane == 0 && m) atomicOr(c.hits + (size_t)si * words + jb, m);
}

// ---- candidate pairs of the MPC step (piadmm_detect.cuh): reach discs, then the symmetric candidate bit matrix
__global__ void k_disc(const __grid_constant__ DevCtx c) {
  const long a = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (a == 0 && c.cand_over) *c.cand_over = 0;              // set by the list kernel that follows
  if (a < (long)c.S * c.V) dev_disc_vehicle(c, a);
  const int words = (c.V + 31) / 32;
  const long wb = a >> 5;                                  // warp -> (scenario, block of 32 vehicles)
  if (wb < (long)c.S * words) dev_block_box(c, (int)(wb / words), (int)(wb % words), threadIdx.x & 31);
}
__global__ void k_cand(const __grid_constant__ DevCtx c) {
  const int words = (c.V + 31) / 32;
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long)c.S * c.V * words) return;
  const int jb = (int)(t % words);
  const long si = t / words;
  c.cand[t] = dev_cand_word(c, (int)(si / c.V), (int)(si % c.V), jb);
}
__global__ void k_cand_list(const __grid_constant__ DevCtx c) {
  const long a = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (a < (long)c.S * c.V) dev_cand_list_vehicle(c, (int)(a / c.V), (int)(a % c.V), threadIdx.x & 31);
}
__global__ void k_cand_direct(const __grid_constant__ DevCtx c) {
  const long a = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (a < (long)c.S * c.V) dev_cand_direct_vehicle(c, (int)(a / c.V), (int)(a % c.V), threadIdx.x & 31);
}

// per vehicle (one warp): exact test of the candidates (EXACT) / walk of the hit matrix, neighbour slots in ascending j
// (this equals the reference's sequential processing of np.where(edge_mat == 1), row-major)
template <bool EXACT>
__global__ void k_slots(const __grid_constant__ DevCtx c) {
  const long a = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (a >= (long)c.S * c.V) return;
  const int s = (int)(a / c.V), i = (int)(a % c.V);
  if (c.b.done[s]) return;
  dev_slots<EXACT>(c, s, i, threadIdx.x & 31, c.i_iter);
}

// per scenario: count the valid edges per vehicle, scan, emit the sorted edge list; flag / early-exit logic of
// casadi/main.py:115-118
__global__ void k_edges(const __grid_constant__ DevCtx c) {
  const int s = blockIdx.x;
  if (c.b.done[s]) return;
  __shared__ int wsum[32];
  int base = 0;
  for (int i0 = 0; i0 < c.V; i0 += blockDim.x) {
    const int i = i0 + threadIdx.x;
    const int cnt = i < c.V ? dev_count_vehicle(c, s, i) : 0;
    int total;
    const int off = cta_excl_scan(cnt, wsum, &total);
    if (cnt > 0) dev_emit_vehicle(c, s, i, base + off, cnt);
    base += total;
  }
  if (threadIdx.x == 0) dev_edges_done(c, s, base);
}

template <int LPQ>
__global__ void __launch_bounds__(CTA_THREADS, LPQ == 16 ? Z16_MIN_CTAS : Z32_MIN_CTAS) k_z_update(const __grid_constant__ DevCtx c) {
  __shared__ __align__(16) double Wsm[PIADMM_NTAB * LPQ * LPQ];
  __shared__ __align__(16) double gbuf[groups_per_cta<LPQ>()][2 * LPQ];
  __shared__ __align__(16) double xch[128];
  Who<LPQ> w;
  const long pairs = (long)c.S * c.E, items = (pairs + pairs_per_cta<LPQ>() - 1) / pairs_per_cta<LPQ>();
  bool loaded = false;
  for (long item = blockIdx.x; item < items; item += gridDim.x) {       // persistent CTAs, grid-stride
    const long pidx = item * pairs_per_cta<LPQ>() + w.pair_in_cta;
    const int s = (int)(pidx / c.E), e = (int)(pidx % c.E);
    bool valid = s < c.S && !c.b.done[s] && e < c.b.edge_n[s];     // uniform per pair
    const int32_t* eij = c.b.edge_ij + ((size_t)s * c.E + e) * 2;
    const int32_t* esl = c.b.edge_slot + ((size_t)s * c.E + e) * 2;
    if (valid) valid = eij[0] >= 0 && ((eij[0] >= c.own_lo && eij[0] < c.own_hi) || (eij[1] >= c.own_lo && eij[1] < c.own_hi));
    if (!__syncthreads_or(valid)) continue;  // empty CTA
    if (!loaded) { load_tabs<LPQ>(c, Wsm, 2, 2); loaded = true; }
    if (!valid) continue;                     // LPQ==16 only: the other pair of the CTA works alone
    dev_z_update<LPQ, false, true>(c, s, e, eij[0], eij[1], esl[0], esl[1], e == c.b.edge_n[s] - 1, true, w.k, w.lane, w.half,
                             Wsm, gbuf[w.group_in_cta], xch, nullptr, true);
    if (w.k == 0) c.veh_dirty[(size_t)s * c.V + eij[w.half]] = 1;       // this endpoint's QP changed
  }
}

// residuals, casadi/main.py:165-173: rank-local partial sums over the edges whose veh1 is owned.
__global__ void k_res_partial(const __grid_constant__ DevCtx c) {
  const int s = blockIdx.x;
  if (c.b.done[s]) return;
  const piadmm_params& p = c.p;
  __shared__ double red[2][128];
  const int n = c.b.edge_n[s];
  const double rho = (p.dual_mode == PIADMM_DUAL_PI_OLD && n > 0) ? c.rho_new[s] : c.b.rho_s[s];
  double sk = 0.0, rk = 0.0;
  for (int e = threadIdx.x; e < n; e += blockDim.x) {
    const int v0 = c.b.edge_ij[((size_t)s * c.E + e) * 2];
    if (v0 < c.own_lo || v0 >= c.own_hi) continue;
    const double* er = c.edge_res + ((size_t)s * c.E + e) * 4;
    if (p.residual_global) { sk += sqrt(rho * rho * (er[0] + er[1])); rk += sqrt(er[2] + er[3]); }
    else { sk += 2.0 * sqrt(rho * rho * er[0]); rk += 2.0 * sqrt(er[2]); }
  }
  red[0][threadIdx.x] = sk; red[1][threadIdx.x] = rk;
  __syncthreads();
  for (int o = blockDim.x / 2; o >= 1; o >>= 1) {
    if ((int)threadIdx.x < o) {
      red[0][threadIdx.x] += red[0][threadIdx.x + o];
      red[1][threadIdx.x] += red[1][threadIdx.x + o];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) { c.b.res_part[(size_t)s * 2] = red[1][0]; c.b.res_part[(size_t)s * 2 + 1] = red[0][0]; }
}

// stop test, alias latch, last_hat copy: casadi/main.py:174-181.  One CTA per scenario; res_part holds
// the (all-reduced) sums.
__global__ void k_res_decide(const __grid_constant__ DevCtx c) {
  const int s = blockIdx.x, N = c.N;
  if (c.b.done[s]) return;
  const piadmm_params& p = c.p;
  __shared__ int decision;
  const int n = c.b.edge_n[s];
  const double rho = (p.dual_mode == PIADMM_DUAL_PI_OLD && n > 0) ? c.rho_new[s] : c.b.rho_s[s];
  if (threadIdx.x == 0) {
    const double rk = c.b.res_part[(size_t)s * 2], sk = c.b.res_part[(size_t)s * 2 + 1];
    c.b.err_rk[(size_t)s * c.I + c.i_iter] = rk;
    c.b.err_sk[(size_t)s * c.I + c.i_iter] = sk;
    c.b.rho_s[s] = rho;
    const bool conv = rk <= p.eps_pri && sk <= p.eps_dual && c.dist_bad[s] == 0;
    c.dist_bad[s] = 0;
    if (conv) { c.b.curr_iter[s] = c.i_iter + 1; c.b.done[s] = 1; }
    else {
      if (p.dual_res_alias) c.b.aliased[s] = 1;
      if (p.curr_iter_latest) c.b.curr_iter[s] = c.i_iter + 1;
      if (c.i_iter + 1 >= c.I) { c.b.done[s] = 1; c.b.status[s] |= PIADMM_ST_OUTER_MAXITER; }      // iteration cap reached
    }
    decision = conv ? 1 : 0;
  }
  __syncthreads();
  if (!decision) {
    // last_iter_hat_pos = hat_pos_old: only the slots of this iteration's edges changed
    const int per = 2 * (N + 1);
    for (int t = threadIdx.x; t < n * 2 * per; t += blockDim.x) {
      const int e = t / (2 * per), side = (t / per) & 1, el = t % per;
      const int v = c.b.edge_ij[((size_t)s * c.E + e) * 2 + side];
      if (v < 0) continue;
      const size_t o = ((((size_t)s * c.V + v) * c.D) + c.b.edge_slot[((size_t)s * c.E + e) * 2 + side]) * per + el;
      c.b.last_hat[o] = c.b.hat[o];
    }
  }
}

// Single-rank form of the three kernels above and below (k_res_partial -> k_dist_test -> k_res_decide) plus the
// active-scenario count, in ONE launch (one CTA per scenario): the phases of a small coupled scene are a few
// microseconds each, so four launches of dependent tiny kernels are most of their time.  The residual sums are formed
// exactly as k_res_partial forms them (128 strided partial sums, the same tree), so the histories are bit-identical.
__global__ void k_res_all(const __grid_constant__ DevCtx c, int* cnt) {
  const int s = blockIdx.x;
  if (c.b.done[s]) return;
  __shared__ double red[2][128];
  __shared__ int ish[2];
  dev_res_all(c, s, c.i_iter, cnt, red, ish);       // piadmm_coop.cuh: the same statements the cooperative kernel runs
}

// MATLAB :202 extra stop condition, evaluated over every allocated neighbour slot (all CTAs of the grid)
__global__ void k_dist_test(const __grid_constant__ DevCtx c) {
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long)c.S * c.V * c.D) return;
  const int s = (int)(t / ((long)c.V * c.D)), i = (int)((t / c.D) % c.V), j = c.b.nbr[t];
  if (j <= i || c.b.done[s]) return;
  const piadmm_params& p = c.p;
  const double* xi = c.b.xt + ((size_t)s * c.V + i) * 3;
  const double* xj = c.b.xt + ((size_t)s * c.V + j) * 3;
  const double vi = c.b.spd[(size_t)s * c.V + i], vj = c.b.spd[(size_t)s * c.V + j];
  const double dx = (xi[0] + p.dt * vi * cos(xi[2])) - (xj[0] + p.dt * vj * cos(xj[2]));
  const double dy = (xi[1] + p.dt * vi * sin(xi[2])) - (xj[1] + p.dt * vj * sin(xj[2]));
  if (!(sqrt(dx * dx + dy * dy) > p.dis_thres)) c.dist_bad[s] = 1;
}

__global__ void k_count_active(const __grid_constant__ DevCtx c, int* out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const bool act = t < c.S && !c.b.done[t];
  const unsigned m = __ballot_sync(PIADMM_FULL, act);
  if ((threadIdx.x & 31) == 0 && m) atomicAdd(out, __popc(m));
}

// refwin[s,v,c,k] = ref[s,v,c,*step + k], k = 0..N: the window cost_function_primal slices out of the resident
// reference (casadi/PI_ADMM_class.py:116-117), scenarios [s_lo,s_hi)
__global__ void k_refwin(const __grid_constant__ DevCtx c, const double* __restrict__ ref, int T, const int* step) {
  const int np1 = c.N + 1, st = *step;
  const size_t rows0 = (size_t)c.s_lo * c.V * 2, n = (size_t)(c.s_hi - c.s_lo) * c.V * 2 * np1;
  double* out = const_cast<double*>(c.b.refwin);
  for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (size_t)gridDim.x * blockDim.x) {
    const size_t row = rows0 + t / np1;
    const int k = (int)(t % np1);
    out[row * np1 + k] = ref[row * T + st + k];
  }
}

// end of one MPC step of the on-device closed loop: append the histories of casadi/main.py:186,199-201, advance
// the plant state (xt <- xt_next) and the step index.  t = position of the step in the history arrays.
__global__ void k_advance(const __grid_constant__ DevCtx c, piadmm_histories o, int t, int* step) {
  const long a = (long)blockIdx.x * blockDim.x + threadIdx.x, SV = (long)c.S * c.V;
  if (a == 0) *step += 1;
  if (a >= SV) return;
  const int s = (int)(a / c.V);
  const double* xn = c.b.xt_next + a * 3;
  double* xt = const_cast<double*>(c.b.xt) + a * 3;
  const double x = xn[0], y = xn[1], th = xn[2];
  if (o.x_vec) { o.x_vec[((size_t)t * SV + a) * 2] = x; o.x_vec[((size_t)t * SV + a) * 2 + 1] = y; }
  if (o.theta_vec) o.theta_vec[(size_t)t * SV + a] = th;
  if (o.u_vec) o.u_vec[(size_t)t * SV + a] = c.b.u[a * c.N];
  if (a % c.V == 0) {
    if (o.iter_his) o.iter_his[(size_t)t * c.S + s] = c.b.curr_iter[s];
    if (o.n_outer_his) o.n_outer_his[(size_t)t * c.S + s] = c.b.n_outer[s];
    if (o.status_his) o.status_his[(size_t)t * c.S + s] = c.b.status[s];
    if (o.err_rk_his) { const int n = c.b.n_outer[s]; o.err_rk_his[(size_t)t * c.S + s] = n > 0 ? c.b.err_rk[(size_t)s * c.I + n - 1] : 0.0; }
  }
  xt[0] = x; xt[1] = y; xt[2] = th;
}

__global__ void k_plant(const __grid_constant__ DevCtx c) {
  const long a = (long)c.s_lo * c.V + (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= (long)c.s_hi * c.V) return;
  const double* xt = c.b.xt + a * 3;
  const double v = c.b.spd[a];
  double* o = c.b.xt_next + a * 3;
  o[0] = xt[0] + v * cos(xt[2]) * c.p.dt;
  o[1] = xt[1] + v * sin(xt[2]) * c.p.dt;
  o[2] = xt[2] + v / c.p.L * c.b.u[a * c.N] * c.p.dt;
}

// ------------------------------------------------------------------ fused V=2 persistent kernel
// One pair per scenario runs casadi/main.py:52-181 start to finish: no host round trips, state
// stays in L1/L2 between phases (each scenario touches ~10 KB).
template <int LPQ, bool KS>
__device__ __forceinline__ void run_scenario_v2(const DevCtx& c, int s, const Who<LPQ>& w, const double* Wsm,
                                                double* gbuf, double* xch, double* krow, int step);

// end of MPC step t of scenario s inside the fused kernel (what k_advance does for the whole batch between launches)
template <int LPQ>
__device__ __forceinline__ void advance_scenario_v2(const DevCtx& c, int s, const Who<LPQ>& w, int t) {
  const piadmm_histories& o = c.hist;
  const size_t SV = (size_t)c.S * 2, a = (size_t)s * 2 + w.half;
  if (w.k == 0) {
    const double* xn = c.b.xt_next + a * 3;
    double* xt = const_cast<double*>(c.b.xt) + a * 3;
    const double x = xn[0], y = xn[1], th = xn[2];
    if (o.x_vec) { o.x_vec[((size_t)t * SV + a) * 2] = x; o.x_vec[((size_t)t * SV + a) * 2 + 1] = y; }
    if (o.theta_vec) o.theta_vec[(size_t)t * SV + a] = th;
    if (o.u_vec) o.u_vec[(size_t)t * SV + a] = c.b.u[a * c.N];
    xt[0] = x; xt[1] = y; xt[2] = th;
    if (w.half == 0) {
      if (o.iter_his) o.iter_his[(size_t)t * c.S + s] = c.b.curr_iter[s];
      if (o.n_outer_his) o.n_outer_his[(size_t)t * c.S + s] = c.b.n_outer[s];
      if (o.status_his) o.status_his[(size_t)t * c.S + s] = c.b.status[s];
      if (o.err_rk_his) { const int n = c.b.n_outer[s]; o.err_rk_his[(size_t)t * c.S + s] = n > 0 ? c.b.err_rk[(size_t)s * c.I + n - 1] : 0.0; }
    }
  }
}

// scenario queue: LPQ==16 one scenario per warp, LPQ==32 one per 64-thread CTA
template <int LPQ>
__device__ __forceinline__ int next_scenario(const DevCtx& c, const Who<LPQ>& w, int* s_next) {
  int s;
  if (LPQ == 16) {
    s = 0;
    if (w.lane == 0) {
      s = c.s_lo + atomicAdd(c.queue, 1);
      if (c.order && s < c.s_hi) s = c.order[s];
    }
    s = __shfl_sync(PIADMM_FULL, s, 0);
  } else {
    if (threadIdx.x == 0) {
      int q = c.s_lo + atomicAdd(c.queue, 1);
      if (c.order && q < c.s_hi) q = c.order[q];
      *s_next = q;
    }
    __syncthreads();
    s = *s_next;
    __syncthreads();
  }
  return s;
}

// Scheduling hint for the NEXT step of the same scenarios: order[s_lo..s_hi) = the scenarios of the range sorted
// by the work they needed in this step, longest first (counting sort on a 512-bucket cost histogram, one CTA).
// Scenario run times differ by up to ~2.7x (outer iterations 1..3, inner iterations 25..225 per solve); with
// about two scenarios per resident warp slot, starting the long ones first removes most of the tail in which
// SMs idle while the last warps finish.  Results never depend on the order.
#define ORDER_BUCKETS 512
__device__ __forceinline__ int order_bucket(const DevCtx& c, int s, float scale) {
  const float cost = (float)c.b.veh_inner[s] + 4.0f * (float)c.b.edge_inner[s];
  const int b = ORDER_BUCKETS - 1 - (int)(cost * scale);                 // bucket 0 = longest
  return b < 0 ? 0 : b;
}
__global__ void __launch_bounds__(1024) k_order(const __grid_constant__ DevCtx c, int* order) {
  __shared__ unsigned hist[ORDER_BUCKETS];
  __shared__ float wmax[32];
  const int n = c.s_hi - c.s_lo, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int b = threadIdx.x; b < ORDER_BUCKETS; b += blockDim.x) hist[b] = 0;
  float m = 1.0f;
  for (int t = threadIdx.x; t < n; t += blockDim.x)
    m = fmaxf(m, (float)c.b.veh_inner[c.s_lo + t] + 4.0f * (float)c.b.edge_inner[c.s_lo + t]);
  for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(PIADMM_FULL, m, o));
  if (lane == 0) wmax[warp] = m;
  __syncthreads();
  m = wmax[lane];
  for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(PIADMM_FULL, m, o));
  const float scale = (ORDER_BUCKETS - 1) / m;
  for (int t = threadIdx.x; t < n; t += blockDim.x) atomicAdd(&hist[order_bucket(c, c.s_lo + t, scale)], 1u);
  __syncthreads();
  if (warp == 0) {                            // exclusive scan of the 512 bucket counts by one warp
    unsigned v[ORDER_BUCKETS / 32], tot = 0;
#pragma unroll
    for (int i = 0; i < ORDER_BUCKETS / 32; i++) { v[i] = hist[lane * (ORDER_BUCKETS / 32) + i]; tot += v[i]; }
    unsigned inc = tot;
    for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(PIADMM_FULL, inc, o); if (lane >= o) inc += t; }
    unsigned run = inc - tot;
#pragma unroll
    for (int i = 0; i < ORDER_BUCKETS / 32; i++) { hist[lane * (ORDER_BUCKETS / 32) + i] = run; run += v[i]; }
  }
  __syncthreads();
  for (int t = threadIdx.x; t < n; t += blockDim.x) {
    const unsigned pos = atomicAdd(&hist[order_bucket(c, c.s_lo + t, scale)], 1u);
    order[c.s_lo + pos] = c.s_lo + t;
  }
  if (threadIdx.x == 0) *c.queue = 0;          // the fused kernel before this one is done: re-arm its work queue
}

// Debug build (-DPIADMM_TRACE, tools/trace_step.py): start / end time and SM of every scenario of the last fused launch
#ifdef PIADMM_TRACE
__device__ long long g_trace[8192 * 4];
__device__ __forceinline__ long long gtime() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
extern "C" int piadmm_debug_read_trace(long long* host, int n) {
  return (int)cudaMemcpyFromSymbol(host, g_trace, sizeof(long long) * 4 * n);
}
#endif
// FEW: the same code compiled for registers instead of occupancy (4 CTAs per SM: ~250 registers, no spills), launched
// when the batch has at most two CTAs per SM -- config 1 is ONE scenario, i.e. one dependent chain whose spills are
// pure latency.  Same PTX arithmetic, so results are bit-identical to the throughput build.
template <int LPQ, bool FEW = false>
#ifdef FUSED16_MAXNREG   // A/B build: cap the registers of the LPQ == 16 throughput build directly (NVCC_EXTRA=-DFUSED16_MAXNREG=112)
__global__ void __launch_bounds__(CTA_THREADS) __maxnreg__(FEW ? 248 : (LPQ == 16 ? FUSED16_MAXNREG : 168)) k_fused_v2(const __grid_constant__ DevCtx c) {
#else
__global__ void __launch_bounds__(CTA_THREADS, FEW ? 4 : (LPQ == 16 ? FUSED16_MIN_CTAS : FUSED32_MIN_CTAS)) k_fused_v2(const __grid_constant__ DevCtx c) {
#endif
  __shared__ __align__(16) double Wsm[PIADMM_NTAB * LPQ * LPQ];
  __shared__ __align__(16) double gbuf[groups_per_cta<LPQ>()][2 * LPQ];
  __shared__ __align__(16) double xch[128];
  __shared__ int s_next;
  __shared__ int step_sh;
  if (c.ref && threadIdx.x == 0) step_sh = *reinterpret_cast<const volatile int*>(c.step);   // overlaps the table load
  load_W<LPQ>(c, Wsm);                       // (ends with a CTA barrier)
  const int step = c.ref ? step_sh : 0;
  Who<LPQ> w;
  // persistent: every pair pulls scenarios from a global counter until the batch is exhausted
  const int nloop = c.loop_steps > 0 ? c.loop_steps : 1;
  for (;;) {
    const int s = next_scenario<LPQ>(c, w, &s_next);
    if (s >= c.s_hi) break;
#ifdef PIADMM_TRACE
    const long long t0 = gtime();
#endif
#pragma unroll 1
    for (int t = 0; t < nloop; t++) {
      run_scenario_v2<LPQ, false>(c, s, w, Wsm, gbuf[w.group_in_cta], xch, nullptr, step + t);
      if (c.loop_steps > 0) {           // histories of casadi/main.py:186,199-201 and xt <- xt_next, by the scenario itself
        pair_sync<LPQ>();
        advance_scenario_v2<LPQ>(c, s, w, t);
        pair_sync<LPQ>();
      }
    }
#ifdef PIADMM_TRACE
    if (w.lane == 0 && s < 8192) {
      unsigned sm; asm("mov.u32 %0, %%smid;" : "=r"(sm));
      g_trace[s * 4] = t0; g_trace[s * 4 + 1] = gtime(); g_trace[s * 4 + 2] = sm; g_trace[s * 4 + 3] = 0;
    }
#endif
  }
}

// Throughput variant: K^-1 rows in shared memory, tables from global memory (see inner_solve).
//   LPQ==16: 128-thread CTAs (4 scenario-warps), 7 per SM = 28 warps per SM (4144 scenario slots on 148 SMs)
//   LPQ==32: 64-thread CTAs (one scenario), up to 12 per SM
#define KS_THREADS(LPQ) ((LPQ) == 16 ? 128 : 64)
#define KS_GROUPS(LPQ) (KS_THREADS(LPQ) / (LPQ))
template <int LPQ>
__global__ void __launch_bounds__(KS_THREADS(LPQ), LPQ == 16 ? 7 : 12) k_fused_ks(const __grid_constant__ DevCtx c) {
  __shared__ __align__(16) double krows[KS_GROUPS(LPQ)][LPQ * PIADMM_KROW_STRIDE(LPQ)];
  __shared__ __align__(16) double gbuf[KS_GROUPS(LPQ)][2 * LPQ];
  __shared__ __align__(16) double xch[128];
  __shared__ int s_next;
  Who<LPQ> w;
  for (;;) {
    const int s = next_scenario<LPQ>(c, w, &s_next);
    if (s >= c.s_hi) break;
    run_scenario_v2<LPQ, true>(c, s, w, c.W, gbuf[w.group_in_cta], xch,
                               krows[w.group_in_cta] + w.k * PIADMM_KROW_STRIDE(LPQ), c.ref ? *c.step : 0);
  }
}

template <int LPQ, bool KS>
__device__ __forceinline__ void run_scenario_v2(const DevCtx& c, int s, const Who<LPQ>& w, const double* Wsm,
                                                double* gbuf, double* xch, double* krow, int step) {
  const piadmm_params& p = c.p;
  const int N = c.N, k = w.k, half = w.half;
  const int i = half, j = half ^ 1;
  const bool lead = half == 0 && k == 0;
  const size_t so = SLOT_OFF(s, i, 0);
  // ---- reference window of vehicle i out of the resident reference (as k_refwin; casadi/PI_ADMM_class.py:116-117)
  if (c.ref && k <= N) {
    const size_t row = ((size_t)s * 2 + i) * 2;
    double* rw = const_cast<double*>(c.b.refwin) + POS_OFF(s, i);
    rw[k] = c.ref[row * c.ref_T + step + k];
    rw[NP1 + k] = c.ref[(row + 1) * c.ref_T + step + k];
  }
  // ---- casadi/main.py:52-72: re-initialise local / edge / dual variables
  if (k <= N) {
    for (int cc = 0; cc < 2; cc++) {
      const size_t o2 = so + (size_t)cc * NP1 + k;
      c.b.hat[o2] = 0.0; c.b.dual[o2] = 0.0; c.b.integ[o2] = 0.0; c.b.diff[o2] = 0.0; c.b.last_hat[o2] = 0.0;
      c.b.pos[POS_OFF(s, i) + (size_t)cc * NP1 + k] = 0.0;
    }
  }
  if (k < N) c.b.u[((size_t)s * 2 + i) * N + k] = 0.0;
  if (k == 0) c.b.nbr[(size_t)s * 2 + i] = p.al_all ? j : -1;
  for (int t = k + half * LPQ; t < c.I; t += 2 * LPQ) {
    c.b.err_rk[(size_t)s * c.I + t] = 0.0;
    c.b.err_sk[(size_t)s * c.I + t] = 0.0;
  }
  if (lead) {
    c.b.aliased[s] = 0; c.b.status[s] = 0; c.b.veh_inner[s] = 0; c.b.edge_inner[s] = 0;
    c.b.n_veh_solves[s] = 0; c.b.n_edge_solves[s] = 0; c.b.done[s] = 0;
    c.rho_new[s] = c.b.rho_s[s];
  }
  pair_sync<LPQ>();
  int flag = 0, curr_iter = 1, n_outer = 0, edge_n = 0, capped = 1;
  const double thr = p.thr_squared ? p.dis_thres * p.dis_thres : p.dis_thres;
  bool dist_ok = true;
  if (p.stop_dist_test) {
    const double* xi = c.b.xt + ((size_t)s * 2) * 3;
    const double* xj = xi + 3;
    const double vi = c.b.spd[(size_t)s * 2], vj = c.b.spd[(size_t)s * 2 + 1];
    const double dx = (xi[0] + p.dt * vi * cos(xi[2])) - (xj[0] + p.dt * vj * cos(xj[2]));
    const double dy = (xi[1] + p.dt * vi * sin(xi[2])) - (xj[1] + p.dt * vj * sin(xj[2]));
    dist_ok = sqrt(dx * dx + dy * dy) > p.dis_thres;
  }
  const bool memo = dev_memo_on(c);
  bool dirty = true;                            // x_memo: both QPs change together (one edge) or not at all
#pragma unroll 1
  for (int i_iter = 0; i_iter < c.I; i_iter++) {
    n_outer++;
    if (dirty || !memo) dev_x_update<LPQ, KS>(c, s, i, true, k, w.lane, half, Wsm, gbuf, xch, krow);
    dirty = false;
    pair_sync<LPQ>();
    // ---- collision test, casadi/main.py:110-118
    bool near = false;
    if (k <= N) {
      const double* pi = c.b.pos + POS_OFF(s, i);
      const double* pj = c.b.pos + POS_OFF(s, j);
      const double dx = pi[k] - pj[k], dy = pi[NP1 + k] - pj[NP1 + k];
      near = dx * dx + dy * dy < thr;
    }
    const bool hit = p.always_edge || group_any<LPQ>(near, w.lane);
    edge_n = hit ? 1 : 0;
    if (!hit && !flag) { capped = 0; break; }
    flag = 1;
    double sk = 0.0, rk = 0.0;
    if (hit) {
      if (k == 0) c.b.nbr[(size_t)s * 2 + i] = j;
      dev_z_update<LPQ, KS>(c, s, 0, 0, 1, 0, 0, true, true, k, w.lane, half, Wsm, gbuf, xch, krow, true);
      dirty = true;
      pair_sync<LPQ>();
      const double* er = c.edge_res + (size_t)s * c.E * 4;
      const double rho = p.dual_mode == PIADMM_DUAL_PI_OLD ? c.rho_new[s] : c.b.rho_s[s];
      if (p.residual_global) { sk = sqrt(rho * rho * (er[0] + er[1])); rk = sqrt(er[2] + er[3]); }
      else { sk = 2.0 * sqrt(rho * rho * er[0]); rk = 2.0 * sqrt(er[2]); }
      pair_sync<LPQ>();                       // everyone has read rho_s / edge_res
      if (lead) c.b.rho_s[s] = rho;
    }
    if (lead) { c.b.err_rk[(size_t)s * c.I + i_iter] = rk; c.b.err_sk[(size_t)s * c.I + i_iter] = sk; }
    if (rk <= p.eps_pri && sk <= p.eps_dual && dist_ok) { curr_iter = i_iter + 1; capped = 0; break; }
    if (p.dual_res_alias && lead) c.b.aliased[s] = 1;
    if (k <= N) {
      c.b.last_hat[so + k] = c.b.hat[so + k];
      c.b.last_hat[so + NP1 + k] = c.b.hat[so + NP1 + k];
    }
    if (p.curr_iter_latest) curr_iter = i_iter + 1;
    pair_sync<LPQ>();
  }
  if (c.host_out) {
    // every lane re-reads what it wrote itself (u_k, the two coordinates of point k): no synchronisation needed
    const size_t uo = ((size_t)s * 2 + i) * N, po = POS_OFF(s, i);
    if (c.ho_u && k < N) c.ho_u[uo + k] = c.b.u[uo + k];
    if (c.ho_pos && k <= N) { c.ho_pos[po + k] = c.b.pos[po + k]; c.ho_pos[po + NP1 + k] = c.b.pos[po + NP1 + k]; }
    if (k == 0) {                              // plant step of vehicle i, casadi/main.py:185-192 (as k_plant)
      const double* xt = c.b.xt + ((size_t)s * 2 + i) * 3;
      const double v = c.b.spd[(size_t)s * 2 + i];
      const double nx = xt[0] + v * cos(xt[2]) * p.dt, ny = xt[1] + v * sin(xt[2]) * p.dt;
      const double nt = xt[2] + v / p.L * c.b.u[uo] * p.dt;
      double* o = c.b.xt_next + ((size_t)s * 2 + i) * 3;
      o[0] = nx; o[1] = ny; o[2] = nt;
      if (c.ho_xt_next) { double* ho = c.ho_xt_next + ((size_t)s * 2 + i) * 3; ho[0] = nx; ho[1] = ny; ho[2] = nt; }
    }
    if (lead) {
      if (c.ho_curr_iter) c.ho_curr_iter[s] = curr_iter;
      if (c.ho_n_outer) c.ho_n_outer[s] = n_outer;
    }
  }
  if (lead) {
    if (capped) atomicOr(c.b.status + s, PIADMM_ST_OUTER_MAXITER);
    c.b.flag[s] = flag; c.b.done[s] = 1; c.b.curr_iter[s] = curr_iter; c.b.n_outer[s] = n_outer;
    c.b.edge_n[s] = edge_n;
    if (edge_n) {
      c.b.edge_ij[(size_t)s * c.E * 2] = 0; c.b.edge_ij[(size_t)s * c.E * 2 + 1] = 1;
      c.b.edge_slot[(size_t)s * c.E * 2] = 0; c.b.edge_slot[(size_t)s * c.E * 2 + 1] = 0;
    }
  }
}

// ------------------------------------------------------------------ FP64 FMA peak micro-benchmark
__global__ void k_fp64_peak(double* out, int iters) {
  double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6,
         a7 = a0 + 7;
  const double b = 1.0000001, cc = 1e-9;
  for (int i = 0; i < iters; i++) {
    a0 = fma(a0, b, cc); a1 = fma(a1, b, cc); a2 = fma(a2, b, cc); a3 = fma(a3, b, cc);
    a4 = fma(a4, b, cc); a5 = fma(a5, b, cc); a6 = fma(a6, b, cc); a7 = fma(a7, b, cc);
  }
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
}

// ------------------------------------------------------------------ host side / C ABI
struct piadmm_handle {
  DevCtx c;
  int device;
  bool bound;
  int lpq;
  double* d_W;
  int* d_count;
  int *d_count2, *h_count2;    // piadmm_step (generic V): double-buffered active counts (device / pinned host)
  cudaEvent_t ev_count[2];
  int* d_queue;
  int fused_grid, x_grid, z_grid;
  int sm_count;
  bool fused_few;             // small batches use the registers-over-occupancy build of the fused kernel (PIADMM_FEW=0 disables)
  bool fused_ks;              // fused kernel variant: K^-1 rows in shared memory (k_fused_ks)
  int* d_order;               // [S] longest-first processing order learnt from the previous step
  bool use_order;             // PIADMM_LPT=0 disables the hint
  bool order_valid[PIADMM_CHUNKS + 1];   // per launch range: [0] whole batch, [1+ch] sub-batch ch
  bool queue_armed[PIADMM_CHUNKS + 1];   // the range's work counter was reset on the device by the last k_order
  cudaStream_t cs[PIADMM_CHUNKS], cs_main;
  cudaEvent_t ev_start, ev_user, ev_done[PIADMM_CHUNKS];
  int n_chunks;               // sub-batches used by piadmm_step_host (PIADMM_NCHUNKS, default 2)
  double* d_ref;              // resident reference trajectories [S,V,2,T] (piadmm_set_reference), library owned
  int ref_T;
  int* h_step;                // pinned host word holding num_step of the current piadmm_step_host_at call
  int* d_step;                //   ... and its device copy (a graph memcpy node refreshes it)
  // cached CUDA graph of the whole host-buffer step (copies + kernels of all sub-batches): one cudaGraphLaunch
  // instead of ~45 stream API calls per step
  bool use_graph;             // PIADMM_GRAPH=0 disables
  bool zero_copy;             // PIADMM_ZEROCOPY=0 disables the zero-copy epilogue of the host-buffer entry points
  cudaGraphExec_t gexec;
  const void* gkey[9];        // host pointers + order-validity bits the graph was captured for
  bool gvalid_after[PIADMM_CHUNKS + 1];
  int64_t glaunches;
  bool have_streams;
  int64_t launches;
  // persistent cooperative kernel for generic V (piadmm_coop.cuh); PIADMM_COOP=0 selects the phase kernels
  bool use_coop;
  bool cand_matrix;           // PIADMM_CAND=matrix: candidate bit matrix + list kernel instead of k_cand_direct (A/B)
  int coop_grid_max;          // co-resident CTAs of k_coop<lpq, COOP_MIN_CTAS>
  int coop_grid_max_hi;       // ... of k_coop<lpq, COOP_MANY_CTAS> (the occupancy build, large scenes)
  int coop_occ;               // PIADMM_COOP_OCC: 0 choose by size, 1 lo, 2 hi
  int* d_nactive;             // [2]
  long long* d_trace;         // piadmm_debug_coop_trace
  // vehicle-sharded multi-GPU exchange (piadmm_peer_export / piadmm_peer_connect)
  char* xblk;                 // library-owned exchange block of THIS rank: pos | u | edge_res | tbox | stamp | need | flags | epoch | aborted
  bool peer_on;
  void* peer_base[PIADMM_MAX_PEERS];   // the peers' exchange blocks as mapped into this process (cudaIpcOpenMemHandle)
  PeerCtx peer;
  std::string err;
};

// layout of an exchange block (S == 1): doubles first, then the barrier words
struct XblkLayout { size_t pos, u, er, tbox, stamp, need, flags, epoch, aborted, bytes; };
static XblkLayout xblk_layout(int V, int N, int E) {
  XblkLayout L;
  L.pos = 0;
  L.u = L.pos + sizeof(double) * (size_t)V * 2 * (N + 1);
  L.er = L.u + sizeof(double) * (size_t)V * N;
  L.tbox = ((L.er + sizeof(double) * (size_t)E * 4 + 31) / 32) * 32;        // double4 accesses
  L.stamp = L.tbox + sizeof(double) * (size_t)V * 4;
  L.need = L.stamp + ((sizeof(int32_t) * (size_t)V + 15) / 16) * 16;
  L.flags = L.need + ((sizeof(int32_t) * (size_t)V + 15) / 16) * 16;
  L.epoch = L.flags + sizeof(unsigned long long) * PIADMM_MAX_PEERS;
  L.aborted = L.epoch + sizeof(unsigned long long);
  L.bytes = L.aborted + 16;
  return L;
}

static thread_local std::string g_create_err;

#define CK(call)                                                                        \
  do {                                                                                  \
    cudaError_t e_ = (call);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                      \
      return PIADMM_E_CUDA;                                                             \
    }                                                                                   \
  } while (0)

static double pow10i(int d) {
  if (d < 0) return 0.0;
  double s = 1.0;
  for (int i = 0; i < d; i++) s *= 10.0;
  return s;
}

static int check_params(const piadmm_params* p, int V, int D, std::string& err) {
  if (!p) { err = "params is NULL"; return PIADMM_E_ARG; }
  if (p->num_ho < 3 || p->num_ho > 31) { err = "num_ho must be in [3,31]"; return PIADMM_E_ARG; }
  if (p->iter_num < 1) { err = "iter_num must be >= 1"; return PIADMM_E_ARG; }
  if (p->in_check < 1 || p->in_max_iter < 1) { err = "in_check / in_max_iter must be >= 1"; return PIADMM_E_ARG; }
  if (!(p->in_rho0 > 0) || !(p->in_sigma > 0) || !(p->in_alpha > 0) || !(p->dt > 0) || !(p->L > 0)) {
    err = "in_rho0, in_sigma, in_alpha, dt, L must be > 0"; return PIADMM_E_ARG;
  }
  if (p->dual_mode < 0 || p->dual_mode > 2) { err = "bad dual_mode"; return PIADMM_E_ARG; }
  if (p->al_all && D < V - 1) { err = "al_all needs D >= V-1"; return PIADMM_E_ARG; }
  return PIADMM_OK;
}

static void drop_graph(piadmm_handle* h) {       // kernel arguments are baked into the captured graph
  if (h->gexec) { cudaGraphExecDestroy(h->gexec); h->gexec = nullptr; }
}

extern "C" {

int piadmm_abi_version(void) { return PIADMM_ABI_VERSION; }

int piadmm_max_edges(int V, int D) { int e = V * D / 2; return e < 1 ? 1 : e; }

void piadmm_default_params(piadmm_params* p) {
  memset(p, 0, sizeof *p);
  p->dt = 0.1; p->L = 1.0; p->dis_thres = 2.0; p->beta = 10.0; p->Pnorm = 1.0; p->Pcost = 1.0;
  p->rho = 2.0; p->eps_pri = 20.0; p->eps_dual = 1.0;
  p->theta1 = 5.0; p->theta2 = 3.0; p->kI = -1.0; p->windup_sat = 30.0; p->backcalc = 1.0;
  p->K_I = 3.0; p->KP_num = 5.0; p->KP_max = 2.5; p->rho_num = 4.0; p->rho_min = 1.0; p->rho_max = 5.0;
  p->in_rho0 = 0.1; p->in_sigma = 1e-6; p->in_alpha = 1.6; p->in_eps_abs = 1e-3; p->in_eps_rel = 1e-3;
  p->in_adapt_tol = 5.0;
  p->num_ho = 15; p->iter_num = 100; p->dual_mode = PIADMM_DUAL_PLAIN;
  p->thr_squared = 0; p->dual_res_alias = 1; p->stop_dist_test = 0; p->pos_nonlinear = 0;
  p->curr_iter_latest = 0; p->round_u = 4; p->round_edge = 4; p->al_all = 1; p->always_edge = 0;
  p->residual_global = 0; p->PI = 1; p->windup = 1; p->trad = 0;
  p->in_max_iter = 4000; p->in_check = 25; p->in_adaptive = 1; p->x_memo = 1;
}

const char* piadmm_last_error(const piadmm_handle* h) { return h ? h->err.c_str() : g_create_err.c_str(); }

int piadmm_create(const piadmm_params* p, int device, int S, int V, int D, piadmm_handle** out) {
  if (!out) return PIADMM_E_ARG;
  *out = nullptr;
  if (S < 1 || V < 1 || D < 1) { g_create_err = "S, V, D must be >= 1"; return PIADMM_E_ARG; }
  int rc = check_params(p, V, D, g_create_err);
  if (rc) return rc;
  int ndev = 0;
  cudaError_t ce = cudaGetDeviceCount(&ndev);
  if (ce != cudaSuccess || ndev < 1 || device < 0 || device >= ndev) {
    g_create_err = std::string("no usable CUDA device (there is no CPU fallback): ") +
                   (ce != cudaSuccess ? cudaGetErrorString(ce) : "bad device index");
    return PIADMM_E_CUDA;
  }
  piadmm_handle* h = new piadmm_handle();
  memset(&h->c, 0, sizeof h->c);
  h->device = device; h->bound = false; h->launches = 0; h->d_W = nullptr; h->d_count = nullptr; h->d_count2 = nullptr; h->h_count2 = nullptr; h->ev_count[0] = h->ev_count[1] = nullptr; h->d_queue = nullptr; h->d_order = nullptr; h->fused_grid = 0; h->have_streams = false; h->gexec = nullptr; h->d_ref = nullptr; h->ref_T = 0; h->h_step = nullptr; h->d_step = nullptr;
  { const char* e_ = getenv("PIADMM_CAND"); h->cand_matrix = e_ && strcmp(e_, "matrix") == 0; }
  h->use_coop = true; h->coop_grid_max = 0; h->coop_grid_max_hi = 0; h->coop_occ = 0; h->d_nactive = nullptr; h->d_trace = nullptr; h->xblk = nullptr; h->peer_on = false;
  memset(&h->peer, 0, sizeof h->peer);
  for (int i = 0; i < PIADMM_MAX_PEERS; i++) h->peer_base[i] = nullptr;
  h->c.p = *p; set_inner_consts(h->c.ik, p->in_alpha); h->c.S = S; h->c.V = V; h->c.D = D; h->c.E = piadmm_max_edges(V, D);
  h->c.N = p->num_ho; h->c.I = p->iter_num;
  h->c.own_lo = 0; h->c.own_hi = V;
  h->c.s_lo = 0; h->c.s_hi = S;
  h->c.round_u_scale = pow10i(p->round_u); h->c.round_edge_scale = pow10i(p->round_edge);
  h->lpq = p->num_ho <= 15 ? 16 : 32;
  auto fail = [&](cudaError_t e, const char* what) {
    g_create_err = std::string(what) + ": " + cudaGetErrorString(e);
    piadmm_destroy(h);
    return PIADMM_E_CUDA;
  };
  if ((ce = cudaSetDevice(device)) != cudaSuccess) return fail(ce, "cudaSetDevice");
  // W = M'M, zero padded to LPQ x LPQ
  const int L = h->lpq, N = p->num_ho;
  double* W = (double*)calloc((size_t)PIADMM_NTAB * L * L, sizeof(double));
  for (int i = 0; i < N; i++)
    for (int j = 0; j < N; j++) {
      double s = 0;
      for (int k = 0; k <= N; k++) {
        int mi = k - 1 - i, mj = k - 1 - j;
        if (mi > 0 && mj > 0) s += (double)mi * (double)mj;
      }
      W[i * L + j] = s;
    }
  for (int r = 0; r + 2 < N; r++) {          // D2'D2, D2 row r = e_r - 2 e_{r+1} + e_{r+2}
    const double w[3] = {1.0, -2.0, 1.0};
    for (int a = 0; a < 3; a++)
      for (int b = 0; b < 3; b++) W[(size_t)L * L + (r + a) * L + (r + b)] += w[a] * w[b];
  }
  {   // position-space tables of the edge problem: B1 = D'D, B2 = D' T D, T = A'A = 2I + 2 D1'D1
    double* Dm = (double*)calloc((size_t)N * N, sizeof(double));
    double* Tm = (double*)calloc((size_t)N * N, sizeof(double));
    double* TD = (double*)calloc((size_t)N * N, sizeof(double));
    for (int j = 0; j < N; j++) {
      Dm[j * N + j] = 1.0;
      if (j <= N - 2) { if (j >= 1) Dm[j * N + j - 1] = -2.0; if (j >= 2) Dm[j * N + j - 2] = 1.0; }
    }
    for (int j = 0; j < N; j++) Tm[j * N + j] = 2.0;
    for (int r = 0; r + 1 < N; r++) {
      Tm[r * N + r] += 2.0; Tm[(r + 1) * N + r + 1] += 2.0; Tm[r * N + r + 1] -= 2.0; Tm[(r + 1) * N + r] -= 2.0;
    }
    for (int i = 0; i < N; i++)
      for (int j = 0; j < N; j++) { double s = 0; for (int m = 0; m < N; m++) s += Tm[i * N + m] * Dm[m * N + j]; TD[i * N + j] = s; }
    for (int i = 0; i < N; i++)
      for (int j = 0; j < N; j++) {
        double s1 = 0, s2 = 0;
        for (int m = 0; m < N; m++) { s1 += Dm[m * N + i] * Dm[m * N + j]; s2 += Dm[m * N + i] * TD[m * N + j]; }
        W[(size_t)2 * L * L + i * L + j] = s1;
        W[(size_t)3 * L * L + i * L + j] = s2;
      }
    free(Dm); free(Tm); free(TD);
  }
  ce = cudaMalloc(&h->d_W, sizeof(double) * PIADMM_NTAB * L * L);
  if (ce == cudaSuccess) ce = cudaMemcpy(h->d_W, W, sizeof(double) * PIADMM_NTAB * L * L, cudaMemcpyHostToDevice);
  free(W);
  if (ce != cudaSuccess) return fail(ce, "W upload");
  h->c.W = h->d_W;
  const int words = (V + 31) / 32;
  if ((ce = cudaMalloc(&h->c.edge_res, sizeof(double) * (size_t)S * h->c.E * 4)) != cudaSuccess) return fail(ce, "edge_res");
  if ((ce = cudaMemset(h->c.edge_res, 0, sizeof(double) * (size_t)S * h->c.E * 4)) != cudaSuccess) return fail(ce, "edge_res");
  if ((ce = cudaMalloc(&h->c.rho_new, sizeof(double) * S)) != cudaSuccess) return fail(ce, "rho_new");
  if ((ce = cudaMalloc(&h->c.hits, sizeof(uint32_t) * (size_t)S * V * words)) != cudaSuccess) return fail(ce, "hits");
  if ((ce = cudaMalloc(&h->c.cand, sizeof(uint32_t) * (size_t)S * V * words)) != cudaSuccess) return fail(ce, "cand");
  if ((ce = cudaMalloc(&h->c.slot_hit, (size_t)S * V * D)) != cudaSuccess) return fail(ce, "slot_hit");
  if ((ce = cudaMemset(h->c.slot_hit, 0, (size_t)S * V * D)) != cudaSuccess) return fail(ce, "slot_hit");
  {   // A/B: PIADMM_DETECT = cand (default) | matrix_warp | matrix_thread ; PIADMM_HITS=1|2 is the round-1 spelling of the last two
    const char* e = getenv("PIADMM_DETECT");
    const char* e2 = getenv("PIADMM_HITS");
    h->c.detect_mode = 0;
    if (e && strcmp(e, "matrix_warp") == 0) h->c.detect_mode = 1;
    else if (e && strcmp(e, "matrix_thread") == 0) h->c.detect_mode = 2;
    else if (e && strcmp(e, "matrix") == 0) h->c.detect_mode = (long)S * V * words <= 9472 ? 1 : 2;
    else if (!e && e2 && atoi(e2) >= 1 && atoi(e2) <= 2) h->c.detect_mode = atoi(e2);
  }
  if ((ce = cudaMalloc(&h->c.cand_list, sizeof(int32_t) * (size_t)S * V * PIADMM_CAND_CAP)) != cudaSuccess) return fail(ce, "cand_list");
  if ((ce = cudaMalloc(&h->c.cand_cnt, sizeof(int32_t) * (size_t)S * V)) != cudaSuccess) return fail(ce, "cand_cnt");
  if ((ce = cudaMemset(h->c.cand_cnt, 0xff, sizeof(int32_t) * (size_t)S * V)) != cudaSuccess) return fail(ce, "cand_cnt");
  if ((ce = cudaMalloc(&h->c.veh_stamp, sizeof(int32_t) * (size_t)S * V)) != cudaSuccess) return fail(ce, "veh_stamp");
  if ((ce = cudaMalloc(&h->c.tbox, sizeof(double) * 4 * (size_t)S * V)) != cudaSuccess) return fail(ce, "tbox");
  if ((ce = cudaMalloc(&h->c.veh_need, sizeof(int32_t) * (size_t)S * V)) != cudaSuccess) return fail(ce, "veh_need");
  if ((ce = cudaMemset(h->c.veh_need, 0, sizeof(int32_t) * (size_t)S * V)) != cudaSuccess) return fail(ce, "veh_need");
  {
    const size_t nct = (size_t)S * ((V + CTA_THREADS - 1) / CTA_THREADS);
    if ((ce = cudaMalloc(&h->c.chunk_tot, sizeof(int32_t) * nct)) != cudaSuccess) return fail(ce, "chunk_tot");
    if ((ce = cudaMemset(h->c.chunk_tot, 0, sizeof(int32_t) * nct)) != cudaSuccess) return fail(ce, "chunk_tot");
  }
  if ((ce = cudaMalloc(&h->c.blockbox, sizeof(double) * 4 * (size_t)S * words)) != cudaSuccess) return fail(ce, "blockbox");
  if ((ce = cudaMalloc(&h->c.veh_dirty, (size_t)S * V)) != cudaSuccess) return fail(ce, "veh_dirty");
  if ((ce = cudaMemset(h->c.veh_dirty, 1, (size_t)S * V)) != cudaSuccess) return fail(ce, "veh_dirty");
  if ((ce = cudaMalloc(&h->c.up_cnt, sizeof(int32_t) * (size_t)S * V)) != cudaSuccess) return fail(ce, "up_cnt");
  if ((ce = cudaMalloc(&h->c.bbox, sizeof(double) * 4 * (size_t)S * V)) != cudaSuccess) return fail(ce, "bbox");
  if ((ce = cudaMalloc(&h->c.dist_bad, sizeof(int) * (size_t)S)) != cudaSuccess) return fail(ce, "dist_bad");
  if ((ce = cudaMemset(h->c.dist_bad, 0, sizeof(int) * (size_t)S)) != cudaSuccess) return fail(ce, "dist_bad");
  if ((ce = cudaMalloc(&h->d_count, sizeof(int))) != cudaSuccess) return fail(ce, "count");
  if ((ce = cudaMalloc(&h->d_count2, 2 * sizeof(int))) != cudaSuccess) return fail(ce, "count2");
  if ((ce = cudaMallocHost(&h->h_count2, 2 * sizeof(int))) != cudaSuccess) return fail(ce, "count2");
  for (int i = 0; i < 2; i++)
    if ((ce = cudaEventCreateWithFlags(&h->ev_count[i], cudaEventDisableTiming)) != cudaSuccess) return fail(ce, "event");
  if ((ce = cudaMalloc(&h->d_queue, sizeof(int) * (PIADMM_CHUNKS + 1))) != cudaSuccess) return fail(ce, "queue");
  if ((ce = cudaMalloc(&h->d_order, sizeof(int) * (size_t)S)) != cudaSuccess) return fail(ce, "order");
  {   // K^-1(rho0) cache (PIADMM_KCACHE=0 disables): zero keys never match (cI > 0)
    const char* e = getenv("PIADMM_KCACHE");
    if (!(e && strcmp(e, "0") == 0)) {
      const size_t nb = sizeof(double) * (size_t)S * V * 2 * PIADMM_KC_STRIDE(h->lpq);
      if ((ce = cudaMalloc(&h->c.kcache, nb)) != cudaSuccess) return fail(ce, "kcache");
      if ((ce = cudaMemset(h->c.kcache, 0, nb)) != cudaSuccess) return fail(ce, "kcache");
    }
  }
  {
    const char* lpt = getenv("PIADMM_LPT");
    h->use_order = !(lpt && strcmp(lpt, "0") == 0);
    for (int i = 0; i <= PIADMM_CHUNKS; i++) { h->order_valid[i] = false; h->queue_armed[i] = false; }
  }
  for (int i = 0; i < PIADMM_CHUNKS; i++) {
    if ((ce = cudaStreamCreateWithFlags(&h->cs[i], cudaStreamNonBlocking)) != cudaSuccess) return fail(ce, "stream");
    if ((ce = cudaEventCreateWithFlags(&h->ev_done[i], cudaEventDisableTiming)) != cudaSuccess) return fail(ce, "event");
  }
  if ((ce = cudaEventCreateWithFlags(&h->ev_start, cudaEventDisableTiming)) != cudaSuccess) return fail(ce, "event");
  if ((ce = cudaEventCreateWithFlags(&h->ev_user, cudaEventDisableTiming)) != cudaSuccess) return fail(ce, "event");
  if ((ce = cudaStreamCreateWithFlags(&h->cs_main, cudaStreamNonBlocking)) != cudaSuccess) return fail(ce, "stream");
  {
    const char* e = getenv("PIADMM_NCHUNKS");
    int n = e ? atoi(e) : 2;
    h->n_chunks = n < 1 ? 1 : (n > PIADMM_CHUNKS ? PIADMM_CHUNKS : n);
    e = getenv("PIADMM_GRAPH");
    h->use_graph = !(e && strcmp(e, "0") == 0);
    e = getenv("PIADMM_ZEROCOPY");
    h->zero_copy = !(e && strcmp(e, "0") == 0);
    h->gexec = nullptr; h->glaunches = 0;
  }
  if ((ce = cudaMallocHost(&h->h_step, sizeof(int))) != cudaSuccess) return fail(ce, "h_step");
  if ((ce = cudaMalloc(&h->d_step, sizeof(int))) != cudaSuccess) return fail(ce, "d_step");
  h->have_streams = true;
  h->c.queue = h->d_queue;
  {
    cudaDeviceProp prop;
    int per_sm = 0;
    if ((ce = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return fail(ce, "props");
    // A/B switch: "ks" = K^-1 rows in shared memory (twice the resident warps, but the LDS traffic of the
    // mat-vec then saturates the shared-memory pipe: measured 0.371 ms vs 0.336 ms per step, profiles/r1_history.md)
    const char* fv = getenv("PIADMM_FUSED");
    h->fused_ks = fv && strcmp(fv, "ks") == 0;
    h->sm_count = prop.multiProcessorCount;
    const char* few = getenv("PIADMM_FEW");
    h->fused_few = !(few && strcmp(few, "0") == 0);
    if (h->fused_ks) {
      if (h->lpq == 16) ce = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_fused_ks<16>, KS_THREADS(16), 0);
      else ce = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_fused_ks<32>, KS_THREADS(32), 0);
    } else if (h->lpq == 16) ce = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_fused_v2<16>, CTA_THREADS, 0);
    else ce = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_fused_v2<32>, CTA_THREADS, 0);
    if (ce != cudaSuccess || per_sm < 1) return fail(ce, "occupancy");
    h->fused_grid = prop.multiProcessorCount * per_sm;     // one resident wave, persistent CTAs
    int px = 0, pz = 0;
    if (h->lpq == 16) {
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&px, k_x_update<16>, CTA_THREADS, 0);
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&pz, k_z_update<16>, CTA_THREADS, 0);
    } else {
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&px, k_x_update<32>, CTA_THREADS, 0);
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&pz, k_z_update<32>, CTA_THREADS, 0);
    }
    h->x_grid = prop.multiProcessorCount * (px < 1 ? 1 : px);
    h->z_grid = prop.multiProcessorCount * (pz < 1 ? 1 : pz);
    int pc = 0;
    int pc_hi = 0;
    if (h->lpq == 16) {
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&pc, k_coop<16, COOP_MIN_CTAS>, CTA_THREADS, 0);
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&pc_hi, k_coop<16, COOP_MANY_CTAS>, CTA_THREADS, 0);
    } else {
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&pc, k_coop<32, COOP_MIN_CTAS>, CTA_THREADS, 0);
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&pc_hi, k_coop<32, COOP_MANY_CTAS>, CTA_THREADS, 0);
    }
    h->coop_grid_max_hi = prop.multiProcessorCount * (pc_hi < 1 ? 0 : pc_hi);
    { const char* e_ = getenv("PIADMM_COOP_OCC"); h->coop_occ = e_ && strcmp(e_, "lo") == 0 ? 1 : (e_ && strcmp(e_, "hi") == 0 ? 2 : 0); }
    int coop_ok = 0;
    cudaDeviceGetAttribute(&coop_ok, cudaDevAttrCooperativeLaunch, device);
    h->coop_grid_max = prop.multiProcessorCount * (pc < 1 ? 0 : pc);
    const char* ce_ = getenv("PIADMM_COOP");
    h->use_coop = coop_ok && h->coop_grid_max > 0 && !(ce_ && strcmp(ce_, "0") == 0);
    if ((ce = cudaMalloc(&h->d_nactive, 4 * sizeof(int))) != cudaSuccess) return fail(ce, "nactive");
    if ((ce = cudaMemset(h->d_nactive, 0, 4 * sizeof(int))) != cudaSuccess) return fail(ce, "nactive");
    h->c.cand_over = h->d_nactive + 3;
  }
  *out = h;
  return PIADMM_OK;
}

void piadmm_destroy(piadmm_handle* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaFree(h->d_W); cudaFree(h->c.edge_res); cudaFree(h->c.rho_new); cudaFree(h->c.hits); cudaFree(h->c.cand); cudaFree(h->c.slot_hit); cudaFree(h->c.veh_dirty); cudaFree(h->c.blockbox); cudaFree(h->c.edge_u); cudaFree(h->c.cand_list); cudaFree(h->c.cand_cnt); cudaFree(h->c.veh_stamp); cudaFree(h->c.veh_need); cudaFree(h->c.chunk_tot); cudaFree(h->c.tbox); cudaFree(h->c.up_cnt); cudaFree(h->c.bbox); cudaFree(h->c.dist_bad); cudaFree(h->c.kcache);
  cudaFree(h->d_count); cudaFree(h->d_count2); if (h->h_count2) cudaFreeHost(h->h_count2);
  for (int i = 0; i < 2; i++) if (h->ev_count[i]) cudaEventDestroy(h->ev_count[i]);
  piadmm_peer_disconnect(h);
  cudaFree(h->xblk); cudaFree(h->d_nactive); cudaFree(h->d_trace);
  cudaFree(h->d_queue); cudaFree(h->d_order); cudaFree(h->d_ref); cudaFree(h->d_step);
  if (h->h_step) cudaFreeHost(h->h_step);
  if (h->have_streams) {
    for (int i = 0; i < PIADMM_CHUNKS; i++) { cudaStreamDestroy(h->cs[i]); cudaEventDestroy(h->ev_done[i]); }
    cudaEventDestroy(h->ev_start); cudaEventDestroy(h->ev_user); cudaStreamDestroy(h->cs_main);
    if (h->gexec) cudaGraphExecDestroy(h->gexec);
  }
  delete h;
}

int piadmm_set_params(piadmm_handle* h, const piadmm_params* p) {
  if (!h) return PIADMM_E_ARG;
  int rc = check_params(p, h->c.V, h->c.D, h->err);
  if (rc) return rc;
  if (p->num_ho != h->c.N || p->iter_num != h->c.I) { h->err = "num_ho / iter_num are fixed at create time"; return PIADMM_E_ARG; }
  h->c.p = *p;
  set_inner_consts(h->c.ik, p->in_alpha);
  h->c.round_u_scale = pow10i(p->round_u); h->c.round_edge_scale = pow10i(p->round_edge);
  drop_graph(h);
  return PIADMM_OK;
}

int piadmm_bind(piadmm_handle* h, const piadmm_buffers* b) {
  if (!h || !b) return PIADMM_E_ARG;
  const void* const* ptrs = reinterpret_cast<const void* const*>(b);
  for (size_t i = 0; i < sizeof(piadmm_buffers) / sizeof(void*); i++)
    if (!ptrs[i]) { h->err = "piadmm_bind: NULL buffer at field index " + std::to_string(i); return PIADMM_E_UNBOUND; }
  h->c.b = *b;
  h->c.slots_dirty = 1;
  h->bound = true;
  drop_graph(h);
  return PIADMM_OK;
}

int64_t piadmm_launch_count(const piadmm_handle* h) { return h ? h->launches : 0; }

#define REQUIRE_BOUND()                                                       \
  do {                                                                        \
    if (!h) return PIADMM_E_ARG;                                              \
    if (!h->bound) { h->err = "buffers not bound"; return PIADMM_E_UNBOUND; } \
    CK(cudaSetDevice(h->device));                                             \
  } while (0)

#define LAUNCHED()              \
  do {                          \
    h->launches++;              \
    CK(cudaGetLastError());     \
  } while (0)

// candidate pairs of this MPC step (detect_mode 0): once per step, from the plant state
static int build_candidates(piadmm_handle* h, cudaStream_t st) {
  const long agents = (long)h->c.S * h->c.V, wtot = agents * ((h->c.V + 31) / 32);
  const long nt_disc = std::max(agents, (long)h->c.S * ((h->c.V + 31) / 32) * 32);
  k_disc<<<(unsigned)((nt_disc + 255) / 256), 256, 0, st>>>(h->c);
  LAUNCHED();
  if (h->c.D <= 32 && !h->cand_matrix) {       // lists straight from the discs (PIADMM_CAND=matrix: the two-kernel form, A/B)
    k_cand_direct<<<(unsigned)((agents * 32 + 127) / 128), 128, 0, st>>>(h->c);
    LAUNCHED();
    return PIADMM_OK;
  }
  k_cand<<<(unsigned)((wtot + 255) / 256), 256, 0, st>>>(h->c);
  LAUNCHED();
  k_cand_list<<<(unsigned)((agents * 32 + 255) / 256), 256, 0, st>>>(h->c);
  LAUNCHED();
  return PIADMM_OK;
}

int piadmm_begin_step(piadmm_handle* h, void* stream) {
  REQUIRE_BOUND();
  cudaStream_t st = (cudaStream_t)stream;
  h->c.i_iter = 0;
  k_begin_step<<<h->sm_count * 4, 256, 0, st>>>(h->c);
  LAUNCHED();
  h->c.slots_dirty = 0;
  if (h->c.detect_mode == 0) return build_candidates(h, st);   // (the fused V=2 kernel never comes through here)
  return PIADMM_OK;
}

int piadmm_x_update(piadmm_handle* h, void* stream) {
  REQUIRE_BOUND();
  cudaStream_t st = (cudaStream_t)stream;
  const long agents = (long)h->c.S * h->c.V;
  if (h->lpq == 16) {
    const int g = groups_per_cta<16>();
    long items = (agents + g - 1) / g;
    k_x_update<16><<<(unsigned)(items < h->x_grid ? items : h->x_grid), CTA_THREADS, 0, st>>>(h->c);
  } else {
    const int g = groups_per_cta<32>();
    long items = (agents + g - 1) / g;
    k_x_update<32><<<(unsigned)(items < h->x_grid ? items : h->x_grid), CTA_THREADS, 0, st>>>(h->c);
  }
  LAUNCHED();
  return PIADMM_OK;
}

int piadmm_detect_edges(piadmm_handle* h, void* stream) {
  REQUIRE_BOUND();
  cudaStream_t st = (cudaStream_t)stream;
  const int words = (h->c.V + 31) / 32;
  const long agents = (long)h->c.S * h->c.V, warps = agents * words;
  if (h->c.detect_mode == 0) {
    k_slots<true><<<(unsigned)((agents * 32 + 255) / 256), 256, 0, st>>>(h->c);
    LAUNCHED();
  } else {                      // round-1 path: bounding boxes + the full pair test in every outer iteration
    k_bbox<<<(unsigned)((agents * 32 + 127) / 128), 128, 0, st>>>(h->c);
    LAUNCHED();
    CK(cudaMemsetAsync(h->c.hits, 0, sizeof(uint32_t) * (size_t)warps, st));
    if (h->c.detect_mode == 1) k_hits_warp<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, st>>>(h->c);   // a lane per pair
    else k_hits<<<(unsigned)((warps + 255) / 256), 256, 0, st>>>(h->c);                                    // a thread per 32 pairs
    LAUNCHED();
    k_slots<false><<<(unsigned)((agents * 32 + 255) / 256), 256, 0, st>>>(h->c);
    LAUNCHED();
  }
  k_edges<<<h->c.S, h->c.V > 512 ? 1024 : (h->c.V > 128 ? 256 : 128), 0, st>>>(h->c);
  LAUNCHED();
  return PIADMM_OK;
}

int piadmm_z_update(piadmm_handle* h, void* stream) {
  REQUIRE_BOUND();
  cudaStream_t st = (cudaStream_t)stream;
  const long pairs = (long)h->c.S * h->c.E;
  if (h->lpq == 16) {
    const int g = pairs_per_cta<16>();
    long items = (pairs + g - 1) / g;
    k_z_update<16><<<(unsigned)(items < h->z_grid ? items : h->z_grid), CTA_THREADS, 0, st>>>(h->c);
  } else {
    k_z_update<32><<<(unsigned)(pairs < h->z_grid ? pairs : h->z_grid), CTA_THREADS, 0, st>>>(h->c);
  }
  LAUNCHED();
  return PIADMM_OK;
}

// casadi/main.py:133-153 for a caller-provided edge list: the edge problems alone (no roll-out of z-hat, no dual update)
int piadmm_solve_edges(piadmm_handle* h, void* stream) {
  REQUIRE_BOUND();
  if (!h->c.edge_u) {
    CK(cudaMalloc(&h->c.edge_u, sizeof(double) * (size_t)h->c.S * h->c.E * 2 * h->c.N));
    CK(cudaMemset(h->c.edge_u, 0, sizeof(double) * (size_t)h->c.S * h->c.E * 2 * h->c.N));
  }
  h->c.z_solve_only = 1;
  const int rc = piadmm_z_update(h, stream);
  h->c.z_solve_only = 0;
  return rc;
}

int piadmm_get_edge_u(piadmm_handle* h, double* host_out, void* stream) {
  REQUIRE_BOUND();
  if (!host_out) return PIADMM_E_ARG;
  if (!h->c.edge_u) { h->err = "piadmm_get_edge_u: no edge solve yet (piadmm_solve_edges)"; return PIADMM_E_UNBOUND; }
  cudaStream_t st = (cudaStream_t)stream;
  CK(cudaMemcpyAsync(host_out, h->c.edge_u, sizeof(double) * (size_t)h->c.S * h->c.E * 2 * h->c.N, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return PIADMM_OK;
}

int piadmm_solve_joint(piadmm_handle* h, const int32_t* pair_n, const int32_t* pair_ij, int P, double hinge_weight,
                       double* u_out, int32_t* iters_out, int32_t* status_out, void* stream) {
  REQUIRE_BOUND();
  if (!pair_n || !pair_ij || P < 1 || !u_out || !iters_out || !status_out || !(hinge_weight >= 0.0)) {
    h->err = "piadmm_solve_joint: NULL argument, P < 1 or a negative hinge weight";
    return PIADMM_E_ARG;
  }
  const int G = JT_THREADS / h->lpq, CL = (h->c.V + G - 1) / G;
  if (CL > JT_MAX_CLUSTER) {
    h->err = "piadmm_solve_joint: the scene does not fit one thread-block cluster (V <= 64 for N > 15, V <= 128 for N <= 15)";
    return PIADMM_E_UNSUPPORTED;
  }
  JointArgs ja;
  ja.pair_n = pair_n; ja.pair_ij = pair_ij; ja.P = P; ja.hinge = hinge_weight;
  ja.u_out = u_out; ja.iters_out = iters_out; ja.status_out = status_out;
  cudaStream_t st = (cudaStream_t)stream;
  CK(h->lpq == 16 ? launch_joint<16>(h->c, ja, CL, st) : launch_joint<32>(h->c, ja, CL, st));
  LAUNCHED();
  return PIADMM_OK;
}

int piadmm_residual_partials(piadmm_handle* h, void* stream) {
  REQUIRE_BOUND();
  cudaStream_t st = (cudaStream_t)stream;
  k_res_partial<<<h->c.S, 128, 0, st>>>(h->c);
  LAUNCHED();
  return PIADMM_OK;
}

int piadmm_residual_decide(piadmm_handle* h, void* stream) {
  REQUIRE_BOUND();
  cudaStream_t st = (cudaStream_t)stream;
  if (h->c.p.stop_dist_test) {
    const long slots = (long)h->c.S * h->c.V * h->c.D;
    k_dist_test<<<(unsigned)((slots + 255) / 256), 256, 0, st>>>(h->c);
    LAUNCHED();
  }
  // the last_hat copy of a dense scene is ~7000 elements behind two dependent index loads: 1024 threads per scenario
  // when there are few scenarios (36 -> ~6 us per launch at config 3), 128 for large batches
  k_res_decide<<<h->c.S, h->c.S <= 64 ? 1024 : 128, 0, st>>>(h->c);
  LAUNCHED();
  h->c.i_iter++;
  return PIADMM_OK;
}

// one launch (k_res_all) when this rank owns every vehicle; cnt (device, zeroed by the caller) receives the number of
// scenarios that are still active, or is NULL
static int residuals_single_rank(piadmm_handle* h, cudaStream_t st, int* cnt) {
  k_res_all<<<h->c.S, h->c.S <= 64 ? 1024 : 128, 0, st>>>(h->c, cnt);
  LAUNCHED();
  h->c.i_iter++;
  return PIADMM_OK;
}

int piadmm_residuals(piadmm_handle* h, void* stream) {
  REQUIRE_BOUND();
  if (h->c.own_lo == 0 && h->c.own_hi == h->c.V) return residuals_single_rank(h, (cudaStream_t)stream, nullptr);
  int rc = piadmm_residual_partials(h, stream);
  return rc ? rc : piadmm_residual_decide(h, stream);
}

int piadmm_set_ownership(piadmm_handle* h, int lo, int hi) {
  if (!h) return PIADMM_E_ARG;
  if (lo < 0 || hi > h->c.V || lo >= hi) { h->err = "ownership range must satisfy 0 <= lo < hi <= V"; return PIADMM_E_ARG; }
  if (h->c.p.dual_mode == PIADMM_DUAL_PI_OLD && (lo != 0 || hi != h->c.V)) {
    h->err = "dual_mode PI_OLD mutates one scene-wide rho and is not supported with vehicle sharding";
    return PIADMM_E_UNSUPPORTED;
  }
  h->c.own_lo = lo; h->c.own_hi = hi;
  return PIADMM_OK;
}

int piadmm_num_active(piadmm_handle* h, void* stream, int32_t* host_out) {
  REQUIRE_BOUND();
  if (!host_out) return PIADMM_E_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  CK(cudaMemsetAsync(h->d_count, 0, sizeof(int), st));
  k_count_active<<<(h->c.S + 255) / 256, 256, 0, st>>>(h->c, h->d_count);
  LAUNCHED();
  CK(cudaMemcpyAsync(host_out, h->d_count, sizeof(int), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return PIADMM_OK;
}

int piadmm_plant_step(piadmm_handle* h, void* stream) {
  REQUIRE_BOUND();
  cudaStream_t st = (cudaStream_t)stream;
  const long agents = (long)h->c.S * h->c.V;
  k_plant<<<(unsigned)((agents + 127) / 128), 128, 0, st>>>(h->c);
  LAUNCHED();
  return PIADMM_OK;
}

// one persistent kernel for the whole outer loop of the scenarios [s_lo, s_hi)
struct HostStepArgs {
  const double *xt, *spd, *refwin;     // spd / refwin NULL: resident (bound spd buffer, piadmm_set_reference)
  double *u_out, *pos_out, *xt_next_out;
  int32_t *curr_iter_out, *n_outer_out;
};

static int launch_fused(piadmm_handle* h, cudaStream_t st, int s_lo, int s_hi, int* queue, int range_id,
                        const HostStepArgs* zc = nullptr, const int* gather_step = nullptr, int loop_steps = 0,
                        const piadmm_histories* hist = nullptr) {
  DevCtx c = h->c;
  c.s_lo = s_lo; c.s_hi = s_hi; c.queue = queue;
  c.loop_steps = loop_steps;
  if (hist) c.hist = *hist;
  if (gather_step) { c.ref = h->d_ref; c.ref_T = h->ref_T; c.step = gather_step; }   // window gather inside the kernel
  if (zc) {      // zero-copy epilogue into mapped host memory (device-visible addresses)
    c.host_out = 1;
    c.ho_u = zc->u_out; c.ho_pos = zc->pos_out; c.ho_xt_next = zc->xt_next_out;
    c.ho_curr_iter = zc->curr_iter_out; c.ho_n_outer = zc->n_outer_out;
  }
  // the whole-batch order and the sub-batch orders share d_order: using one invalidates the other
  c.order = (h->use_order && h->order_valid[range_id]) ? h->d_order : nullptr;
  if (!h->queue_armed[range_id]) CK(cudaMemsetAsync(queue, 0, sizeof(int), st));   // else k_order of the last launch did it
  const int pairs = h->lpq == 16 ? (h->fused_ks ? KS_THREADS(16) / 32 : 2) : 1;
  int grid = (s_hi - s_lo + pairs - 1) / pairs;
  if (grid > h->fused_grid) grid = h->fused_grid;
  if (h->fused_ks) {
    if (h->lpq == 16) k_fused_ks<16><<<grid, KS_THREADS(16), 0, st>>>(c);
    else k_fused_ks<32><<<grid, KS_THREADS(32), 0, st>>>(c);
  } else if (grid <= 2 * h->sm_count && h->fused_few) {
    if (h->lpq == 16) k_fused_v2<16, true><<<grid, CTA_THREADS, 0, st>>>(c);
    else k_fused_v2<32, true><<<grid, CTA_THREADS, 0, st>>>(c);
  } else if (h->lpq == 16) k_fused_v2<16><<<grid, CTA_THREADS, 0, st>>>(c);
  else k_fused_v2<32><<<grid, CTA_THREADS, 0, st>>>(c);
  LAUNCHED();
  if (h->use_order && s_hi - s_lo >= 2 * grid) {      // learn the order for the next step (needs >1 scenario per slot)
    k_order<<<1, 1024, 0, st>>>(c, h->d_order);
    LAUNCHED();
    if (range_id == 0) { for (int i = 1; i <= PIADMM_CHUNKS; i++) h->order_valid[i] = false; }
    else h->order_valid[0] = false;
    h->order_valid[range_id] = true;
    h->queue_armed[range_id] = true;
  } else { h->order_valid[range_id] = false; h->queue_armed[range_id] = false; }
  return PIADMM_OK;
}

// ---- vehicle-sharded multi-GPU exchange: library-owned exchange block, exported / mapped with CUDA IPC
int piadmm_peer_export(piadmm_handle* h, void* handle_out) {
  if (!h || !handle_out) return PIADMM_E_ARG;
  if (h->c.S != 1) { h->err = "the peer exchange shards ONE coupled scene (S == 1)"; return PIADMM_E_UNSUPPORTED; }
  static_assert(sizeof(cudaIpcMemHandle_t) == PIADMM_PEER_HANDLE_BYTES, "handle size");
  CK(cudaSetDevice(h->device));
  const XblkLayout L = xblk_layout(h->c.V, h->c.N, h->c.E);
  if (!h->xblk) {
    CK(cudaMalloc(&h->xblk, L.bytes));
    CK(cudaMemset(h->xblk, 0, L.bytes));
    CK(cudaDeviceSynchronize());
  }
  cudaIpcMemHandle_t ih;
  CK(cudaIpcGetMemHandle(&ih, h->xblk));
  memcpy(handle_out, &ih, sizeof ih);
  return PIADMM_OK;
}

int piadmm_peer_disconnect(piadmm_handle* h) {
  if (!h) return PIADMM_E_ARG;
  if (h->peer_on) cudaSetDevice(h->device);
  for (int r = 0; r < PIADMM_MAX_PEERS; r++) {
    if (h->peer_base[r]) { cudaIpcCloseMemHandle(h->peer_base[r]); h->peer_base[r] = nullptr; }
  }
  h->peer_on = false;
  memset(&h->peer, 0, sizeof h->peer);
  return PIADMM_OK;
}

int piadmm_peer_connect(piadmm_handle* h, int rank, int world, const void* handles) {
  if (!h || !handles) return PIADMM_E_ARG;
  if (world < 2 || world > PIADMM_MAX_PEERS || rank < 0 || rank >= world) { h->err = "piadmm_peer_connect: 2 <= world <= 8, 0 <= rank < world"; return PIADMM_E_ARG; }
  if (!h->xblk) { h->err = "piadmm_peer_connect: call piadmm_peer_export first"; return PIADMM_E_UNBOUND; }
  if (!h->use_coop || h->c.detect_mode != 0) { h->err = "the peer exchange needs the cooperative kernel (PIADMM_COOP / PIADMM_DETECT defaults)"; return PIADMM_E_UNSUPPORTED; }
  CK(cudaSetDevice(h->device));
  piadmm_peer_disconnect(h);
  const XblkLayout L = xblk_layout(h->c.V, h->c.N, h->c.E);
  const cudaIpcMemHandle_t* ih = reinterpret_cast<const cudaIpcMemHandle_t*>(handles);
  PeerCtx pc;
  memset(&pc, 0, sizeof pc);
  pc.rank = rank; pc.world = world;
  for (int r = 0; r < world; r++) {
    char* base = h->xblk;
    if (r != rank) {
      void* q = nullptr;
      cudaError_t e = cudaIpcOpenMemHandle(&q, ih[r], cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) {
        h->err = std::string("cudaIpcOpenMemHandle (rank ") + std::to_string(r) + "): " + cudaGetErrorString(e);
        cudaGetLastError();
        piadmm_peer_disconnect(h);
        return PIADMM_E_CUDA;
      }
      h->peer_base[r] = q;
      base = (char*)q;
    }
    pc.pos[r] = (double*)(base + L.pos); pc.u[r] = (double*)(base + L.u); pc.edge_res[r] = (double*)(base + L.er);
    pc.flags[r] = (unsigned long long*)(base + L.flags);
    pc.stamp[r] = (int32_t*)(base + L.stamp);
    pc.need[r] = (int32_t*)(base + L.need);
    pc.tbox[r] = (double*)(base + L.tbox);
  }
  pc.epoch = (unsigned long long*)(h->xblk + L.epoch);
  pc.aborted = (int*)(h->xblk + L.aborted);
  const char* to = getenv("PIADMM_PEER_TIMEOUT_MS");
  pc.timeout_ns = (unsigned long long)(to ? atoll(to) : 20000) * 1000000ull;
  h->peer = pc;
  h->peer_on = true;
  return PIADMM_OK;
}

int piadmm_debug_coop_trace(piadmm_handle* h, int64_t* host_out, int n) {
  if (!h) return PIADMM_E_ARG;
  CK(cudaSetDevice(h->device));
  if (!h->d_trace) {
    CK(cudaMalloc(&h->d_trace, sizeof(long long) * PIADMM_COOP_TRACE_MAX));
    CK(cudaMemset(h->d_trace, 0, sizeof(long long) * PIADMM_COOP_TRACE_MAX));
  }
  if (host_out && n > 0) {
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(host_out, h->d_trace, sizeof(long long) * (size_t)std::min(n, PIADMM_COOP_TRACE_MAX), cudaMemcpyDeviceToHost));
  }
  return PIADMM_OK;
}

// the whole outer loop in ONE cooperative launch (piadmm_coop.cuh)
static int launch_coop(piadmm_handle* h, cudaStream_t st) {
  DevCtx c = h->c;
  CoopArgs a;
  memset(&a, 0, sizeof a);
  a.n_active = h->d_nactive;
  { const char* e_ = getenv("PIADMM_SLOT_GATE"); a.slot_gate = !(e_ && strcmp(e_, "0") == 0); }
  c.any_overflow = h->d_nactive + 2;
  c.use_chunk_tot = 1;
  { const char* e_ = getenv("PIADMM_TBOX"); c.use_tbox = !(e_ && strcmp(e_, "0") == 0); }
  a.trace = h->d_trace;
  if (h->d_trace) CK(cudaMemsetAsync(h->d_trace, 0, sizeof(long long), st));
  if (h->peer_on) {
    const XblkLayout L = xblk_layout(c.V, c.N, c.E);
    a.peer = h->peer;
    a.peer.user_pos = h->c.b.pos; a.peer.user_u = h->c.b.u;
    c.b.pos = (double*)(h->xblk + L.pos); c.b.u = (double*)(h->xblk + L.u); c.edge_res = (double*)(h->xblk + L.er);
    c.veh_stamp = (int32_t*)(h->xblk + L.stamp);
    c.veh_need = (int32_t*)(h->xblk + L.need);
    c.tbox = (double*)(h->xblk + L.tbox);
    c.stamps_global = 1;
  }
  const long agents = (long)c.S * c.V, pairs = (long)c.S * c.E;
  long want;
  if (h->lpq == 16) want = std::max((agents + 3) / 4, (pairs + 1) / 2);
  else want = std::max((agents + 1) / 2, pairs);
  // the occupancy build when this rank's vehicle QPs are more than two waves of the register build's group slots
  const long own_agents = (long)c.S * (c.own_hi - c.own_lo);
  const long slots_lo = (long)h->coop_grid_max * (h->lpq == 16 ? groups_per_cta<16>() : groups_per_cta<32>());
  // (not at tight inner tolerances: every solve is then hundreds of inner iterations, the step is the slowest chains again
  //  and the spills are on them -- config 5 with inner=tight: 2.71 ms with the register build, 3.33 ms with this one)
  const bool hi = h->coop_grid_max_hi > h->coop_grid_max &&
                  (h->coop_occ == 2 || (h->coop_occ == 0 && own_agents > 2 * slots_lo && c.p.in_eps_abs >= 1e-6));
  int grid = (int)std::min<long>(hi ? h->coop_grid_max_hi : h->coop_grid_max, std::max<long>(want, 1));
  int rc = build_candidates(h, st);          // reads xt / spd only
  if (rc) return rc;
  void* args[2] = {&c, &a};
  void* kern = h->lpq == 16 ? (hi ? (void*)k_coop<16, COOP_MANY_CTAS> : (void*)k_coop<16, COOP_MIN_CTAS>)
                            : (hi ? (void*)k_coop<32, COOP_MANY_CTAS> : (void*)k_coop<32, COOP_MIN_CTAS>);
  CK(cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(CTA_THREADS), args, 0, st));
  LAUNCHED();
  h->c.slots_dirty = 0;
  return PIADMM_OK;
}

int piadmm_step(piadmm_handle* h, void* stream) {
  REQUIRE_BOUND();
  cudaStream_t st = (cudaStream_t)stream;
  if (h->c.V == 2 && h->c.D == 1) return launch_fused(h, st, 0, h->c.S, h->d_queue, 0);
  const bool partial = h->c.own_lo != 0 || h->c.own_hi != h->c.V;
  if (partial && !h->peer_on) {
    h->err = "piadmm_step on a vehicle-sharded handle needs the peer exchange (piadmm_peer_connect); without it drive the "
             "phase entry points and reduce res_part between piadmm_residual_partials and piadmm_residual_decide";
    return PIADMM_E_UNSUPPORTED;
  }
  if (h->use_coop && h->c.detect_mode == 0) return launch_coop(h, st);
  if (h->peer_on) { h->err = "peer exchange without the cooperative kernel"; return PIADMM_E_UNSUPPORTED; }
  // generic V: phase kernels.  The active-scenario count of every outer iteration is copied to a pinned word
  // asynchronously and looked at ONE ITERATION LATE: the host has iteration it+1 enqueued before it waits for the
  // count of iteration it, so the GPU never idles for a host round trip.  At most one superfluous iteration is
  // enqueued after the last scenario finished; every kernel of it is a no-op (all phases are gated by done[s]).
  int rc = piadmm_begin_step(h, stream);
  for (int it = 0; it < h->c.I && rc == PIADMM_OK; it++) {
    if ((rc = piadmm_x_update(h, stream))) break;
    if ((rc = piadmm_detect_edges(h, stream))) break;
    if ((rc = piadmm_z_update(h, stream))) break;
    int* cnt = h->d_count2 + (it & 1);
    CK(cudaMemsetAsync(cnt, 0, sizeof(int), st));
    if (h->c.own_lo == 0 && h->c.own_hi == h->c.V) {
      if ((rc = residuals_single_rank(h, st, cnt))) break;      // residuals + stop test + active count in one launch
    } else {
      if ((rc = piadmm_residuals(h, stream))) break;
      k_count_active<<<(h->c.S + 255) / 256, 256, 0, st>>>(h->c, cnt);
      LAUNCHED();
    }
    CK(cudaMemcpyAsync(h->h_count2 + (it & 1), cnt, sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(h->ev_count[it & 1], st));
    if (it >= 1) {
      CK(cudaEventSynchronize(h->ev_count[(it - 1) & 1]));
      if (h->h_count2[(it - 1) & 1] == 0) break;
    }
  }
  return rc;
}

// Independent scenarios: the batch is split into n_chunks sub-batches on internal streams (forked from and
// joined back into `ms`), so that the H2D / D2H copies of one sub-batch overlap the kernels of the others.
// zc != NULL (zero-copy): the kernels store the results themselves into the caller's mapped pinned buffers (zc =
// their device-visible addresses) while they run: no D2H copies, no plant kernel; with a resident reference the
// inputs are a few 100 KB and one launch over the whole batch is best.
static int enqueue_chunked_host_step(piadmm_handle* h, cudaStream_t ms, const HostStepArgs& a, const HostStepArgs* zc) {
  const size_t N = h->c.N, np1 = N + 1;
  int nch = (zc && !a.refwin) ? 1 : h->n_chunks;
  if (h->c.S < 64 * nch) nch = 1;
  // zero-copy step on the resident reference: the fused kernel gathers the windows itself (no gather kernel).  The
  // step index still travels as a 4-byte copy: 1184 CTAs reading one pinned host word over PCIe took ~0.9 ms.
  const int* gather = (zc && !a.refwin && nch == 1) ? h->d_step : nullptr;
  if (!a.refwin) CK(cudaMemcpyAsync(h->d_step, h->h_step, sizeof(int), cudaMemcpyHostToDevice, ms));
  CK(cudaEventRecord(h->ev_start, ms));
  for (int ch = 0; ch < nch; ch++) {
    cudaStream_t cs = h->cs[ch];
    const int s0 = (int)((long)h->c.S * ch / nch), s1 = (int)((long)h->c.S * (ch + 1) / nch);
    const size_t a0 = (size_t)s0 * 2, na = (size_t)(s1 - s0) * 2;
    CK(cudaStreamWaitEvent(cs, h->ev_start, 0));
    CK(cudaMemcpyAsync((double*)h->c.b.xt + a0 * 3, a.xt + a0 * 3, sizeof(double) * na * 3, cudaMemcpyHostToDevice, cs));
    if (a.spd) CK(cudaMemcpyAsync((double*)h->c.b.spd + a0, a.spd + a0, sizeof(double) * na, cudaMemcpyHostToDevice, cs));
    DevCtx c = h->c;
    c.s_lo = s0; c.s_hi = s1;
    if (a.refwin) {
      CK(cudaMemcpyAsync((double*)h->c.b.refwin + a0 * 2 * np1, a.refwin + a0 * 2 * np1, sizeof(double) * na * 2 * np1, cudaMemcpyHostToDevice, cs));
    } else if (!gather) {
      k_refwin<<<(unsigned)((na * 2 * np1 + 255) / 256), 256, 0, cs>>>(c, h->d_ref, h->ref_T, h->d_step);
      LAUNCHED();
    }
    int rc = nch == 1 ? launch_fused(h, cs, s0, s1, h->d_queue, 0, zc, gather) : launch_fused(h, cs, s0, s1, h->d_queue + 1 + ch, 1 + ch, zc);
    if (rc) return rc;
    if (zc) {
      CK(cudaEventRecord(h->ev_done[ch], cs));
      CK(cudaStreamWaitEvent(ms, h->ev_done[ch], 0));
      continue;
    }
    k_plant<<<(unsigned)((na + 127) / 128), 128, 0, cs>>>(c);
    LAUNCHED();
    if (a.u_out) CK(cudaMemcpyAsync(a.u_out + a0 * N, h->c.b.u + a0 * N, sizeof(double) * na * N, cudaMemcpyDeviceToHost, cs));
    if (a.pos_out) CK(cudaMemcpyAsync(a.pos_out + a0 * 2 * np1, h->c.b.pos + a0 * 2 * np1, sizeof(double) * na * 2 * np1, cudaMemcpyDeviceToHost, cs));
    if (a.xt_next_out) CK(cudaMemcpyAsync(a.xt_next_out + a0 * 3, h->c.b.xt_next + a0 * 3, sizeof(double) * na * 3, cudaMemcpyDeviceToHost, cs));
    if (a.curr_iter_out) CK(cudaMemcpyAsync(a.curr_iter_out + s0, h->c.b.curr_iter + s0, sizeof(int32_t) * (s1 - s0), cudaMemcpyDeviceToHost, cs));
    if (a.n_outer_out) CK(cudaMemcpyAsync(a.n_outer_out + s0, h->c.b.n_outer + s0, sizeof(int32_t) * (s1 - s0), cudaMemcpyDeviceToHost, cs));
    CK(cudaEventRecord(h->ev_done[ch], cs));
    CK(cudaStreamWaitEvent(ms, h->ev_done[ch], 0));
  }
  return PIADMM_OK;
}

// device-visible address of a mapped pinned host pointer (NULL stays NULL); false if p is not mapped pinned memory
static bool mapped_dev_ptr(void* p, void** out) {
  *out = nullptr;
  if (!p) return true;
  if (cudaHostGetDevicePointer(out, p, 0) != cudaSuccess) { cudaGetLastError(); *out = nullptr; return false; }
  return true;
}

static bool pinned_or_null(const void* p) {
  if (!p) return true;
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged;
}

static int step_host_impl(piadmm_handle* h, const double* xt, const double* spd, const double* refwin,
                          double* u_out, double* pos_out, double* xt_next_out, int32_t* curr_iter_out,
                          int32_t* n_outer_out, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const size_t SV = (size_t)h->c.S * h->c.V, N = h->c.N;
  const HostStepArgs a = {xt, spd, refwin, u_out, pos_out, xt_next_out, curr_iter_out, n_outer_out};
  HostStepArgs zc = a;
  const bool fused = h->c.V == 2 && h->c.D == 1;
  const bool zero_copy = fused && h->zero_copy && mapped_dev_ptr(u_out, (void**)&zc.u_out) &&
                         mapped_dev_ptr(pos_out, (void**)&zc.pos_out) && mapped_dev_ptr(xt_next_out, (void**)&zc.xt_next_out) &&
                         mapped_dev_ptr(curr_iter_out, (void**)&zc.curr_iter_out) &&
                         mapped_dev_ptr(n_outer_out, (void**)&zc.n_outer_out);
  if (fused && (zero_copy || h->c.S >= 64 * h->n_chunks)) {
    cudaStream_t ms = h->cs_main;                 // internal origin stream (the caller's may be the legacy stream,
    CK(cudaEventRecord(h->ev_user, st));          //  which cannot be captured): ordered after the caller's work
    CK(cudaStreamWaitEvent(ms, h->ev_user, 0));
    bool graph = h->use_graph && pinned_or_null(xt) && pinned_or_null(spd) && pinned_or_null(refwin) &&
                 pinned_or_null(u_out) && pinned_or_null(pos_out) && pinned_or_null(xt_next_out) &&
                 pinned_or_null(curr_iter_out) && pinned_or_null(n_outer_out);
    auto enqueue = [&]() { return enqueue_chunked_host_step(h, ms, a, zero_copy ? &zc : nullptr); };
    if (graph) {
      uintptr_t bits = zero_copy ? (uintptr_t)1 << 20 : 0;
      for (int i = 0; i <= PIADMM_CHUNKS; i++) bits |= (uintptr_t)(h->order_valid[i] ? 1 : 0) << i;
      const void* key[9] = {xt, spd, refwin, u_out, pos_out, xt_next_out, curr_iter_out, n_outer_out, (const void*)bits};
      if (!h->gexec || memcmp(key, h->gkey, sizeof key) != 0) {
        drop_graph(h);
        const int64_t l0 = h->launches;
        CK(cudaStreamBeginCapture(ms, cudaStreamCaptureModeThreadLocal));
        const int rc = enqueue();
        cudaGraph_t g = nullptr;
        cudaError_t ce = cudaStreamEndCapture(ms, &g);
        if (rc == PIADMM_OK && ce == cudaSuccess) ce = cudaGraphInstantiate(&h->gexec, g, 0);
        if (g) cudaGraphDestroy(g);
        h->glaunches = h->launches - l0;
        h->launches = l0;
        if (rc != PIADMM_OK || ce != cudaSuccess) {      // capture unavailable: plain stream path from now on
          cudaGetLastError();
          h->gexec = nullptr; h->use_graph = false; graph = false;
          for (int i = 0; i <= PIADMM_CHUNKS; i++) { h->order_valid[i] = false; h->queue_armed[i] = false; }
        } else {
          memcpy(h->gkey, key, sizeof key);
          for (int i = 0; i <= PIADMM_CHUNKS; i++) h->gvalid_after[i] = h->order_valid[i];
        }
      }
      if (graph) {
        CK(cudaGraphLaunch(h->gexec, ms));
        h->launches += h->glaunches;
        for (int i = 0; i <= PIADMM_CHUNKS; i++) h->order_valid[i] = h->gvalid_after[i];
      }
    }
    if (!graph) {
      const int rc = enqueue();
      if (rc) return rc;
    }
    CK(cudaStreamSynchronize(ms));
    return PIADMM_OK;
  }
  CK(cudaMemcpyAsync((void*)h->c.b.xt, xt, sizeof(double) * SV * 3, cudaMemcpyHostToDevice, st));
  if (spd) CK(cudaMemcpyAsync((void*)h->c.b.spd, spd, sizeof(double) * SV, cudaMemcpyHostToDevice, st));
  if (refwin) {
    CK(cudaMemcpyAsync((void*)h->c.b.refwin, refwin, sizeof(double) * SV * 2 * (N + 1), cudaMemcpyHostToDevice, st));
  } else {
    CK(cudaMemcpyAsync(h->d_step, h->h_step, sizeof(int), cudaMemcpyHostToDevice, st));
    k_refwin<<<(unsigned)((SV * 2 * (N + 1) + 255) / 256), 256, 0, st>>>(h->c, h->d_ref, h->ref_T, h->d_step);
    LAUNCHED();
  }
  int rc = piadmm_step(h, stream);
  if (rc) return rc;
  rc = piadmm_plant_step(h, stream);
  if (rc) return rc;
  if (u_out) CK(cudaMemcpyAsync(u_out, h->c.b.u, sizeof(double) * SV * N, cudaMemcpyDeviceToHost, st));
  if (pos_out) CK(cudaMemcpyAsync(pos_out, h->c.b.pos, sizeof(double) * SV * 2 * (N + 1), cudaMemcpyDeviceToHost, st));
  if (xt_next_out) CK(cudaMemcpyAsync(xt_next_out, h->c.b.xt_next, sizeof(double) * SV * 3, cudaMemcpyDeviceToHost, st));
  if (curr_iter_out) CK(cudaMemcpyAsync(curr_iter_out, h->c.b.curr_iter, sizeof(int32_t) * h->c.S, cudaMemcpyDeviceToHost, st));
  if (n_outer_out) CK(cudaMemcpyAsync(n_outer_out, h->c.b.n_outer, sizeof(int32_t) * h->c.S, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return PIADMM_OK;
}

int piadmm_step_host(piadmm_handle* h, const double* xt, const double* spd, const double* refwin,
                     double* u_out, double* pos_out, double* xt_next_out, int32_t* curr_iter_out,
                     int32_t* n_outer_out, void* stream) {
  REQUIRE_BOUND();
  if (!xt || !spd || !refwin) return PIADMM_E_ARG;
  return step_host_impl(h, xt, spd, refwin, u_out, pos_out, xt_next_out, curr_iter_out, n_outer_out, stream);
}

int piadmm_set_reference(piadmm_handle* h, const double* ref, int T, void* stream) {
  if (!h) return PIADMM_E_ARG;
  if (!ref || T < h->c.N + 1) { h->err = "piadmm_set_reference: ref is NULL or T < num_ho + 1"; return PIADMM_E_ARG; }
  CK(cudaSetDevice(h->device));
  const size_t n = (size_t)h->c.S * h->c.V * 2 * T;
  if (T != h->ref_T || !h->d_ref) {
    cudaFree(h->d_ref); h->d_ref = nullptr; h->ref_T = 0;
    CK(cudaMalloc(&h->d_ref, sizeof(double) * n));
    drop_graph(h);
  }
  CK(cudaMemcpyAsync(h->d_ref, ref, sizeof(double) * n, cudaMemcpyDefault, (cudaStream_t)stream));
  CK(cudaStreamSynchronize((cudaStream_t)stream));
  h->ref_T = T;
  return PIADMM_OK;
}

int piadmm_step_host_at(piadmm_handle* h, int num_step, const double* xt, double* u_out, double* pos_out,
                        double* xt_next_out, int32_t* curr_iter_out, int32_t* n_outer_out, void* stream) {
  REQUIRE_BOUND();
  if (!xt) return PIADMM_E_ARG;
  if (!h->d_ref) { h->err = "piadmm_step_host_at: no resident reference (call piadmm_set_reference first)"; return PIADMM_E_UNBOUND; }
  if (num_step < 0 || num_step + h->c.N + 1 > h->ref_T) {
    h->err = "piadmm_step_host_at: window [num_step, num_step + num_ho] leaves the reference (casadi/main.py:43)";
    return PIADMM_E_ARG;
  }
  *h->h_step = num_step;
  return step_host_impl(h, xt, nullptr, nullptr, u_out, pos_out, xt_next_out, curr_iter_out, n_outer_out, stream);
}

int piadmm_run_closed_loop(piadmm_handle* h, int first_step, int n_steps, const piadmm_histories* out, void* stream) {
  REQUIRE_BOUND();
  if (!h->d_ref) { h->err = "piadmm_run_closed_loop: no resident reference (call piadmm_set_reference first)"; return PIADMM_E_UNBOUND; }
  if (n_steps < 1 || first_step < 0 || first_step + n_steps - 1 + h->c.N + 1 > h->ref_T) {
    h->err = "piadmm_run_closed_loop: the windows [step, step + num_ho] leave the reference (casadi/main.py:43)";
    return PIADMM_E_ARG;
  }
  cudaStream_t st = (cudaStream_t)stream;
  piadmm_histories o;
  memset(&o, 0, sizeof o);
  if (out) o = *out;
  const size_t SV = (size_t)h->c.S * h->c.V, np1 = h->c.N + 1;
  const bool fused = h->c.V == 2 && h->c.D == 1;
  *h->h_step = first_step;
  CK(cudaMemcpyAsync(h->d_step, h->h_step, sizeof(int), cudaMemcpyHostToDevice, st));
  HostStepArgs none;
  memset(&none, 0, sizeof none);
  if (fused && !h->fused_ks) {
    // V == 2: ONE launch.  Scenarios are independent, so each one runs its n_steps MPC steps back to back inside the
    // persistent kernel (window gather at first_step + t, outer loop, plant step, histories, xt <- xt_next): no launch
    // and no batch-wide tail per MPC step -- a step of the batch no longer lasts as long as its slowest scenario.
    return launch_fused(h, st, 0, h->c.S, h->d_queue, 0, &none, h->d_step, n_steps, &o);
  }
  for (int t = 0; t < n_steps; t++) {
    int rc;
    if (fused) {     // host_out with NULL targets: inline plant step; the kernel gathers the windows at *d_step itself
      rc = launch_fused(h, st, 0, h->c.S, h->d_queue, 0, &none, h->d_step);
    } else {
      k_refwin<<<(unsigned)((SV * 2 * np1 + 255) / 256), 256, 0, st>>>(h->c, h->d_ref, h->ref_T, h->d_step);
      LAUNCHED();
      rc = piadmm_step(h, stream);
      if (!rc) rc = piadmm_plant_step(h, stream);
    }
    if (rc) return rc;
    k_advance<<<(unsigned)((SV + 127) / 128), 128, 0, st>>>(h->c, o, t, h->d_step);
    LAUNCHED();
  }
  return PIADMM_OK;
}

int piadmm_fp64_peak(int device, int iters, double* tflops, double* ms) {
  if (!tflops || iters < 1) return PIADMM_E_ARG;
  cudaDeviceProp prop;
  if (cudaSetDevice(device) != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess) return PIADMM_E_CUDA;
  const int blocks = prop.multiProcessorCount * 8, threads = 256;
  double* d = nullptr;
  if (cudaMalloc(&d, sizeof(double) * (size_t)blocks * threads) != cudaSuccess) return PIADMM_E_CUDA;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 6; rep++) {       // first 3 are warm-up
    cudaEventRecord(e0);
    k_fp64_peak<<<blocks, threads>>>(d, iters);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(d); return PIADMM_E_CUDA; }
    float t; cudaEventElapsedTime(&t, e0, e1);
    if (rep >= 3 && t < best) best = t;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
  const double flops = 2.0 * 8.0 * (double)iters * (double)blocks * threads;
  *tflops = flops / (best * 1e-3) / 1e12;
  if (ms) *ms = best;
  return PIADMM_OK;
}

}  // extern "C"
