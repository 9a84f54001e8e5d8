// piadmm_dev.cuh -- device-side building blocks of the PI-ADMM hot path (sm_100a).
//
// Execution model
//   * a "group" is LPQ (16 or 32) consecutive lanes of a warp; lane k of a group owns time index k
//     of one vehicle: the k-th steering variable, the k-th trajectory point and the four constraint
//     rows (I, -I, -D1, D1 of casadi/PI_ADMM_class.py:172-192) that belong to index k.
//   * a "pair" is two groups: the two vehicles of a collision edge.  LPQ==16: both groups live in
//     one warp (exchange by shuffle).  LPQ==32: the pair is a 64-thread CTA of two warps (exchange
//     through shared memory + __syncthreads()).
//   * the N x N matrix K^-1 = (P + sigma I + rho A'A)^-1 of the OSQP-style inner ADMM lives in
//     REGISTERS, one row per lane; the right-hand side is broadcast through shared memory.
//   * M (double summation, casadi/PI_ADMM_class.py:59-69) and M' are never formed: they are two
//     chained warp scans.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/piadmm.h"

#define PIADMM_FULL 0xffffffffu

// constants of the inner iteration derived from in_alpha on the host (set_inner_consts): as members of the kernel
// parameter block they are constant-bank operands of the loop's DFMAs, not registers
struct InnerK {
  double ha, b1;        // alpha/2, 1 - alpha/2
  double alu, ald;      // alpha * l_u, alpha * l_d
};
#define PIADMM_L_U (-0.52359877559829887308)   // -pi/6: steering bound rows
#define PIADMM_L_D (-0.34906585039886591538)   // -pi/9: steering rate rows
inline void set_inner_consts(InnerK& ik, double alpha) {
  ik.ha = 0.5 * alpha; ik.b1 = 1.0 - 0.5 * alpha; ik.alu = alpha * PIADMM_L_U; ik.ald = alpha * PIADMM_L_D;
}

#define PIADMM_CAND_CAP 64
struct DevCtx {
  piadmm_params p;
  InnerK ik;
  piadmm_buffers b;
  int S, V, D, E, N, I;
  double round_u_scale, round_edge_scale;  // 10^decimals, 0 = no rounding
  const double* W;                         // [2][LPQ*LPQ]: M'M and D2'D2, zero padded, row stride LPQ
  double* edge_res;                        // [S,E,4] ss0, ss1, rr0, rr1 of the current outer iteration
  double* rho_new;                         // [S] penalty chosen by the last edge (dual_mode PI_OLD)
  uint32_t* hits;                          // [S,V,ceil(V/32)] collision bit matrix (matrix detection modes)
  uint32_t* cand;                          // [S,V,ceil(V/32)] candidate pairs of this MPC step (piadmm_detect.cuh)
  uint8_t* slot_hit;                       // [S,V,D] the partner in this slot collides in the current outer iteration
  int32_t* up_cnt;                         // [S,V] edges (i,j>i) of the current outer iteration per vehicle
  double* bbox;                            // [S,V,4] trajectory bounding boxes (xmin,xmax,ymin,ymax) / reach discs (x0,y0,R,0)
  int32_t* cand_list;                      // [S,V,PIADMM_CAND_CAP] candidate partners of every vehicle, ascending
  int32_t* cand_cnt;                       // [S,V] their number, or -1: more than the cap, walk the bit row instead
  double* edge_u;                          // [S,E,2,N] solution of every edge problem of the last z phase (piadmm_solve_edges), or NULL
  int z_solve_only;                        // edge problems only: no z-hat roll-out, no dual / PI update (the caller's script does both)
  int stamps_global;                       // veh_stamp is exchanged between the ranks (peer mode): no rank-local guessing
  int* any_overflow;                       // cooperative kernel: a neighbour table overflowed in this step (or NULL)
  int* cand_over;                          // some vehicle has more than PIADMM_CAND_CAP candidates in this MPC step (set by the list kernels)
  int32_t* veh_need;                       // [S,V] cooperative kernel: outer iteration (>= 1) in which the vehicle or one of its candidates
                                           // was re-solved, i.e. its slot phase has something to re-test (stale values only cost work)
  int32_t* chunk_tot;                      // [S,ceil(V/CTA_THREADS)] cooperative kernel: sum of up_cnt over each chunk of vehicles
  int use_chunk_tot;                       // up_cnt is kept together with chunk_tot (set in the cooperative launch only)
  double* tbox;                            // [S,V,4] cooperative kernel: (xmin, xmax, ymin, ymax) of the vehicle's pos_old row, written by
                                           // whoever solves its QP (x phase); the slot phase rejects a candidate on the two boxes
                                           // (32 B per candidate) before it loads the 2(N+1) points of the exact test
  int use_tbox;                            // tbox is maintained (set in the cooperative launch only)
  int32_t* veh_stamp;                      // [S,V] outer iteration in which the vehicle's trajectory was last recomputed
  double* blockbox;                        // [S,ceil(V/32),4] bounding box of the reach discs of 32 consecutive vehicles
  uint8_t* veh_dirty;                      // [S,V] x_memo: the vehicle's QP changed since its last solve in this step
  int slots_dirty;                         // the slot arrays may hold anything (first step after piadmm_bind): zero all of them
  int detect_mode;                         // 0 candidate rows (default), 1 full pair test lane-per-pair, 2 thread-per-word
  int* dist_bad;                           // [S] MATLAB :202 distance stop test failed
  int i_iter;                              // outer iteration index (phase API)
  int own_lo, own_hi;                      // vehicles x-updated by this rank (multi-GPU vehicle sharding)
  int* queue;                              // work counter of the persistent fused kernel
  int s_lo, s_hi;                          // scenario range of this launch (fused kernel / plant step)
  const int* order;                        // fused kernel: processing order of [s_lo,s_hi) (longest first), or NULL
  double* kcache;                          // [S*V][2][PIADMM_KC_STRIDE]: K^-1(rho0) of every vehicle's two problems, or NULL
  // fused kernel, host-buffer entry points: zero-copy epilogue.  When set, every finished scenario also takes the
  // plant step and stores its results straight into the caller's pinned (mapped) host buffers, so the D2H traffic
  // streams out during the kernel instead of as copies after it.  Any pointer may be NULL.
  // fused kernel with a resident reference (piadmm_set_reference): every scenario gathers its own window
  // refwin[s,v,c,k] = ref[s,v,c,*step + k] when it starts (no separate gather kernel); `step` is device memory
  const double* ref;
  int ref_T;
  const int* step;
  int host_out;
  double *ho_u, *ho_pos, *ho_xt_next;
  int32_t *ho_curr_iter, *ho_n_outer;
  // fused kernel, on-device closed loop (piadmm_run_closed_loop, V == 2): every scenario runs ALL loop_steps MPC steps
  // back to back -- its step t+1 needs nothing but its own step t -- and appends the script's histories itself
  int loop_steps;
  piadmm_histories hist;
};

// ------------------------------------------------------------------ group primitives
// v += t on the lanes where k >= o (one predicated DADD instead of a DADD + two FSEL)
__device__ __forceinline__ void add_if_ge(double& v, double t, int k, int o) {
  asm("{\n\t.reg .pred p;\n\tsetp.ge.s32 p, %2, %3;\n\t@p add.f64 %0, %0, %1;\n\t}" : "+d"(v) : "d"(t), "r"(k), "r"(o));
}
template <int LPQ>
__device__ __forceinline__ double incl_prefix(double v, int k) {
#pragma unroll
  for (int o = 1; o < LPQ; o <<= 1) {
    double t = __shfl_up_sync(PIADMM_FULL, v, o, LPQ);
    add_if_ge(v, t, k, o);
  }
  return v;
}
template <int LPQ>
__device__ __forceinline__ double incl_suffix(double v, int k) {
#pragma unroll
  for (int o = 1; o < LPQ; o <<= 1) {
    double t = __shfl_down_sync(PIADMM_FULL, v, o, LPQ);
    add_if_ge(v, t, LPQ - 1 - k, o);
  }
  return v;
}
template <int LPQ>
__device__ __forceinline__ double shift_up1(double v, int k) {  // lane k <- lane k-1, lane 0 <- 0
  double t = __shfl_up_sync(PIADMM_FULL, v, 1, LPQ);
  return k >= 1 ? t : 0.0;
}
template <int LPQ>
__device__ __forceinline__ double shift_down1(double v, int k) {  // lane k <- lane k+1, last <- 0
  double t = __shfl_down_sync(PIADMM_FULL, v, 1, LPQ);
  return k + 1 < LPQ ? t : 0.0;
}
template <int LPQ>
__device__ __forceinline__ double excl_prefix(double v, int k) { return shift_up1<LPQ>(incl_prefix<LPQ>(v, k), k); }
template <int LPQ>
__device__ __forceinline__ double excl_suffix(double v, int k) { return shift_down1<LPQ>(incl_suffix<LPQ>(v, k), k); }

template <int LPQ>
__device__ __forceinline__ double group_max(double v) {
#pragma unroll
  for (int o = LPQ / 2; o >= 1; o >>= 1) { double t = __shfl_xor_sync(PIADMM_FULL, v, o); v = v > t ? v : t; }
  return v;
}
// group maximum of NON-NEGATIVE finite doubles (norms): their IEEE bit patterns order like unsigned integers, so the
// maximum is two warp reductions (redux.sync: high words, then the low words of the lanes that hold the maximal
// high word) instead of log2(LPQ) rounds of two shuffles, a compare and two selects (LPQ == 32: N=30 0.604 -> 0.593 ms)
template <int LPQ>
__device__ __forceinline__ double group_max_nn(double v) {
  // half-warp member masks make REDUX slower than the four shuffle rounds (N=15: 0.2121 vs 0.2093 ms): full warps only
  if (LPQ == 16) return group_max<LPQ>(v);
  const unsigned mask = 0xffffffffu;
  const unsigned hi = (unsigned)__double2hiint(v), lo = (unsigned)__double2loint(v);
  const unsigned mh = __reduce_max_sync(mask, hi);
  const unsigned ml = __reduce_max_sync(mask, hi == mh ? lo : 0u);
  return __hiloint2double((int)mh, (int)ml);
}
// maxima of FOUR quantities over a 16-lane group in one transposed butterfly: after the xor-8 exchange a lane keeps
// two of them, after xor-4 one; 9 exchanges + 4 broadcasts instead of 16 exchanges (each = two SHFL + compare + two
// selects).  Lanes 0-3 / 4-7 / 8-11 / 12-15 of the group end up owning a / b / c / d.
// pair = true: maxima over both groups of the warp (the two halves of an edge problem), one more exchange.
__device__ __forceinline__ void group_max4_16(double& a, double& b, double& c, double& d, int k, bool pair) {
  const bool h8 = k & 8, h4 = k & 4;
  double m0 = h8 ? c : a, m1 = h8 ? d : b;
  const double s0 = h8 ? a : c, s1 = h8 ? b : d;
  const double r0 = __shfl_xor_sync(PIADMM_FULL, s0, 8), r1 = __shfl_xor_sync(PIADMM_FULL, s1, 8);
  m0 = m0 > r0 ? m0 : r0; m1 = m1 > r1 ? m1 : r1;
  double m = h4 ? m1 : m0;
  const double s = h4 ? m0 : m1;
  const double r = __shfl_xor_sync(PIADMM_FULL, s, 4);
  m = m > r ? m : r;
  double t = __shfl_xor_sync(PIADMM_FULL, m, 2); m = m > t ? m : t;
  t = __shfl_xor_sync(PIADMM_FULL, m, 1); m = m > t ? m : t;
  if (pair) { t = __shfl_xor_sync(PIADMM_FULL, m, 16); m = m > t ? m : t; }
  a = __shfl_sync(PIADMM_FULL, m, 0, 16); b = __shfl_sync(PIADMM_FULL, m, 4, 16);
  c = __shfl_sync(PIADMM_FULL, m, 8, 16); d = __shfl_sync(PIADMM_FULL, m, 12, 16);
}
template <int LPQ>
__device__ __forceinline__ double group_min(double v) {
#pragma unroll
  for (int o = LPQ / 2; o >= 1; o >>= 1) { double t = __shfl_xor_sync(PIADMM_FULL, v, o); v = v < t ? v : t; }
  return v;
}
template <int LPQ>
__device__ __forceinline__ double group_sum(double v) {
#pragma unroll
  for (int o = LPQ / 2; o >= 1; o >>= 1) v += __shfl_xor_sync(PIADMM_FULL, v, o);
  return v;
}
template <int LPQ>
__device__ __forceinline__ bool group_any(bool pred, int lane) {
  unsigned m = __ballot_sync(PIADMM_FULL, pred);
  if (LPQ == 16) m = (m >> (lane & 16)) & 0xffffu;
  return m != 0u;
}

// (M x)_k at lane k (k = 0..LPQ-1); *next = (M x)_{k+1}.  x must be 0 on padded lanes.
template <int LPQ>
__device__ __forceinline__ double apply_M(double x, int k, double* next) {
  double c1 = excl_prefix<LPQ>(x, k);   // sum_{j<k} x_j
  double m = excl_prefix<LPQ>(c1, k);   // sum_{i<k} c1_i
  *next = m + c1;
  return m;
}
// (M' v)_j at lane j for v given at lane k = trajectory index k (0..N; 0 on padded lanes)
template <int LPQ>
__device__ __forceinline__ double apply_Mt(double v, int k) {
  double t = excl_suffix<LPQ>(v, k);    // t_i = sum_{k>i} v_k
  return excl_suffix<LPQ>(t, k);        // sum_{i>j} t_i
}
// (M' v)_j for v given in the "s layout": lane L holds the entry of trajectory index L+1
template <int LPQ>
__device__ __forceinline__ double apply_Mt_s(double ws, int k) {
  double t = incl_suffix<LPQ>(ws, k);   // t_i = sum_{L>=i} ws_L = sum_{k>i} v_k
  return excl_suffix<LPQ>(t, k);
}

// exchange one double with the partner group of the pair.  LPQ==32: through shared memory, DOUBLE BUFFERED (xch =
// 2 x 64 doubles, `par` selects the buffer and flips): one CTA barrier per exchange instead of two -- the buffer
// written now is not written again before the barrier of the NEXT exchange, which every thread only reaches after it
// has read this one.
template <int LPQ>
__device__ __forceinline__ double partner(double mine, double* xch, int k, int half, int& par) {
  if (LPQ == 16) {
    return __shfl_xor_sync(PIADMM_FULL, mine, 16);
  } else {
    double* b = xch + par * 64;
    par ^= 1;
    b[half * 32 + k] = mine;
    __syncthreads();
    return b[(half ^ 1) * 32 + k];
  }
}
// the four group-uniform termination quantities of the partner in one exchange
template <int LPQ>
__device__ __forceinline__ void partner_max4(double& a0, double& a1, double& a2, double& a3, double* xch, int k, int half,
                                             int& par) {
  if (LPQ == 16) {
    a0 = fmax(a0, __shfl_xor_sync(PIADMM_FULL, a0, 16)); a1 = fmax(a1, __shfl_xor_sync(PIADMM_FULL, a1, 16));
    a2 = fmax(a2, __shfl_xor_sync(PIADMM_FULL, a2, 16)); a3 = fmax(a3, __shfl_xor_sync(PIADMM_FULL, a3, 16));
  } else {
    double* b = xch + par * 64;
    par ^= 1;
    if (k == 0) { double* o = b + half * 32; o[0] = a0; o[1] = a1; o[2] = a2; o[3] = a3; }
    __syncthreads();
    const double* o = b + (half ^ 1) * 32;
    a0 = fmax(a0, o[0]); a1 = fmax(a1, o[1]); a2 = fmax(a2, o[2]); a3 = fmax(a3, o[3]);
  }
}

// max/min without the NaN plumbing of fmax/fmin (inputs are finite): DSETP + 2 SEL
// (written in PTX so that the compiler does not pattern-match it back into the 7-instruction fmax sequence)
__device__ __forceinline__ double dmax(double a, double b) {
  double r;
  asm("{\n\t.reg .pred p;\n\tsetp.gt.f64 p, %1, %2;\n\tselp.f64 %0, %1, %2, p;\n\t}" : "=d"(r) : "d"(a), "d"(b));
  return r;
}
__device__ __forceinline__ double dmin(double a, double b) {
  double r;
  asm("{\n\t.reg .pred p;\n\tsetp.lt.f64 p, %1, %2;\n\tselp.f64 %0, %1, %2, p;\n\t}" : "=d"(r) : "d"(a), "d"(b));
  return r;
}

__device__ __forceinline__ double round_scale(double x, double scale) {
  return scale > 0.0 ? rint(x * scale) / scale : x;   // np.around: rint(x*10^d)/10^d
}

// ------------------------------------------------------------------ K^-1 in registers
// acc += a*b on the lanes where pred != 0 (one predicated DFMA)
__device__ __forceinline__ void fma_if(double& acc, double a, double b, int pred) {
  asm("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %3, 0;\n\t@p fma.rn.f64 %0, %1, %2, %0;\n\t}" : "+d"(acc) : "d"(a), "d"(b), "r"(pred));
}

// 1/d for a positive, normal d (the determinant of an SPD pivot block): hardware seed (2^-20) + two Newton steps,
// without the special-case plumbing of the IEEE division (about 25 of the ~140 instructions of a block step)
__device__ __forceinline__ double fast_rcp(double d) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  r = fma(fma(-d, r, 1.0), r, r);
  r = fma(fma(-d, r, 1.0), r, r);
  return r;
}

// One 2x2-BLOCK step of the in-place Gauss-Jordan inversion of an SPD matrix held one row per lane.
//
// Two pivots (rows/columns p, p+1; p even) are eliminated per step, so a build costs LPQ/2 instead of LPQ
// rounds of "publish pivot rows -> warp barrier -> broadcast read -> reciprocal -> update": the rounds are
// latency chains (the builds were ~20 % of the fused kernel's time with scalar pivots), the flop count is
// unchanged.  The loop stays ROLLED with compile-time register indices: every step rotates the row by two
// positions, so the pivot columns are always a[0], a[1]; after LPQ/2 steps the rotation has come full circle.
// Padded rows/columns (>= N) are identity and decoupled (a 2x2 block [[a,0],[0,1]] is handled exactly).
//   pivot lanes    : publish their raw rows and clear their registers, so that the common update
//                    a[j-2] = c0*rowp[j] + c1*rowq[j] + a[j] yields (B^-1 [rowp; rowq])_j with (c0,c1) = B^-1 row;
//   other lanes    : (c0,c1) = -[a0 a1] B^-1; the new inverse columns a[LPQ-2], a[LPQ-1] are (c0,c1) for everyone.
// rowbuf: 2*LPQ doubles of shared memory per group.
template <int LPQ>
__device__ __forceinline__ void gj_step2(double (&a)[LPQ], int k, int p, double* rowbuf) {
  const double f0 = a[0], f1 = a[1];
  const bool piv = (k >> 1) == (p >> 1);
  if (piv) {
    double* rb = rowbuf + (k & 1) * LPQ;
#pragma unroll
    for (int j = 0; j < LPQ; j += 2) *reinterpret_cast<double2*>(rb + j) = make_double2(a[j], a[j + 1]);
#pragma unroll
    for (int j = 0; j < LPQ; j++) a[j] = 0.0;
  }
  __syncwarp();
  const double2 bp = *reinterpret_cast<const double2*>(rowbuf);         // b00 b01
  const double2 bq = *reinterpret_cast<const double2*>(rowbuf + LPQ);   // b10 b11
  const double idet = fast_rcp(fma(bp.x, bq.y, -bp.y * bq.x));
  // non-pivot: (c0,c1) = -[f0 f1] B^-1 = idet * (f1*b10 - f0*b11 , f0*b01 - f1*b00)
  double c0 = fma(f1, bq.x, -f0 * bq.y), c1 = fma(f0, bp.y, -f1 * bp.x);
  if (piv) {                                                             // rows of B^-1 = idet*[[b11,-b01],[-b10,b00]]
    c0 = (k & 1) ? -bq.x : bq.y;
    c1 = (k & 1) ? bp.x : -bp.y;
  }
  c0 *= idet; c1 *= idet;
#pragma unroll
  for (int j = 2; j < LPQ; j += 2) {             // in place, increasing j: a[j-2] <- a[j] + c0*rowp[j] + c1*rowq[j]
    const double2 rp = *reinterpret_cast<const double2*>(rowbuf + j);
    const double2 rq = *reinterpret_cast<const double2*>(rowbuf + LPQ + j);
    a[j - 2] = fma(c0, rp.x, fma(c1, rq.x, a[j]));
    a[j - 1] = fma(c0, rp.y, fma(c1, rq.y, a[j + 1]));
  }
  a[LPQ - 2] = c0;
  a[LPQ - 1] = c1;
  __syncwarp();
}

// Shared-memory tables (zero padded LPQ x LPQ, symmetric, read column-wise = conflict free):
//   0: W  = M'M           1: D2'D2          (vehicle problem, steering space)
//   2: B1 = D'D           3: B2 = D'(A'A)D  (edge problem, position space; D = second difference map x = D t)
#define PIADMM_NTAB 4

// Vehicle problem: K = cI*I + c2*D2'D2 + cW*W + cT*T, T = A'A = 2I + 2 D1'D1 (tridiagonal: diag 4 at the
// ends / 6 inside, off-diagonal -2); inverted in place.  rowbuf: 2*LPQ doubles of shared memory per group.
template <bool GT>
__device__ __forceinline__ double tab_ld(const double* p) { return GT ? __ldg(p) : *p; }

// GT: the tables are read from global memory (L1-resident, read-only path) instead of shared memory
template <int LPQ, bool GT>
__device__ __forceinline__ void build_Kinv_vehicle(double (&a)[LPQ], int N, int k, double cI, double c2, double cW,
                                                   double cT, const double* __restrict__ tabs, double* rowbuf) {
  const double* Wsm = tabs;
  const double* D2sm = tabs + LPQ * LPQ;
  const double tdiag = (k == 0 || k == N - 1) ? 4.0 : 6.0;
#pragma unroll
  for (int j = 0; j < LPQ; j++) {
    double v = fma(cW, tab_ld<GT>(Wsm + j * LPQ + k), c2 * tab_ld<GT>(D2sm + j * LPQ + k));
    if (j == k) v += fma(cT, tdiag, cI);
    if ((j == k + 1 || j == k - 1) && j < N) v = fma(cT, -2.0, v);
    if (k >= N) v = (j == k) ? 1.0 : 0.0;
    a[j] = v;
  }
#pragma unroll 1
  for (int p = 0; p < LPQ; p += 2) gj_step2<LPQ>(a, k, p, rowbuf);
}

// Edge half problem in position space: K_t = c1*B1 + cT*B2 + cS*I_s, I_s = identity on lanes <= N-2.
template <int LPQ, bool GT>
__device__ __forceinline__ void build_Kinv_edge(double (&a)[LPQ], int N, int k, double c1, double cT, double cS,
                                                const double* __restrict__ tabs, double* rowbuf) {
  const double* B1 = tabs + 2 * LPQ * LPQ;
  const double* B2 = tabs + 3 * LPQ * LPQ;
#pragma unroll
  for (int j = 0; j < LPQ; j++) {
    double v = fma(c1, tab_ld<GT>(B1 + j * LPQ + k), cT * tab_ld<GT>(B2 + j * LPQ + k));
    if (j == k && k <= N - 2) v += cS;
    if (k >= N) v = (j == k) ? 1.0 : 0.0;
    a[j] = v;
  }
#pragma unroll 1
  for (int p = 0; p < LPQ; p += 2) gj_step2<LPQ>(a, k, p, rowbuf);
}

// sum_{L>=j} (L-j+1) v_L  (= M'(.) for a vector given on the position rows), v = 0 on masked lanes
template <int LPQ>
__device__ __forceinline__ double suffix2(double v, int k) { return incl_suffix<LPQ>(incl_suffix<LPQ>(v, k), k); }

// ------------------------------------------------------------------ inner OSQP-style ADMM
struct InnerOut {
  double x;
  int iters;
  int status;
};

// One (edge=false) or two coupled (edge=true) half problems per group; see oracle/solver.py.
//   P = pI*I + p2*D2'D2 + pW*W ; box rows A = [I;-I;-D1;D1] >= -[pi/6,pi/6,pi/9,pi/9]
//   edge: extra position rows s = (M x)_{1..N}, coupled with the partner by the prox of
//         beta*max(0, c - a0*s0 - a1*s1).  `edge` must be uniform over the pair.
//
// The edge half problem is iterated in POSITION SPACE: unknown t with x = D t,
//   t_L = (M x)_{L+2} (L <= N-2),  t_{N-1} = x_{N-1};   x_j = t_j - 2 t_{j-1} + t_{j-2} (j <= N-2).
// It is the same iteration in other coordinates (K_t = D' K_x D, rhs_t = D' rhs_x, x~ = D t~): identical
// iterates up to rounding, but M and M' (two chained warp scans each) drop out of the loop: the position
// rows become the identity (lane L <-> trajectory index L+2; lane N-1 holds the all-zero row of index 1)
// and D, D' are 3-point stencils.  Termination is still evaluated on the steering-space residuals.
// cvec follows that lane layout.  rhsbuf: 2*LPQ doubles of shared memory per group (the mat-vec uses the first
// LPQ, the K^-1 builds all of it), xch: 2 x 64 doubles per pair.
//
// KS (throughput variant of the fused kernel): the row of K^-1 is kept in SHARED memory (krow: this lane's
// row, LPQ doubles, rows padded to LPQ+2 so that a quarter-warp LDS.128 is conflict free) instead of
// registers, and the tables come from global memory.  That halves the register footprint, so that twice
// as many scenario-warps are resident to hide the dependent FP64 / LDS / SHFL latencies.  The Gauss-Jordan
// build still runs in registers; around a rebuild the ADMM state is parked in the (then free) row.
#define PIADMM_KROW_STRIDE(LPQ) ((LPQ) + 2)
#define PIADMM_KC_KEY 8                                        // doubles reserved for the key of a K^-1 cache entry
#define PIADMM_KC_STRIDE(LPQ) (PIADMM_KC_KEY + (LPQ) * (LPQ))  // doubles per cache entry
// Independent accumulator chains of the mat-vec K^-1 rhs (LPQ DFMA per lane): NA chains cost NA-1 extra DADD per
// iteration but shorten the dependent chain to LPQ/NA DFMA + log2(NA) DADD.  The batch kernels are issue bound
// (FP64 instructions hold the dispatch port for two cycles): fewest instructions win, NA = 4 (N=15: 0.2154 -> 0.2105 ms
// against 8; N=30: 0.626 -> 0.605 ms against 16).  The generic-V edge kernel is latency bound (few concurrent edge
// problems, config 5: 7.99 / 8.40 / 8.91 ms with 16 / 8 / 4 chains at N=30): most chains win.
template <int LPQ, bool LATENCY> __host__ __device__ constexpr int nacc() { return LATENCY ? (LPQ == 16 ? 8 : 16) : 4; }
template <int LPQ, bool KS, bool LATENCY = false>
__device__ __forceinline__ InnerOut inner_solve(const piadmm_params& ip, const InnerK& ik, int N, int k, int half, bool edge,
                                                double pI, double p2, double pW, double q, double a0, double a1,
                                                double cvec, const double* __restrict__ tabs, double* rhsbuf,
                                                double* xch, double* krow, double* kc) {
  const double al = ip.in_alpha, sg = ip.in_sigma, oma = 1.0 - al;
  // LPQ == 32 problems (N = 16..31) are latency critical -- few concurrent edge problems, the slowest one of a coupled
  // scene bounds the outer iteration -- so their edge iteration uses the forms with the shortest loop-carried chain
  // (one shuffle level less in the right-hand side and in A x~, at ~15 % more instructions); the LPQ == 16 batch
  // kernels are issue bound and keep the forms with the fewest instructions (measured both ways, r1_history.md).
  constexpr bool LAT = LPQ == 32;
  const double ha = ik.ha, b1 = ik.b1, alu = ik.alu, ald = ik.ald;
  const double l_u = PIADMM_L_U, l_d = PIADMM_L_D;
  double Kinv[LPQ];
  double rho = ip.in_rho0, irho = 1.0 / rho;
  // ADMM state per constraint row r: z_r and c_r = (1-alpha) z_r + y_r/rho.  The dual y_r is implicit,
  // y_r = rho (c_r - (1-alpha) z_r); the recurrences below are OSQP's written in (z, c):
  //   v = alpha (A x~)_r + c_r,  z' = Pi(v),  y' = rho (v - z')  =>  c' = v - alpha z',  g' = z' - y'/rho = 2 z' - v.
  // Per iteration and row that is 1 FMA + projection + 2 FMA (instead of relaxation, projection and a separate
  // dual update), and the loop-carried dependency chain is three FP64 operations shorter.
  // The bound rows are carried SHIFTED by their bound l: a_r = v_r - l and ch_r = c_r - l, so that
  //   a = alpha (A x~)_r + ch,   z' = l + max(a, 0) = l + (a + |a|)/2   (never formed in the loop),
  //   g' = 2 z' - v = l + |a|    (the l cancel in g0 - g1 and g3 - g2),
  //   ch' = v - alpha z' - l = (1 - alpha/2) a - (alpha/2) |a| - alpha l:
  // three DFMA and half a DADD per row, no compare / select, and |a| is one operation after the mat-vec result.
  double x = 0.0, a0r = -l_u, a1r = -l_u, a2r = -l_d, a3r = -l_d, ch0 = -l_u, ch1 = -l_u, ch2 = -l_d, ch3 = -l_d;
  double t = 0.0, zs = 0.0, cs = 0.0;
  double g01 = 0.0, ge = 0.0, gs = 0.0;       // (g0 - g1), (g3 - g2), g_s of the rows of this lane
  const double na = edge ? fma(a1, a1, a0 * a0) : 1.0;
  const double a_me = half == 0 ? a0 : a1;
  const bool inner_row = k < N - 1;
  const int srow = k <= N - 2;               // lane carries a real position row / a second-difference row of D
  // LAT (168 registers): per-lane coefficients instead of selects in the loop; at 128 registers (LPQ == 16) the six
  // extra registers cost more than the eight selects save (N=15: 0.2120 vs 0.2093 ms; N=30: 0.5925 -> 0.5839 ms)
  const double al_d = (LAT && !inner_row) ? 0.0 : al, al_s = srow ? al : 0.0;
  const int d_up1 = srow && k >= 1, d_up2 = srow && k >= 2;          // x_j = t_j - 2 t_{j-1} + t_{j-2}
  const int srow_next = k <= N - 3;                                  // the same three predicates for lane k+1
  const int n_up1 = srow_next, n_up2 = srow_next && k >= 1;
  // K^-1(rho0) depends only on the vehicle (speed), the outer penalty and the inner settings, not on the iterate:
  // it is the same in every outer iteration and every MPC step, so it is cached in global memory (kc: 8 doubles of
  // key = the four coefficients that define K, then the matrix stored so that lane k reads/writes consecutive
  // doubles).  A miss (first use, changed rho / parameters) builds and refreshes the entry; the decision is warp
  // uniform because the build synchronises the warp.
  {
    const double kI = pI + sg, kB = edge ? rho : p2, kW = edge ? pW + rho : pW, kT = edge ? -1.0 : rho;
    bool hit = false;
    if (kc) hit = kc[0] == kI && kc[1] == kB && kc[2] == kW && kc[3] == kT;
    if (__all_sync(PIADMM_FULL, hit)) {
#pragma unroll
      for (int j = 0; j < LPQ; j++) Kinv[j] = kc[PIADMM_KC_KEY + j * LPQ + k];
    } else {
      if (edge) build_Kinv_edge<LPQ, KS>(Kinv, N, k, pI + sg, rho, pW + rho, tabs, rhsbuf);
      else build_Kinv_vehicle<LPQ, KS>(Kinv, N, k, pI + sg, p2, pW, rho, tabs, rhsbuf);
      // kc is uniform per GROUP only: at LPQ == 16 a warp of the generic-V x-update can hold one group with an entry
      // and one masked group (finished scenario / vehicle of another rank) without, so the barrier between the
      // matrix stores and the key store must sit outside the `if (kc)` (a barrier only half the warp reaches pairs
      // up with the next barrier of the other half, and the halves then read each other's rhs one iteration off)
      if (kc) {
#pragma unroll
        for (int j = 0; j < LPQ; j++) kc[PIADMM_KC_KEY + j * LPQ + k] = Kinv[j];
      }
      // an edge entry is shared by every edge problem of the vehicle (K_t depends on the vehicle's speed and the outer
      // penalty only), and in the generic-V edge kernel those run concurrently in different CTAs: concurrent writers
      // store identical bits, and a reader that sees the new key must see the new matrix -> fence between the two
      // (keys and rows are only ever read after a key match of the same launch, so no stale L1 line can pair up)
      __threadfence();
      __syncwarp();
      if (kc && k == 0) { kc[0] = kI; kc[1] = kB; kc[2] = kW; kc[3] = kT; }
    }
  }
  if (KS) {
#pragma unroll
    for (int j = 0; j < LPQ; j += 2) *reinterpret_cast<double2*>(krow + j) = make_double2(Kinv[j], Kinv[j + 1]);
  }
  InnerOut out;
  out.x = 0.0; out.iters = ip.in_max_iter; out.status = 1;
  bool done = false;
  // everything the loop body reads is copied out of *ip first: ip points into the kernel parameter block
  // through a generic pointer, and a load from it inside the loop is a real memory access on the chain
  const int check_every = ip.in_check;
  int xpar = 0;                              // buffer parity of the LPQ==32 partner exchange
  int left = ip.in_max_iter;                 // iterations still allowed
  const double inv_na = 1.0 / na;
  double tb = ip.beta * irho;                // hinge prox step bound beta / rho
  double rho_s = srow ? rho : 0.0;           // rho on the position rows, 0 elsewhere
  // two nested loops: the hot one counts a block of in_check iterations down in ONE register; `left`, the
  // termination test and the K^-1 rebuild are touched once per block
#pragma unroll 1
  while (left > 0) {
   const int blk = left < check_every ? left : check_every;
#pragma unroll 1
   for (int cnt = blk; cnt > 0; cnt--) {
    // ---- steering-space right-hand side  v = sigma x - q + A'(rho z - y)
    // A'(rho z - y) = rho A'g.  e_{k-1}: lane 0 reads the last lane of the group, a padded lane (N <= LPQ-1)
    // whose g is exactly 0.
    double rhs;
    if (!edge || !LAT) {
      rhs = fma(rho, (g01 - ge) + __shfl_sync(PIADMM_FULL, ge, (k - 1) & (LPQ - 1), LPQ), fma(sg, x, -q));
      if (edge) {
        const double vm = srow ? rhs : 0.0;
        const double v1 = __shfl_down_sync(PIADMM_FULL, vm, 1, LPQ), v2 = __shfl_down_sync(PIADMM_FULL, vm, 2, LPQ);
        rhs = fma(-2.0, v1, rhs) + v2;
        fma_if(rhs, rho, gs, srow);
      }
    } else {
      // rhs_t = D' v + rho g_s on the position rows, with v_k = e_k + f_{k-1} split into its lane-local part
      // e = rho (g0 - g1 - ge) + sigma x - q and the neighbour part f = rho ge: all four shuffles of the composed
      // stencil start from lane-local values (one shuffle level on the loop-carried chain instead of two).
      // Lanes beyond N-2 must not contribute: em / fm are zeroed at the source, and the clamped shuffles of the
      // last lanes return their own, zeroed, value.
      const double e = fma(rho, g01 - ge, fma(sg, x, -q)), f = rho * ge;
      const double em = srow ? e : 0.0, fm = srow_next ? f : 0.0;        // masks of the rows k and k+1
      const double fu = __shfl_sync(PIADMM_FULL, f, (k - 1) & (LPQ - 1), LPQ);
      const double e1 = __shfl_down_sync(PIADMM_FULL, em, 1, LPQ), e2 = __shfl_down_sync(PIADMM_FULL, em, 2, LPQ);
      const double f1 = __shfl_down_sync(PIADMM_FULL, fm, 1, LPQ);
      rhs = fma(-2.0, e1 + fm, e + fu) + (e2 + f1);
      rhs = fma(rho_s, gs, rhs);
    }
    rhsbuf[k] = rhs;
    __syncwarp();
    // ---- K^-1 rhs  (row k in registers, rhs broadcast from shared memory)
    // independent accumulator chains of the mat-vec: the dependent-DFMA depth is LPQ / NACC (+ log2 NACC adds)
    constexpr int PIADMM_NACC = nacc<LPQ, LATENCY>();
    double acc[PIADMM_NACC];
#pragma unroll
    for (int u = 0; u < PIADMM_NACC; u++) acc[u] = 0.0;
#pragma unroll
    for (int j = 0; j < LPQ; j += 4) {
      const double2 r0 = *reinterpret_cast<const double2*>(rhsbuf + j);
      const double2 r1 = *reinterpret_cast<const double2*>(rhsbuf + j + 2);
      double k0x, k0y, k1x, k1y;
      if (KS) {
        const double2 k0 = *reinterpret_cast<const double2*>(krow + j);
        const double2 k1 = *reinterpret_cast<const double2*>(krow + j + 2);
        k0x = k0.x; k0y = k0.y; k1x = k1.x; k1y = k1.y;
      } else {
        k0x = Kinv[j]; k0y = Kinv[j + 1]; k1x = Kinv[j + 2]; k1y = Kinv[j + 3];
      }
      acc[(j + 0) % PIADMM_NACC] = fma(k0x, r0.x, acc[(j + 0) % PIADMM_NACC]);
      acc[(j + 1) % PIADMM_NACC] = fma(k0y, r0.y, acc[(j + 1) % PIADMM_NACC]);
      acc[(j + 2) % PIADMM_NACC] = fma(k1x, r1.x, acc[(j + 2) % PIADMM_NACC]);
      acc[(j + 3) % PIADMM_NACC] = fma(k1y, r1.y, acc[(j + 3) % PIADMM_NACC]);
    }
    __syncwarp();
#pragma unroll
    for (int w = PIADMM_NACC / 2; w >= 1; w >>= 1) {
#pragma unroll
      for (int u = 0; u < w; u++) acc[u] += acc[u + w];
    }
    double xt = acc[0];
    double tt = 0.0;
    double d;
    if (!edge || !LAT) {
      if (edge) {
        tt = xt;
        const double u1 = __shfl_up_sync(PIADMM_FULL, tt, 1, LPQ), u2 = __shfl_up_sync(PIADMM_FULL, tt, 2, LPQ);
        fma_if(xt, -2.0, u1, d_up1);
        fma_if(xt, 1.0, u2, d_up2);
      }
      d = __shfl_down_sync(PIADMM_FULL, xt, 1, LPQ) - xt;              // the last lane is padded (xt = 0): no clamp needed
    } else {                                                           // x~ = D t~ and its forward difference,
      tt = xt;                                                         // both straight from the shuffles of t~
      const double u1 = __shfl_up_sync(PIADMM_FULL, tt, 1, LPQ), u2 = __shfl_up_sync(PIADMM_FULL, tt, 2, LPQ);
      const double dn = __shfl_down_sync(PIADMM_FULL, tt, 1, LPQ);
      xt = fma(-2.0, d_up1 ? u1 : 0.0, tt) + (d_up2 ? u2 : 0.0);
      const double xn = fma(-2.0, n_up1 ? tt : 0.0, dn) + (n_up2 ? u1 : 0.0);   // x~_{k+1}
      d = xn - xt;
    }
    if (!LAT && !inner_row) d = 0.0;
    x = fma(al, xt, oma * x);
    {   // LAT: al_d = 0 on the lanes without a rate row, their (meaningless) d never enters
      a0r = fma(al, xt, ch0); a1r = fma(-al, xt, ch1); a2r = fma(-al_d, d, ch2); a3r = fma(al_d, d, ch3);
      const double m0 = fabs(a0r), m1 = fabs(a1r), m2 = fabs(a2r), m3 = fabs(a3r);
      g01 = m0 - m1;
      ge = m3 - m2;
      ch0 = fma(b1, a0r, fma(-ha, m0, -alu)); ch1 = fma(b1, a1r, fma(-ha, m1, -alu));
      ch2 = fma(b1, a2r, fma(-ha, m2, -ald)); ch3 = fma(b1, a3r, fma(-ha, m3, -ald));
    }
    if (edge) {
      t = fma(al, tt, oma * t);
      // position row value (lane N-1: the zero row); LAT: al_s = 0 there instead of a select
      const double vs = LAT ? fma(al_s, tt, cs) : fma(al, srow ? tt : 0.0, cs);
      double tau;
      if (LAT) {                                    // the halves exchange their products a_me s_me: the sum is
        const double pm = a_me * vs;                // bitwise the same on both sides, no selects
        tau = pm + partner<LPQ>(pm, xch, k, half, xpar);
      } else {
        const double vo = partner<LPQ>(vs, xch, k, half, xpar);
        const double s0 = half == 0 ? vs : vo, s1 = half == 0 ? vo : vs;
        tau = fma(a1, s1, a0 * s0);
      }
      // prox of beta*max(0, c - a's): step = clamp((c - tau)/|a|^2, 0, beta/rho), branch free
      const double step = dmin(dmax((cvec - tau) * inv_na, 0.0), tb);
      zs = fma(step, a_me, vs);
      gs = fma(2.0, zs, -vs);
      cs = fma(-al, zs, vs);
    }
   }
   left -= blk;
    // ---- termination / adaptive rho every in_check iterations (steering-space residuals)
    if (blk == check_every) {
      const double z0 = l_u + dmax(a0r, 0.0), z1 = l_u + dmax(a1r, 0.0), z2 = l_d + dmax(a2r, 0.0),
                   z3 = l_d + dmax(a3r, 0.0);
      const double c0 = ch0 + l_u, c1 = ch1 + l_u, c2 = ch2 + l_d, c3 = ch3 + l_d;
      const double y0 = rho * fma(-oma, z0, c0), y1 = rho * fma(-oma, z1, c1), y2 = rho * fma(-oma, z2, c2),
                   y3 = rho * fma(-oma, z3, c3), ys = rho * fma(-oma, zs, cs);      // the implicit duals
      double dd = shift_down1<LPQ>(x, k) - x;
      if (!inner_row) dd = 0.0;
      double rp = dmax(dmax(fabs(x - z0), fabs(-x - z1)), dmax(fabs(-dd - z2), fabs(dd - z3)));
      double np = dmax(dmax(fabs(x), fabs(dd)), dmax(dmax(fabs(z0), fabs(z1)), dmax(fabs(z2), fabs(z3))));
      double Px, Aty;
      const double ey = y3 - y2;
      Aty = (y0 - y1) + shift_up1<LPQ>(ey, k) - ey;
      if (edge) {
        const double st = srow ? t : 0.0;           // (M x) on the position rows
        rp = dmax(rp, fabs(st - zs));
        np = dmax(np, dmax(fabs(st), fabs(zs)));
        Px = fma(pW, suffix2<LPQ>(st, k), pI * x);   // P x = pI x + pW M'(M x)
        Aty += suffix2<LPQ>(srow ? ys : 0.0, k);     // + M' y_s
      } else {
        // (M'M x)_k as a dot product with column k of the W table (the one K was built from) instead of four warp
        // scans (~150 instructions at LPQ == 16).  Deliberately a ROLLED loop: unrolled, its loads next to the live
        // K^-1 row make ptxas spill (measured: slower than the scans).
        rhsbuf[k] = x;
        __syncwarp();
        double w0 = 0.0, w1 = 0.0;
#pragma unroll 1
        for (int j = 0; j < LPQ; j += 2) {
          const double2 xj = *reinterpret_cast<const double2*>(rhsbuf + j);
          w0 = fma(tab_ld<KS>(tabs + j * LPQ + k), xj.x, w0);
          w1 = fma(tab_ld<KS>(tabs + (j + 1) * LPQ + k), xj.y, w1);
        }
        __syncwarp();
        Px = fma(pW, w0 + w1, pI * x);
        const double x1 = shift_down1<LPQ>(x, k), x2 = shift_down1<LPQ>(x1, k);
        const double ee = (k < N - 2) ? (x - 2.0 * x1 + x2) : 0.0;
        const double e1 = shift_up1<LPQ>(ee, k), e2 = shift_up1<LPQ>(e1, k);
        Px = fma(p2, ee - 2.0 * e1 + e2, Px);
      }
      double rd = fabs(Px + q + Aty);
      double nd = dmax(fabs(Px), dmax(fabs(Aty), fabs(q)));
      if (LPQ == 16) group_max4_16(rp, np, rd, nd, k, edge);
      else { rp = group_max_nn<LPQ>(rp); np = group_max_nn<LPQ>(np); rd = group_max_nn<LPQ>(rd); nd = group_max_nn<LPQ>(nd); }
      if (edge && LPQ == 32) {
        partner_max4<LPQ>(rp, np, rd, nd, xch, k, half, xpar);
      }
      const bool conv = rp <= ip.in_eps_abs + ip.in_eps_rel * np && rd <= ip.in_eps_abs + ip.in_eps_rel * nd;
      if (conv && !done) {
        done = true;
        out.x = x; out.iters = ip.in_max_iter - left; out.status = 0;
      }
      bool refac = false;
      double rn = rho;
      if (!conv && !done && ip.in_adaptive) {
        const double a = rp / dmax(np, 1e-30);
        const double b = dmax(rd / dmax(nd, 1e-30), 1e-30);
        rn = dmin(dmax(rho * sqrt(a / b), 1e-6), 1e6);
        refac = rn > ip.in_adapt_tol * rho || rn < rho / ip.in_adapt_tol;
      }
      // loop exit must be uniform over everything that synchronises inside the loop:
      // the warp (both groups when LPQ==16) and, for an edge with LPQ==32, the CTA pair
      // (conv is identical in both halves of an edge by construction).
      if (__all_sync(PIADMM_FULL, done)) break;
      if (__any_sync(PIADMM_FULL, refac)) {
        if (refac) {       // OSQP keeps (z, y) and changes rho: re-express c and g in the new rho
          rho = rn; irho = 1.0 / rho; tb = ip.beta * irho; rho_s = srow ? rho : 0.0;
          ch0 = fma(y0, irho, oma * z0) - l_u; ch1 = fma(y1, irho, oma * z1) - l_u;
          ch2 = fma(y2, irho, oma * z2) - l_d; ch3 = fma(y3, irho, oma * z3) - l_d;
          cs = fma(ys, irho, oma * zs);
          g01 = fma(-y0, irho, z0) - fma(-y1, irho, z1);
          ge = fma(-y3, irho, z3) - fma(-y2, irho, z2);
          gs = fma(-ys, irho, zs);
        }
        if (KS) {      // park the ADMM state in the row that is about to be rebuilt (frees 24 registers for the build)
          krow[0] = x; krow[1] = a0r; krow[2] = a1r; krow[3] = a2r; krow[4] = a3r; krow[5] = ch0; krow[6] = ch1;
          krow[7] = ch2; krow[8] = ch3; krow[9] = t; krow[10] = zs; krow[11] = cs; krow[12] = g01; krow[13] = ge;
          krow[14] = gs;
        }
        if (edge) build_Kinv_edge<LPQ, KS>(Kinv, N, k, pI + sg, rho, pW + rho, tabs, rhsbuf);
        else build_Kinv_vehicle<LPQ, KS>(Kinv, N, k, pI + sg, p2, pW, rho, tabs, rhsbuf);
        if (KS) {
          x = krow[0]; a0r = krow[1]; a1r = krow[2]; a2r = krow[3]; a3r = krow[4]; ch0 = krow[5]; ch1 = krow[6];
          ch2 = krow[7]; ch3 = krow[8]; t = krow[9]; zs = krow[10]; cs = krow[11]; g01 = krow[12]; ge = krow[13];
          gs = krow[14];
#pragma unroll
          for (int j = 0; j < LPQ; j += 2) *reinterpret_cast<double2*>(krow + j) = make_double2(Kinv[j], Kinv[j + 1]);
        }
      }
    }
  }
  if (!done) out.x = x;
  return out;
}
