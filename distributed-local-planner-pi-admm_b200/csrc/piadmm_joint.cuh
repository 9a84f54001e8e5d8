// piadmm_joint.cuh -- the joint N-vehicle RSU edge problem: cost_function_edge of
// matlab_old_files/ADMM_fmincon_multi_veh_intesection_static.m:250-291 (= casadi_old_PI_ADMM/PI_ADMM_class.py:141-178) in
// its convex statement, ONE problem over all vehicles of a scene:
//
//   min  sum_i [ Pcost |u_i|^2 + rho/2 |pos_old_i - P_i(u_i) + lambda_i|^2 ]
//        + hinge * sum_{pairs (i<j)} sum_{k=1..N} max(0, d^2 - (2 l_ij'(P_j,k - P_i,k) - |l_ij|^2))      s.t. box / rate rows
//
// with ONE consensus copy and ONE multiplier per vehicle (neighbour slot 0 of the bound buffers) and l_ij = seed_j - seed_i.
// The reference sums over ORDERED pairs, i.e. hinge = 2 beta per pair (the host passes the weight).
//
// Splitting (oracle/joint.py): z_i = [A u_i ; one copy of (M u_i)_{1..N} per incident pair], so the linear system stays one
// N x N matrix per vehicle, K_i = H_i + sigma I + rho (A'A + deg_i M'M), iterated in position space exactly as the edge
// half problem of inner_solve (piadmm_dev.cuh), and the pairs couple through the closed-form 2-D prox of their hinge.
//
// Execution: one THREAD-BLOCK CLUSTER per scene.  A group of LPQ lanes owns a vehicle (K^-1 row per lane in registers,
// per-pair row state in registers), 256-thread CTAs hold 8 (N <= 31) or 16 (N <= 15) vehicles, a cluster of up to 8 CTAs
// holds the scene (V <= 64 / 128).  Per inner iteration every vehicle publishes its position-space iterate t~ (LPQ
// doubles) in its CTA's shared memory, the cluster meets at ONE hardware cluster barrier, and every vehicle reads the t~ of
// its <= 8 partners through distributed shared memory; both endpoints of a pair evaluate the pair's prox redundantly from
// identical inputs in identical order (bitwise equal results, so no second exchange).  Termination and the adaptive rho
// use the residuals of the WHOLE problem: four maxima per vehicle, exchanged the same way every in_check iterations.
#pragma once
#include <cooperative_groups.h>
#include "piadmm_dev.cuh"
namespace cg = cooperative_groups;

#define JT_THREADS 256
#define JT_D 8                         // incident pairs per vehicle (per-pair row state lives in registers)
#define JT_MAX_CLUSTER 8               // portable cluster size

struct JointArgs {
  const int32_t* pair_n;               // [S] pairs of every scenario
  const int32_t* pair_ij;              // [S,P,2] (veh_lo < veh_hi)
  int P;
  double hinge;                        // weight of one pair's hinge (2 beta for the reference's ordered-pair sum)
  double* u_out;                       // [S,V,N] rounded joint controls
  int32_t* iters_out;                  // [S] inner iterations of the joint solve
  int32_t* status_out;                 // [S] 0 ok, PIADMM_ST_EDGE_MAXITER, PIADMM_ST_SLOT_OVERFLOW (more than JT_D pairs at a vehicle)
};

template <int LPQ> __host__ __device__ constexpr int jt_groups() { return JT_THREADS / LPQ; }
// dynamic shared memory of one CTA, in doubles
template <int LPQ> __host__ __device__ constexpr size_t jt_smem_doubles() {
  return (size_t)PIADMM_NTAB * LPQ * LPQ            // tables
         + (size_t)jt_groups<LPQ>() * 2 * LPQ       // gbuf: rhs broadcast / Gauss-Jordan rows
         + (size_t)2 * jt_groups<LPQ>() * LPQ       // tbuf: published t~, double buffered
         + (size_t)2 * jt_groups<LPQ>() * 4         // red: four maxima per vehicle, double buffered
         + (size_t)jt_groups<LPQ>() * JT_D * LPQ    // cvec of every pair slot
         + (size_t)jt_groups<LPQ>() * JT_D * LPQ    // zs of the partner's side (needed when rho changes)
         + (size_t)jt_groups<LPQ>() * JT_D * 4      // a_me, a_pt, 1/|a|^2, me-is-lo per slot
         + (size_t)jt_groups<LPQ>() * JT_D;         // partner ids (as int32 pairs) + degree
}

template <int LPQ>
__global__ void __launch_bounds__(JT_THREADS, 1) k_joint(const __grid_constant__ DevCtx c, const JointArgs ja, int CL) {
  extern __shared__ __align__(16) double jsm[];
  constexpr int G = jt_groups<LPQ>();
  double* Wsm = jsm;
  double* gbuf_all = Wsm + PIADMM_NTAB * LPQ * LPQ;
  double* tbuf = gbuf_all + G * 2 * LPQ;            // [2][G][LPQ]
  double* red = tbuf + 2 * G * LPQ;                 // [2][G][4]
  double* pcv = red + 2 * G * 4;                    // [G][JT_D][LPQ]
  double* zpt = pcv + G * JT_D * LPQ;               // [G][JT_D][LPQ]
  double* psc = zpt + G * JT_D * LPQ;               // [G][JT_D][4]
  int32_t* pnb = reinterpret_cast<int32_t*>(psc + G * JT_D * 4);   // [G][JT_D] partner vehicle, then [G] degree at +G*JT_D
  int32_t* pdeg = pnb + G * JT_D;

  cg::cluster_group cl = cg::this_cluster();
  const int rank = (int)cl.block_rank();
  const int s = blockIdx.x / CL;
  const piadmm_params& p = c.p;
  const int N = c.N, V = c.V;
  const int lane = threadIdx.x & 31, k = lane & (LPQ - 1);
  const int g = threadIdx.x / LPQ;                  // group in the CTA
  const int i = rank * G + g;                       // vehicle
  const bool valid = i < V;
  double* gbuf = gbuf_all + g * 2 * LPQ;

  load_tabs<LPQ>(c, Wsm, 0, PIADMM_NTAB);           // (ends with a CTA barrier)

  // ---- set-up: this vehicle's data, its incident pairs in list order
  const double rho_o = c.b.rho_s[s];
  double x0 = 0.0, y0 = 0.0, th0 = 0.0, v = 1.0;
  if (valid) {
    const double* xt = c.b.xt + ((size_t)s * V + i) * 3;
    x0 = xt[0]; y0 = xt[1]; th0 = xt[2];
    v = c.b.spd[(size_t)s * V + i];
  }
  double sn, cs;
  sincos(th0, &sn, &cs);
  const double gi = v * v * p.dt * p.dt / p.L;
  if (k == 0) {
    int deg = 0;
    bool over = false;
    if (valid) {
      const int np_ = ja.pair_n[s];
      const int32_t* pl = ja.pair_ij + (size_t)s * ja.P * 2;
      for (int pi = 0; pi < np_; pi++) {
        const int a = pl[2 * pi], b = pl[2 * pi + 1];
        if (a == i || b == i) {
          if (deg < JT_D) pnb[g * JT_D + deg++] = a == i ? b : a; else over = true;
        }
      }
    }
    for (int d = deg; d < JT_D; d++) pnb[g * JT_D + d] = valid ? i : 0;   // unused slots point at a readable row
    pdeg[g] = over ? -1 : deg;
  }
  __syncwarp();
  int deg = pdeg[g];
  // a vehicle with more than JT_D pairs: nobody solves (the decision must be uniform over the cluster)
  {
    if (k == 0) red[g * 4] = deg < 0 ? 1.0 : 0.0;
    cl.sync();
    double any = 0.0;
    for (int j = lane; j < CL * G; j += 32) any = fmax(any, cl.map_shared_rank(red, j / G)[(j % G) * 4]);
    any = fmax(any, __shfl_xor_sync(PIADMM_FULL, any, 16)); any = fmax(any, __shfl_xor_sync(PIADMM_FULL, any, 8));
    any = fmax(any, __shfl_xor_sync(PIADMM_FULL, any, 4)); any = fmax(any, __shfl_xor_sync(PIADMM_FULL, any, 2));
    any = fmax(any, __shfl_xor_sync(PIADMM_FULL, any, 1));
    cl.sync();                                       // red is reused below
    if (any > 0.0) {
      if (rank == 0 && threadIdx.x == 0) { ja.status_out[s] = PIADMM_ST_SLOT_OVERFLOW; ja.iters_out[s] = 0; }
      return;
    }
  }
  // AL term: q = rho g M'(s ex - c ey) with the vehicle's ONE consensus copy (slot 0)
  double vec = 0.0;
  if (valid && k <= N) {
    const double* pos = c.b.pos + POS_OFF(s, i);
    const double* du = c.b.dual + SLOT_OFF(s, i, 0);
    const double xb = x0 + k * p.dt * v * cs, yb = y0 + k * p.dt * v * sn;
    vec = sn * (pos[k] + du[k] - xb) - cs * (pos[NP1 + k] + du[NP1 + k] - yb);
  }
  const double q = rho_o * gi * apply_Mt<LPQ>(vec, k);
  const double sdx = rint((x0 + p.dt * v * cs) * 1e4) / 1e4, sdy = rint((y0 + p.dt * v * sn) * 1e4) / 1e4;   // casadi/main.py:48-49
  for (int d = 0; d < JT_D; d++) {
    double cfull = 0.0, a_me = 0.0, a_pt = 0.0, ina = 0.0, melo = 1.0;
    if (d < deg) {                                   // group uniform
      const int j = pnb[g * JT_D + d];
      const double* xj = c.b.xt + ((size_t)s * V + j) * 3;
      const double vj = c.b.spd[(size_t)s * V + j];
      double sj, cj;
      sincos(xj[2], &sj, &cj);
      const double gj = vj * vj * p.dt * p.dt / p.L;
      const double sjx = rint((xj[0] + p.dt * vj * cj) * 1e4) / 1e4, sjy = rint((xj[1] + p.dt * vj * sj) * 1e4) / 1e4;
      const bool me_lo = i < j;
      // lo / hi endpoint of the pair: every quantity is formed from (lo, hi) in the same order on both sides
      const double lx = me_lo ? sjx - sdx : sdx - sjx, ly = me_lo ? sjy - sdy : sdy - sjy;       // l = seed_hi - seed_lo
      const double s_lo = me_lo ? sn : sj, c_lo = me_lo ? cs : cj, g_lo = me_lo ? gi : gj;
      const double s_hi = me_lo ? sj : sn, c_hi = me_lo ? cj : cs, g_hi = me_lo ? gj : gi;
      const double a_lo = -2.0 * g_lo * (-lx * s_lo + ly * c_lo), a_hi = 2.0 * g_hi * (-lx * s_hi + ly * c_hi);
      const double na = fma(a_hi, a_hi, a_lo * a_lo);
      a_me = me_lo ? a_lo : a_hi; a_pt = me_lo ? a_hi : a_lo; ina = na > 0.0 ? 1.0 / na : 0.0; melo = me_lo ? 1.0 : 0.0;
      if (k <= N) {
        const double xlo0 = me_lo ? x0 : xj[0], ylo0 = me_lo ? y0 : xj[1], vlo = me_lo ? v : vj;
        const double xhi0 = me_lo ? xj[0] : x0, yhi0 = me_lo ? xj[1] : y0, vhi = me_lo ? vj : v;
        const double xb0 = xlo0 + k * p.dt * vlo * c_lo, yb0 = ylo0 + k * p.dt * vlo * s_lo;
        const double xb1 = xhi0 + k * p.dt * vhi * c_hi, yb1 = yhi0 + k * p.dt * vhi * s_hi;
        const double base0 = -2.0 * (lx * xb0 + ly * yb0), base1 = 2.0 * (lx * xb1 + ly * yb1);
        cfull = p.dis_thres * p.dis_thres - (base0 + base1 - (lx * lx + ly * ly));
      }
    }
    // position-row layout: lane L <= N-2 <-> trajectory index L+2, lane N-1 <-> index 1
    double cv = __shfl_down_sync(PIADMM_FULL, cfull, 2, LPQ);
    const double c_one = __shfl_sync(PIADMM_FULL, cfull, 1, LPQ);
    if (k == N - 1) cv = c_one;
    if (!valid || k >= N || d >= deg) cv = -1e300;
    pcv[(g * JT_D + d) * LPQ + k] = cv;
    zpt[(g * JT_D + d) * LPQ + k] = 0.0;
    if (k == 0) { double* o = psc + (g * JT_D + d) * 4; o[0] = a_me; o[1] = a_pt; o[2] = ina; o[3] = melo; }
  }
  __syncwarp();

  // ---- inner ADMM (OSQP iteration, (z, c) state form of inner_solve; edge half problem in position space)
  const double al = p.in_alpha, sg = p.in_sigma, oma = 1.0 - al;
  const double ha = c.ik.ha, b1 = c.ik.b1, alu = c.ik.alu, ald = c.ik.ald;
  const double l_u = PIADMM_L_U, l_d = PIADMM_L_D;
  const double pI = 2.0 * p.Pcost, pW = valid ? rho_o * gi * gi : 0.0;
  double Kinv[LPQ];
  double rho = p.in_rho0, irho = 1.0 / rho;
  build_Kinv_edge<LPQ, false>(Kinv, N, k, pI + sg, rho, pW + deg * rho, Wsm, gbuf);
  double x = 0.0, t = 0.0, a0r = -l_u, a1r = -l_u, a2r = -l_d, a3r = -l_d, ch0 = -l_u, ch1 = -l_u, ch2 = -l_d, ch3 = -l_d;
  double g01 = 0.0, ge = 0.0, gsum = 0.0;
  double cs_me[JT_D], cs_pt[JT_D], zs_me[JT_D];
#pragma unroll
  for (int d = 0; d < JT_D; d++) { cs_me[d] = 0.0; cs_pt[d] = 0.0; zs_me[d] = 0.0; }
  const bool inner_row = k < N - 1;
  const int srow = k <= N - 2;
  const int d_up1 = srow && k >= 1, d_up2 = srow && k >= 2;
  const int check_every = p.in_check;
  int left = p.in_max_iter, iters = p.in_max_iter, status = PIADMM_ST_EDGE_MAXITER;
  double tb = ja.hinge * irho;
  int par = 0, rpar = 0;
  bool done = false;
  double x_out = 0.0;
#pragma unroll 1
  while (left > 0) {
    const int blk = left < check_every ? left : check_every;
#pragma unroll 1
    for (int cnt = blk; cnt > 0; cnt--) {
      double rhs = fma(rho, (g01 - ge) + __shfl_sync(PIADMM_FULL, ge, (k - 1) & (LPQ - 1), LPQ), fma(sg, x, -q));
      {
        const double vm = srow ? rhs : 0.0;
        const double v1 = __shfl_down_sync(PIADMM_FULL, vm, 1, LPQ), v2 = __shfl_down_sync(PIADMM_FULL, vm, 2, LPQ);
        rhs = fma(-2.0, v1, rhs) + v2;
        fma_if(rhs, rho, gsum, srow);
      }
      gbuf[k] = rhs;
      __syncwarp();
      double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
      for (int j = 0; j < LPQ; j += 4) {
        const double2 r0 = *reinterpret_cast<const double2*>(gbuf + j);
        const double2 r1 = *reinterpret_cast<const double2*>(gbuf + j + 2);
        acc[0] = fma(Kinv[j], r0.x, acc[0]); acc[1] = fma(Kinv[j + 1], r0.y, acc[1]);
        acc[2] = fma(Kinv[j + 2], r1.x, acc[2]); acc[3] = fma(Kinv[j + 3], r1.y, acc[3]);
      }
      __syncwarp();
      const double tt = (acc[0] + acc[2]) + (acc[1] + acc[3]);
      // ---- the pairs: publish t~ (position rows only) and ARRIVE at the cluster barrier; the bound rows of this vehicle
      // are updated while the other CTAs arrive, the partners' rows are read after the WAIT
      const double tme = srow ? tt : 0.0;
      tbuf[(par * G + g) * LPQ + k] = tme;
      cl.barrier_arrive();
      double xt = tt;
      {
        const double u1 = __shfl_up_sync(PIADMM_FULL, tt, 1, LPQ), u2 = __shfl_up_sync(PIADMM_FULL, tt, 2, LPQ);
        fma_if(xt, -2.0, u1, d_up1);
        fma_if(xt, 1.0, u2, d_up2);
      }
      double dd = __shfl_down_sync(PIADMM_FULL, xt, 1, LPQ) - xt;
      if (!inner_row) dd = 0.0;
      x = fma(al, xt, oma * x);
      a0r = fma(al, xt, ch0); a1r = fma(-al, xt, ch1); a2r = fma(-al, dd, ch2); a3r = fma(al, dd, ch3);
      const double m0 = fabs(a0r), m1 = fabs(a1r), m2 = fabs(a2r), m3 = fabs(a3r);
      g01 = m0 - m1;
      ge = m3 - m2;
      ch0 = fma(b1, a0r, fma(-ha, m0, -alu)); ch1 = fma(b1, a1r, fma(-ha, m1, -alu));
      ch2 = fma(b1, a2r, fma(-ha, m2, -ald)); ch3 = fma(b1, a3r, fma(-ha, m3, -ald));
      t = fma(al, tt, oma * t);
      cl.barrier_wait();
      gsum = 0.0;
      // all partner rows first (independent distributed-shared-memory loads in flight together), then the proxes
      double tjv[JT_D];
#pragma unroll
      for (int d = 0; d < JT_D; d++) {
        const int j = pnb[g * JT_D + d];
        tjv[d] = cl.map_shared_rank(tbuf, j / G)[(par * G + (j % G)) * LPQ + k];
      }
#pragma unroll
      for (int d = 0; d < JT_D; d++) {
        if (d < deg) {                               // group uniform
          const double* sc = psc + (g * JT_D + d) * 4;
          const double tj = tjv[d];
          const double a_me = sc[0], a_pt = sc[1];
          const double vs_me = fma(al, tme, cs_me[d]), vs_pt = fma(al, tj, cs_pt[d]);
          const bool me_lo = sc[3] != 0.0;
          const double v_lo = me_lo ? vs_me : vs_pt, v_hi = me_lo ? vs_pt : vs_me;
          const double a_lo = me_lo ? a_me : a_pt, a_hi = me_lo ? a_pt : a_me;
          const double tau = fma(a_hi, v_hi, a_lo * v_lo);
          const double step = dmin(dmax((pcv[(g * JT_D + d) * LPQ + k] - tau) * sc[2], 0.0), tb);
          const double zm = fma(step, a_me, vs_me), zp = fma(step, a_pt, vs_pt);
          zs_me[d] = zm;
          zpt[(g * JT_D + d) * LPQ + k] = zp;
          gsum += fma(2.0, zm, -vs_me);
          cs_me[d] = fma(-al, zm, vs_me);
          cs_pt[d] = fma(-al, zp, vs_pt);
        }
      }
      par ^= 1;
    }
    left -= blk;
    if (blk == check_every) {
      // ---- termination / adaptive rho on the residuals of the WHOLE problem (steering-space residuals)
      const double z0 = l_u + dmax(a0r, 0.0), z1 = l_u + dmax(a1r, 0.0), z2 = l_d + dmax(a2r, 0.0), z3 = l_d + dmax(a3r, 0.0);
      const double c0 = ch0 + l_u, c1 = ch1 + l_u, c2 = ch2 + l_d, c3 = ch3 + l_d;
      const double y0_ = rho * fma(-oma, z0, c0), y1_ = rho * fma(-oma, z1, c1), y2_ = rho * fma(-oma, z2, c2),
                   y3_ = rho * fma(-oma, z3, c3);
      double dx = shift_down1<LPQ>(x, k) - x;
      if (!inner_row) dx = 0.0;
      double rp = dmax(dmax(fabs(x - z0), fabs(-x - z1)), dmax(fabs(-dx - z2), fabs(dx - z3)));
      double np = dmax(dmax(fabs(x), fabs(dx)), dmax(dmax(fabs(z0), fabs(z1)), dmax(fabs(z2), fabs(z3))));
      const double ey = y3_ - y2_;
      double Aty = (y0_ - y1_) + shift_up1<LPQ>(ey, k) - ey;
      const double st = srow ? t : 0.0;
      double ysum = 0.0;
#pragma unroll
      for (int d = 0; d < JT_D; d++) {
        if (d < deg) {
          rp = dmax(rp, fabs(st - zs_me[d]));
          np = dmax(np, dmax(fabs(st), fabs(zs_me[d])));
          ysum += rho * fma(-oma, zs_me[d], cs_me[d]);
        }
      }
      const double Px = fma(pW, suffix2<LPQ>(st, k), pI * x);
      Aty += suffix2<LPQ>(srow ? ysum : 0.0, k);
      double rd = fabs(Px + q + Aty);
      double nd = dmax(fabs(Px), dmax(fabs(Aty), fabs(q)));
      rp = group_max<LPQ>(rp); np = group_max<LPQ>(np); rd = group_max<LPQ>(rd); nd = group_max<LPQ>(nd);
      if (k == 0) { double* o = red + (rpar * G + g) * 4; o[0] = rp; o[1] = np; o[2] = rd; o[3] = nd; }
      cl.sync();
      {
        double m0_ = 0.0, m1_ = 0.0, m2_ = 0.0, m3_ = 0.0;
        for (int j = lane; j < CL * G; j += 32) {
          const double* o = cl.map_shared_rank(red, j / G) + (rpar * G + (j % G)) * 4;
          m0_ = fmax(m0_, o[0]); m1_ = fmax(m1_, o[1]); m2_ = fmax(m2_, o[2]); m3_ = fmax(m3_, o[3]);
        }
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
          m0_ = fmax(m0_, __shfl_xor_sync(PIADMM_FULL, m0_, o)); m1_ = fmax(m1_, __shfl_xor_sync(PIADMM_FULL, m1_, o));
          m2_ = fmax(m2_, __shfl_xor_sync(PIADMM_FULL, m2_, o)); m3_ = fmax(m3_, __shfl_xor_sync(PIADMM_FULL, m3_, o));
        }
        rp = m0_; np = m1_; rd = m2_; nd = m3_;
      }
      rpar ^= 1;
      const bool conv = rp <= p.in_eps_abs + p.in_eps_rel * np && rd <= p.in_eps_abs + p.in_eps_rel * nd;
      if (conv && !done) { done = true; x_out = x; iters = p.in_max_iter - left; status = 0; }
      if (done) break;                               // uniform over the cluster: the four maxima are global
      if (p.in_adaptive) {
        const double a = rp / dmax(np, 1e-30);
        const double b = dmax(rd / dmax(nd, 1e-30), 1e-30);
        const double rn = dmin(dmax(rho * sqrt(a / b), 1e-6), 1e6);
        if (rn > p.in_adapt_tol * rho || rn < rho / p.in_adapt_tol) {
          // OSQP keeps (z, y) and changes rho: re-express c and g in the new rho
          const double ro = rho;
          rho = rn; irho = 1.0 / rho; tb = ja.hinge * irho;
          ch0 = fma(y0_, irho, oma * z0) - l_u; ch1 = fma(y1_, irho, oma * z1) - l_u;
          ch2 = fma(y2_, irho, oma * z2) - l_d; ch3 = fma(y3_, irho, oma * z3) - l_d;
          g01 = fma(-y0_, irho, z0) - fma(-y1_, irho, z1);
          ge = fma(-y3_, irho, z3) - fma(-y2_, irho, z2);
          gsum = 0.0;
#pragma unroll
          for (int d = 0; d < JT_D; d++) {
            if (d < deg) {
              const double ym = ro * fma(-oma, zs_me[d], cs_me[d]);
              const double zp = zpt[(g * JT_D + d) * LPQ + k];
              const double yp = ro * fma(-oma, zp, cs_pt[d]);
              cs_me[d] = fma(ym, irho, oma * zs_me[d]);
              cs_pt[d] = fma(yp, irho, oma * zp);
              gsum += fma(-ym, irho, zs_me[d]);
            }
          }
          build_Kinv_edge<LPQ, false>(Kinv, N, k, pI + sg, rho, pW + deg * rho, Wsm, gbuf);
        }
      }
    }
  }
  if (!done) x_out = x;
  if (valid && k < N) ja.u_out[((size_t)s * V + i) * N + k] = round_scale(x_out, c.round_edge_scale);
  if (rank == 0 && threadIdx.x == 0) { ja.iters_out[s] = iters; ja.status_out[s] = status; }
  cl.sync();                                         // nobody leaves while a peer may still read its shared memory
}

// host side: one cluster of CL CTAs per scenario
template <int LPQ>
static cudaError_t launch_joint(const DevCtx& c, const JointArgs& ja, int CL, cudaStream_t st) {
  const size_t bytes = jt_smem_doubles<LPQ>() * sizeof(double);
  cudaError_t ce = cudaFuncSetAttribute(k_joint<LPQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (ce != cudaSuccess) return ce;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(c.S * CL)); cfg.blockDim = dim3(JT_THREADS); cfg.dynamicSmemBytes = bytes; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = (unsigned)CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, k_joint<LPQ>, c, ja, CL);
}
