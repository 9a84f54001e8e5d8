// piadmm_phases.cuh -- group-level statements of the outer-loop phases (x-update, edge step with
// dual/PI update).  State lives in the caller-owned piadmm_buffers; every function is executed by
// whole warps (all lanes reach every shuffle), `valid` masks groups that have no work.
#pragma once
#include "piadmm_dev.cuh"

#define NP1 (c.N + 1)
#define SLOT_OFF(s, i, d) ((((size_t)(s) * c.V + (i)) * c.D + (d)) * 2 * NP1)
#define POS_OFF(s, i) (((size_t)(s) * c.V + (i)) * 2 * NP1)

// non-linear roll-out (casadi/PI_ADMM_class.py:91-104): lane k gets point k.  u is 0 on lanes >= N.
template <int LPQ>
__device__ __forceinline__ void rollout_nonlinear(const piadmm_params& p, double x0, double y0, double th0,
                                                  double v, double u, int k, double* px, double* py,
                                                  double* pth) {
  const double th = th0 + (v / p.L * p.dt) * excl_prefix<LPQ>(u, k);
  double sn, cs;
  sincos(th, &sn, &cs);
  *px = x0 + (v * p.dt) * excl_prefix<LPQ>(cs, k);
  *py = y0 + (v * p.dt) * excl_prefix<LPQ>(sn, k);
  *pth = th;
}

// ---- x-update of agent (s,i): casadi/main.py:81-106 --------------------------------------------
template <int LPQ, bool KS>
__device__ __forceinline__ void dev_x_update(const DevCtx& c, int s, int i, bool valid, int k, int lane,
                                             int half, const double* Wsm, double* gbuf, double* xch, double* krow,
                                             double* px_out = nullptr, double* py_out = nullptr, double* u_out = nullptr) {
  const piadmm_params& p = c.p;
  const int N = c.N;
  double x0 = 0.0, y0 = 0.0, th0 = 0.0, v = 1.0, rho = 1.0;
  double tx = 0.0, ty = 0.0;
  int deg = 0;
  if (valid) {
    const double* xt = c.b.xt + ((size_t)s * c.V + i) * 3;
    x0 = xt[0]; y0 = xt[1]; th0 = xt[2];
    v = c.b.spd[(size_t)s * c.V + i];
    rho = c.b.rho_s[s];
    if (k <= N) {
      const double* rw = c.b.refwin + POS_OFF(s, i);
      tx = p.Pnorm * rw[k];
      ty = p.Pnorm * rw[NP1 + k];
    }
  }
  // AL term over the allocated neighbour slots, ascending d.  The slot table is read by the lanes in parallel (lane d =
  // slot d) and turned into a bit mask per group: D serial dependent loads (2.4 us at D = 8) become one, and a vehicle
  // without neighbours -- almost every vehicle of a large scene -- skips the loop altogether.
  if (c.D <= LPQ) {
    const int nbv = (valid && k < c.D) ? c.b.nbr[((size_t)s * c.V + i) * c.D + k] : -1;
    unsigned m = __ballot_sync(PIADMM_FULL, nbv >= 0);
    if (LPQ == 16) m = (m >> (lane & 16)) & 0xffffu;
    deg = __popc(m);
    while (m) {
      const int d = __ffs(m) - 1;
      m &= m - 1;
      if (k <= N) {
        const double* hat = c.b.hat + SLOT_OFF(s, i, d);
        const double* du = c.b.dual + SLOT_OFF(s, i, d);
        tx += 0.5 * rho * (hat[k] - du[k]);
        ty += 0.5 * rho * (hat[NP1 + k] - du[NP1 + k]);
      }
    }
  } else if (valid) {
    for (int d = 0; d < c.D; d++) {
      if (c.b.nbr[((size_t)s * c.V + i) * c.D + d] < 0) continue;
      deg++;
      if (k <= N) {
        const double* hat = c.b.hat + SLOT_OFF(s, i, d);
        const double* du = c.b.dual + SLOT_OFF(s, i, d);
        tx += 0.5 * rho * (hat[k] - du[k]);
        ty += 0.5 * rho * (hat[NP1 + k] - du[NP1 + k]);
      }
    }
  }
  double sn, cs;
  sincos(th0, &sn, &cs);
  const double g = v * v * p.dt * p.dt / p.L;
  const double w = p.Pnorm + 0.5 * rho * deg;
  const double e0 = -sn * x0 + cs * y0;
  const double vec = (valid && k <= N) ? (w * e0 + sn * tx - cs * ty) : 0.0;
  const double q = 2.0 * g * apply_Mt<LPQ>(vec, k);
  const double pI = 2.0 * p.Pcost, p2 = valid ? 2.0 : 0.0, pW = valid ? 2.0 * w * g * g : 0.0;
  double* kc = (c.kcache && valid) ? c.kcache + (((size_t)s * c.V + i) * 2) * PIADMM_KC_STRIDE(LPQ) : nullptr;
  InnerOut o = inner_solve<LPQ, KS>(p, c.ik, N, k, half, false, pI, p2, pW, q, 0.0, 0.0, 0.0, Wsm, gbuf, xch, krow, kc);
  const double u = (k < N) ? round_scale(o.x, c.round_u_scale) : 0.0;
  double px, py;
  if (p.pos_nonlinear) {
    double pth;
    rollout_nonlinear<LPQ>(p, x0, y0, th0, v, u, k, &px, &py, &pth);
  } else {
    double nxt;
    const double m = apply_M<LPQ>(u, k, &nxt);
    px = x0 + k * p.dt * v * cs - g * sn * m;
    py = y0 + k * p.dt * v * sn + g * cs * m;
  }
  if (px_out) { *px_out = px; *py_out = py; *u_out = u; }      // this lane's point k / control k of the new row (k <= N / k < N)
  if (valid) {
    if (k < N) c.b.u[((size_t)s * c.V + i) * N + k] = u;
    if (k <= N) {
      double* pos = c.b.pos + POS_OFF(s, i);
      pos[k] = px;
      pos[NP1 + k] = py;
    }
    if (k == 0) {
      atomicAdd(reinterpret_cast<unsigned long long*>(c.b.veh_inner + s), (unsigned long long)o.iters);
      atomicAdd(c.b.n_veh_solves + s, 1);
      if (o.status) atomicOr(c.b.status + s, PIADMM_ST_VEH_MAXITER);
    }
  }
}

// ---- edge step of edge e = (v0 < v1) of scenario s, half = which endpoint this group owns -------
// casadi/main.py:121-162 with the convex edge problem (MATLAB ...PI_antiwindup.m:276-304,383-396)
// and the three dual-update laws.  Writes hat, dual, integ, diff (own side), edge_res, rho_new.
template <int LPQ, bool KS, bool LATENCY = false>
__device__ __forceinline__ void dev_z_update(const DevCtx& c, int s, int e, int v0, int v1, int sl0, int sl1,
                                             bool last_edge, bool valid, int k, int lane, int half,
                                             const double* Wsm, double* gbuf, double* xch, double* krow,
                                             bool edge_cache) {
  const piadmm_params& p = c.p;
  const int N = c.N;
  int vv[2] = {v0, v1}, sl[2] = {sl0, sl1};
  double xs[2][3] = {{0, 0, 0}, {0, 0, 0}}, spd[2] = {1.0, 1.0};
  double rho = 1.0;
  if (valid) {
    for (int r = 0; r < 2; r++) {
      const double* xt = c.b.xt + ((size_t)s * c.V + vv[r]) * 3;
      xs[r][0] = xt[0]; xs[r][1] = xt[1]; xs[r][2] = xt[2];
      spd[r] = c.b.spd[(size_t)s * c.V + vv[r]];
    }
    rho = c.b.rho_s[s];
  }
  double sn[2], cs[2], seed[2][2];
  for (int r = 0; r < 2; r++) {
    sincos(xs[r][2], &sn[r], &cs[r]);
    seed[r][0] = rint((xs[r][0] + p.dt * spd[r] * cs[r]) * 1e4) / 1e4;   // casadi/main.py:48-49
    seed[r][1] = rint((xs[r][1] + p.dt * spd[r] * sn[r]) * 1e4) / 1e4;
  }
  const double lx = seed[1][0] - seed[0][0], ly = seed[1][1] - seed[0][1];
  const int me = half, ot = half ^ 1;
  const double vme = spd[me], g = vme * vme * p.dt * p.dt / p.L;
  const size_t so = SLOT_OFF(s, vv[me], sl[me]);
  double pos_x = 0.0, pos_y = 0.0, du_x = 0.0, du_y = 0.0, opos_x = 0.0, opos_y = 0.0;
  if (valid && k <= N) {
    const double* pos = c.b.pos + POS_OFF(s, vv[me]);
    const double* opos = c.b.pos + POS_OFF(s, vv[ot]);
    pos_x = pos[k]; pos_y = pos[NP1 + k];
    opos_x = opos[k]; opos_y = opos[NP1 + k];
    du_x = c.b.dual[so + k]; du_y = c.b.dual[so + NP1 + k];
  }
  // q = rho g M'(s ex - c ey), hinge offsets c_k = dth^2 - dis_k(u=0)
  double vec = 0.0, cfull = 0.0;
  if (valid && k <= N) {
    const double xb = xs[me][0] + k * p.dt * vme * cs[me], yb = xs[me][1] + k * p.dt * vme * sn[me];
    vec = sn[me] * (pos_x + du_x - xb) - cs[me] * (pos_y + du_y - yb);
    const double xb0 = xs[0][0] + k * p.dt * spd[0] * cs[0], yb0 = xs[0][1] + k * p.dt * spd[0] * sn[0];
    const double xb1 = xs[1][0] + k * p.dt * spd[1] * cs[1], yb1 = xs[1][1] + k * p.dt * spd[1] * sn[1];
    const double base0 = -2.0 * (lx * xb0 + ly * yb0), base1 = 2.0 * (lx * xb1 + ly * yb1);
    cfull = p.dis_thres * p.dis_thres - (base0 + base1 - (lx * lx + ly * ly));
  }
  const double q = rho * g * apply_Mt<LPQ>(vec, k);
  // position-row layout of inner_solve: lane L <= N-2 <-> trajectory index L+2, lane N-1 <-> index 1
  double cvec = __shfl_down_sync(PIADMM_FULL, cfull, 2, LPQ);
  const double c_one = __shfl_sync(PIADMM_FULL, cfull, 1, LPQ);
  if (k == N - 1) cvec = c_one;
  if (!valid || k >= N) cvec = -1e300;
  const double g0 = spd[0] * spd[0] * p.dt * p.dt / p.L, g1 = spd[1] * spd[1] * p.dt * p.dt / p.L;
  const double a0 = valid ? -2.0 * g0 * (-lx * sn[0] + ly * cs[0]) : 0.0;
  const double a1 = valid ? 2.0 * g1 * (-lx * sn[1] + ly * cs[1]) : 0.0;
  const double pI = 2.0 * p.Pcost, pW = valid ? rho * g * g : 0.0;
  // a vehicle can be an endpoint of several edges solved concurrently (generic V): they share the vehicle's edge
  // entry -- same key, same matrix, identical bits from every writer (see the fence in inner_solve)
  double* kc = (c.kcache && valid && edge_cache) ? c.kcache + (((size_t)s * c.V + vv[me]) * 2 + 1) * PIADMM_KC_STRIDE(LPQ) : nullptr;
  InnerOut o = inner_solve<LPQ, KS, LATENCY>(p, c.ik, N, k, half, true, pI, 0.0, pW, q, a0, a1, cvec, Wsm, gbuf, xch, krow, kc);
  // z-hat: non-linear roll-out of the rounded edge controls (casadi/main.py:153-158)
  const double uh = (k < N) ? round_scale(o.x, c.round_edge_scale) : 0.0;
  if (c.edge_u && valid && k < N) c.edge_u[(((size_t)s * c.E + e) * 2 + half) * N + k] = uh;
  if (c.z_solve_only) {            // casadi/main.py:133-153 alone; uniform over the pair (kernel parameter)
    if (valid && k == 0 && half == 0) {
      atomicAdd(reinterpret_cast<unsigned long long*>(c.b.edge_inner + s), (unsigned long long)o.iters);
      atomicAdd(c.b.n_edge_solves + s, 1);
      if (o.status) atomicOr(c.b.status + s, PIADMM_ST_EDGE_MAXITER);
    }
    return;
  }
  double hx, hy, hth;
  rollout_nonlinear<LPQ>(p, xs[me][0], xs[me][1], xs[me][2], vme, uh, k, &hx, &hy, &hth);
  // gains
  double kP = 0.0, kI = p.kI < 0.0 ? rho : p.kI, rho_dual = rho;
  if (p.dual_mode == PIADMM_DUAL_PI) {
    const double dx = pos_x - opos_x, dy = pos_y - opos_y;
    const double dmin = group_min<LPQ>(k <= N ? sqrt(dx * dx + dy * dy) : 1e300);
    kP = p.theta1 - p.theta2 / (1.0 + exp(-dmin));
  } else if (p.dual_mode == PIADMM_DUAL_PI_OLD) {
    // distance along the non-linear roll-out of primal_u (casadi_old_PI_ADMM/main.py:127-137)
    double um = 0.0, uo = 0.0;
    if (valid && k < N) {
      um = c.b.u[((size_t)s * c.V + vv[me]) * N + k];
      uo = c.b.u[((size_t)s * c.V + vv[ot]) * N + k];
    }
    double ax, ay, bx, by, t;
    rollout_nonlinear<LPQ>(p, xs[me][0], xs[me][1], xs[me][2], vme, um, k, &ax, &ay, &t);
    rollout_nonlinear<LPQ>(p, xs[ot][0], xs[ot][1], xs[ot][2], spd[ot], uo, k, &bx, &by, &t);
    const double dx = ax - bx, dy = ay - by;
    const double dmin = group_min<LPQ>(k <= N ? sqrt(dx * dx + dy * dy) : 1e300);
    kP = fmin(p.KP_num / dmin, p.KP_max);
    rho_dual = fmax(p.rho_min, fmin(p.rho_max, p.rho_num / dmin));
  }
  // dual / PI update, own side (both coordinates of index k)
  double ss = 0.0, rr = 0.0;
  if (valid && k <= N) {
    const bool alias = p.dual_res_alias && c.b.aliased[s];
    for (int cc = 0; cc < 2; cc++) {
      const size_t o2 = so + (size_t)cc * NP1 + k;
      const double hv = cc == 0 ? hx : hy, pv = cc == 0 ? pos_x : pos_y;
      const double err = pv - hv;
      double du = cc == 0 ? du_x : du_y, raw;
      if (p.dual_mode == PIADMM_DUAL_PLAIN || (p.dual_mode == PIADMM_DUAL_PI && !p.PI)) {
        raw = du + rho_dual * err;
      } else if (p.dual_mode == PIADMM_DUAL_PI) {
        const double in = c.b.integ[o2] + kI * err + p.backcalc * c.b.diff[o2];
        c.b.integ[o2] = in;
        raw = in + kP * err;
      } else if (p.trad) {
        raw = du + rho_dual * err + c.b.diff[o2];
      } else {
        const double in = c.b.integ[o2];
        raw = in + kP * err;
        c.b.integ[o2] = in + p.K_I * err + p.backcalc * c.b.diff[o2];
      }
      if (p.dual_mode != PIADMM_DUAL_PLAIN && p.windup) {
        const double cl = fmin(p.windup_sat, fmax(raw, -p.windup_sat));
        c.b.diff[o2] = cl - raw;
        raw = cl;
      }
      c.b.dual[o2] = raw;
      const double lh = alias ? hv : c.b.last_hat[o2];
      if (alias) c.b.last_hat[o2] = hv;
      c.b.hat[o2] = hv;
      ss += (lh - hv) * (lh - hv);
      rr += err * err;
    }
  }
  ss = group_sum<LPQ>(ss);
  rr = group_sum<LPQ>(rr);
  if (valid && k == 0) {
    double* er = c.edge_res + ((size_t)s * c.E + e) * 4;
    er[half] = ss;
    er[2 + half] = rr;
    if (half == 0) {
      atomicAdd(reinterpret_cast<unsigned long long*>(c.b.edge_inner + s), (unsigned long long)o.iters);
      atomicAdd(c.b.n_edge_solves + s, 1);
      if (o.status) atomicOr(c.b.status + s, PIADMM_ST_EDGE_MAXITER);
      if (p.dual_mode == PIADMM_DUAL_PI_OLD && last_edge) c.rho_new[s] = rho_dual;
    }
  }
}
