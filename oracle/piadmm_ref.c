/* CPU ORACLE, C twin (test infrastructure and timed CPU baseline -- never a product path).
 *
 * Plain-C fp64 restatement of the PI-ADMM hot path of
 * KaigeTan/Distributed-Local-Planner-PI-ADMM, the same algorithm as oracle/driver.py +
 * oracle/solver.py (which are the readable statements; this file exists because the numpy
 * version is too slow for the 4096-scenario parity cases and for a fair CPU baseline):
 *
 *   vehicle QP        casadi/PI_ADMM_class.py:114-135,172-192 ; casadi/main.py:81-106
 *   collision test    casadi/main.py:110-118
 *   edge problem      matlab_old_files/ADMM_CVX_two_veh_intesection_PI_antiwindup.m:276-304,383-396
 *                     (convex statement of casadi/PI_ADMM_class.py:145-169) ; casadi/main.py:121-158
 *   dual / PI update  casadi/main.py:161-162 ; MATLAB :152-188 ; casadi_old_PI_ADMM/main.py:127-151
 *   residuals / stop  casadi/main.py:165-180
 *   plant step        casadi/main.py:185-192
 *   inner solver      OSQP ADMM iteration (third party, not under /root/reference; see oracle/solver.py)
 *
 * PARITY UNPINNED: the reference has no tests/golden vectors and CasADi+OSQP cannot be installed
 * here, so this restatement was never compared with outputs of the reference itself.
 *
 * Build: see oracle/Makefile (gcc -O3 -fopenmp -shared).  Scenarios are independent and are
 * spread over OpenMP threads.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
static int omp_thread_id(void) { return omp_get_thread_num(); }
#else
static int omp_thread_id(void) { return 0; }
#endif
#include "../include/piadmm.h"

#define NMAX 32

typedef struct {
  int N;
  double W[NMAX * NMAX];   /* M'M */
  double D2[NMAX * NMAX];  /* D2'D2 */
  double T[NMAX * NMAX];   /* A'A = 2I + 2 D1'D1 for A = [I;-I;-D1;D1] */
} consts_t;

static void build_consts(consts_t* c, int N) {
  c->N = N;
  memset(c->W, 0, sizeof c->W); memset(c->D2, 0, sizeof c->D2); memset(c->T, 0, sizeof c->T);
  for (int i = 0; i < N; i++)
    for (int j = 0; j < N; j++) {
      double s = 0;
      for (int k = 0; k <= N; k++) {
        int mi = k - 1 - i, mj = k - 1 - j;
        if (mi > 0 && mj > 0) s += (double)mi * (double)mj;
      }
      c->W[i * NMAX + j] = s;
    }
  for (int k = 0; k + 2 < N; k++) {
    const double w[3] = {1.0, -2.0, 1.0};
    for (int a = 0; a < 3; a++)
      for (int b = 0; b < 3; b++) c->D2[(k + a) * NMAX + (k + b)] += w[a] * w[b];
  }
  for (int i = 0; i < N; i++) c->T[i * NMAX + i] = 2.0;
  for (int k = 0; k + 1 < N; k++) {
    c->T[k * NMAX + k] += 2.0; c->T[(k + 1) * NMAX + k + 1] += 2.0;
    c->T[k * NMAX + k + 1] -= 2.0; c->T[(k + 1) * NMAX + k] -= 2.0;
  }
}

/* out[k] = (M x)_k, k = 0..N ;  M[k][j] = max(k-1-j,0) */
static void apply_M(int N, const double* x, double* out) {
  double c1 = 0.0, acc = 0.0;
  out[0] = 0.0;
  for (int k = 1; k <= N; k++) {
    acc += c1;          /* acc = sum_{i<k} c1_i , c1_i = sum_{j<i} x_j */
    out[k] = acc;
    c1 += x[k - 1];
  }
}

/* out[j] = (M' v)_j, j = 0..N-1, v has N+1 entries */
static void apply_Mt(int N, const double* v, double* out) {
  double t = 0.0, acc = 0.0;   /* t_i = sum_{k>i} v_k */
  for (int j = N - 1; j >= 0; j--) {
    /* t_{j+1} = sum_{k>j+1} v_k */
    acc += t;                  /* acc = sum_{i=j+1..N} t_i */
    out[j] = acc;
    t += v[j + 1];
  }
}

/* K = cI*I + c2*D2'D2 + cW*W  ->  in-place inverse (Gauss-Jordan, SPD so no pivoting) */
static void build_Kinv(const consts_t* c, double cI, double c2, double cW, double cT, double* K) {
  int N = c->N;
  for (int i = 0; i < N; i++)
    for (int j = 0; j < N; j++)
      K[i * NMAX + j] = (i == j ? cI : 0.0) + c2 * c->D2[i * NMAX + j] + cW * c->W[i * NMAX + j] +
                        cT * c->T[i * NMAX + j];
  for (int p = 0; p < N; p++) {
    double piv = 1.0 / K[p * NMAX + p];
    for (int j = 0; j < N; j++) if (j != p) K[p * NMAX + j] *= piv;
    K[p * NMAX + p] = piv;
    for (int i = 0; i < N; i++) {
      if (i == p) continue;
      double f = K[i * NMAX + p];
      for (int j = 0; j < N; j++) if (j != p) K[i * NMAX + j] -= f * K[p * NMAX + j];
      K[i * NMAX + p] = -f * piv;
    }
  }
}

/* One "half problem": N unknowns, 4N-2 box rows (+ N position rows when with_s). */
typedef struct {
  double x[NMAX], q[NMAX];
  double z[4][NMAX], y[4][NMAX];   /* rows: I, -I, -D1 (k<N-1), D1 (k<N-1) */
  double zs[NMAX], ys[NMAX];       /* position rows s_k = (M x)_k, k=1..N stored at k-1 */
  double pI, p2, pW;               /* P = pI*I + p2*D2'D2 + pW*W */
  double Kinv[NMAX * NMAX];
  /* scratch of the current iteration */
  double zr[4][NMAX], v[4][NMAX], zrs[NMAX], vs[NMAX];
} half_t;

static void half_reset(half_t* h, int N) {
  memset(h->x, 0, sizeof h->x);
  memset(h->z, 0, sizeof h->z); memset(h->y, 0, sizeof h->y);
  memset(h->zs, 0, sizeof h->zs); memset(h->ys, 0, sizeof h->ys);
  (void)N;
}

/* A' w for w = (w1..w4 [, ws]) */
static void apply_At(int N, double w[4][NMAX], const double* ws, int with_s, double* out) {
  for (int k = 0; k < N; k++) {
    double d = w[0][k] - w[1][k];
    double a = (k < N - 1) ? (w[3][k] - w[2][k]) : 0.0;       /* D1' e : e_{k-1} - e_k */
    double am = (k > 0) ? (w[3][k - 1] - w[2][k - 1]) : 0.0;
    out[k] = d + am - a;
  }
  if (with_s) {
    double v[NMAX + 1], t[NMAX];
    v[0] = 0.0;
    for (int k = 1; k <= N; k++) v[k] = ws[k - 1];
    apply_Mt(N, v, t);
    for (int k = 0; k < N; k++) out[k] += t[k];
  }
}

static void apply_P(int N, const half_t* h, const double* x, double* out) {
  double m[NMAX + 1], t[NMAX];
  apply_M(N, x, m);
  apply_Mt(N, m, t);
  for (int k = 0; k < N; k++) out[k] = h->pI * x[k] + h->pW * t[k];
  if (h->p2 != 0.0) {
    for (int k = 0; k + 2 < N; k++) {
      double e = x[k] - 2.0 * x[k + 1] + x[k + 2];
      out[k] += h->p2 * e; out[k + 1] -= 2.0 * h->p2 * e; out[k + 2] += h->p2 * e;
    }
  }
}

/* first part of an iteration: linear solve, relaxation, v = zr + y/rho */
static void half_iter_a(const piadmm_params* p, int N, half_t* h, double rho, int with_s) {
  double w[4][NMAX], ws[NMAX], rhs[NMAX], xt[NMAX], ms[NMAX + 1];
  const double al = p->in_alpha, sg = p->in_sigma;
  for (int r = 0; r < 4; r++)
    for (int k = 0; k < N; k++) w[r][k] = rho * h->z[r][k] - h->y[r][k];
  if (with_s) for (int k = 0; k < N; k++) ws[k] = rho * h->zs[k] - h->ys[k];
  apply_At(N, w, ws, with_s, rhs);
  for (int k = 0; k < N; k++) rhs[k] += sg * h->x[k] - h->q[k];
  for (int i = 0; i < N; i++) {
    double s = 0.0;
    for (int j = 0; j < N; j++) s += h->Kinv[i * NMAX + j] * rhs[j];
    xt[i] = s;
  }
  for (int k = 0; k < N; k++) {
    double zt0 = xt[k], zt1 = -xt[k];
    h->zr[0][k] = al * zt0 + (1.0 - al) * h->z[0][k];
    h->zr[1][k] = al * zt1 + (1.0 - al) * h->z[1][k];
    if (k < N - 1) {
      double d = xt[k + 1] - xt[k];
      h->zr[2][k] = al * (-d) + (1.0 - al) * h->z[2][k];
      h->zr[3][k] = al * d + (1.0 - al) * h->z[3][k];
    } else { h->zr[2][k] = 0.0; h->zr[3][k] = 0.0; }
    for (int r = 0; r < 4; r++) h->v[r][k] = h->zr[r][k] + h->y[r][k] / rho;
  }
  if (with_s) {
    apply_M(N, xt, ms);
    for (int k = 0; k < N; k++) {
      h->zrs[k] = al * ms[k + 1] + (1.0 - al) * h->zs[k];
      h->vs[k] = h->zrs[k] + h->ys[k] / rho;
    }
  }
  for (int k = 0; k < N; k++) h->x[k] = al * xt[k] + (1.0 - al) * h->x[k];
}

/* second part: projections of the box rows and dual update; position rows get zs_new from caller */
static void half_iter_b(int N, half_t* h, double rho, int with_s, const double* zs_new) {
  const double lu = -M_PI / 6.0, ld = -M_PI / 9.0;
  for (int k = 0; k < N; k++)
    for (int r = 0; r < 4; r++) {
      if (r >= 2 && k >= N - 1) continue;
      double l = (r < 2) ? lu : ld;
      double zn = h->v[r][k] > l ? h->v[r][k] : l;
      h->y[r][k] += rho * (h->zr[r][k] - zn);
      h->z[r][k] = zn;
    }
  if (with_s)
    for (int k = 0; k < N; k++) {
      h->ys[k] += rho * (h->zrs[k] - zs_new[k]);
      h->zs[k] = zs_new[k];
    }
}

/* residual pieces of one half (inf-norms) */
static void half_norms(int N, const half_t* h, int with_s, double* rp, double* rd, double* n_p, double* n_d) {
  double Ax[4][NMAX], ms[NMAX + 1], Px[NMAX], Aty[NMAX];
  double yy[4][NMAX];
  for (int k = 0; k < N; k++) {
    Ax[0][k] = h->x[k]; Ax[1][k] = -h->x[k];
    if (k < N - 1) { double d = h->x[k + 1] - h->x[k]; Ax[2][k] = -d; Ax[3][k] = d; }
    else { Ax[2][k] = 0.0; Ax[3][k] = 0.0; }
  }
  for (int r = 0; r < 4; r++)
    for (int k = 0; k < N; k++) {
      if (r >= 2 && k >= N - 1) { yy[r][k] = 0.0; continue; }
      double a = fabs(Ax[r][k] - h->z[r][k]); if (a > *rp) *rp = a;
      a = fabs(Ax[r][k]); if (a > *n_p) *n_p = a;
      a = fabs(h->z[r][k]); if (a > *n_p) *n_p = a;
      yy[r][k] = h->y[r][k];
    }
  if (with_s) {
    apply_M(N, h->x, ms);
    for (int k = 0; k < N; k++) {
      double a = fabs(ms[k + 1] - h->zs[k]); if (a > *rp) *rp = a;
      a = fabs(ms[k + 1]); if (a > *n_p) *n_p = a;
      a = fabs(h->zs[k]); if (a > *n_p) *n_p = a;
    }
  }
  apply_P(N, h, h->x, Px);
  apply_At(N, yy, h->ys, with_s, Aty);
  for (int k = 0; k < N; k++) {
    double a = fabs(Px[k] + h->q[k] + Aty[k]); if (a > *rd) *rd = a;
    a = fabs(Px[k]); if (a > *n_d) *n_d = a;
    a = fabs(Aty[k]); if (a > *n_d) *n_d = a;
    a = fabs(h->q[k]); if (a > *n_d) *n_d = a;
  }
}

static double new_rho(double rho, double rp, double rd, double n_p, double n_d) {
  double a = rp / (n_p > 1e-30 ? n_p : 1e-30);
  double b = rd / (n_d > 1e-30 ? n_d : 1e-30);
  if (b < 1e-30) b = 1e-30;
  double rn = rho * sqrt(a / b);
  if (rn < 1e-6) rn = 1e-6;
  if (rn > 1e6) rn = 1e6;
  return rn;
}

static double round_dec(double x, int dec) {
  if (dec < 0) return x;
  double s = 1.0;
  for (int i = 0; i < dec; i++) s *= 10.0;
  return rint(x * s) / s;      /* np.around: rint(x*10^d)/10^d, ties to even */
}

/* vehicle QP: returns inner iterations, *status = 0 converged / 1 max_iter */
static int solve_vehicle(const piadmm_params* p, const consts_t* c, half_t* h, int* status) {
  const int N = c->N;
  double rho = p->in_rho0;
  half_reset(h, N);
  build_Kinv(c, h->pI + p->in_sigma, h->p2, h->pW, rho, h->Kinv);
  *status = 1;
  int it;
  for (it = 1; it <= p->in_max_iter; it++) {
    half_iter_a(p, N, h, rho, 0);
    half_iter_b(N, h, rho, 0, NULL);
    if (it % p->in_check == 0) {
      double rp = 0, rd = 0, n_p = 0, n_d = 0;
      half_norms(N, h, 0, &rp, &rd, &n_p, &n_d);
      if (rp <= p->in_eps_abs + p->in_eps_rel * n_p && rd <= p->in_eps_abs + p->in_eps_rel * n_d) {
        *status = 0;
        return it;
      }
      if (p->in_adaptive) {
        double rn = new_rho(rho, rp, rd, n_p, n_d);
        if (rn > p->in_adapt_tol * rho || rn < rho / p->in_adapt_tol) {
          rho = rn;
          build_Kinv(c, h->pI + p->in_sigma, h->p2, h->pW, rho, h->Kinv);
        }
      }
    }
  }
  return p->in_max_iter;
}

/* edge problem on two halves coupled by the hinge prox on (s1_k, s2_k) */
static int solve_edge(const piadmm_params* p, const consts_t* c, half_t* h0, half_t* h1, double a0,
                      double a1, const double* cvec, int* status) {
  const int N = c->N;
  double rho = p->in_rho0;
  half_t* hh[2] = {h0, h1};
  const double na = a0 * a0 + a1 * a1;
  for (int r = 0; r < 2; r++) {
    half_reset(hh[r], N);
    /* K = P + sigma I + rho (T + W) */
    build_Kinv(c, hh[r]->pI + p->in_sigma, hh[r]->p2, hh[r]->pW + rho, rho, hh[r]->Kinv);
  }
  *status = 1;
  for (int it = 1; it <= p->in_max_iter; it++) {
    double zs0[NMAX], zs1[NMAX];
    half_iter_a(p, N, h0, rho, 1);
    half_iter_a(p, N, h1, rho, 1);
    const double t = p->beta / rho;
    for (int k = 0; k < N; k++) {
      double s1 = h0->vs[k], s2 = h1->vs[k];
      double tau = a0 * s1 + a1 * s2, step;
      if (!(na > 0.0) || tau >= cvec[k]) step = 0.0;
      else if (tau + t * na <= cvec[k]) step = t;
      else step = (cvec[k] - tau) / na;
      zs0[k] = s1 + step * a0; zs1[k] = s2 + step * a1;
    }
    half_iter_b(N, h0, rho, 1, zs0);
    half_iter_b(N, h1, rho, 1, zs1);
    if (it % p->in_check == 0) {
      double rp = 0, rd = 0, n_p = 0, n_d = 0;
      half_norms(N, h0, 1, &rp, &rd, &n_p, &n_d);
      half_norms(N, h1, 1, &rp, &rd, &n_p, &n_d);
      if (rp <= p->in_eps_abs + p->in_eps_rel * n_p && rd <= p->in_eps_abs + p->in_eps_rel * n_d) {
        *status = 0;
        return it;
      }
      if (p->in_adaptive) {
        double rn = new_rho(rho, rp, rd, n_p, n_d);
        if (rn > p->in_adapt_tol * rho || rn < rho / p->in_adapt_tol) {
          rho = rn;
          for (int r = 0; r < 2; r++)
            build_Kinv(c, hh[r]->pI + p->in_sigma, hh[r]->p2, hh[r]->pW + rho, rho, hh[r]->Kinv);
        }
      }
    }
  }
  return p->in_max_iter;
}

/* non-linear roll-out (casadi/PI_ADMM_class.py:91-104; the three-term x_dot there equals v*cos(theta_k)) */
static void rollout_nonlinear(const piadmm_params* p, int N, const double* xt, double v, const double* u,
                              double* x, double* y, double* th) {
  x[0] = xt[0]; y[0] = xt[1];
  double t = xt[2];
  if (th) th[0] = t;
  for (int k = 0; k < N; k++) {
    x[k + 1] = x[k] + v * cos(t) * p->dt;
    y[k + 1] = y[k] + v * sin(t) * p->dt;
    t = t + v / p->L * u[k] * p->dt;
    if (th) th[k + 1] = t;
  }
}

/* linearised roll-out about xt[2] (casadi/PI_ADMM_class.py:56-69) */
static void rollout_linear(const piadmm_params* p, int N, const double* xt, double v, const double* u,
                           double* x, double* y) {
  double m[NMAX + 1];
  const double s = sin(xt[2]), c = cos(xt[2]), g = v * v * p->dt * p->dt / p->L;
  apply_M(N, u, m);
  for (int k = 0; k <= N; k++) {
    x[k] = xt[0] + k * p->dt * v * c - g * s * m[k];
    y[k] = xt[1] + k * p->dt * v * s + g * c * m[k];
  }
}


int piadmm_ref_max_edges(int V, int D) {
  int e = V * D / 2;
  return e < 1 ? 1 : e;
}

/* casadi/PI_ADMM_class.py:15-28 + the semantics of casadi/main.py as written */
void piadmm_ref_default_params(piadmm_params* p) {
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

/* ---- indexing helpers over the piadmm_buffers layout (host pointers here) ---- */
#define NP1 (N + 1)
#define SLOT_OFF(s, i, d) ((((size_t)(s) * V + (i)) * D + (d)) * 2 * NP1)
#define POS_OFF(s, i) (((size_t)(s) * V + (i)) * 2 * NP1)

static int find_or_alloc_slot(int32_t* nbr_i, int D, int j) {
  for (int d = 0; d < D; d++) if (nbr_i[d] == j) return d;
  for (int d = 0; d < D; d++) if (nbr_i[d] < 0) { nbr_i[d] = j; return d; }
  return -1;
}

static void begin_scenario(const piadmm_params* p, const piadmm_buffers* b, int s, int V, int D, int E) {
  const int N = p->num_ho, I = p->iter_num;
  memset(b->u + (size_t)s * V * N, 0, sizeof(double) * V * N);
  memset(b->pos + POS_OFF(s, 0), 0, sizeof(double) * V * 2 * NP1);
  size_t so = SLOT_OFF(s, 0, 0), sn = (size_t)V * D * 2 * NP1;
  memset(b->hat + so, 0, sizeof(double) * sn); memset(b->dual + so, 0, sizeof(double) * sn);
  memset(b->integ + so, 0, sizeof(double) * sn); memset(b->diff + so, 0, sizeof(double) * sn);
  memset(b->last_hat + so, 0, sizeof(double) * sn);
  for (int i = 0; i < V; i++)
    for (int d = 0; d < D; d++) {
      int j = -1;
      if (p->al_all && d < V - 1) j = d < i ? d : d + 1;      /* slots = all j != i, ascending */
      b->nbr[((size_t)s * V + i) * D + d] = j;
    }
  b->flag[s] = 0; b->done[s] = 0; b->aliased[s] = 0; b->curr_iter[s] = 1; b->n_outer[s] = 0; b->status[s] = 0;
  b->veh_inner[s] = 0; b->edge_inner[s] = 0; b->n_veh_solves[s] = 0; b->n_edge_solves[s] = 0;
  b->edge_n[s] = 0;
  for (int k = 0; k < I; k++) { b->err_rk[(size_t)s * I + k] = 0.0; b->err_sk[(size_t)s * I + k] = 0.0; }
  (void)E;
}

static void x_update_vehicle(const piadmm_params* p, const consts_t* c, const piadmm_buffers* b, int s, int i, int V,
                             int D, half_t* h, int* it_out, int* st_out) {
  const int N = p->num_ho;
  const double rho = b->rho_s[s];
  const double* xt = b->xt + ((size_t)s * V + i) * 3;
  const double v = b->spd[(size_t)s * V + i];
  const double* rw = b->refwin + POS_OFF(s, i);
  const double sn = sin(xt[2]), cs = cos(xt[2]), g = v * v * p->dt * p->dt / p->L;
  double tx[NMAX + 1], ty[NMAX + 1], vec[NMAX + 1];
  int deg = 0;
  for (int k = 0; k <= N; k++) { tx[k] = p->Pnorm * rw[k]; ty[k] = p->Pnorm * rw[NP1 + k]; }
  for (int d = 0; d < D; d++) {
    if (b->nbr[((size_t)s * V + i) * D + d] < 0) continue;
    deg++;
    const double* hat = b->hat + SLOT_OFF(s, i, d);
    const double* du = b->dual + SLOT_OFF(s, i, d);
    for (int k = 0; k <= N; k++) {
      tx[k] += 0.5 * rho * (hat[k] - du[k]);
      ty[k] += 0.5 * rho * (hat[NP1 + k] - du[NP1 + k]);
    }
  }
  const double w = p->Pnorm + 0.5 * rho * deg;
  const double e0 = -sn * xt[0] + cs * xt[1];
  for (int k = 0; k <= N; k++) vec[k] = w * e0 + sn * tx[k] - cs * ty[k];
  apply_Mt(N, vec, h->q);
  for (int k = 0; k < N; k++) h->q[k] *= 2.0 * g;
  h->pI = 2.0 * p->Pcost; h->p2 = 2.0; h->pW = 2.0 * w * g * g;
  *it_out = solve_vehicle(p, c, h, st_out);
  double* u = b->u + ((size_t)s * V + i) * N;
  for (int k = 0; k < N; k++) u[k] = round_dec(h->x[k], p->round_u);
  double* pos = b->pos + POS_OFF(s, i);
  if (p->pos_nonlinear) rollout_nonlinear(p, N, xt, v, u, pos, pos + NP1, NULL);
  else rollout_linear(p, N, xt, v, u, pos, pos + NP1);
}

/* `pool` = one half_t pair per OpenMP thread; inner != 0: the vehicles of this ONE scenario are spread over the
 * threads (coupled scenes, S < threads), else the caller already runs scenarios in parallel */
/* x_memo: dirty[i] != 0 <=> the QP of vehicle i changed since its last solve in this MPC step (set for everybody at the
 * start of the step and by every edge problem for its two endpoints); clean vehicles keep u / pos_old as they are */
static int memo_on(const piadmm_params* p) { return p->x_memo && p->dual_mode != PIADMM_DUAL_PI_OLD; }

static void x_update_scenario(const piadmm_params* p, const consts_t* c, const piadmm_buffers* b, int s, int V,
                              int D, half_t** pool, int inner, uint8_t* dirty) {
  long tot_it = 0; int any_st = 0, max_it = 0, n_solved = 0;
  const int memo = memo_on(p);
  if (inner) {
#pragma omp parallel for schedule(dynamic, 8) reduction(+ : tot_it, n_solved) reduction(| : any_st) reduction(max : max_it)
    for (int i = 0; i < V; i++) {
      int it, st;
      if (memo && !dirty[i]) continue;
      dirty[i] = 0;
      x_update_vehicle(p, c, b, s, i, V, D, pool[2 * omp_thread_id()], &it, &st);
      tot_it += it; any_st |= st; n_solved++;
      if (it > max_it) max_it = it;
    }
  } else {
    for (int i = 0; i < V; i++) {
      int it, st;
      if (memo && !dirty[i]) continue;
      dirty[i] = 0;
      x_update_vehicle(p, c, b, s, i, V, D, pool[2 * omp_thread_id()], &it, &st);
      tot_it += it; any_st |= st; n_solved++;
      if (it > max_it) max_it = it;
    }
  }
  if (getenv("PIADMM_REF_DEBUG"))
    fprintf(stderr, "[ref] scenario %d outer %d: %d vehicle QPs, inner iterations sum %ld max %d\n", s, (int)b->n_outer[s], n_solved, tot_it, max_it);
  b->veh_inner[s] += tot_it; b->n_veh_solves[s] += n_solved;
  if (any_st) b->status[s] |= PIADMM_ST_VEH_MAXITER;
}

/* casadi/main.py:110-118.  The pair test fills a V x V bitmap (in parallel over rows when `inner`), the slot allocation
 * then walks it in row-major order = the order of np.where(edge_mat == 1). */
static void detect_scenario(const piadmm_params* p, const piadmm_buffers* b, int s, int V, int D, int E,
                            uint8_t* hitmap, int inner) {
  const int N = p->num_ho;
  const double thr = p->thr_squared ? p->dis_thres * p->dis_thres : p->dis_thres;
  const size_t rowb = ((size_t)V + 7) / 8;
#pragma omp parallel for schedule(dynamic, 16) if (inner)
  for (int i = 0; i < V; i++) {
    uint8_t* row = hitmap + (size_t)i * rowb;
    memset(row, 0, rowb);
    const double* pi = b->pos + POS_OFF(s, i);
    for (int j = i + 1; j < V; j++) {
      const double* pj = b->pos + POS_OFF(s, j);
      int hit = p->always_edge;
      for (int k = 0; k <= N && !hit; k++) {
        double dx = pi[k] - pj[k], dy = pi[NP1 + k] - pj[NP1 + k];
        if (dx * dx + dy * dy < thr) hit = 1;
      }
      if (hit) row[j >> 3] |= (uint8_t)(1u << (j & 7));
    }
  }
  int n = 0;
  for (int i = 0; i < V; i++) {
    const uint8_t* row = hitmap + (size_t)i * rowb;
    for (size_t w = (size_t)(i + 1) >> 3; w < rowb; w++) {
      uint8_t m = row[w];
      while (m) {
        const int bit = __builtin_ctz(m);
        m &= (uint8_t)(m - 1);
        const int j = (int)(w * 8) + bit;
        int32_t* ni = b->nbr + ((size_t)s * V + i) * D;
        int32_t* nj = b->nbr + ((size_t)s * V + j) * D;
        int di = find_or_alloc_slot(ni, D, j), dj = find_or_alloc_slot(nj, D, i);
        if (di < 0 || dj < 0 || n >= E) { b->status[s] |= PIADMM_ST_SLOT_OVERFLOW; continue; }
        b->edge_ij[((size_t)s * E + n) * 2 + 0] = i; b->edge_ij[((size_t)s * E + n) * 2 + 1] = j;
        b->edge_slot[((size_t)s * E + n) * 2 + 0] = di; b->edge_slot[((size_t)s * E + n) * 2 + 1] = dj;
        n++;
      }
    }
  }
  b->edge_n[s] = n;
}

/* dual / PI update of one directed side; e = pos - hat */
static void dual_side(const piadmm_params* p, int N, double rho, double kP, double kI, const double* pos,
                      const double* hat, double* dual, double* integ, double* diff) {
  for (int k = 0; k < 2 * NP1; k++) {
    const double e = pos[k] - hat[k];
    double raw;
    if (p->dual_mode == PIADMM_DUAL_PLAIN || (p->dual_mode == PIADMM_DUAL_PI && !p->PI)) {
      raw = dual[k] + rho * e;
    } else if (p->dual_mode == PIADMM_DUAL_PI) {
      integ[k] = integ[k] + kI * e + p->backcalc * diff[k];
      raw = integ[k] + kP * e;
    } else {
      if (p->trad) raw = dual[k] + rho * e + diff[k];
      else {
        raw = integ[k] + kP * e;
        integ[k] = integ[k] + p->K_I * e + p->backcalc * diff[k];
      }
    }
    if (p->dual_mode != PIADMM_DUAL_PLAIN && p->windup) {
      double cl = raw > p->windup_sat ? p->windup_sat : (raw < -p->windup_sat ? -p->windup_sat : raw);
      diff[k] = cl - raw;
      raw = cl;
    }
    dual[k] = raw;
  }
}

/* edge problem e of scenario s + its dual / PI update (edges of one outer iteration are independent of each other:
 * every directed slot belongs to exactly one edge, all edges see the same penalty) */
static void z_update_edge(const piadmm_params* p, const consts_t* c, const piadmm_buffers* b, int s, int e, int V,
                          int D, int E, half_t* h0, half_t* h1, int* it_out, int* st_out, double* rho_next) {
  const int N = p->num_ho;
  half_t* hh[2] = {h0, h1};
  const double rho = b->rho_s[s];   /* all edges of one outer iteration see the same penalty */
  int vv[2] = {b->edge_ij[((size_t)s * E + e) * 2], b->edge_ij[((size_t)s * E + e) * 2 + 1]};
  int sl[2] = {b->edge_slot[((size_t)s * E + e) * 2], b->edge_slot[((size_t)s * E + e) * 2 + 1]};
  double seed[2][2], a[2], base[2][NMAX + 1], cvec[NMAX];
  const double* xt[2]; double spd[2];
  for (int r = 0; r < 2; r++) {
    xt[r] = b->xt + ((size_t)s * V + vv[r]) * 3;
    spd[r] = b->spd[(size_t)s * V + vv[r]];
    seed[r][0] = round_dec(xt[r][0] + p->dt * spd[r] * cos(xt[r][2]), 4);   /* casadi/main.py:48-49 */
    seed[r][1] = round_dec(xt[r][1] + p->dt * spd[r] * sin(xt[r][2]), 4);
  }
  const double lx = seed[1][0] - seed[0][0], ly = seed[1][1] - seed[0][1];
  for (int r = 0; r < 2; r++) {
    const double sn = sin(xt[r][2]), cs = cos(xt[r][2]), v = spd[r], g = v * v * p->dt * p->dt / p->L;
    const double* pos = b->pos + POS_OFF(s, vv[r]);
    const double* du = b->dual + SLOT_OFF(s, vv[r], sl[r]);
    double vec[NMAX + 1];
    const double sign = r == 0 ? -1.0 : 1.0;
    for (int k = 0; k <= N; k++) {
      const double xb = xt[r][0] + k * p->dt * v * cs, yb = xt[r][1] + k * p->dt * v * sn;
      const double ex = pos[k] + du[k] - xb, ey = pos[NP1 + k] + du[NP1 + k] - yb;
      vec[k] = sn * ex - cs * ey;
      base[r][k] = sign * 2.0 * (lx * xb + ly * yb);
    }
    apply_Mt(N, vec, hh[r]->q);
    for (int k = 0; k < N; k++) hh[r]->q[k] *= rho * g;
    hh[r]->pI = 2.0 * p->Pcost; hh[r]->p2 = 0.0; hh[r]->pW = rho * g * g;
    a[r] = sign * 2.0 * g * (-lx * sn + ly * cs);
  }
  for (int k = 1; k <= N; k++)
    cvec[k - 1] = p->dis_thres * p->dis_thres - (base[0][k] + base[1][k] - (lx * lx + ly * ly));
  int st;
  *it_out = solve_edge(p, c, h0, h1, a[0], a[1], cvec, &st);
  *st_out = st;
  /* z-hat: non-linear roll-out of the rounded edge controls (casadi/main.py:153-158) */
  double uh[2][NMAX];
  for (int r = 0; r < 2; r++) {
    for (int k = 0; k < N; k++) uh[r][k] = round_dec(hh[r]->x[k], p->round_edge);
    double* hat = b->hat + SLOT_OFF(s, vv[r], sl[r]);
    rollout_nonlinear(p, N, xt[r], spd[r], uh[r], hat, hat + NP1, NULL);
    if (p->dual_res_alias && b->aliased[s])
      memcpy(b->last_hat + SLOT_OFF(s, vv[r], sl[r]), hat, sizeof(double) * 2 * NP1);
  }
  /* gains */
  double kP = 0.0, kI = p->kI < 0.0 ? rho : p->kI, rho_dual = rho;
  if (p->dual_mode == PIADMM_DUAL_PI) {
    const double* p0 = b->pos + POS_OFF(s, vv[0]);
    const double* p1 = b->pos + POS_OFF(s, vv[1]);
    double dmin = INFINITY;
    for (int k = 0; k <= N; k++) {
      double dx = p0[k] - p1[k], dy = p0[NP1 + k] - p1[NP1 + k];
      double d = sqrt(dx * dx + dy * dy);
      if (d < dmin) dmin = d;
    }
    kP = p->theta1 - p->theta2 / (1.0 + exp(-dmin));
  } else if (p->dual_mode == PIADMM_DUAL_PI_OLD) {
    double px[2][NMAX + 1], py[2][NMAX + 1];
    for (int r = 0; r < 2; r++)
      rollout_nonlinear(p, N, xt[r], spd[r], b->u + ((size_t)s * V + vv[r]) * N, px[r], py[r], NULL);
    double dmin = INFINITY;
    for (int k = 0; k <= N; k++) {
      double dx = px[0][k] - px[1][k], dy = py[0][k] - py[1][k];
      double d = sqrt(dx * dx + dy * dy);
      if (d < dmin) dmin = d;
    }
    kP = p->KP_num / dmin; if (kP > p->KP_max) kP = p->KP_max;
    double rn = p->rho_num / dmin; if (rn > p->rho_max) rn = p->rho_max; if (rn < p->rho_min) rn = p->rho_min;
    rho_dual = rn;      /* old main.py:137: rho is updated before the dual update */
    *rho_next = rn;
  }
  for (int r = 0; r < 2; r++) {
    size_t o = SLOT_OFF(s, vv[r], sl[r]);
    dual_side(p, N, rho_dual, kP, kI, b->pos + POS_OFF(s, vv[r]), b->hat + o, b->dual + o, b->integ + o,
              b->diff + o);
  }
}

static void z_update_scenario(const piadmm_params* p, const consts_t* c, const piadmm_buffers* b, int s, int V,
                              int D, int E, half_t** pool, int inner, uint8_t* dirty) {
  const int n = b->edge_n[s];
  for (int e = 0; e < n; e++) {
    dirty[b->edge_ij[((size_t)s * E + e) * 2]] = 1; dirty[b->edge_ij[((size_t)s * E + e) * 2 + 1]] = 1;
  }
  double rho_next = b->rho_s[s];
  long tot_it = 0; int any_st = 0, max_it = 0;
  const int outer_tid = omp_thread_id();      /* !inner: the region below is serialised and would report thread 0 */
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : tot_it) reduction(| : any_st) reduction(max : max_it) if (inner)
  for (int e = 0; e < n; e++) {
    int it, st;
    double rn = 0.0;
    half_t** hp = pool + 2 * (inner ? omp_thread_id() : outer_tid);
    z_update_edge(p, c, b, s, e, V, D, E, hp[0], hp[1], &it, &st, &rn);
    tot_it += it; any_st |= st;
    if (it > max_it) max_it = it;
    if (e == n - 1 && p->dual_mode == PIADMM_DUAL_PI_OLD) rho_next = rn;   /* old main.py:137: the last edge's rho stays */
  }
  if (getenv("PIADMM_REF_DEBUG"))
    fprintf(stderr, "[ref] scenario %d outer %d: %d edges, inner iterations sum %ld max %d\n", s, (int)b->n_outer[s], n, tot_it, max_it);
  b->edge_inner[s] += tot_it; b->n_edge_solves[s] += n;
  if (any_st) b->status[s] |= PIADMM_ST_EDGE_MAXITER;
  b->rho_s[s] = rho_next;
}

static void residual_scenario(const piadmm_params* p, const piadmm_buffers* b, int s, int V, int D, int E,
                              int i_iter) {
  const int N = p->num_ho, I = p->iter_num;
  const int n = b->edge_n[s];
  const double rho = b->rho_s[s];
  double sk = 0.0, rk = 0.0;
  for (int e = 0; e < n; e++) {
    int vv[2] = {b->edge_ij[((size_t)s * E + e) * 2], b->edge_ij[((size_t)s * E + e) * 2 + 1]};
    int sl[2] = {b->edge_slot[((size_t)s * E + e) * 2], b->edge_slot[((size_t)s * E + e) * 2 + 1]};
    double ss[2] = {0, 0}, rr[2] = {0, 0};
    for (int r = 0; r < 2; r++) {
      const double* hat = b->hat + SLOT_OFF(s, vv[r], sl[r]);
      const double* last = b->last_hat + SLOT_OFF(s, vv[r], sl[r]);
      const double* pos = b->pos + POS_OFF(s, vv[r]);
      for (int k = 0; k < 2 * NP1; k++) {
        double a = rho * (last[k] - hat[k]), c = pos[k] - hat[k];
        ss[r] += a * a; rr[r] += c * c;
      }
    }
    if (p->residual_global) { sk += sqrt(ss[0] + ss[1]); rk += sqrt(rr[0] + rr[1]); }
    else { sk += 2.0 * sqrt(ss[0]); rk += 2.0 * sqrt(rr[0]); }
  }
  b->err_rk[(size_t)s * I + i_iter] = rk; b->err_sk[(size_t)s * I + i_iter] = sk;
  int dist_ok = 1;
  if (p->stop_dist_test) {
    for (int i = 0; i < V; i++)
      for (int d = 0; d < D; d++) {
        int j = b->nbr[((size_t)s * V + i) * D + d];
        if (j <= i) continue;
        const double* xi = b->xt + ((size_t)s * V + i) * 3;
        const double* xj = b->xt + ((size_t)s * V + j) * 3;
        const double vi = b->spd[(size_t)s * V + i], vj = b->spd[(size_t)s * V + j];
        double dx = (xi[0] + p->dt * vi * cos(xi[2])) - (xj[0] + p->dt * vj * cos(xj[2]));
        double dy = (xi[1] + p->dt * vi * sin(xi[2])) - (xj[1] + p->dt * vj * sin(xj[2]));
        if (!(sqrt(dx * dx + dy * dy) > p->dis_thres)) dist_ok = 0;
      }
  }
  if (rk <= p->eps_pri && sk <= p->eps_dual && dist_ok) {
    b->curr_iter[s] = i_iter + 1;
    b->done[s] = 1;
    return;
  }
  if (p->dual_res_alias) b->aliased[s] = 1;
  memcpy(b->last_hat + SLOT_OFF(s, 0, 0), b->hat + SLOT_OFF(s, 0, 0), sizeof(double) * V * D * 2 * NP1);
  if (p->curr_iter_latest) b->curr_iter[s] = i_iter + 1;
}

static void plant_scenario(const piadmm_params* p, const piadmm_buffers* b, int s, int V) {
  const int N = p->num_ho;
  for (int i = 0; i < V; i++) {
    const double* xt = b->xt + ((size_t)s * V + i) * 3;
    const double v = b->spd[(size_t)s * V + i];
    const double* u = b->u + ((size_t)s * V + i) * N;
    double* o = b->xt_next + ((size_t)s * V + i) * 3;
    o[0] = xt[0] + v * cos(xt[2]) * p->dt;
    o[1] = xt[1] + v * sin(xt[2]) * p->dt;
    o[2] = xt[2] + v / p->L * u[0] * p->dt;
  }
}

int piadmm_ref_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* One MPC step for S scenarios: begin_step + outer loop + plant step.  All pointers in *b are HOST
 * pointers.  nthreads <= 0 -> OpenMP default.  Batches (S >= threads) run scenarios in parallel; coupled scenes
 * (S < threads) run the vehicles / pair-test rows / edges of one scenario in parallel.  Results do not depend on the
 * thread count (no floating-point reduction crosses threads). */
int piadmm_ref_step(const piadmm_params* p, const piadmm_buffers* b, int S, int V, int D, int nthreads) {
  if (!p || !b || p->num_ho < 3 || p->num_ho > NMAX || V < 1 || D < 1) return PIADMM_E_ARG;
  if (p->al_all && D < V - 1) return PIADMM_E_ARG;
  const int E = piadmm_ref_max_edges(V, D);
  consts_t* c = (consts_t*)malloc(sizeof(consts_t));
  build_consts(c, p->num_ho);
  int nt = 1;
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
  nt = omp_get_max_threads();
#endif
  const int inner = S < nt && V >= 16;   /* tiny scenes: not worth a parallel region per phase */
  half_t** pool = (half_t**)malloc(sizeof(half_t*) * 2 * nt);
  for (int t = 0; t < 2 * nt; t++) pool[t] = (half_t*)malloc(sizeof(half_t));
  const size_t mapb = (size_t)V * (((size_t)V + 7) / 8);
  uint8_t* hitmaps = (uint8_t*)malloc(mapb * (inner ? 1 : nt));
  uint8_t* dirty_all = (uint8_t*)malloc((size_t)S * V);
#pragma omp parallel for schedule(dynamic, 4) if (!inner)
  for (int s = 0; s < S; s++) {
    uint8_t* hitmap = hitmaps + (inner ? 0 : mapb * omp_thread_id());
    uint8_t* dirty = dirty_all + (size_t)s * V;
    memset(dirty, 1, V);
    begin_scenario(p, b, s, V, D, E);
    for (int i_iter = 0; i_iter < p->iter_num; i_iter++) {
      b->n_outer[s] += 1;
      x_update_scenario(p, c, b, s, V, D, pool, inner, dirty);
      detect_scenario(p, b, s, V, D, E, hitmap, inner);
      if (b->edge_n[s] == 0 && b->flag[s] == 0) break;
      b->flag[s] = 1;
      z_update_scenario(p, c, b, s, V, D, E, pool, inner, dirty);
      residual_scenario(p, b, s, V, D, E, i_iter);
      if (b->done[s]) break;
    }
    if (!b->done[s] && b->n_outer[s] == p->iter_num && b->flag[s]) b->status[s] |= PIADMM_ST_OUTER_MAXITER;
    b->done[s] = 1;
    plant_scenario(p, b, s, V);
  }
  for (int t = 0; t < 2 * nt; t++) free(pool[t]);
  free(pool); free(hitmaps); free(dirty_all);
  free(c);
  return PIADMM_OK;
}
