/* obca_consensus.h -- C ABI of the consensus pieces of the OBCA overtaking ADMM (SURVEY.md 8(a) row a13,
 * BASELINE config 2), part of libpiadmm.so.
 *
 * The local / RSU problems of that driver are IPOPT NLPs (out of scope); what sits between
 * them on the per-iteration path is plain array arithmetic, batched here over S independent scenarios:
 *
 *   obca_bar_state_update     Distributed_planner/decentralized/optimizer.py:205-222  (bar_state_update, util.py half-spaces)
 *   obca_lambda_update        Distributed_planner/decentralized/optimizer.py:330-335  (lambda_update)
 *   obca_residuals            Distributed_planner/decentralized_overtaking_ADMM.py:82-85 (L1 residuals)
 *   obca_check_converge       Distributed_planner/decentralized/optimizer.py:225-235  (check_converge)
 *   obca_iterate_next_state   Distributed_planner/decentralized/optimizer.py:337-344  (iterate_next_state)
 *   obca_consensus_tail       Distributed_planner/decentralized_overtaking_ADMM.py:79-97 (lambda update,
 *                             residuals, stop test, then either horizon shift of the previous state or
 *                             previous <- current), one kernel, one CTA per scenario
 *
 * Layouts (row-major, fp64): S scenarios, V vehicles, H = N_horz-1 predicted steps, ns states (5),
 * nl local multipliers per step (4), nz = ns+nl (9), C = control entries per vehicle ((N_horz-1)*n_controls).
 * The fields of struct obca_state are the members of the reference's `mid_state`
 * (Distributed_planner/decentralized/optimizer.py:351-373) with a leading scenario axis.
 *
 * Conventions as in piadmm.h: plain C, raw device pointers borrowed for the call, asynchronous on
 * `stream` (cudaStream_t as void*), 0 = ok / negative = PIADMM_E_*, no CPU fallback.
 */
#ifndef OBCA_CONSENSUS_H
#define OBCA_CONSENSUS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct obca_dims {
  int32_t S, V, H, ns, nl, C;
} obca_dims;

typedef struct obca_state {
  double* Z_bar;     /* [S,V,H,ns+nl] */
  double* A;         /* [S,V,H,nl,2]  */
  double* b;         /* [S,V,H,nl]    */
  double* lamb_bar;  /* [S,V,H,ns+nl] */
  double* lamb_ij;   /* [S,V,H,nl]    */
  double* local_x;   /* [S,V,H,ns]    */
} obca_state;

/* vehicle rectangle + communication-delay model: Distributed_planner/decentralized/veh_config.py:9-27 */
typedef struct obca_vehicle {
  double length, width;          /* 3.5, 2 */
  double avg_delay, var_delay;   /* 0.05, 0.025 */
  double prob;                   /* 0.95 */
} obca_vehicle;

/* bar_state_update (optimizer.py:205-222): for every vehicle and predicted step, A [4,2] and b [4] of the vehicle
 * rectangle's half-spaces A x <= b from the local solutions fullx [S,V,H,ns+nl] (state = x, y, v, heading, steer), and
 * local_x <- the states.  prob != 0: compute_square_halfspaces_ca_prob (util.py:76-101, delay-tightened, what the driver
 * runs: if_comm_delay = 1, decentralized_overtaking_ADMM.py:21-23); prob == 0: generate_vehicle_vertices +
 * compute_square_halfspaces_ca (util.py:12-74).  lamb_ij is NOT touched (the reference's update of it is commented out,
 * optimizer.py:219-220). */
int obca_bar_state_update(const obca_dims* d, obca_state* st, const double* fullx, const obca_vehicle* cfg, int prob, void* stream);

/* lamb_bar += rho * ([local_x, lamb_ij] - Z_bar) */
int obca_lambda_update(const obca_dims* d, obca_state* st, double rho, void* stream);

/* primal[s] = sum |ctrl - ctrl_prev| ; dual[s] = sum |lamb_bar - lamb_bar_prev| (each term sqrt(x*x)) */
int obca_residuals(const obca_dims* d, const double* ctrl, const double* ctrl_prev, const double* lamb_bar,
                   const double* lamb_bar_prev, double* primal, double* dual, void* stream);

/* ok[s] = all |sum_v A_v' lamb_v| <= thres  and  all (-sum_v b_v' lamb_v) >= min_dis */
int obca_check_converge(const obca_dims* d, const obca_state* st, double thres, double min_dis, int32_t* ok,
                        void* stream);

/* every member: row h <- row h+1 along H, last row repeated (in place) */
int obca_iterate_next_state(const obca_dims* d, obca_state* st, void* stream);

/* Tail of one ADMM iteration for all scenarios that are still iterating (stop[s] == 0 on entry):
 *   cur.lamb_bar += rho*(local_z - cur.Z_bar); residuals -> primal[s], dual[s]; n_iter[s] += 1;
 *   if (primal <= primal_thres && dual <= dual_thres) || n_iter > max_iter:
 *        prev <- iterate_next_state(prev); stop[s] = 1
 *   else prev <- cur (all six members), ctrl_prev <- ctrl
 */
int obca_consensus_tail(const obca_dims* d, obca_state* cur, obca_state* prev, const double* ctrl,
                        double* ctrl_prev, double rho, double primal_thres, double dual_thres, int32_t max_iter,
                        double* primal, double* dual, int32_t* n_iter, int32_t* stop, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* OBCA_CONSENSUS_H */
