/* piadmm.h -- C ABI of the B200-native PI-ADMM consensus solver (libpiadmm.so).
 *
 * Drop-in boundary for the hot path of KaigeTan/Distributed-Local-Planner-PI-ADMM.
 * The reference has no FFI of its own (pure Python -> CasADi -> OSQP); every entry
 * point below cites the reference statement block it replaces (paths relative to the
 * reference repository root):
 *
 *   piadmm_x_update      casadi/main.py:81-106   (vehicle QPs: cost_function_primal +
 *                        nonlcon_function -> ca.qpsol("osqp") -> np.around -> dynamic_update_local)
 *   piadmm_detect_edges  casadi/main.py:110-118  (edge_mat, flag / early exit)
 *   piadmm_z_update      casadi/main.py:121-162  (edge problem, non-linear roll-out of z-hat,
 *                        dual update; PI forms: casadi_old_PI_ADMM/main.py:127-151 and
 *                        matlab_old_files/ADMM_CVX_two_veh_intesection_PI_antiwindup.m:152-188)
 *   piadmm_residuals     casadi/main.py:165-180  (error_rk / error_sk, stop test, alias quirk)
 *   piadmm_step          casadi/main.py:78-181   (the whole outer loop, one persistent kernel)
 *   piadmm_plant_step    casadi/main.py:185-192  (dynamic_update_edge numeric, xt <- column 1)
 *
 * Conventions
 *   - plain C, no torch types; all array arguments are raw pointers; fp64 = double, i32 = int32_t.
 *   - return value: 0 = ok, negative = error (PIADMM_E_*); nothing throws across the boundary.
 *   - device entry points are asynchronous on `stream` (a cudaStream_t passed as void*, NULL =
 *     legacy default stream).  The library BORROWS the device buffers bound with piadmm_bind();
 *     the caller (PyTorch on the Python side) owns and frees them.
 *   - a handle is not thread-safe; one handle per (process, device).
 *   - there is NO CPU fallback: every compute entry point fails with PIADMM_E_CUDA when no
 *     CUDA device is usable.
 *
 * Array layouts (row-major, contiguous).  S scenarios, V vehicles per scenario, N = num_ho,
 * D = neighbour slots per vehicle, I = iter_num, E = max edges per scenario = V*D/2 (>=1).
 */
#ifndef PIADMM_H
#define PIADMM_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PIADMM_ABI_VERSION 3

enum {
  PIADMM_OK = 0,
  PIADMM_E_ARG = -1,      /* bad argument / shape */
  PIADMM_E_CUDA = -2,     /* CUDA runtime error (see piadmm_last_error) */
  PIADMM_E_UNBOUND = -3,  /* piadmm_bind() not called / NULL buffer */
  PIADMM_E_UNSUPPORTED = -4
};

/* dual_mode */
enum { PIADMM_DUAL_PLAIN = 0, PIADMM_DUAL_PI = 1, PIADMM_DUAL_PI_OLD = 2 };

/* bits of status[s] */
enum {
  PIADMM_ST_VEH_MAXITER = 1,   /* a vehicle QP hit in_max_iter (reference: CasADi raises) */
  PIADMM_ST_EDGE_MAXITER = 2,  /* an edge problem hit in_max_iter (reference: error_on_fail=False) */
  PIADMM_ST_SLOT_OVERFLOW = 4, /* a vehicle needed more than D neighbour slots (the pair is dropped) */
  PIADMM_ST_OUTER_MAXITER = 8, /* the outer loop ended at iter_num without passing the stop test (casadi/main.py:78
                                  just falls out of its for loop; curr_iter then keeps its last value) */
  PIADMM_ST_PEER_TIMEOUT = 16  /* multi-GPU persistent kernel: a peer did not arrive at a barrier in time */
};

typedef struct piadmm_params {
  /* --- problem constants: casadi/PI_ADMM_class.py:15-28 --- */
  double dt, L, dis_thres, beta, Pnorm, Pcost, rho, eps_pri, eps_dual;
  /* --- PI law --- */
  double theta1, theta2;   /* MATLAB :22-23 sigmoid gain kP = theta1 - theta2/(1+exp(-min d)) */
  double kI;               /* MATLAB :21 ; < 0 means "use rho" */
  double windup_sat;       /* MATLAB :43 (30) ; old main.py:24 (20) */
  double backcalc;         /* weight of the back-calculation term: 1 (MATLAB :163) or 2 (old main.py:142) */
  double K_I, KP_num, KP_max, rho_num, rho_min, rho_max; /* old main.py:135-137 */
  /* --- inner OSQP-style ADMM (OSQP defaults: 0.1, 1e-6, 1.6, 1e-3, 1e-3, 5) --- */
  double in_rho0, in_sigma, in_alpha, in_eps_abs, in_eps_rel, in_adapt_tol;
  int32_t num_ho;          /* horizon N (3..31) */
  int32_t iter_num;        /* outer iteration cap */
  int32_t dual_mode;       /* PIADMM_DUAL_* */
  int32_t thr_squared;     /* 0: d^2 < dis_thres (casadi/main.py:113) ; 1: d^2 < dis_thres^2 (MATLAB :109) */
  int32_t dual_res_alias;  /* casadi/main.py:180 aliasing: dual residual == 0 from outer iteration 2 */
  int32_t stop_dist_test;  /* MATLAB :202 extra stop condition dis_vec(2) > dis_thres */
  int32_t pos_nonlinear;   /* MATLAB :312-330: numeric pos_old roll-out is the non-linear one */
  int32_t curr_iter_latest;/* MATLAB :209 */
  int32_t round_u;         /* decimals for np.around of the vehicle solution (casadi/main.py:103); <0 = none */
  int32_t round_edge;      /* decimals for the edge solution (casadi/main.py:153); <0 = none */
  int32_t al_all;          /* 1: AL term over every j != i (PI_ADMM_class.py:126-129; needs D >= V-1) */
  int32_t always_edge;     /* old main.py: edge step every iteration, no collision test */
  int32_t residual_global; /* old main.py:154-155 */
  int32_t PI, windup, trad;/* MATLAB :24-25 ; old main.py:18 */
  int32_t in_max_iter, in_check, in_adaptive;
  int32_t x_memo;          /* 1: a vehicle QP whose inputs did not change since its last solve in this MPC step (no slot of
                              the vehicle was touched by an edge problem) is not solved again -- the reference re-solves
                              every vehicle in every outer iteration (casadi/main.py:81) and gets the identical u; exact.
                              n_veh_solves / veh_inner then count the solves that were executed.  Off with dual_mode PI_OLD
                              (rho changes between iterations). */
} piadmm_params;

/* Device buffers owned by the caller.  Shapes in comments. */
typedef struct piadmm_buffers {
  /* inputs of one MPC step */
  const double* xt;      /* [S,V,3]  (x, y, theta) */
  const double* spd;     /* [S,V] */
  const double* refwin;  /* [S,V,2,N+1] reference window ref_traj[2i:2i+2, step:step+N+1] */
  /* primal state */
  double* u;             /* [S,V,N]     primal_u */
  double* pos;           /* [S,V,2,N+1] pos_old */
  /* per directed neighbour slot */
  int32_t* nbr;          /* [S,V,D] neighbour vehicle id, -1 = empty */
  double* hat;           /* [S,V,D,2,N+1] hat_pos_old[i,j] */
  double* dual;          /* [S,V,D,2,N+1] dual_var_old[i,j] */
  double* integ;         /* [S,V,D,2,N+1] PI integrator (sum_err) */
  double* diff;          /* [S,V,D,2,N+1] back-calculation term (diff_val) */
  double* last_hat;      /* [S,V,D,2,N+1] last_iter_hat_pos */
  /* per scenario */
  double* rho_s;         /* [S] current penalty (mutated only by dual_mode PI_OLD) */
  int32_t* flag;         /* [S] "a collision was seen" latch (casadi/main.py:71,115-118) */
  int32_t* done;         /* [S] 1 once the scenario left the outer loop */
  int32_t* aliased;      /* [S] alias latch of casadi/main.py:180 */
  int32_t* curr_iter;    /* [S] iter_his entry */
  int32_t* n_outer;      /* [S] outer iterations executed */
  int32_t* status;       /* [S] PIADMM_ST_* bits */
  double* err_rk;        /* [S,I] primal residual history */
  double* err_sk;        /* [S,I] dual residual history */
  int64_t* veh_inner;    /* [S] inner ADMM iterations summed over vehicle QPs */
  int64_t* edge_inner;   /* [S] inner ADMM iterations summed over edge problems */
  int32_t* n_veh_solves; /* [S] */
  int32_t* n_edge_solves;/* [S] */
  /* edge list of the current outer iteration */
  int32_t* edge_n;       /* [S] */
  int32_t* edge_ij;      /* [S,E,2] (veh1 < veh2), sorted */
  int32_t* edge_slot;    /* [S,E,2] slot of veh2 in veh1's table, slot of veh1 in veh2's table */
  double* res_part;      /* [S,2] rank-local partial residual sums (rk, sk) of the current outer iteration */
  /* plant */
  double* xt_next;       /* [S,V,3] */
} piadmm_buffers;

typedef struct piadmm_handle piadmm_handle;

/* version / introspection (no GPU needed) */
int piadmm_abi_version(void);
void piadmm_default_params(piadmm_params* p);          /* casadi/PI_ADMM_class.py:15-28 + casadi/main.py semantics */
int piadmm_max_edges(int V, int D);

/* lifecycle */
int piadmm_create(const piadmm_params* p, int device, int S, int V, int D, piadmm_handle** out);
int piadmm_set_params(piadmm_handle* h, const piadmm_params* p);   /* num_ho / iter_num must not change */
int piadmm_bind(piadmm_handle* h, const piadmm_buffers* b);
void piadmm_destroy(piadmm_handle* h);
const char* piadmm_last_error(const piadmm_handle* h);             /* h may be NULL: last create error */

/* one MPC step */
int piadmm_begin_step(piadmm_handle* h, void* stream);   /* casadi/main.py:52-72 re-initialisation */
int piadmm_x_update(piadmm_handle* h, void* stream);
int piadmm_detect_edges(piadmm_handle* h, void* stream);
int piadmm_z_update(piadmm_handle* h, void* stream);     /* edge problems + dual / PI update */
int piadmm_residuals(piadmm_handle* h, void* stream);    /* histories, stop flags, alias latch (= partials + decide) */
/* The edge problems ALONE, for scripts that keep casadi/main.py's own statements around the solve (:133-153: build the
 * problem, F = ca.qpsol("osqp"), result = F(lbg=0)) and do the z-hat roll-out (:156-158) and the dual update (:161-162)
 * themselves: solves the edges listed in the bound edge_n / edge_ij / edge_slot from the bound xt, spd, pos, dual, rho_s;
 * writes nothing but the library-owned solution buffer, read back with piadmm_get_edge_u (host_out [S,E,2,N], edge
 * controls rounded to round_edge decimals; synchronises). */
int piadmm_solve_edges(piadmm_handle* h, void* stream);
int piadmm_get_edge_u(piadmm_handle* h, double* host_out, void* stream);
/* The joint N-vehicle RSU edge problem: cost_function_edge of
 * matlab_old_files/ADMM_fmincon_multi_veh_intesection_static.m:250-291 (= casadi_old_PI_ADMM/PI_ADMM_class.py:141-178) in its
 * convex statement -- ONE edge problem over all V vehicles of every scenario, with ONE consensus copy / multiplier per
 * vehicle (neighbour slot 0 of the bound pos_old / dual buffers), the hinge summed over the pairs of the given list
 * (pair_n [S], pair_ij [S,P,2] with veh_lo < veh_hi; DEVICE pointers; at most 8 pairs per vehicle, else the scenario's
 * status is PIADMM_ST_SLOT_OVERFLOW and nothing is solved), bounds / rate rows of nonlcon_function on every vehicle.
 * hinge_weight = the weight of ONE pair's hinge: the reference sums over ordered pairs (:265-276), i.e. 2 * beta.
 * The reference forms last_dis from linear indices (:268), the pair's own offset only for V = 2; here every pair is
 * linearised about its own seed offset l_ij (as every two-vehicle statement does).  Outputs (device): u_out [S,V,N]
 * (rounded to round_edge decimals), iters_out [S], status_out [S] (0 / PIADMM_ST_EDGE_MAXITER / PIADMM_ST_SLOT_OVERFLOW).
 * One thread-block cluster per scenario (V <= 64 for N > 15, V <= 128 for N <= 15; PIADMM_E_UNSUPPORTED beyond). */
int piadmm_solve_joint(piadmm_handle* h, const int32_t* pair_n, const int32_t* pair_ij, int P, double hinge_weight,
                       double* u_out, int32_t* iters_out, int32_t* status_out, void* stream);
/* vehicle-sharded (multi-GPU) form: this rank x-updates vehicles [lo,hi), solves every edge with an owned
 * endpoint, and sums the residual terms of the edges whose veh1 it owns into res_part; the caller
 * all-gathers pos/u rows after x_update and all-reduces (sum) res_part between the two residual calls. */
int piadmm_set_ownership(piadmm_handle* h, int lo, int hi);
int piadmm_residual_partials(piadmm_handle* h, void* stream);
int piadmm_residual_decide(piadmm_handle* h, void* stream);
int piadmm_num_active(piadmm_handle* h, void* stream, int32_t* host_out); /* scenarios with done==0 (syncs) */
int piadmm_step(piadmm_handle* h, void* stream);         /* begin_step + all outer iterations, persistent kernel */
int piadmm_plant_step(piadmm_handle* h, void* stream);

/* Vehicle-sharded multi-GPU step without NCCL in the data path (config 5; there is no reference counterpart: the
 * reference's "communication" is Python variable passing, casadi/main.py:75-77).  Every rank creates a handle for the
 * whole scene (S == 1), sets its ownership range, exports a library-owned exchange block as a CUDA IPC handle
 * (PIADMM_PEER_HANDLE_BYTES bytes), the caller all-gathers the handles of all ranks (torch.distributed) and every rank
 * maps its peers' blocks.  piadmm_step then runs the persistent cooperative kernel on every rank: owned pos_old /
 * primal_u rows and per-edge residual terms are stored straight into the peers' blocks over NVLink, two of the grid
 * barriers per outer iteration span all ranks (flags in peer memory).  All ranks must call piadmm_step collectively.
 * A peer that does not arrive within PIADMM_PEER_TIMEOUT_MS (default 20000) sets PIADMM_ST_PEER_TIMEOUT. */
#define PIADMM_PEER_HANDLE_BYTES 64
int piadmm_peer_export(piadmm_handle* h, void* handle_out);
int piadmm_peer_connect(piadmm_handle* h, int rank, int world, const void* handles /* world x PIADMM_PEER_HANDLE_BYTES */);
int piadmm_peer_disconnect(piadmm_handle* h);

/* host-buffer convenience path (what a ctypes caller without torch uses; copies inside):
 * runs begin_step + the whole outer loop + plant step for host inputs and returns host outputs. */
int piadmm_step_host(piadmm_handle* h, const double* xt, const double* spd, const double* refwin,
                     double* u_out, double* pos_out, double* xt_next_out, int32_t* curr_iter_out,
                     int32_t* n_outer_out, void* stream);

/* Resident reference trajectories.  The reference builds ref_traj once in PI_ADMM_CASADI.__init__
 * (casadi/PI_ADMM_class.py:31-37) and every MPC step only slices the window [num_step, num_step+num_ho] out of
 * it (cost_function_primal, :116-117; casadi/main.py:43,84).  piadmm_set_reference copies ref [S,V,2,T]
 * (host or device pointer; row (s,v,0) = x reference, (s,v,1) = y reference) into library-owned device memory;
 * piadmm_step_host_at then needs only the plant state xt [S,V,3] per step: the window is gathered on the device,
 * spd is the bound (device-resident) buffer.  Outputs as piadmm_step_host.  Errors: PIADMM_E_UNBOUND without a
 * reference, PIADMM_E_ARG when the window leaves the reference (the reference script stops at Nt/dt - num_ho). */
int piadmm_set_reference(piadmm_handle* h, const double* ref, int T, void* stream);
int piadmm_step_host_at(piadmm_handle* h, int num_step, const double* xt, double* u_out, double* pos_out,
                        double* xt_next_out, int32_t* curr_iter_out, int32_t* n_outer_out, void* stream);

/* Whole receding-horizon run on the device: n_steps MPC steps of casadi/main.py:43-201 without a
 * host round trip.  Per step: window gather from the resident reference at num_step = first_step + t, the outer
 * PI-ADMM loop, the plant step (:185-192), the histories the script appends (:186,199-201) and xt <- xt_next.  The
 * initial plant state is the bound xt buffer (overwritten: on return it holds the final state).  History arguments are
 * DEVICE pointers, any may be NULL:
 *   x_vec [n_steps,S,V,2], theta_vec [n_steps,S,V], u_vec [n_steps,S,V] (first control of the step),
 *   iter_his [n_steps,S] (curr_iter), n_outer_his [n_steps,S], err_rk_his [n_steps,S] (last primal residual).
 * Asynchronous on `stream` when V == 2 (fused kernel); the generic-V path synchronises once per outer iteration. */
typedef struct piadmm_histories {
  double *x_vec, *theta_vec, *u_vec, *err_rk_his;
  int32_t *iter_his, *n_outer_his;
  int32_t *status_his;   /* [n_steps,S] PIADMM_ST_* bits of every step (vehicle / edge solver failures, slot overflow) */
} piadmm_histories;
int piadmm_run_closed_loop(piadmm_handle* h, int first_step, int n_steps, const piadmm_histories* out, void* stream);

/* Debug: phase timeline of the LAST cooperative launch (generic V).  The first call enables the trace; later calls copy
 * out[0] = number of stamps, out[1..] = %globaltimer (ns) at the kernel start, after the re-initialisation, and then
 * seven per outer iteration: after the x phase (+ exchange barrier), slots, count, edges, z phase, its exchange
 * barrier, decide.  n <= 1024. */
int piadmm_debug_coop_trace(piadmm_handle* h, int64_t* host_out, int n);

/* Allocation-free collision-edge test.  NO COUNTERPART IN THE REFERENCE as an entry point: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  What it computes is casadi/main.py:110-113 (edge_mat[i,j] = 1 when any of the N+1
 * trajectory points of vehicles i < j are closer than the threshold) and the row-major pair list of :121
 * (np.where(edge_mat == 1)) -- the matrix form of piadmm_detect_edges (k_bbox + k_hits + emit) without a handle and
 * without slot tables: its scratch (trajectory bounding boxes + the hit bit matrix) lives in ONE file-scope __device__
 * array of fixed size (csrc/piadmm_static.cu, 2.25 MiB, held by every GPU that loads the library until exit).
 *   pos [S,V,2,N+1] device, fp64 (pos_old of piadmm_buffers); N = 1..31;
 *   threshold: d^2 < dis_thres (thr_squared = 0, casadi/main.py:113) or d^2 < dis_thres^2 (thr_squared = 1, MATLAB :109);
 *   edge_n [S] device: number of colliding pairs of the scenario (may exceed e_cap);
 *   edge_ij [S,e_cap,2] device: the first e_cap pairs (veh1 < veh2), sorted by (veh1, veh2); rows >= edge_n untouched.
 * Limits: S*V <= PIADMM_STATIC_MAX_VEH and S*V*ceil(V/32) <= PIADMM_STATIC_MAX_WORDS (one 4096-vehicle scene, or 4096
 * two-vehicle scenarios); anything larger returns PIADMM_E_UNSUPPORTED (text: piadmm_static_last_error).
 * Calls no cudaMalloc / cudaFree / allocating routine; asynchronous on `stream`; runs on the device that owns `pos` and
 * leaves the caller's current device as it was.  The first call on a device looks up the array's device address
 * (cudaGetSymbolAddress) -- make it before starting a graph capture.  All calls on a device share the array: calls
 * ordered on the GPU are fine, calls that OVERLAP on the GPU (different unordered streams) race on it.  Nothing has to
 * be cleared between calls: every scratch word a call reads was written by the same call. */
#define PIADMM_STATIC_MAX_VEH 8192
#define PIADMM_STATIC_MAX_WORDS 524288
int piadmm_edge_list_static(const double* pos, int S, int V, int N, double dis_thres, int thr_squared, int e_cap,
                            int32_t* edge_n, int32_t* edge_ij, void* stream);
const char* piadmm_static_last_error(void);

/* statistics for the roofline: kernels launched by this handle since creation */
int64_t piadmm_launch_count(const piadmm_handle* h);

/* FP64 FMA micro-benchmark (independent DFMA chains on every SM); returns TFLOP/s in *tflops */
int piadmm_fp64_peak(int device, int iters, double* tflops, double* ms);

#ifdef __cplusplus
}
#endif
#endif /* PIADMM_H */
