// obca_consensus.cu -- consensus pieces of the OBCA overtaking ADMM (include/obca_consensus.h), sm_100a.
// HBM-bound elementwise / reduction work, batched over scenarios: one CTA per scenario for everything that
// needs a per-scenario decision, flat grids for the pure elementwise entry points.
#include <cuda_runtime.h>
#include "../../include/obca_consensus.h"
#include "../../include/piadmm.h"

#define OBCA_THREADS 128

struct ObcaCtx {
  obca_dims d;
  obca_state cur, prev;
  const double* ctrl;
  double* ctrl_prev;
  double rho, primal_thres, dual_thres;
  int max_iter;
  double *primal, *dual;
  int32_t *n_iter, *stop;
};

// lamb_bar[t] + rho*(local_z - Z_bar[t]) with numpy's rounding sequence (no FMA contraction):
// Distributed_planner/decentralized/optimizer.py:333-335
__device__ __forceinline__ double lambda_elem(const obca_dims& d, const obca_state& st, size_t row, int w, double rho) {
  const int nz = d.ns + d.nl;
  const double lz = w < d.ns ? st.local_x[row * d.ns + w] : st.lamb_ij[row * d.nl + (w - d.ns)];
  const size_t t = row * nz + w;
  return __dadd_rn(st.lamb_bar[t], __dmul_rn(rho, __dsub_rn(lz, st.Z_bar[t])));
}

__device__ __forceinline__ double abs_via_sqrt(double x) { return sqrt(__dmul_rn(x, x)); }   // np.sqrt(x*x)

__device__ __forceinline__ double block_sum(double v, double* red) {
  red[threadIdx.x] = v;
  __syncthreads();
  for (int o = blockDim.x / 2; o >= 1; o >>= 1) {
    if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  const double r = red[0];
  __syncthreads();
  return r;
}

// rows h <- h+1 of one member ([V,H,W] block of one scenario), last row repeated; executed by a whole CTA.
// Every thread owns whole columns, so the in-place shift has no cross-thread hazard.
__device__ __forceinline__ void shift_block(double* a, int V, int H, int W) {
  for (int col = threadIdx.x; col < V * W; col += blockDim.x) {
    double* p = a + (size_t)(col / W) * H * W + (col % W);
    for (int h = 0; h + 1 < H; h++) p[(size_t)h * W] = p[(size_t)(h + 1) * W];
  }
}
__device__ __forceinline__ void copy_block(double* dst, const double* src, int n) {
  for (int t = threadIdx.x; t < n; t += blockDim.x) dst[t] = src[t];
}

__global__ void k_obca_lambda(const obca_dims d, obca_state st, double rho) {
  const int nz = d.ns + d.nl;
  const size_t n = (size_t)d.S * d.V * d.H * nz;
  for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (size_t)gridDim.x * blockDim.x)
    st.lamb_bar[t] = lambda_elem(d, st, t / nz, (int)(t % nz), rho);
}

__global__ void k_obca_res(const obca_dims d, const double* ctrl, const double* ctrl_prev, const double* lb,
                           const double* lb_prev, double* primal, double* dual) {
  __shared__ double red[OBCA_THREADS];
  const int s = blockIdx.x, nz = d.ns + d.nl;
  const size_t nc = (size_t)d.V * d.C, nl = (size_t)d.V * d.H * nz;
  double p = 0.0, q = 0.0;
  for (size_t t = threadIdx.x; t < nc; t += blockDim.x) p += abs_via_sqrt(ctrl[s * nc + t] - ctrl_prev[s * nc + t]);
  for (size_t t = threadIdx.x; t < nl; t += blockDim.x) q += abs_via_sqrt(lb[s * nl + t] - lb_prev[s * nl + t]);
  p = block_sum(p, red);
  q = block_sum(q, red);
  if (threadIdx.x == 0) { primal[s] = p; dual[s] = q; }
}

__global__ void k_obca_check(const obca_dims d, obca_state st, double thres, double min_dis, int32_t* ok) {
  const int s = blockIdx.x;
  int bad = 0;
  for (int h = threadIdx.x; h < d.H; h += blockDim.x) {
    double e0 = 0.0, e1 = 0.0, ue = 0.0;
    for (int v = 0; v < d.V; v++) {
      const size_t row = ((size_t)s * d.V + v) * d.H + h;
      for (int r = 0; r < d.nl; r++) {
        const double lam = st.lamb_ij[row * d.nl + r];
        e0 += st.A[(row * d.nl + r) * 2] * lam;
        e1 += st.A[(row * d.nl + r) * 2 + 1] * lam;
        ue -= st.b[row * d.nl + r] * lam;
      }
    }
    if (!(fabs(e0) <= thres && fabs(e1) <= thres && ue >= min_dis)) bad = 1;
  }
  bad = __syncthreads_or(bad);
  if (threadIdx.x == 0) ok[s] = bad ? 0 : 1;
}

__device__ __forceinline__ void shift_state(const obca_dims& d, const obca_state& st, int s) {
  const int nz = d.ns + d.nl;
  const size_t r = (size_t)s * d.V * d.H;
  shift_block(st.Z_bar + r * nz, d.V, d.H, nz);
  shift_block(st.A + r * d.nl * 2, d.V, d.H, d.nl * 2);
  shift_block(st.b + r * d.nl, d.V, d.H, d.nl);
  shift_block(st.lamb_bar + r * nz, d.V, d.H, nz);
  shift_block(st.lamb_ij + r * d.nl, d.V, d.H, d.nl);
  shift_block(st.local_x + r * d.ns, d.V, d.H, d.ns);
}

__global__ void k_obca_shift(const obca_dims d, obca_state st) { shift_state(d, st, blockIdx.x); }

__global__ void k_obca_tail(const ObcaCtx c) {
  __shared__ double red[OBCA_THREADS];
  const obca_dims& d = c.d;
  const int s = blockIdx.x, nz = d.ns + d.nl;
  if (c.stop[s]) return;
  const size_t rows = (size_t)d.V * d.H, r0 = (size_t)s * rows, nc = (size_t)d.V * d.C;
  // lambda update + dual residual against the previous iteration's multipliers
  double q = 0.0;
  for (size_t t = threadIdx.x; t < rows * nz; t += blockDim.x) {
    const double v = lambda_elem(d, c.cur, r0 + t / nz, (int)(t % nz), c.rho);
    c.cur.lamb_bar[r0 * nz + t] = v;
    q += abs_via_sqrt(v - c.prev.lamb_bar[r0 * nz + t]);
  }
  double p = 0.0;
  for (size_t t = threadIdx.x; t < nc; t += blockDim.x) p += abs_via_sqrt(c.ctrl[s * nc + t] - c.ctrl_prev[s * nc + t]);
  p = block_sum(p, red);
  q = block_sum(q, red);
  const int it = c.n_iter[s] + 1;
  const bool stop = (p <= c.primal_thres && q <= c.dual_thres) || it > c.max_iter;
  __syncthreads();
  if (threadIdx.x == 0) { c.primal[s] = p; c.dual[s] = q; c.n_iter[s] = it; c.stop[s] = stop ? 1 : 0; }
  if (stop) {
    shift_state(d, c.prev, s);      // bar_state_prev = iterate_next_state(bar_state_prev)
  } else {                          // bar_state_prev = deepcopy(optimizer.bar_state); bar_ctrl_prev = bar_ctrl
    copy_block(c.prev.Z_bar + r0 * nz, c.cur.Z_bar + r0 * nz, (int)(rows * nz));
    copy_block(c.prev.A + r0 * d.nl * 2, c.cur.A + r0 * d.nl * 2, (int)(rows * d.nl * 2));
    copy_block(c.prev.b + r0 * d.nl, c.cur.b + r0 * d.nl, (int)(rows * d.nl));
    copy_block(c.prev.lamb_bar + r0 * nz, c.cur.lamb_bar + r0 * nz, (int)(rows * nz));
    copy_block(c.prev.lamb_ij + r0 * d.nl, c.cur.lamb_ij + r0 * d.nl, (int)(rows * d.nl));
    copy_block(c.prev.local_x + r0 * d.ns, c.cur.local_x + r0 * d.ns, (int)(rows * d.ns));
    copy_block(c.ctrl_prev + s * nc, c.ctrl + s * nc, (int)nc);
  }
}

// bar_state_update (optimizer.py:205-222): half-spaces A x <= b of every vehicle rectangle at every predicted step from
// the local solutions, and local_x <- the states.  prob != 0: compute_square_halfspaces_ca_prob (util.py:76-101, the
// delay-tightened form: A = [R'; -R'], b = b0 + A (pos + avg_delay v (c,s) + sqrt(p/(1-p)) (var_delay v (c,s))^2));
// prob == 0: generate_vehicle_vertices + compute_square_halfspaces_ca (util.py:12-74).  One thread per (s, v, h).
__global__ void k_obca_halfspaces(const obca_dims d, obca_state st, const double* __restrict__ fullx, const obca_vehicle cfg, int prob) {
  const size_t n = (size_t)d.S * d.V * d.H;
  const int nz = d.ns + d.nl;
  for (size_t row = (size_t)blockIdx.x * blockDim.x + threadIdx.x; row < n; row += (size_t)gridDim.x * blockDim.x) {
    const double* fx = fullx + row * nz;
    const double x = fx[0], y = fx[1], v = fx[2], th = fx[3];
    for (int w = 0; w < d.ns; w++) st.local_x[row * d.ns + w] = fx[w];
    double sn, cs;
    sincos(th, &sn, &cs);
    double* A = st.A + row * d.nl * 2;
    double* b = st.b + row * d.nl;
    const double hl = 0.5 * cfg.length, hw = 0.5 * cfg.width;
    if (prob) {
      const double a[4][2] = {{cs, sn}, {-sn, cs}, {-cs, -sn}, {sn, -cs}};
      const double b0[4] = {hl, hw, hl, hw};
      const double kq = sqrt(cfg.prob / (1.0 - cfg.prob));
      const double vx = cfg.var_delay * v * cs, vy = cfg.var_delay * v * sn;
      const double px = x + cfg.avg_delay * v * cs + kq * (vx * vx), py = y + cfg.avg_delay * v * sn + kq * (vy * vy);
      for (int r = 0; r < 4; r++) {
        A[2 * r] = a[r][0]; A[2 * r + 1] = a[r][1];
        b[r] = __dadd_rn(b0[r], __dadd_rn(__dmul_rn(a[r][0], px), __dmul_rn(a[r][1], py)));
      }
    } else {
      const double vxs[4] = {x + hl * cs - hw * sn, x + hl * cs + hw * sn, x - hl * cs + hw * sn, x - hl * cs - hw * sn};
      const double vys[4] = {y + hl * sn + hw * cs, y + hl * sn - hw * cs, y - hl * sn - hw * cs, y - hl * sn + hw * cs};
      for (int r = 0; r < 4; r++) {
        const double p1x = vxs[r], p1y = vys[r], p2x = vxs[(r + 1) & 3], p2y = vys[(r + 1) & 3];
        const double nx = p2y - p1y, ny = p1x - p2x, nrm = sqrt(__dadd_rn(__dmul_rn(nx, nx), __dmul_rn(ny, ny)));
        A[2 * r] = (p1y - p2y) / nrm; A[2 * r + 1] = (p2x - p1x) / nrm;
        b[r] = __dsub_rn(__dmul_rn(p2x, p1y), __dmul_rn(p2y, p1x)) / nrm;
      }
    }
  }
}

// ------------------------------------------------------------------ C ABI
static int dims_ok(const obca_dims* d) {
  return d && d->S >= 1 && d->V >= 1 && d->H >= 1 && d->ns >= 1 && d->nl >= 1 && d->C >= 0;
}
static int state_ok(const obca_state* st) {
  return st && st->Z_bar && st->A && st->b && st->lamb_bar && st->lamb_ij && st->local_x;
}
static int have_device() {
  int n = 0;
  return cudaGetDeviceCount(&n) == cudaSuccess && n > 0;
}
#define OBCA_ENTER(cond)                          \
  do {                                            \
    if (!(cond)) return PIADMM_E_ARG;             \
    if (!have_device()) return PIADMM_E_CUDA;     \
  } while (0)
#define OBCA_LEAVE() return cudaGetLastError() == cudaSuccess ? PIADMM_OK : PIADMM_E_CUDA

// run on the device that owns the caller's buffers (the entry points carry no handle) and size grids by ITS SM count
static int use_device_of(const void* p) {
  cudaPointerAttributes at;
  int dev = 0;
  if (p && cudaPointerGetAttributes(&at, p) == cudaSuccess && at.type == cudaMemoryTypeDevice) dev = at.device;
  else { cudaGetLastError(); cudaGetDevice(&dev); }
  cudaSetDevice(dev);
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms < 1) sms = 1;
  return sms;
}

extern "C" {

int obca_lambda_update(const obca_dims* d, obca_state* st, double rho, void* stream) {
  OBCA_ENTER(dims_ok(d) && state_ok(st));
  const size_t n = (size_t)d->S * d->V * d->H * (d->ns + d->nl);
  const size_t cap = (size_t)use_device_of(st->lamb_bar) * 16;
  size_t grid = (n + 255) / 256;
  if (grid > cap) grid = cap;
  k_obca_lambda<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(*d, *st, rho);
  OBCA_LEAVE();
}

int obca_residuals(const obca_dims* d, const double* ctrl, const double* ctrl_prev, const double* lamb_bar,
                   const double* lamb_bar_prev, double* primal, double* dual, void* stream) {
  OBCA_ENTER(dims_ok(d) && ctrl && ctrl_prev && lamb_bar && lamb_bar_prev && primal && dual);
  use_device_of(ctrl);
  k_obca_res<<<d->S, OBCA_THREADS, 0, (cudaStream_t)stream>>>(*d, ctrl, ctrl_prev, lamb_bar, lamb_bar_prev, primal, dual);
  OBCA_LEAVE();
}

int obca_check_converge(const obca_dims* d, const obca_state* st, double thres, double min_dis, int32_t* ok,
                        void* stream) {
  OBCA_ENTER(dims_ok(d) && state_ok(st) && ok);
  use_device_of(ok);
  k_obca_check<<<d->S, 32, 0, (cudaStream_t)stream>>>(*d, *st, thres, min_dis, ok);
  OBCA_LEAVE();
}

int obca_iterate_next_state(const obca_dims* d, obca_state* st, void* stream) {
  OBCA_ENTER(dims_ok(d) && state_ok(st));
  use_device_of(st->lamb_bar);
  k_obca_shift<<<d->S, OBCA_THREADS, 0, (cudaStream_t)stream>>>(*d, *st);
  OBCA_LEAVE();
}

int obca_bar_state_update(const obca_dims* d, obca_state* st, const double* fullx, const obca_vehicle* cfg, int prob, void* stream) {
  OBCA_ENTER(dims_ok(d) && state_ok(st) && fullx && cfg && d->nl == 4 && d->ns >= 4 && cfg->length > 0.0 && cfg->width > 0.0 &&
             (!prob || (cfg->prob > 0.0 && cfg->prob < 1.0)));
  const size_t n = (size_t)d->S * d->V * d->H;
  const size_t cap = (size_t)use_device_of(fullx) * 16;
  size_t grid = (n + 127) / 128;
  if (grid > cap) grid = cap;
  k_obca_halfspaces<<<(unsigned)grid, 128, 0, (cudaStream_t)stream>>>(*d, *st, fullx, *cfg, prob);
  OBCA_LEAVE();
}

int obca_consensus_tail(const obca_dims* d, obca_state* cur, obca_state* prev, const double* ctrl,
                        double* ctrl_prev, double rho, double primal_thres, double dual_thres, int32_t max_iter,
                        double* primal, double* dual, int32_t* n_iter, int32_t* stop, void* stream) {
  OBCA_ENTER(dims_ok(d) && state_ok(cur) && state_ok(prev) && ctrl && ctrl_prev && primal && dual && n_iter && stop);
  use_device_of(ctrl);
  ObcaCtx c;
  c.d = *d; c.cur = *cur; c.prev = *prev; c.ctrl = ctrl; c.ctrl_prev = ctrl_prev; c.rho = rho;
  c.primal_thres = primal_thres; c.dual_thres = dual_thres; c.max_iter = max_iter;
  c.primal = primal; c.dual = dual; c.n_iter = n_iter; c.stop = stop;
  k_obca_tail<<<d->S, OBCA_THREADS, 0, (cudaStream_t)stream>>>(c);
  OBCA_LEAVE();
}

}  // extern "C"
