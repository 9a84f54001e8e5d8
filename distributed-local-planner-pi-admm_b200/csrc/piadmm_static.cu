// piadmm_static.cu -- piadmm_edge_list_static (include/piadmm.h): the collision-edge test of casadi/main.py:110-113 and
// the row-major edge list of :121 as ONE allocation-free entry point, for callers that cannot call cudaMalloc (CUDA
// graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the
// reference.
//
// It is the matrix form of the detection (k_bbox + k_hits + the emit of k_edges in piadmm.cu) with the scratch those
// kernels keep in cudaMalloc'ed, handle-owned buffers (trajectory bounding boxes, the hit bit matrix) moved into a
// file-scope __device__ array of fixed size.  Every GPU that loads libpiadmm.so holds that array (2.25 MiB) until exit.
//
// Sharing: all calls on a device use the SAME array.  Calls that are ordered on the GPU (one stream, or streams ordered
// by events) are fine; calls that overlap on the GPU race on it.  Nothing relies on what an earlier call left behind:
// every box and every word of the bit matrix that the emit kernel reads is written by the box / hit kernels of the same
// call (no atomics into a pre-zeroed matrix as k_hits does), so there is nothing to clear.
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/piadmm.h"

// limits of one call (stated in include/piadmm.h next to the declaration)
#define ST_MAX_WORDS PIADMM_STATIC_MAX_WORDS   // S * V * ceil(V/32) words of the hit bit matrix (2 MiB)
#define ST_MAX_VEH PIADMM_STATIC_MAX_VEH       // S * V bounding boxes of 32 bytes (256 KiB)
#define ST_FULL 0xffffffffu

// layout of the array: [0, 32 * ST_MAX_VEH) bounding boxes (xmin, xmax, ymin, ymax) of every trajectory, then the hit
// bit matrix: word (s, i, jb), bit jj <=> vehicle j = 32 jb + jj > i collides with i (upper triangle)
#define ST_BBOX_BYTES (sizeof(double4) * (size_t)ST_MAX_VEH)
#define ST_BYTES (ST_BBOX_BYTES + sizeof(uint32_t) * (size_t)ST_MAX_WORDS)
static __device__ __align__(256) unsigned char g_static_scratch[ST_BYTES];

struct StaticScratch {             // the two views of the array, from its DEVICE address
  double4* bbox;
  uint32_t* hits;
};

// one warp per vehicle, lane k = trajectory point k (N + 1 <= 32; padded lanes repeat point 0)
__global__ void k_static_bbox(const double* __restrict__ pos, long agents, int N, StaticScratch sc) {
  const long a = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (a >= agents) return;
  const int k = lane <= N ? lane : 0;
  const double* p = pos + (size_t)a * 2 * (N + 1);
  double x0 = p[k], x1 = x0, y0 = p[N + 1 + k], y1 = y0;
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    x0 = fmin(x0, __shfl_xor_sync(ST_FULL, x0, o)); x1 = fmax(x1, __shfl_xor_sync(ST_FULL, x1, o));
    y0 = fmin(y0, __shfl_xor_sync(ST_FULL, y0, o)); y1 = fmax(y1, __shfl_xor_sync(ST_FULL, y1, o));
  }
  if (lane == 0) sc.bbox[a] = make_double4(x0, x1, y0, y1);
}

// one thread per (s, i, word jb): the exact point-by-point test of casadi/main.py:110-113 behind a box reject, for the
// j > i of the word; EVERY word is stored (zero below the diagonal), so the matrix needs no clearing
__global__ void k_static_hits(const double* __restrict__ pos, int S, int V, int N, double thr, StaticScratch sc) {
  const int words = (V + 31) / 32;
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long)S * V * words) return;
  const int jb = (int)(t % words);
  const long si = t / words;
  const int i = (int)(si % V), s = (int)(si / V);
  uint32_t m = 0;
  const int j0 = jb * 32, j1 = min(j0 + 32, V);
  if (j1 - 1 > i) {
    const double4* bb = sc.bbox + (size_t)s * V;
    const double4 bi = bb[i];
    const double* pi = pos + (size_t)si * 2 * (N + 1);
    for (int j = max(j0, i + 1); j < j1; j++) {
      const double4 bj = bb[j];
      const double gx = fmax(0.0, fmax(bi.x - bj.y, bj.x - bi.y)), gy = fmax(0.0, fmax(bi.z - bj.w, bj.z - bi.w));
      if (gx * gx + gy * gy < thr) {            // boxes closer than the threshold: exact test
        const double* pj = pos + ((size_t)s * V + j) * 2 * (N + 1);
        bool hit = false;
        for (int k = 0; k <= N && !hit; k++) {
          const double dx = pi[k] - pj[k], dy = pi[N + 1 + k] - pj[N + 1 + k];
          hit = dx * dx + dy * dy < thr;
        }
        if (hit) m |= 1u << (j - j0);
      }
    }
  }
  sc.hits[t] = m;
}

// one CTA per scenario: pairs (i, j > i) in row-major order (np.where(edge_mat == 1), casadi/main.py:121);
// edge_n[s] = number of colliding pairs, the first e_cap of them are stored
__global__ void k_static_emit(int V, int e_cap, StaticScratch sc, int32_t* edge_n, int32_t* edge_ij) {
  __shared__ int wsum[32];
  const int s = blockIdx.x, words = (V + 31) / 32, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  int base = 0;
  for (int i0 = 0; i0 < V; i0 += blockDim.x) {       // uniform trip count: barriers inside
    const int i = i0 + threadIdx.x;
    const uint32_t* row = sc.hits + ((size_t)s * V + (i < V ? i : 0)) * words;
    int cnt = 0;
    if (i < V) for (int w = i >> 5; w < words; w++) cnt += __popc(row[w]);
    int inc = cnt;
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(ST_FULL, inc, o); if (lane >= o) inc += v; }
    __syncthreads();                                  // wsum of the previous chunk has been read
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      int w = lane < nw ? wsum[lane] : 0;
      for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(ST_FULL, w, o); if (lane >= o) w += v; }
      wsum[lane] = w;
    }
    __syncthreads();
    int n = base + (warp ? wsum[warp - 1] : 0) + inc - cnt;
    base += wsum[nw - 1];
    if (cnt) {
      for (int w = i >> 5; w < words; w++) {
        uint32_t m = row[w];
        while (m) {
          const int j = w * 32 + __ffs(m) - 1;
          m &= m - 1;
          if (n < e_cap) { int32_t* e = edge_ij + ((size_t)s * e_cap + n) * 2; e[0] = i; e[1] = j; }
          n++;
        }
      }
    }
  }
  if (threadIdx.x == 0) edge_n[s] = base;
}

// address of the scratch array on every device, looked up once per device with that device current (in host code the
// symbol's name is a host-side stand-in, not a device address)
#define ST_MAX_DEVICES 64
static unsigned char* g_scratch_ptr[ST_MAX_DEVICES];
static const char* g_static_err = "";

extern "C" const char* piadmm_static_last_error(void) { return g_static_err; }

extern "C" int piadmm_edge_list_static(const double* pos, int S, int V, int N, double dis_thres, int thr_squared,
                                       int e_cap, int32_t* edge_n, int32_t* edge_ij, void* stream) {
  if (!pos || !edge_n || (!edge_ij && e_cap > 0) || S < 1 || V < 1 || N < 1 || N > 31 || e_cap < 0) {
    g_static_err = "piadmm_edge_list_static: bad argument (NULL buffer, S/V < 1, N outside 1..31 or e_cap < 0)";
    return PIADMM_E_ARG;
  }
  const long agents = (long)S * V, words = (V + 31) / 32;
  if (agents > ST_MAX_VEH || agents * words > ST_MAX_WORDS) {
    g_static_err = "piadmm_edge_list_static: scene larger than the static scratch (S*V <= PIADMM_STATIC_MAX_VEH and "
                   "S*V*ceil(V/32) <= PIADMM_STATIC_MAX_WORDS); use piadmm_detect_edges on a handle instead";
    return PIADMM_E_UNSUPPORTED;
  }
  int dev = -1;
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, pos) == cudaSuccess && at.type == cudaMemoryTypeDevice) dev = at.device;
  else cudaGetLastError();
  if (dev < 0 && cudaGetDevice(&dev) != cudaSuccess) { g_static_err = "piadmm_edge_list_static: no CUDA device"; return PIADMM_E_CUDA; }
  if (dev >= ST_MAX_DEVICES) { g_static_err = "piadmm_edge_list_static: device index out of range"; return PIADMM_E_UNSUPPORTED; }
  int prev = dev;
  cudaGetDevice(&prev);
  if (prev != dev && cudaSetDevice(dev) != cudaSuccess) { g_static_err = "piadmm_edge_list_static: cudaSetDevice failed"; return PIADMM_E_CUDA; }
  if (!g_scratch_ptr[dev]) {
    unsigned char* p = nullptr;
    const cudaError_t ce = cudaGetSymbolAddress((void**)&p, g_static_scratch);
    if (ce != cudaSuccess || !p) { g_static_err = cudaGetErrorString(ce); if (prev != dev) cudaSetDevice(prev); return PIADMM_E_CUDA; }
    g_scratch_ptr[dev] = p;
  }
  StaticScratch sc;
  sc.bbox = reinterpret_cast<double4*>(g_scratch_ptr[dev]);
  sc.hits = reinterpret_cast<uint32_t*>(g_scratch_ptr[dev] + ST_BBOX_BYTES);
  const cudaStream_t st = (cudaStream_t)stream;
  const double thr = thr_squared ? dis_thres * dis_thres : dis_thres;
  k_static_bbox<<<(unsigned)((agents * 32 + 255) / 256), 256, 0, st>>>(pos, agents, N, sc);
  k_static_hits<<<(unsigned)((agents * words + 127) / 128), 128, 0, st>>>(pos, S, V, N, thr, sc);
  const int threads = V >= 1024 ? 1024 : ((V + 31) / 32) * 32;
  k_static_emit<<<S, threads, 0, st>>>(V, e_cap, sc, edge_n, edge_ij);
  const cudaError_t ce = cudaGetLastError();
  if (prev != dev) cudaSetDevice(prev);            // the caller's current device is left as it was
  if (ce != cudaSuccess) { g_static_err = cudaGetErrorString(ce); return PIADMM_E_CUDA; }
  return PIADMM_OK;
}
