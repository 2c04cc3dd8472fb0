// d4b_static.cu -- d4b_vec_dot_noalloc: the dot product of d4b_solver.cu (d4b_vec_dot) for
// callers that must not allocate: code under CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc.  No counterpart in the
// reference.
//
// d4b_vec_dot keeps the per-block partial sums of its two-pass reduction in a cudaMalloc'ed
// buffer of the mesh (d4b_mesh::d_red).  Here the partial sums live in a file-scope device array
// of fixed size, the caller passes inputs, output and stream, and nothing on the path calls
// cudaMalloc / cudaFree / an allocating library routine.  Same grid, same block size, same
// reduction trees as dot_partial / dot_final of d4b_solver.cu, so one vector pair gives the
// bit-identical sum (single rank: there is no allreduce here).
//
// Limits and hazards (also in include/d4est_b200.h):
//   * at most D4B_DOT_NOALLOC_MAX_VECS vector pairs per call (scratch = MAX_VECS x 592 doubles
//     = 303 104 bytes, held by every GPU that loads the library until the process exits);
//   * every call on a device shares that one array: calls that overlap on the GPU (different
//     streams, or graphs replayed concurrently) race.  Serialise them.
//   * nothing has to start at zero: every block of the first pass overwrites its partial sum
//     (an empty chunk writes 0.0), and the second pass reads only what this call wrote.
#include <cstdarg>
#include <cstdio>
#include <mutex>
#include <cuda_runtime.h>
#include "d4est_b200.h"
#include "d4b_reduce_cfg.h"

namespace {

constexpr int kBlocks = kD4bRedBlocks;     // the geometry of d4b_vec_dot (d4b_reduce_cfg.h)
constexpr int kThreads = kD4bRedThreads;
constexpr int kMaxDevices = 64;

// The scratch.  In host code this name is only a stand-in for the symbol: its device address
// comes from cudaGetSymbolAddress (scratch_for_current_device), never from &g_dot_scratch.
static __device__ __align__(256) double g_dot_scratch[(size_t)D4B_DOT_NOALLOC_MAX_VECS * kBlocks];

// pass 1: block (b, v) reduces chunk b of vector pair v into part[v * kBlocks + b]
__global__ void __launch_bounds__(kThreads) dot_partial_static(const double *__restrict__ x, const double *__restrict__ y,
                                                               int64_t n, double *__restrict__ part) {
  __shared__ double sh[kThreads];
  const double *xv = x + (int64_t)blockIdx.y * n, *yv = y + (int64_t)blockIdx.y * n;
  const int64_t chunk = (n + gridDim.x - 1) / gridDim.x;
  const int64_t lo = (int64_t)blockIdx.x * chunk, hi = lo + chunk < n ? lo + chunk : n;
  double s = 0.;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kThreads) s += xv[i] * yv[i];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int w = kThreads / 2; w > 0; w >>= 1) {
    if (threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) part[(size_t)blockIdx.y * kBlocks + blockIdx.x] = sh[0];
}

// pass 2: block v reduces the kBlocks partial sums of vector pair v
__global__ void __launch_bounds__(1024) dot_final_static(const double *__restrict__ part, double *__restrict__ out) {
  __shared__ double sh[1024];
  sh[threadIdx.x] = (threadIdx.x < kBlocks) ? part[(size_t)blockIdx.x * kBlocks + threadIdx.x] : 0.;
  __syncthreads();
  for (int w = 512; w > 0; w >>= 1) {
    if (threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[blockIdx.x] = sh[0];
}

std::mutex g_mu;
double *g_scratch_addr[kMaxDevices] = {};   // device address of g_dot_scratch, per device

int fail(char *err, int64_t err_len, int code, const char *fmt, ...) {
  if (err && err_len > 0) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err, (size_t)err_len, fmt, ap);
    va_end(ap);
  }
  return code;
}

// set-up, once per device, with that device current
int scratch_for_current_device(double **out, char *err, int64_t err_len) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess)
    return fail(err, err_len, D4B_NOALLOC_E_CUDA, "d4b_vec_dot_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices)
    return fail(err, err_len, D4B_NOALLOC_E_CUDA, "d4b_vec_dot_noalloc: device ordinal %d outside 0..%d", dev, kMaxDevices - 1);
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_addr[dev]) {
    double *p = nullptr;
    e = cudaGetSymbolAddress((void **)&p, g_dot_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return fail(err, err_len, D4B_NOALLOC_E_CUDA, "d4b_vec_dot_noalloc: cudaGetSymbolAddress on device %d: %s", dev,
                  cudaGetErrorString(e));
    }
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return D4B_NOALLOC_OK;
}

}  // namespace

extern "C" int32_t d4b_vec_dot_noalloc(const double *x, const double *y, int64_t n, int32_t nvec, double *out,
                                       void *cuda_stream, char *err, int64_t err_len) {
  if (err && err_len > 0) err[0] = '\0';
  if (nvec < 0 || n < 0)
    return fail(err, err_len, D4B_NOALLOC_E_ARG, "d4b_vec_dot_noalloc: negative size (n = %lld, nvec = %d)", (long long)n,
                (int)nvec);
  if (nvec > D4B_DOT_NOALLOC_MAX_VECS)
    return fail(err, err_len, D4B_NOALLOC_E_TOO_LARGE,
                "d4b_vec_dot_noalloc: nvec = %d exceeds D4B_DOT_NOALLOC_MAX_VECS = %d (the static scratch holds %d x %d "
                "partial sums); split the batch or use d4b_vec_dot",
                (int)nvec, D4B_DOT_NOALLOC_MAX_VECS, D4B_DOT_NOALLOC_MAX_VECS, kBlocks);
  double *scratch = nullptr;
  const int rc = scratch_for_current_device(&scratch, err, err_len);
  if (rc != D4B_NOALLOC_OK) return rc;
  if (nvec == 0) return D4B_NOALLOC_OK;                 // set-up only (call once before a capture)
  if (!out || (n > 0 && (!x || !y)))
    return fail(err, err_len, D4B_NOALLOC_E_ARG, "d4b_vec_dot_noalloc: NULL pointer (x = %p, y = %p, out = %p)", (const void *)x,
                (const void *)y, (void *)out);
  cudaStream_t st = (cudaStream_t)cuda_stream;
  dot_partial_static<<<dim3(kBlocks, nvec), kThreads, 0, st>>>(x, y, n, scratch);
  const cudaError_t e1 = cudaGetLastError();
  if (e1 != cudaSuccess)   // nothing else has been enqueued
    return fail(err, err_len, D4B_NOALLOC_E_CUDA, "d4b_vec_dot_noalloc: kernel launch: %s", cudaGetErrorString(e1));
  dot_final_static<<<nvec, 1024, 0, st>>>(scratch, out);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess)
    return fail(err, err_len, D4B_NOALLOC_E_CUDA, "d4b_vec_dot_noalloc: kernel launch: %s", cudaGetErrorString(e));
  return D4B_NOALLOC_OK;
}
