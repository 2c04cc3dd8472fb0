// d4b_krylov.cu -- device-resident outer Krylov loops around d4b_apply.
//
// Replaces, for device-resident vectors, the reference's in-house solvers
//   d4est_solver_cg_solve     src/Solver/d4est_solver_cg.c:75-197            (unpreconditioned CG; also the
//                             multigrid bottom solver, ..._bottom_solver_cg.c:48-200, same loop)
//   d4est_solver_fcg_solve    src/Solver/d4est_solver_fcg_improved.c:103-346 (flexible CG, truncation 1,
//                             after Notay & Napov) with the multigrid V-cycle as preconditioner
//                             (d4est_krylov_pc_multigrid_apply, d4est_krylov_pc_multigrid.c:41-79)
// Same recurrences, same stopping tests, hence the same iteration counts; the vectors never
// leave HBM and every iteration makes one fused pass over them.
#include <cmath>
#include <vector>
#include "d4b_internal.h"

namespace {

bool is_dev(const void *p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}
int grid_for(int64_t n) {
  int64_t b = (n + 255) / 256;
  return (int)(b < 148 * 8 ? (b > 0 ? b : 1) : 148 * 8);
}

// ---------------------------------------------------------------- CG ------------------
// The stopping test (delta_new > atol^2 + delta_0 rtol^2, d4est_solver_cg.c:134) is evaluated on
// the device: every iteration is enqueued behind a `done` flag the vector kernels honour, and
// the host only looks at the flag once per chunk of iterations.
struct CgScal { double delta_new, dAd, delta_0, alpha, beta, rr; int32_t done, iters; };
constexpr int kCgChunk = 16;

__global__ void cg_begin_k(CgScal *s, double atol, double rtol, int imax) {
  s->delta_new = s->rr; s->delta_0 = s->rr; s->iters = 0;
  s->done = !(imax > 0 && s->delta_new > atol * atol + s->delta_0 * rtol * rtol);
}
__global__ void cg_alpha_k(CgScal *s) { if (!s->done) s->alpha = s->delta_new / s->dAd; }
__global__ void cg_beta_k(CgScal *s, double atol, double rtol, int imax) {
  if (s->done) return;
  s->beta = s->rr / s->delta_new;
  s->delta_new = s->rr;
  s->iters++;
  // 2: finished by this iteration -- its d-update still runs, like the reference's
  if (!(s->iters < imax && s->delta_new > atol * atol + s->delta_0 * rtol * rtol)) s->done = 2;
}
__global__ void cg_init_k(const double *__restrict__ rhs, const double *__restrict__ Au, double *__restrict__ r,
                          double *__restrict__ d, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double v = rhs[i] - Au[i];
    r[i] = v; d[i] = v;
  }
}
__global__ void cg_step_k(const CgScal *__restrict__ s, const double *__restrict__ d, const double *__restrict__ Au,
                          double *__restrict__ u, double *__restrict__ r, int64_t n) {
  if (s->done) return;
  const double alpha = s->alpha;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    u[i] += alpha * d[i];
    r[i] -= alpha * Au[i];
  }
}
__global__ void cg_xpby_k(const CgScal *__restrict__ s, const double *__restrict__ r, double *__restrict__ d,
                          int64_t n) {
  if (s->done == 1) return;
  const double beta = s->beta;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    d[i] = r[i] + beta * d[i];
}
__global__ void cg_settle_k(CgScal *s) { if (s->done == 2) s->done = 1; }

// ---------------------------------------------------------------- FCG -----------------
constexpr int kDotBlocks = 256, kDotThreads = 256;

// partial sums of v.r, v.w, v.q, r.r in one pass (the four dots of d4est_solver_fcg_improved.c:232-252)
__global__ void __launch_bounds__(kDotThreads) fcg_dots_k(const double *__restrict__ v, const double *__restrict__ r,
                                                          const double *__restrict__ w, const double *__restrict__ q,
                                                          int with_q, int64_t n, double *__restrict__ part) {
  __shared__ double sh[4][kDotThreads];
  const int64_t chunk = (n + gridDim.x - 1) / gridDim.x;
  const int64_t lo = (int64_t)blockIdx.x * chunk, hi = lo + chunk < n ? lo + chunk : n;
  double s0 = 0., s1 = 0., s2 = 0., s3 = 0.;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kDotThreads) {
    const double vi = v[i], ri = r[i];
    s0 += vi * ri; s1 += vi * w[i]; s3 += ri * ri;
    if (with_q) s2 += vi * q[i];
  }
  sh[0][threadIdx.x] = s0; sh[1][threadIdx.x] = s1; sh[2][threadIdx.x] = s2; sh[3][threadIdx.x] = s3;
  __syncthreads();
  for (int wd = kDotThreads / 2; wd > 0; wd >>= 1) {
    if (threadIdx.x < wd)
      for (int c = 0; c < 4; ++c) sh[c][threadIdx.x] += sh[c][threadIdx.x + wd];
    __syncthreads();
  }
  if (threadIdx.x < 4) part[threadIdx.x * kDotBlocks + blockIdx.x] = sh[threadIdx.x][0];
}
__global__ void __launch_bounds__(kDotBlocks) fcg_dots_final_k(const double *__restrict__ part, double *__restrict__ out) {
  __shared__ double sh[kDotBlocks];
  for (int c = 0; c < 4; ++c) {
    sh[threadIdx.x] = part[c * kDotBlocks + threadIdx.x];
    __syncthreads();
    for (int wd = kDotBlocks / 2; wd > 0; wd >>= 1) {
      if (threadIdx.x < wd) sh[threadIdx.x] += sh[threadIdx.x + wd];
      __syncthreads();
    }
    if (threadIdx.x == 0) out[c] = sh[0];
    __syncthreads();
  }
}
// direction update, then u += c2 d ; r -= c2 q (d4est_solver_fcg_improved.c:258-283), one pass.
// mode 0: first iteration, d = v, q = w
// mode 1: what the reference executes for k > 0.  It writes d4est_linalg_vec_axpyeqz(c1, d, v, d)
//         (:260-262) and axpyeqz copies y into z before adding c1 x (d4est_linalg.c:217-221); with z
//         aliasing x the result is d = v + c1 v, q = w + c1 w.  Reproduced as is: parity with the
//         reference's iterates and iteration counts requires it.
// mode 2: the recurrence the reference's comments state, d = v + c1 d_{k-1}, q = w + c1 q_{k-1}
//         (D4B_FCG_TEXTBOOK)
__global__ void fcg_update_k(int mode, double c1, double c2, const double *__restrict__ v,
                             const double *__restrict__ w, double *__restrict__ d, double *__restrict__ q,
                             double *__restrict__ u, double *__restrict__ r, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double vi = v[i], wi = w[i];
    const double di = mode == 0 ? vi : c1 * (mode == 1 ? vi : d[i]) + vi;
    const double qi = mode == 0 ? wi : c1 * (mode == 1 ? wi : q[i]) + wi;
    d[i] = di; q[i] = qi;
    u[i] = c2 * di + u[i];
    r[i] = -c2 * qi + r[i];
  }
}
__global__ void resid_k(const double *__restrict__ rhs, const double *__restrict__ Au, double *__restrict__ r, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    r[i] = -1. * Au[i] + rhs[i];
}
__global__ void copy_k(const double *__restrict__ a, double *__restrict__ b, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    b[i] = a[i];
}

struct Scratch {   // device vectors of one solve
  d4b_mesh *m;
  std::vector<double *> v;
  double *get(int64_t n) { double *p = d4b_dev_alloc<double>(m, (size_t)n); v.push_back(p); return p; }
  ~Scratch() { for (double *p : v) cudaFree(p); }
};

}  // namespace

// CG on device vectors; d_scal: >= 8 doubles of device memory, h_scal: pinned.  r receives the
// recursively updated residual.  Returns the number of iterations.
int d4b_cg_run(d4b_mesh *m, double *u, const double *rhs, double *r, double *d, int imax, double rtol, double atol,
               double *d_scal, double *h_scal) {
  const int64_t n = m->local_nodes;
  const int gr = grid_for(n);
  double *Au = m->d_Au_stage;
  CgScal *s = reinterpret_cast<CgScal *>(d_scal);
  CgScal *h = reinterpret_cast<CgScal *>(h_scal);
  d4b_apply(m, u, Au);
  cg_init_k<<<gr, 256, 0, m->stream>>>(rhs, Au, r, d, n);
  d4b_dot_async(m, r, r, n, &s->rr);
  cg_begin_k<<<1, 1, 0, m->stream>>>(s, atol, rtol, imax);
  h->iters = 0;
  // iterations are enqueued in chunks between two looks at the done flag; the chunk grows 1, 2, 4, ... kCgChunk so that
  // a solve that converges in a few iterations (multigrid bottom solver) does not run a whole chunk of dead
  // applies, halo exchanges and reductions past its end
  for (int i0 = 0, chunk = 1; i0 < imax; i0 += chunk, chunk = chunk < kCgChunk ? 2 * chunk : kCgChunk) {
    for (int i = i0; i < imax && i < i0 + chunk; ++i) {
      d4b_apply(m, d, Au);
      d4b_dot_async(m, d, Au, n, &s->dAd);
      cg_alpha_k<<<1, 1, 0, m->stream>>>(s);
      cg_step_k<<<gr, 256, 0, m->stream>>>(s, d, Au, u, r, n);
      d4b_dot_async(m, r, r, n, &s->rr);
      cg_beta_k<<<1, 1, 0, m->stream>>>(s, atol, rtol, imax);
      cg_xpby_k<<<gr, 256, 0, m->stream>>>(s, r, d, n);
      cg_settle_k<<<1, 1, 0, m->stream>>>(s);
    }
    D4B_CUDA(cudaMemcpyAsync(h, s, sizeof(CgScal), cudaMemcpyDeviceToHost, m->stream));
    D4B_CUDA(cudaStreamSynchronize(m->stream));
    if (h->done) break;
  }
  return h->iters;
}

extern "C" int32_t d4b_cg_solve(d4b_mesh_t *m, double *u, const double *rhs, int32_t imax, double rtol, double atol) {
  D4B_CHECK(m && u && rhs, "d4b_cg_solve: NULL argument");
  const int64_t n = m->local_nodes;
  const size_t bytes = (size_t)n * sizeof(double);
  Scratch S{m};
  const bool u_dev = is_dev(u), rhs_dev = is_dev(rhs);
  double *du = u_dev ? u : S.get(n);
  const double *drhs = rhs;
  if (!rhs_dev) {
    double *t = S.get(n);
    D4B_CUDA(cudaMemcpyAsync(t, rhs, bytes, cudaMemcpyHostToDevice, m->stream));
    drhs = t;
  }
  if (!u_dev) D4B_CUDA(cudaMemcpyAsync(du, u, bytes, cudaMemcpyHostToDevice, m->stream));
  double *r = S.get(n), *d = S.get(n), *d_scal = S.get(16);
  const int its = d4b_cg_run(m, du, drhs, r, d, imax, rtol, atol, d_scal, m->h_red + 2048);
  if (!u_dev) D4B_CUDA(cudaMemcpyAsync(u, du, bytes, cudaMemcpyDeviceToHost, m->stream));
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  return its;
}

extern "C" int32_t d4b_fcg_solve(d4b_mesh_t *m, d4b_mg_t *pc, double *u, const double *rhs, int32_t imax,
                                 double rtol, double atol, int32_t flags, double *r2_history) {
  D4B_CHECK(m && u && rhs, "d4b_fcg_solve: NULL argument");
  D4B_CHECK(!pc || d4b_mg_top_mesh(pc) == m, "d4b_fcg_solve: the preconditioner's finest level is another mesh");
  const int64_t n = m->local_nodes;
  const size_t bytes = (size_t)n * sizeof(double);
  const int gr = grid_for(n);
  Scratch S{m};
  const bool u_dev = is_dev(u), rhs_dev = is_dev(rhs);
  double *du = u_dev ? u : S.get(n);
  const double *drhs = rhs;
  if (!rhs_dev) {
    double *t = S.get(n);
    D4B_CUDA(cudaMemcpyAsync(t, rhs, bytes, cudaMemcpyHostToDevice, m->stream));
    drhs = t;
  }
  if (!u_dev) D4B_CUDA(cudaMemcpyAsync(du, u, bytes, cudaMemcpyHostToDevice, m->stream));
  double *r = S.get(n), *w = S.get(n), *d = S.get(n), *v = S.get(n), *q = S.get(n);
  double *part = S.get(4 * kDotBlocks + 8), *dots = part + 4 * kDotBlocks;
  double *h = m->h_red + 2048;

  // r_0 = rhs - A u_0, tol = atol + rtol |r_0|   (:181-213)
  d4b_apply(m, du, w);
  resid_k<<<gr, 256, 0, m->stream>>>(drhs, w, r, n);
  const double r0 = d4b_vec_dot(m, r, r, n);
  const double tol = atol + rtol * sqrt(r0);
  double rho = 0.;
  int k = 0, its = 0;
  for (k = 0; k < imax; ++k) {
    // v_k = B(r_k): V-cycles from a zero guess, or the identity (:224-236)
    if (pc) {
      D4B_CUDA(cudaMemsetAsync(v, 0, bytes, m->stream));
      d4b_mg_solve(pc, v, r, nullptr);
    } else {
      copy_k<<<gr, 256, 0, m->stream>>>(r, v, n);
    }
    d4b_apply(m, v, w);
    fcg_dots_k<<<kDotBlocks, kDotThreads, 0, m->stream>>>(v, r, w, q, k > 0, n, part);
    fcg_dots_final_k<<<1, kDotBlocks, 0, m->stream>>>(part, dots);
    d4b_comm_allreduce_sum(m, dots, 4);
    D4B_CUDA(cudaMemcpyAsync(h, dots, 4 * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    D4B_CUDA(cudaStreamSynchronize(m->stream));
    const double alpha = h[0], beta = h[1], gamma = h[2], rr = h[3];
    if (r2_history) r2_history[k] = rr;
    ++its;
    double c1 = 0.;
    if (k > 0) {
      c1 = -gamma / rho;
      rho = beta - (gamma * gamma) / rho;
    } else {
      rho = beta;
    }
    const int mode = k == 0 ? 0 : ((flags & D4B_FCG_TEXTBOOK) ? 2 : 1);
    fcg_update_k<<<gr, 256, 0, m->stream>>>(mode, c1, alpha / rho, v, w, d, q, du, r, n);
    // the test looks at |r_k|, computed before this iteration's update (:285)
    if (k > 0 && sqrt(rr) <= tol) break;
  }
  if (!u_dev) D4B_CUDA(cudaMemcpyAsync(u, du, bytes, cudaMemcpyDeviceToHost, m->stream));
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  return its;
}
