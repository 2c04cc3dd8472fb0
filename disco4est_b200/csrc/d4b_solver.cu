// d4b_solver.cu -- device-resident BLAS-1, cg_eigs and the Chebyshev smoother.
//
// Replaces, for device-resident vectors:
//   d4est_linalg_vec_axpy/xpby/scale/dot   src/LinearAlgebra/d4est_linalg.c:179-236
//   cg_eigs                                src/Solver/d4est_solver_cg_eigs.c:115-276
//   ..._smoother_cheby_iterate_aux         src/Solver/d4est_solver_multigrid_smoother_cheby.c:81-176
// Reductions are deterministic: fixed grid, fixed-shape tree, no atomics.
#include <cmath>
#include <algorithm>
#include <cstdlib>
#include <vector>
#include "d4b_internal.h"
#include "d4b_reduce_cfg.h"

namespace {

constexpr int kRedBlocks = kD4bRedBlocks, kRedThreads = kD4bRedThreads;

__global__ void __launch_bounds__(kRedThreads) dot_partial(const double *__restrict__ x,
                                                           const double *__restrict__ y, int64_t n,
                                                           double *__restrict__ part) {
  __shared__ double sh[kRedThreads];
  int64_t chunk = (n + gridDim.x - 1) / gridDim.x;
  int64_t lo = (int64_t)blockIdx.x * chunk, hi = lo + chunk < n ? lo + chunk : n;
  double s = 0.;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kRedThreads) s += x[i] * y[i];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int w = kRedThreads / 2; w > 0; w >>= 1) {
    if (threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) part[blockIdx.x] = sh[0];
}

__global__ void __launch_bounds__(1024) dot_final(const double *__restrict__ part, int np,
                                                   double *__restrict__ out) {
  __shared__ double sh[1024];
  sh[threadIdx.x] = (threadIdx.x < np) ? part[threadIdx.x] : 0.;
  __syncthreads();
  for (int w = 512; w > 0; w >>= 1) {
    if (threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = sh[0];
}

// sum of n partial sums (the energy dot of the apply kernels): stage 1 of 2 for large n
__global__ void __launch_bounds__(kRedThreads) sum_partial(const double *__restrict__ x, int64_t n,
                                                           double *__restrict__ part) {
  __shared__ double sh[kRedThreads];
  int64_t chunk = (n + gridDim.x - 1) / gridDim.x;
  int64_t lo = (int64_t)blockIdx.x * chunk, hi = lo + chunk < n ? lo + chunk : n;
  double s = 0.;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kRedThreads) s += x[i];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int w = kRedThreads / 2; w > 0; w >>= 1) {
    if (threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) part[blockIdx.x] = sh[0];
}

__global__ void axpy_k(double a, const double *__restrict__ x, double *__restrict__ y, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] += a * x[i];
}
// y = x + b y   (d4est_linalg_vec_xpby, d4est_linalg.c:200-215)
__global__ void xpby_k(const double *__restrict__ x, double b, double *__restrict__ y, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] = x[i] + b * y[i];
}
__global__ void scale_k(double a, double *__restrict__ x, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    x[i] *= a;
}
// r = rhs - Au
__global__ void residual_k(const double *__restrict__ rhs, const double *__restrict__ Au,
                           double *__restrict__ r, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    r[i] = rhs[i] - Au[i];
}
// r = rhs - Au ; d = r   (cg_eigs start)
__global__ void residual2_k(const double *__restrict__ rhs, const double *__restrict__ Au,
                            double *__restrict__ r, double *__restrict__ d, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    double v = rhs[i] - Au[i];
    r[i] = v; d[i] = v;
  }
}
// Chebyshev update: p = alpha (rhs - Au) + beta p ; u += p
__global__ void cheby_update_k(double alpha, double beta, const double *__restrict__ rhs,
                               const double *__restrict__ Au, double *__restrict__ p,
                               double *__restrict__ u, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    double r = (rhs[i] - Au[i]) * alpha;
    double pn = r + beta * p[i];
    p[i] = pn;
    u[i] += pn;
  }
}

inline int grid_for(int64_t n) {
  int64_t b = (n + 255) / 256;
  return (int)(b < 148 * 8 ? (b > 0 ? b : 1) : 148 * 8);
}

double *work(d4b_mesh *m, int i) {
  if (!m->d_work[i]) m->d_work[i] = d4b_dev_alloc<double>(m, (size_t)m->local_nodes);
  return m->d_work[i];
}

bool is_dev(const void *p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

}  // namespace

extern "C" double d4b_vec_dot(d4b_mesh_t *m, const double *x, const double *y, int64_t n) {
  dot_partial<<<kRedBlocks, kRedThreads, 0, m->stream>>>(x, y, n, m->d_red);
  dot_final<<<1, 1024, 0, m->stream>>>(m->d_red, kRedBlocks, m->d_red + kRedBlocks);
  d4b_comm_allreduce_sum(m, m->d_red + kRedBlocks, 1);   // sc_allreduce of the reference; no-op on one rank
  D4B_CUDA(cudaMemcpyAsync(m->h_red, m->d_red + kRedBlocks, sizeof(double), cudaMemcpyDeviceToHost, m->stream));
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  return m->h_red[0];
}
extern "C" void d4b_vec_axpy(d4b_mesh_t *m, double a, const double *x, double *y, int64_t n) {
  axpy_k<<<grid_for(n), 256, 0, m->stream>>>(a, x, y, n);
}
extern "C" void d4b_vec_xpby(d4b_mesh_t *m, const double *x, double b, double *y, int64_t n) {
  xpby_k<<<grid_for(n), 256, 0, m->stream>>>(x, b, y, n);
}
extern "C" void d4b_vec_scale(d4b_mesh_t *m, double a, double *x, int64_t n) {
  scale_k<<<grid_for(n), 256, 0, m->stream>>>(a, x, n);
}

namespace {

// Gershgorin bounds of the Lanczos tridiagonal, d4est_solver_cg_eigs.c:9-65.  Runs on the
// device (one thread) so that the estimate never has to visit the host; the explicitly
// rounded intrinsics keep nvcc from contracting into FMAs, i.e. the result is the one the
// reference's scalar code produces from the same dot products.
__device__ double gershgorin(int use_new, int i, int64_t n, double a0, double b0, double a1, double b1) {
  double diag, off;
  if (!use_new) {
    if (i != 0 && i < n - 1) { diag = __dadd_rn(1. / a1, b0 / a0); off = __dadd_rn(fabs(sqrt(b1) / a1), fabs(sqrt(b0) / a0)); }
    else if (i == 0) { diag = 1. / a1; off = sqrt(b1) / a1; }
    else { diag = __dadd_rn(1. / a1, b0 / a0); off = fabs(sqrt(b0) / a0); }
  } else {
    if (i != 0) { diag = __dadd_rn(1. / a1, b0 / a0); off = fabs(sqrt(b0) / a0); }
    else { diag = 1. / a1; off = sqrt(b1) / a1; }
  }
  return __dadd_rn(diag, off);
}

// d_red layout: [0,1024) partial sums | delta[kMaxCgIter+1] | dAd[kMaxCgIter+1] | eig | coef[2*kMaxCheby]
constexpr int kScalOff = 1024, kMaxCgIter = 500;
constexpr int kEigOff = kScalOff + 2 * (kMaxCgIter + 1) + 2, kCoefOff = kEigOff + 2, kMaxCheby = 500;
static_assert(kCoefOff + 2 * kMaxCheby <= 4096, "d_red is 4096 doubles (d4b_mesh.cu)");

// One thread: (imax > 0) the bound of cg_eigs from its CG scalars, times `mult`, into *eig;
// then the Chebyshev coefficients {alpha_i, beta_i} for the spectrum [*eig / ratio, *eig]
// (d4est_solver_multigrid_smoother_cheby.c:100-128) into coef[2 i], coef[2 i + 1].
__global__ void eig_plan_k(const double *__restrict__ delta, const double *__restrict__ dAd, int imax, int use_new,
                           int64_t n, double mult, double ratio, int iter, double *__restrict__ eig,
                           double *__restrict__ coef) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  if (imax > 0) {
    double alpha = -1., beta = -1., bound = 0.;
    for (int i = 0; i < imax; ++i) {
      const double alpha_old = alpha, beta_old = beta;
      alpha = delta[i] / dAd[i];
      beta = delta[i + 1] / delta[i];
      const double t = gershgorin(use_new, i, n, alpha_old, beta_old, alpha, beta);
      bound = (i > 0) ? fmax(bound, t) : t;
    }
    *eig = __dmul_rn(bound, mult);
  }
  const double lmax = *eig, lmin = lmax / ratio;
  const double d = __dmul_rn(__dadd_rn(lmax, lmin), .5), c = __dmul_rn(__dadd_rn(lmax, -lmin), .5);
  const double cc = __dmul_rn(c, c);
  double alpha = 0.;
  for (int i = 0; i < iter; ++i) {
    if (i == 0) alpha = 1. / d;
    else if (i == 1) alpha = __dmul_rn(2., d) / __dadd_rn(__dmul_rn(__dmul_rn(2., d), d), -cc);
    else alpha = 1. / __dadd_rn(d, -(__dmul_rn(alpha, cc) / 4.));
    coef[2 * i] = alpha;
    coef[2 * i + 1] = __dadd_rn(__dmul_rn(alpha, d), -1.);
  }
}

void dot_to(d4b_mesh *m, const double *x, const double *y, int64_t n, double *d_out) {
  dot_partial<<<kRedBlocks, kRedThreads, 0, m->stream>>>(x, y, n, m->d_red);
  dot_final<<<1, 1024, 0, m->stream>>>(m->d_red, kRedBlocks, d_out);
  d4b_comm_allreduce_sum(m, d_out, 1);
}

// u += alpha d ; r -= alpha Ad with alpha = *delta / *dAd ; block partials of r.r
// (six streams, 48 B per node: 6 blocks of 256 threads per SM and two independent nodes per thread and trip
//  keep enough loads in flight; kUpdBlocks <= 1024 = what dot_final sums)
constexpr int kUpdBlocks = 888;
__global__ void __launch_bounds__(kRedThreads) cg_update_dot_k(const double *__restrict__ delta,
                                                               const double *__restrict__ dAd,
                                                               const double *__restrict__ d,
                                                               const double *__restrict__ Ad, double *__restrict__ u,
                                                               double *__restrict__ r, int64_t n,
                                                               double *__restrict__ part) {
  __shared__ double sh[kRedThreads];
  const double alpha = *delta / *dAd;
  int64_t chunk = (n + gridDim.x - 1) / gridDim.x;
  int64_t lo = (int64_t)blockIdx.x * chunk, hi = lo + chunk < n ? lo + chunk : n;
  double s0 = 0., s1 = 0.;
  int64_t i = lo + threadIdx.x;
  for (; i + kRedThreads < hi; i += 2 * kRedThreads) {
    const int64_t j = i + kRedThreads;
    const double di = d[i], dj = d[j], ai = Ad[i], aj = Ad[j], ui = u[i], uj = u[j], ri0 = r[i], rj0 = r[j];
    u[i] = ui + alpha * di;
    u[j] = uj + alpha * dj;
    const double ri = ri0 - alpha * ai, rj = rj0 - alpha * aj;
    r[i] = ri;
    r[j] = rj;
    s0 += ri * ri;
    s1 += rj * rj;
  }
  if (i < hi) {
    u[i] += alpha * d[i];
    const double ri = r[i] - alpha * Ad[i];
    r[i] = ri;
    s0 += ri * ri;
  }
  sh[threadIdx.x] = s0 + s1;
  __syncthreads();
  for (int w = kRedThreads / 2; w > 0; w >>= 1) {
    if (threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) part[blockIdx.x] = sh[0];
}

// d = r + beta d with beta = *delta_new / *delta_old
__global__ void xpby_dev_k(const double *__restrict__ r, const double *__restrict__ delta_new,
                           const double *__restrict__ delta_old, double *__restrict__ d, int64_t n) {
  const double beta = *delta_new / *delta_old;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    d[i] = r[i] + beta * d[i];
}
// Chebyshev update with the coefficients in device memory
__global__ void cheby_update_dev_k(const double *__restrict__ coef, const double *__restrict__ rhs,
                                   const double *__restrict__ Au, double *__restrict__ p,
                                   double *__restrict__ u, int64_t n) {
  const double alpha = coef[0], beta = coef[1];
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    double r = (rhs[i] - Au[i]) * alpha;
    double pn = r + beta * p[i];
    p[i] = pn;
    u[i] += pn;
  }
}

// the partial sums epi.dot_part collected during the last apply -> one device scalar (all-reduced)
void energy_dot_finish(d4b_mesh *m, double *d_out) {
  const int64_t n = m->dot_count;
  if (n <= 1024) {
    dot_final<<<1, 1024, 0, m->stream>>>(m->d_dot_part, (int)n, d_out);
  } else {
    const int blocks = (int)std::min<int64_t>((n + 4 * kRedThreads - 1) / (4 * kRedThreads), 256);
    sum_partial<<<blocks, kRedThreads, 0, m->stream>>>(m->d_dot_part, n, m->d_red);
    dot_final<<<1, 1024, 0, m->stream>>>(m->d_red, blocks, d_out);
  }
  d4b_comm_allreduce_sum(m, d_out, 1);
}

void energy_dot_reserve(d4b_mesh *m) {
  if (m->d_dot_part) return;
  m->dot_capacity = 2 * (int64_t)m->n_local + 64;   // at most two warps per element, one launch tail per class and part
  for (const auto &c : m->fast) m->dot_capacity += 8;
  m->d_dot_part = d4b_dev_alloc<double>(m, (size_t)m->dot_capacity);
}

// D4B_FUSED_SOLVERS=0: cg_eigs and the smoother as separate apply + vector kernels
bool fused_solvers_allowed(const d4b_mesh *m) {
  const char *e = getenv("D4B_FUSED_SOLVERS");
  return m->all_flat && !m->pre_apply && !m->d_W && !(e && e[0] == '0');
}

// cg_eigs on device vectors, enqueued without a host round trip: kernel i reads
// alpha_i = delta_i / (d.Ad)_i and beta_i = delta_{i+1} / delta_i from the slots earlier
// reductions filled.  Leaves delta[0..imax], dAd[0..imax) in d_red.
//
// On meshes whose elements all run in the tangential-free kernel an iteration is three passes instead of six:
//   interpolation kernel   d = r + beta d (written back), dh = (I x I x I) d
//   apply kernel           Ad, and d.Ad in its energy form from the registers of the kernel (D4bEpilogue::dot_part)
//   cg_update_dot_k        u += alpha d, r -= alpha Ad, r.r
// (the same recurrences as d4est_solver_cg_eigs.c:181-257; d.Ad differs from the dot product of the two vectors
//  in the last bits only).
void cg_eigs_enqueue(d4b_mesh *m, double *du, const double *drhs, int imax, int use_zero_guess) {
  D4B_CHECK(imax >= 1 && imax <= kMaxCgIter, "cg_eigs: imax %d out of range", imax);
  const int64_t n = m->local_nodes;
  if (use_zero_guess) {
    // the reference iterates on a scratch zero vector and leaves u untouched
    // (d4est_solver_multigrid_smoother_cheby.c:281-305)
    du = work(m, 3);
    D4B_CUDA(cudaMemsetAsync(du, 0, (size_t)n * sizeof(double), m->stream));
  }
  double *Au = m->d_Au_stage, *d = work(m, 0), *r = work(m, 1);
  double *delta = m->d_red + kScalOff, *dAd = delta + kMaxCgIter + 1;
  const int g = grid_for(n);
  const bool fused = fused_solvers_allowed(m);
  // the packed halo (NCCL send/recv of nodal values) is taken from d before the interpolation kernel runs:
  // the direction update can only ride on that kernel when no nodal values are sent
  const bool fuse_xpby = fused && (m->mpisize == 1 || d4b_comm_peer_ready(m));
  if (fused) energy_dot_reserve(m);
  m->epi = D4bEpilogue();
  d4b_apply(m, du, Au);
  residual2_k<<<g, 256, 0, m->stream>>>(drhs, Au, r, d, n);
  dot_to(m, r, r, n, delta);
  for (int i = 0; i < imax; ++i) {
    if (fused) {
      m->epi = D4bEpilogue();
      m->epi.dot_part = m->d_dot_part;
      m->dot_count = 0;
      if (i > 0) {
        if (fuse_xpby) { m->epi.xpby_r = r; m->epi.xpby_num = delta + i; m->epi.xpby_den = delta + i - 1; }
        else xpby_dev_k<<<g, 256, 0, m->stream>>>(r, delta + i, delta + i - 1, d, n);
      }
      d4b_apply_device(m, d, Au);
      m->epi = D4bEpilogue();
      energy_dot_finish(m, dAd + i);
    } else {
      d4b_apply(m, d, Au);
      dot_to(m, d, Au, n, dAd + i);
    }
    cg_update_dot_k<<<kUpdBlocks, kRedThreads, 0, m->stream>>>(delta + i, dAd + i, d, Au, du, r, n, m->d_red);
    dot_final<<<1, 1024, 0, m->stream>>>(m->d_red, kUpdBlocks, delta + i + 1);
    d4b_comm_allreduce_sum(m, delta + i + 1, 1);
    // (the reference also forms the next direction after its last iteration; nothing reads it)
    if (!fused) xpby_dev_k<<<g, 256, 0, m->stream>>>(r, delta + i + 1, delta + i, d, n);
  }
}

// Chebyshev recurrence on device vectors.  Coefficients: host arrays (alpha/beta as kernel
// arguments) or, when d_coef is given, pairs in device memory.
//
// Three forms, fastest applicable first:
//  (a) Gauss-resident (one-rank meshes whose elements all run in the tangential-free kernel): the recurrence
//      is carried in the Gauss basis, uh <- uh + ph, ph <- alpha (I3 rhs - (B x B x B) rh) + beta ph with B = I I^T,
//      so a sweep is ONE kernel (no interpolation kernel, no nodal vectors); I3 rhs before the first sweep,
//      u = I3^-1 uh after the last, and the closing residual apply starts from the uh already there.
//      Same numbers as (b) up to round-off of the basis change.
//  (b) the recurrence in the apply's epilogue on nodal vectors (every element in a fast kernel)
//  (c) apply + one vector kernel
void cheby_enqueue(d4b_mesh *m, double *du, const double *drhs, double *dr, int iter, const double *h_alpha,
                   const double *h_beta, const double *d_coef) {
  const int64_t n = m->local_nodes;
  const int g = grid_for(n);
  double *Au = m->d_Au_stage, *p = work(m, 0);
  D4B_CUDA(cudaMemsetAsync(p, 0, (size_t)n * sizeof(double), m->stream));
  // every element in a fast kernel: the recurrence runs in the apply's epilogue
  // (u updated in place, no Au vector); otherwise apply + one vector kernel
  const char *nf = getenv("D4B_NO_FUSED_CHEBY");
  const bool fused = m->n_generic == 0 && !m->fast.empty() && !(nf && nf[0] == '1');
  // (multi-rank: only with the peer-memory halo -- the neighbours read each other's uh, nodal values never exist)
  const bool multi = m->mpisize > 1;
  const bool hat = fused && iter > 0 && fused_solvers_allowed(m) && (!multi || d4b_comm_peer_ready(m));
  if (hat) {
    if (!m->d_rhs_hat) m->d_rhs_hat = d4b_dev_alloc<double>(m, (size_t)n);
    if (!multi && !m->d_uhat_alt) m->d_uhat_alt = d4b_dev_alloc<double>(m, (size_t)n);
    double *const uh_own = m->d_uhat;
    // one rank: two private buffers.  Peer halo: the two copies of uhat the neighbours have mapped; every sweep is
    // one epoch of the flag protocol (signal = "the copy of this epoch is complete", sent when the previous sweep is)
    if (multi) d4b_comm_peer_begin(m);
    double *cur = m->d_uhat;
    d4b_launch_basis_change(m, du, cur, true);
    d4b_launch_basis_change(m, drhs, m->d_rhs_hat, true);
    for (int i = 0; i < iter; ++i) {
      double *nxt = multi ? d4b_comm_peer_next_copy(m) : (cur == uh_own ? m->d_uhat_alt : uh_own);
      m->epi = D4bEpilogue();
      m->epi.mode = 3; m->epi.rhs = m->d_rhs_hat; m->epi.p = p; m->epi.uh_out = nxt;
      if (d_coef) m->epi.coef = d_coef + 2 * i;
      else { m->epi.alpha = h_alpha[i]; m->epi.beta = h_beta[i]; }
      m->d_uhat = cur;
      d4b_apply_device(m, nullptr, nullptr);   // the apply kernels only: uh is already there
      if (multi) { d4b_comm_peer_begin(m); cur = m->d_uhat; }
      else cur = nxt;
    }
    d4b_launch_basis_change(m, cur, du, false);
    // r = rhs - A u from the uh of the last sweep
    m->epi = D4bEpilogue();
    m->epi.mode = 2; m->epi.rhs = drhs;
    m->d_uhat = cur;
    d4b_apply_device(m, nullptr, dr);
    m->epi = D4bEpilogue();
    if (!multi) m->d_uhat = uh_own;
    return;
  }
  for (int i = 0; i < iter; ++i) {
    if (fused) {
      m->epi = D4bEpilogue();
      m->epi.mode = 1; m->epi.rhs = drhs; m->epi.p = p; m->epi.u = du;
      if (d_coef) m->epi.coef = d_coef + 2 * i;
      else { m->epi.alpha = h_alpha[i]; m->epi.beta = h_beta[i]; }
      d4b_apply_device(m, du, nullptr);
    } else {
      d4b_apply(m, du, Au);
      if (d_coef) cheby_update_dev_k<<<g, 256, 0, m->stream>>>(d_coef + 2 * i, drhs, Au, p, du, n);
      else cheby_update_k<<<g, 256, 0, m->stream>>>(h_alpha[i], h_beta[i], drhs, Au, p, du, n);
    }
  }
  if (fused) {
    m->epi = D4bEpilogue();
    m->epi.mode = 2; m->epi.rhs = drhs;
    d4b_apply_device(m, du, dr);
    m->epi = D4bEpilogue();
  } else {
    d4b_apply(m, du, Au);
    residual_k<<<g, 256, 0, m->stream>>>(drhs, Au, dr, n);
  }
}

}  // namespace

// One smoothing of a multigrid level (d4est_solver_multigrid_smoother_cheby.c:264-376) on
// device vectors, stream-ordered from the first kernel to the last: the spectrum estimate,
// its multiplier and the recurrence coefficients stay in device memory (m->d_red).
void d4b_smooth_async(d4b_mesh *m, double *du, const double *drhs, double *dr, const d4b_cheby_params_t &c,
                      bool compute_eigs, double *d_eig) {
  D4B_CHECK(c.cheby_imax >= 0 && c.cheby_imax <= kMaxCheby, "cheby_imax %d out of range", c.cheby_imax);
  // with a zero guess the estimate runs on a scratch vector and u is untouched (:281-305)
  if (compute_eigs) cg_eigs_enqueue(m, du, drhs, c.cheby_eigs_cg_imax, c.cheby_use_zero_guess_for_eigs);
  // (:322-349 re-runs cg_eigs on a zero vector into a dummy when the estimate is
  //  re-used with the zero-guess option; it changes no state and is skipped here)
  double *delta = m->d_red + kScalOff, *dAd = delta + kMaxCgIter + 1;
  eig_plan_k<<<1, 32, 0, m->stream>>>(delta, dAd, compute_eigs ? c.cheby_eigs_cg_imax : 0, c.cheby_use_new_cg_eigs,
                                      m->local_nodes, c.cheby_eigs_max_multiplier, c.cheby_eigs_lmax_lmin_ratio,
                                      c.cheby_imax, d_eig, m->d_red + kCoefOff);
  cheby_enqueue(m, du, drhs, dr, c.cheby_imax, nullptr, nullptr, m->d_red + kCoefOff);
}

// global dot product into a device scalar, stream-ordered
void d4b_dot_async(d4b_mesh *m, const double *x, const double *y, int64_t n, double *d_out) { dot_to(m, x, y, n, d_out); }

static double d4b_smooth_eig(d4b_mesh *m) {
  D4B_CUDA(cudaMemcpyAsync(m->h_red, m->d_red + kEigOff, sizeof(double), cudaMemcpyDeviceToHost, m->stream));
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  return m->h_red[0];
}

extern "C" double d4b_cg_eigs(d4b_mesh_t *m, double *u, const double *rhs, int32_t imax,
                              int32_t use_new, int32_t use_zero_guess) {
  d4b_check_device(m);
  const int64_t n = m->local_nodes;
  const size_t bytes = (size_t)n * sizeof(double);
  bool u_dev = is_dev(u), rhs_dev = is_dev(rhs);
  double *du = u_dev ? u : work(m, 2);
  const double *drhs = rhs;
  if (!rhs_dev) {
    D4B_CUDA(cudaMemcpyAsync(m->d_u_stage, rhs, bytes, cudaMemcpyHostToDevice, m->stream));
    drhs = m->d_u_stage;
  }
  if (!u_dev) D4B_CUDA(cudaMemcpyAsync(du, u, bytes, cudaMemcpyHostToDevice, m->stream));
  cg_eigs_enqueue(m, du, drhs, imax, use_zero_guess);
  double *delta = m->d_red + kScalOff, *dAd = delta + kMaxCgIter + 1;
  eig_plan_k<<<1, 32, 0, m->stream>>>(delta, dAd, imax, use_new, n, 1., 1., 0, m->d_red + kEigOff,
                                      m->d_red + kCoefOff);
  if (!u_dev && !use_zero_guess)
    D4B_CUDA(cudaMemcpyAsync(u, du, bytes, cudaMemcpyDeviceToHost, m->stream));
  return d4b_smooth_eig(m);
}

extern "C" void d4b_cheby_smooth(d4b_mesh_t *m, double *u, const double *rhs, double *r, int32_t iter,
                                 double lmin, double lmax) {
  D4B_CHECK(iter >= 0 && iter <= kMaxCheby, "d4b_cheby_smooth: iter %d out of range", iter);
  d4b_check_device(m);
  const int64_t n = m->local_nodes;
  const size_t bytes = (size_t)n * sizeof(double);
  bool u_dev = is_dev(u), rhs_dev = is_dev(rhs), r_dev = is_dev(r);
  double *du = u_dev ? u : work(m, 2);
  const double *drhs = rhs;
  if (!rhs_dev) {
    D4B_CUDA(cudaMemcpyAsync(m->d_u_stage, rhs, bytes, cudaMemcpyHostToDevice, m->stream));
    drhs = m->d_u_stage;
  }
  if (!u_dev) D4B_CUDA(cudaMemcpyAsync(du, u, bytes, cudaMemcpyHostToDevice, m->stream));
  double *dr = r_dev ? r : work(m, 1);
  const double d = (lmax + lmin) * .5, c = (lmax - lmin) * .5;
  std::vector<double> al(iter + 1), be(iter + 1);
  double alpha = 0.;
  for (int i = 0; i < iter; ++i) {
    if (i == 0) alpha = 1. / d;
    else if (i == 1) alpha = 2. * d / (2 * d * d - c * c);
    else alpha = 1. / (d - (alpha * c * c / 4.));
    al[i] = alpha; be[i] = alpha * d - 1.;
  }
  cheby_enqueue(m, du, drhs, dr, iter, al.data(), be.data(), nullptr);
  if (!u_dev) D4B_CUDA(cudaMemcpyAsync(u, du, bytes, cudaMemcpyDeviceToHost, m->stream));
  if (!r_dev) D4B_CUDA(cudaMemcpyAsync(r, dr, bytes, cudaMemcpyDeviceToHost, m->stream));
  if (!u_dev || !r_dev) D4B_CUDA(cudaStreamSynchronize(m->stream));
}
