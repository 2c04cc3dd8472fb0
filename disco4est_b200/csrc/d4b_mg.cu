// d4b_mg.cu -- device-resident multigrid V-cycle over a caller-described
// hierarchy of meshes (include/d4est_b200.h, "multigrid").
//
// Follows d4est_solver_multigrid_vcycle (src/Solver/d4est_solver_multigrid.c:751-1349):
//   down, level top..1 : u_l = 0 below the top; smooth -> r_l = rhs_l - A u_l;
//                        rhs_{l-1} = P^T r_l            (callbacks.h:110-203)
//   bottom             : u_0 = 0; bottom solver
//   up, level 0..top-1 : u_{l+1} += P u_l; smooth       (callbacks.h:237-329, :1250)
// with the Chebyshev smoother of d4est_solver_multigrid_smoother_cheby.c:222-376
// (eigenvalue re-use rules included) and the CG / Chebyshev bottom solvers.
// All level vectors stay in HBM; the only host traffic is the scalars of the
// dot products that steer CG and the stopping tests.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <cstdlib>
#include <cstdio>
#include "d4b_internal.h"

struct d4b_mg {
  int n_levels = 0;
  std::vector<d4b_mesh *> mesh;
  std::vector<int32_t *> d_child_first, d_child_count;   // [n_levels-1]
  std::vector<double *> d_u, d_rhs, d_r;                 // per level (top: only d_r; u/rhs are the caller's)
  double *d_top_u = nullptr, *d_top_rhs = nullptr;       // staging for host-pointer calls
  double *d_P = nullptr;                                 // 1-D prolongation tables
  int32_t *d_P_off = nullptr;                            // ((degH-1)*7 + (degh-1))*3 + half -> offset
  d4b_mg_desc_t prm;
  double *d_eigs = nullptr;      // [n_levels + 1] smoother estimates per level, then the bottom solver's
  double *d_scal = nullptr;      // bottom CG scalars (BcgScal)
  double *h_scal = nullptr;      // pinned
  int vcycle_num_finished = 0;
  // CUDA graphs of whole V-cycles (one rank, Chebyshev bottom solver: the kernel sequence is then
  // fixed and all its scalars live on the device).  Keyed by the caller's vectors and the two
  // eigenvalue-reuse decisions; a key is captured the second time it is seen.
  struct CycleGraph {
    double *u = nullptr; const double *rhs = nullptr;
    bool eigs_down = false, eigs_up = false;
    int64_t epoch = 0;             // sum of the levels' config_epoch when the key was created
    uint64_t parity = 0;           // multi-rank: bit l = which copy of uhat level l writes next (the captured kernels
                                   // carry the pointers of the copies; d4b_comm_peer_parity)
    std::vector<unsigned long long> advance;   // multi-rank: applies per level in one cycle (host-side epoch bookkeeping)
    int seen = 0;
    cudaGraphExec_t exec = nullptr;
  };
  std::vector<CycleGraph> graphs;
  int64_t n_replays = 0;
  double *d_r2 = nullptr;        // squared residual of the last cycle
  cudaStream_t cap_stream = nullptr;
  // Galerkin mass term of the coarse levels (d4b_mg_set_coarse_mass): per level l >= 1 a prolonged copy of the
  // input and the weighted / restricted result on the way back
  int coarse_mass = 0;
  std::vector<double *> d_mw, d_mt;
  double *d_W0 = nullptr;        // mode 1: w J f_u(x, 0) on the finest level (see d4b_mg_set_coarse_mass)
  // the same operator as dense per-element matrices of the coarse levels (uniform degree, what the reference's
  // PT_mat_P stores): column-major, Mc[l][(e n3 + k) n3 + i]; built level by level from the one above whenever the
  // fine weights change (build_coarse_matrices), applied by one kernel instead of the prolong / weight / restrict chain
  std::vector<double *> d_Mc;
  std::vector<int> mc_n3;        // 0: level not available in dense form (chain)
  int mc_mode = 0;               // what d_Mc was built for
  int64_t mc_version = -1;       // ... and for which contents of the finest level's W (mode 2)
  double *d_unit = nullptr;      // unit vectors of the build
  // mode 1, levels whose families all have 8 children of one degree: the slices Q_i as a table
  // Qt[(i nh3 + b) nH3 + a] (transposed for coalesced reads), per coarse level, else nullptr
  std::vector<double *> d_Qt;
  std::vector<int> q_nH3, q_nh3;
};

namespace {

bool is_dev(const void *p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

constexpr int kMaxN = D4B_MAXN;

// One block per coarse element.  dir = 0: fine[child] (+)= P_c coarse;  dir = 1: coarse = sum_c P_c^T fine[child]
// (children in z-order, summed in that order like d4est_operators_apply_hp_prolong_transpose).
__global__ void __launch_bounds__(128) transfer_kernel(int dir, int accumulate, int n_coarse,
                                                       const int32_t *__restrict__ child_first,
                                                       const int32_t *__restrict__ child_count,
                                                       const int32_t *__restrict__ deg_c,
                                                       const int64_t *__restrict__ stride_c,
                                                       const int32_t *__restrict__ deg_f,
                                                       const int64_t *__restrict__ stride_f,
                                                       const double *__restrict__ P, const int32_t *__restrict__ P_off,
                                                       const double *__restrict__ in, double *__restrict__ out) {
  __shared__ double sA[kMaxN * kMaxN * kMaxN], sB[kMaxN * kMaxN * kMaxN], sAcc[kMaxN * kMaxN * kMaxN];
  const int e = blockIdx.x;
  if (e >= n_coarse) return;
  const int nH = deg_c[e] + 1, nH3 = nH * nH * nH;
  const int64_t sH = stride_c[e];
  const int first = child_first[e], count = child_count[e];
  if (dir == 2 && count == 8) {
    // What the reference's multigrid matrix operator EXECUTES for an h-coarsened family
    // (d4est_operators_compute_PT_mat_P, src/dGMath/d4est_operators.c:608-661): it transposes the stacked
    // prolongation [sum_i n_i x nH3] as a whole (:638) and then takes, for child i, the CONTIGUOUS slice
    // PT[off_i nH3 ...) of that transpose as if it were P_i^T (:650) -- which it is only for a single child.
    // The slice read as an [nH3 x n_i] row-major matrix is  Q_i[a][b] = Pstack[k][j]  with
    // f = off_i nH3 + a n_i + b, j = f / total, k = f % total.  out_H = sum_i Q_i in_i.
    int nh3[8], off[9];
    off[0] = 0;
    for (int c = 0; c < 8; ++c) {
      const int nh = deg_f[first + c] + 1;
      nh3[c] = nh * nh * nh;
      off[c + 1] = off[c] + nh3[c];
    }
    const int total = off[8];
    for (int a = threadIdx.x; a < nH3; a += blockDim.x) {
      double acc = 0.;
      for (int i = 0; i < 8; ++i) {
        const double *y = in + stride_f[first + i];
        const int64_t f0 = (int64_t)off[i] * nH3 + (int64_t)a * nh3[i];
        for (int b = 0; b < nh3[i]; ++b) {
          const int64_t f = f0 + b;
          const int j = (int)(f / total), k = (int)(f % total);
          int c = 0;
          while (k >= off[c + 1]) ++c;
          const int m = k - off[c];
          const int nh = deg_f[first + c] + 1;
          const int tb = ((nH - 2) * 7 + (nh - 2)) * 3;
          const double *Px = P + P_off[tb + 1 + (c & 1)], *Py = P + P_off[tb + 1 + ((c >> 1) & 1)],
                       *Pz = P + P_off[tb + 1 + ((c >> 2) & 1)];
          const int mx = m % nh, my = (m / nh) % nh, mz = m / (nh * nh);
          const int jx = j % nH, jy = (j / nH) % nH, jz = j / (nH * nH);
          acc += Px[mx * nH + jx] * Py[my * nH + jy] * Pz[mz * nH + jz] * y[b];
        }
      }
      out[sH + a] = acc;
    }
    return;
  }
  if (dir == 2) dir = 1;
  if (dir == 0) {
    for (int i = threadIdx.x; i < nH3; i += blockDim.x) sAcc[i] = in[sH + i];
  } else {
    for (int i = threadIdx.x; i < nH3; i += blockDim.x) sAcc[i] = 0.;
  }
  __syncthreads();
  for (int c = 0; c < count; ++c) {
    const int ef = first + c;
    const int nh = deg_f[ef] + 1;
    const int64_t sh = stride_f[ef];
    const int hx = count == 1 ? 0 : 1 + (c & 1), hy = count == 1 ? 0 : 1 + ((c >> 1) & 1),
              hz = count == 1 ? 0 : 1 + ((c >> 2) & 1);
    const int tb = ((nH - 2) * 7 + (nh - 2)) * 3;
    const double *Px = P + P_off[tb + hx], *Py = P + P_off[tb + hy], *Pz = P + P_off[tb + hz];   // [nh x nH]
    if (dir == 0) {
      // x: [nH,nH,nH] -> [nH,nH,nh]
      for (int o = threadIdx.x; o < nH * nH * nh; o += blockDim.x) {
        int i2 = o % nh, jk = o / nh;
        double s = 0.;
        for (int i = 0; i < nH; ++i) s += Px[i2 * nH + i] * sAcc[jk * nH + i];
        sA[o] = s;
      }
      __syncthreads();
      // y: [nH,nH,nh] -> [nH,nh,nh]
      for (int o = threadIdx.x; o < nH * nh * nh; o += blockDim.x) {
        int i2 = o % nh, j2 = (o / nh) % nh, k = o / (nh * nh);
        double s = 0.;
        for (int j = 0; j < nH; ++j) s += Py[j2 * nH + j] * sA[(k * nH + j) * nh + i2];
        sB[o] = s;
      }
      __syncthreads();
      // z: [nH,nh,nh] -> [nh,nh,nh]
      for (int o = threadIdx.x; o < nh * nh * nh; o += blockDim.x) {
        int ij = o % (nh * nh), k2 = o / (nh * nh);
        double s = 0.;
        for (int k = 0; k < nH; ++k) s += Pz[k2 * nH + k] * sB[k * nh * nh + ij];
        if (accumulate) out[sh + o] += s; else out[sh + o] = s;
      }
      __syncthreads();
    } else {
      for (int i = threadIdx.x; i < nh * nh * nh; i += blockDim.x) sA[i] = in[sh + i];
      __syncthreads();
      // x^T: [nh,nh,nh] -> [nh,nh,nH]
      for (int o = threadIdx.x; o < nh * nh * nH; o += blockDim.x) {
        int i = o % nH, jk = o / nH;
        double s = 0.;
        for (int i2 = 0; i2 < nh; ++i2) s += Px[i2 * nH + i] * sA[jk * nh + i2];
        sB[o] = s;
      }
      __syncthreads();
      // y^T: [nh,nh,nH] -> [nh,nH,nH]
      for (int o = threadIdx.x; o < nh * nH * nH; o += blockDim.x) {
        int i = o % nH, j = (o / nH) % nH, k2 = o / (nH * nH);
        double s = 0.;
        for (int j2 = 0; j2 < nh; ++j2) s += Py[j2 * nH + j] * sB[(k2 * nh + j2) * nH + i];
        sA[o] = s;
      }
      __syncthreads();
      // z^T: [nh,nH,nH] -> [nH,nH,nH], added to the parent
      for (int o = threadIdx.x; o < nH3; o += blockDim.x) {
        int ij = o % (nH * nH), k = o / (nH * nH);
        double s = 0.;
        for (int k2 = 0; k2 < nh; ++k2) s += Pz[k2 * nH + k] * sA[k2 * nH * nH + ij];
        sAcc[o] += s;
      }
      __syncthreads();
    }
  }
  if (dir == 1)
    for (int i = threadIdx.x; i < nH3; i += blockDim.x) out[sH + i] = sAcc[i];
}

// out_H[a] = sum_i sum_b Qt[(i nh3 + b) nH3 + a] in_i[b] for uniform families (see transfer_kernel dir 2 for what
// Q is); kQEpb coarse elements per block share one pass over the table
constexpr int kQEpb = 8;
__global__ void __launch_bounds__(128) qtable_restrict_kernel(int n_coarse, int nH3, int nh3,
                                                              const int32_t *__restrict__ child_first,
                                                              const int64_t *__restrict__ stride_c,
                                                              const int64_t *__restrict__ stride_f,
                                                              const double *__restrict__ Qt,
                                                              const double *__restrict__ in, double *__restrict__ out) {
  extern __shared__ double sy[];   // [kQEpb][8 nh3]
  const int e0 = blockIdx.x * kQEpb;
  const int ne = min(kQEpb, n_coarse - e0);
  const int tot = 8 * nh3;
  for (int k = 0; k < ne; ++k) {
    const int first = child_first[e0 + k];
    for (int j = threadIdx.x; j < tot; j += blockDim.x) sy[k * tot + j] = in[stride_f[first + j / nh3] + j % nh3];
  }
  __syncthreads();
  for (int a = threadIdx.x; a < nH3; a += blockDim.x) {
    double acc[kQEpb];
#pragma unroll
    for (int k = 0; k < kQEpb; ++k) acc[k] = 0.;
    for (int j = 0; j < tot; ++j) {
      const double q = Qt[(size_t)j * nH3 + a];
#pragma unroll
      for (int k = 0; k < kQEpb; ++k)
        if (k < ne) acc[k] += q * sy[k * tot + j];
    }
#pragma unroll
    for (int k = 0; k < kQEpb; ++k)
      if (k < ne) out[stride_c[e0 + k] + a] = acc[k];
  }
}

void transfer(d4b_mg *g, int dir, int accumulate, int lc, const double *in, double *out) {
  if (dir == 2 && lc < (int)g->d_Qt.size() && g->d_Qt[lc]) {
    d4b_mesh *mc = g->mesh[lc], *mf = g->mesh[lc + 1];
    const size_t smem = (size_t)kQEpb * 8 * g->q_nh3[lc] * sizeof(double);
    static size_t configured = 48 * 1024;
    if (smem > configured) {
      D4B_CUDA(cudaFuncSetAttribute(qtable_restrict_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      configured = smem;
    }
    qtable_restrict_kernel<<<(mc->n_local + kQEpb - 1) / kQEpb, 128, smem, mf->stream>>>(
        mc->n_local, g->q_nH3[lc], g->q_nh3[lc], g->d_child_first[lc], mc->d_nodal_stride, mf->d_nodal_stride,
        g->d_Qt[lc], in, out);
    D4B_CUDA(cudaGetLastError());
    return;
  }
  d4b_mesh *mc = g->mesh[lc], *mf = g->mesh[lc + 1];
  transfer_kernel<<<mc->n_local, 128, 0, mf->stream>>>(dir, accumulate, mc->n_local, g->d_child_first[lc],
                                                       g->d_child_count[lc], mc->d_deg, mc->d_nodal_stride,
                                                       mf->d_deg, mf->d_nodal_stride, g->d_P, g->d_P_off, in, out);
  D4B_CUDA(cudaGetLastError());
}

__global__ void resid_k(const double *__restrict__ rhs, const double *__restrict__ Au, double *__restrict__ r,
                        int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    r[i] = rhs[i] - Au[i];
}
int grid_for(int64_t n) {
  int64_t b = (n + 255) / 256;
  return (int)(b < 148 * 8 ? (b > 0 ? b : 1) : 148 * 8);
}

// out_e = Mc_e in_e, one block per element, thread i one row (column-major storage: coalesced)
__global__ void dense_elem_matvec_k(int n3, const double *__restrict__ Mc, const double *__restrict__ in,
                                    double *__restrict__ out) {
  extern __shared__ double su[];
  const size_t e = blockIdx.x;
  for (int k = threadIdx.x; k < n3; k += blockDim.x) su[k] = in[e * n3 + k];
  __syncthreads();
  const double *M = Mc + e * (size_t)n3 * n3;
  for (int i = threadIdx.x; i < n3; i += blockDim.x) {
    double acc = 0.;
    for (int k = 0; k < n3; ++k) acc += M[(size_t)k * n3 + i] * su[k];
    out[e * n3 + i] = acc;
  }
}
__global__ void unit_vector_k(int64_t n, int n3, int k, double *__restrict__ u) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    u[i] = (i % n3 == k) ? 1. : 0.;
}
__global__ void store_column_k(int64_t n, int n3, int k, const double *__restrict__ col, double *__restrict__ Mc) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    Mc[((i / n3) * n3 + k) * (int64_t)n3 + i % n3] = col[i];
}

void dense_coarse_apply(d4b_mg *g, int l, const double *in, double *out) {
  d4b_mesh *m = g->mesh[l];
  const int n3 = g->mc_n3[l];
  dense_elem_matvec_k<<<m->n_local, n3 <= 64 ? 64 : 128, (size_t)n3 * sizeof(double), m->stream>>>(n3, g->d_Mc[l], in, out);
  D4B_CUDA(cudaGetLastError());
}

void transfer(d4b_mg *g, int dir, int accumulate, int lc, const double *in, double *out);

// The coarse-level mass term as dense per-element matrices, M_l = L_l M_{l+1} P_l with L = P^T (Galerkin) or the
// reference's slices, column by column: the chain of ONE level applied to the unit vector k of every element at once
// gives column k of every element's matrix (the operator is block diagonal over the elements).  Levels must have a
// uniform degree N <= 5 and fit D4B_MG_DENSE_COARSE_MB (default 4096) together; D4B_MG_DENSE_COARSE=0 keeps the chain.
// Not inside a stream capture: called at the start of every V-cycle and of every stand-alone coarse apply.
void build_coarse_matrices(d4b_mg *g) {
  const int top = g->n_levels - 1;
  d4b_mesh *mt = g->mesh[top];
  if (!g->coarse_mass || !mt->d_W) return;
  if (g->mc_mode == g->coarse_mass && (g->coarse_mass == 1 || g->mc_version == mt->w_version)) return;
  static const bool off = getenv("D4B_MG_DENSE_COARSE") && atoi(getenv("D4B_MG_DENSE_COARSE")) == 0;
  static const double budget_mb = getenv("D4B_MG_DENSE_COARSE_MB") ? atof(getenv("D4B_MG_DENSE_COARSE_MB")) : 4096.;
  g->mc_mode = g->coarse_mass;
  g->mc_version = mt->w_version;
  if (g->d_Mc.empty()) { g->d_Mc.assign(g->n_levels, nullptr); g->mc_n3.assign(g->n_levels, 0); }
  cudaStream_t s = mt->stream;
  std::vector<cudaStream_t> saved(g->n_levels);
  for (int j = 0; j <= top; ++j) {
    saved[j] = g->mesh[j]->stream;
    if (saved[j] != s) { D4B_CUDA(cudaStreamSynchronize(saved[j])); g->mesh[j]->stream = s; }
  }
  const double *Wfine = g->coarse_mass == 1 ? g->d_W0 : mt->d_W;
  const int back = g->coarse_mass == 1 ? 2 : 1;
  double used_mb = 0.;
  bool chain_from_here = off;
  for (int l = top - 1; l >= 0; --l) {
    d4b_mesh *ml = g->mesh[l];
    const int N = ml->max_deg + 1, n3 = N * N * N;
    const double mb = (double)ml->n_local * n3 * n3 * 8. / 1048576.;
    const bool uniform = ml->n_local > 0 && ml->local_nodes == (int64_t)ml->n_local * n3 && N <= 5 && ml->mpisize >= 1;
    if (chain_from_here || !uniform || used_mb + mb > budget_mb) {   // this level and everything below: the chain
      chain_from_here = true;
      g->mc_n3[l] = 0;
      continue;
    }
    used_mb += mb;
    if (!g->d_Mc[l]) g->d_Mc[l] = d4b_dev_alloc<double>(ml, (size_t)ml->n_local * n3 * n3);
    if (!g->d_unit) {
      int64_t mx = 1;
      for (int j = 0; j < top; ++j) mx = std::max(mx, g->mesh[j]->local_nodes);
      g->d_unit = d4b_dev_alloc<double>(mt, (size_t)mx);
    }
    const int64_t n = ml->local_nodes;
    for (int k = 0; k < n3; ++k) {
      unit_vector_k<<<grid_for(n), 256, 0, s>>>(n, n3, k, g->d_unit);
      transfer(g, 0, 0, l, g->d_unit, g->d_mw[l + 1]);
      if (l + 1 == top) d4b_launch_quad_op(mt, 4, g->d_mw[top], Wfine, g->d_mt[top]);
      else dense_coarse_apply(g, l + 1, g->d_mw[l + 1], g->d_mt[l + 1]);
      transfer(g, back, 0, l, g->d_mt[l + 1], ml->d_addterm);
      store_column_k<<<grid_for(n), 256, 0, s>>>(n, n3, k, ml->d_addterm, g->d_Mc[l]);
    }
    D4B_CUDA(cudaGetLastError());
    g->mc_n3[l] = n3;
  }
  for (int j = 0; j <= top; ++j) g->mesh[j]->stream = saved[j];
}

// t = P_l^T ... P_{top-1}^T (V^T diag(W) V) P_{top-1} ... P_l u on coarse level l: what the reference's dense
// per-element matrices PT_mat_P of the multigrid matrix operator apply
// (d4est_solver_multigrid_matrix_operator.c:5-47; d4est_operators_compute_PT_mat_P), matrix-free.
// Returns false (nothing to add) while the finest level is not linearised.
bool coarse_mass_term(d4b_mg *g, int l, const double *u) {
  const int top = g->n_levels - 1;
  d4b_mesh *mt = g->mesh[top], *ml = g->mesh[l];
  if (!g->coarse_mass || !mt->d_W) return false;
  if (!g->d_Mc.empty() && g->mc_n3[l] && g->mc_mode == g->coarse_mass &&
      (g->coarse_mass == 1 || g->mc_version == mt->w_version)) {
    dense_coarse_apply(g, l, u, ml->d_addterm);
    return true;
  }
  // every level shares the top stream inside a V-cycle; a stand-alone coarse apply runs on its own
  // stream and the level vectors are ordered through it
  cudaStream_t s = ml->stream;
  std::vector<cudaStream_t> saved(g->n_levels);
  for (int j = l; j <= top; ++j) {
    saved[j] = g->mesh[j]->stream;
    if (saved[j] != s) { D4B_CUDA(cudaStreamSynchronize(saved[j])); g->mesh[j]->stream = s; }
  }
  const double *in = u;
  for (int j = l; j < top; ++j) {
    transfer(g, 0, 0, j, in, g->d_mw[j + 1]);
    in = g->d_mw[j + 1];
  }
  d4b_launch_quad_op(mt, 4, in, g->coarse_mass == 1 ? g->d_W0 : mt->d_W, g->d_mt[top]);
  // mode 1: the slices the reference executes (transfer_kernel dir 2); mode 2: the Galerkin product P^T M P
  const int back = g->coarse_mass == 1 ? 2 : 1;
  for (int j = top - 1; j >= l; --j) transfer(g, back, 0, j, g->d_mt[j + 1], j == l ? ml->d_addterm : g->d_mt[j]);
  for (int j = l; j <= top; ++j) g->mesh[j]->stream = saved[j];
  return true;
}

// Chebyshev smoother of one level, d4est_solver_multigrid_smoother_cheby.c:264-376; the
// estimate of the level lives in g->d_eigs[level] and never visits the host
void smooth(d4b_mg *g, int level, double *u, const double *rhs, double *r, bool compute_eigs) {
  d4b_smooth_async(g->mesh[level], u, rhs, r, g->prm.smoother, compute_eigs, g->d_eigs + level);
}

// Bottom CG, d4est_solver_multigrid_bottom_solver_cg.c:48-200: the loop of d4est_solver_cg.c,
// stopping test on the device (d4b_krylov.cu)
void bottom_cg(d4b_mg *g, double *u, const double *rhs, double *r) {
  d4b_mesh *m = g->mesh[0];
  if (!m->d_work[3]) m->d_work[3] = d4b_dev_alloc<double>(m, (size_t)m->local_nodes);
  d4b_cg_run(m, u, rhs, r, m->d_work[3], g->prm.bottom_iter, g->prm.bottom_rtol, g->prm.bottom_atol, g->d_scal,
             g->h_scal);
}

// d4est_solver_multigrid_bottom_solver_cheby.c:60-115
void bottom_cheby(d4b_mg *g, double *u, const double *rhs, double *r) {
  d4b_cheby_params_t c = g->prm.bottom_cheby;
  c.cheby_use_zero_guess_for_eigs = 0;
  d4b_smooth_async(g->mesh[0], u, rhs, r, c, true, g->d_eigs + g->n_levels);
}

// one V-cycle on device pointers; returns the global squared residual of the top level.
// Everything between the first smoother kernel and the final residual norm is enqueued on
// one stream without a host synchronisation (a CG bottom solver polls its flag once per
// kBcgChunk iterations): while the fine level runs, the host is already queueing the coarse
// levels, so their launch latencies are hidden.
double vcycle_device(d4b_mg *g, double *u_top, const double *rhs_top) {
  const int top = g->n_levels - 1;
  d4b_check_device(g->mesh[top]);
  build_coarse_matrices(g);   // (a no-op unless the coarse-level mass term is on and the fine weights changed)
  const d4b_cheby_params_t &c = g->prm.smoother;
  const int vc = g->vcycle_num_finished;
  // ..._smoother_cheby_update (:222-262): decided at PRE_V for the whole down leg,
  // at UPV_PRE_SMOOTH for every smoothing of the up leg
  const bool eigs_down = !(c.cheby_eigs_reuse_fromlastvcycle == 1 && vc != 0);
  const bool eigs_up = !(c.cheby_eigs_reuse_fromdownvcycle == 1 || (c.cheby_eigs_reuse_fromlastvcycle == 1 && vc != 0));
  D4B_CHECK(!(c.cheby_use_zero_guess_for_eigs == 1 && c.cheby_eigs_reuse_fromdownvcycle != 1),
            "cheby_use_zero_guess_for_eigs = 1 needs cheby_eigs_reuse_fromdownvcycle = 1 "
            "(d4est_solver_multigrid_smoother_cheby.c:315-320)");
  // all levels on the top level's stream for the duration of the cycle
  cudaStream_t s = g->mesh[top]->stream;
  std::vector<cudaStream_t> saved(g->n_levels);
  for (int l = 0; l <= top; ++l) {
    saved[l] = g->mesh[l]->stream;
    if (saved[l] != s) D4B_CUDA(cudaStreamSynchronize(saved[l]));
    g->mesh[l]->stream = s;
  }
  d4b_mesh *mt = g->mesh[top];
  if (!g->d_r2) g->d_r2 = d4b_dev_alloc<double>(mt, 2);
  auto enqueue = [&](cudaStream_t s) {
    for (int l = top; l > 0; --l) {
      d4b_mesh *m = g->mesh[l];
      double *u = (l == top) ? u_top : g->d_u[l];
      const double *rhs = (l == top) ? rhs_top : g->d_rhs[l];
      if (l != top) D4B_CUDA(cudaMemsetAsync(u, 0, (size_t)m->local_nodes * sizeof(double), s));
      smooth(g, l, u, rhs, g->d_r[l], eigs_down);
      // restricted residual is the coarse right-hand side
      transfer(g, 1, 0, l - 1, g->d_r[l], g->d_rhs[l - 1]);
    }
    {
      d4b_mesh *m = g->mesh[0];
      D4B_CUDA(cudaMemsetAsync(g->d_u[0], 0, (size_t)m->local_nodes * sizeof(double), s));
      if (g->prm.bottom_solver == 0) bottom_cg(g, g->d_u[0], g->d_rhs[0], g->d_r[0]);
      else bottom_cheby(g, g->d_u[0], g->d_rhs[0], g->d_r[0]);
    }
    for (int l = 0; l < top; ++l) {
      const int f = l + 1;
      double *u = (f == top) ? u_top : g->d_u[f];
      const double *rhs = (f == top) ? rhs_top : g->d_rhs[f];
      transfer(g, 0, 1, l, g->d_u[l], u);   // u_fine += P u_coarse
      smooth(g, f, u, rhs, g->d_r[f], eigs_up);
    }
    d4b_dot_async(mt, g->d_r[top], g->d_r[top], mt->local_nodes, g->d_r2);
  };
  // graph replay: only when the sequence is static (no host polling).  Multi-rank meshes qualify when the halo
  // and the all-reduces run over peer memory: no NCCL call in the sequence, the epochs of both flag protocols are
  // device counters, and the second stream of the large levels is forked from and joined to the captured one
  bool graphable = g->prm.bottom_solver == 1 && !(getenv("D4B_MG_GRAPH") && atoi(getenv("D4B_MG_GRAPH")) == 0);
  for (int l = 0; l <= top && graphable; ++l) graphable = d4b_comm_graph_safe(g->mesh[l]);
  bool multi = false;
  for (int l = 0; l <= top; ++l) multi = multi || g->mesh[l]->mpisize > 1;
  if (multi && getenv("D4B_MG_GRAPH_MULTI") && atoi(getenv("D4B_MG_GRAPH_MULTI")) == 0) graphable = false;
  d4b_mg::CycleGraph *cg = nullptr;
  int64_t epoch = 0;
  uint64_t parity = 0;
  for (int l = 0; l <= top; ++l) {
    epoch += g->mesh[l]->config_epoch;
    parity |= (uint64_t)d4b_comm_peer_parity(g->mesh[l]) << l;
  }
  if (graphable) {
    for (auto &c2 : g->graphs)
      if (c2.u == u_top && c2.rhs == rhs_top && c2.eigs_down == eigs_down && c2.eigs_up == eigs_up &&
          c2.epoch == epoch && c2.parity == parity) cg = &c2;
    if (!cg) {
      if (g->graphs.size() >= 8) {   // vectors keep changing: stop caching
        for (auto &c2 : g->graphs) if (c2.exec) cudaGraphExecDestroy(c2.exec);
        g->graphs.clear();
      }
      g->graphs.emplace_back();
      cg = &g->graphs.back();
      cg->u = u_top; cg->rhs = rhs_top; cg->eigs_down = eigs_down; cg->eigs_up = eigs_up; cg->epoch = epoch;
      cg->parity = parity;
    }
  }
  if (getenv("D4B_MG_DEBUG"))
    fprintf(stderr, "[d4b mg] cycle %d: graphable %d key parity %llx config %lld -> %s (cached keys %zu)\n", vc, (int)graphable,
            (unsigned long long)parity, (long long)epoch,
            !cg ? "eager (not graphable)" : cg->exec ? "replay" : cg->seen >= 1 ? "record" : "eager (first time)", g->graphs.size());
  std::vector<unsigned long long> e0(g->n_levels);
  for (int l = 0; l <= top; ++l) e0[l] = d4b_comm_peer_epoch(g->mesh[l]);
  if (cg && cg->exec) {
    D4B_CUDA(cudaGraphLaunch(cg->exec, s));
    ++g->n_replays;
    for (int l = 0; l <= top; ++l) d4b_comm_peer_advance(g->mesh[l], cg->advance[l]);   // what the eager cycle does on the host
  } else if (cg && cg->seen >= 1) {
    // every lazy allocation and kernel attribute was set by the first, eager cycle with this key
    // captured on a private stream (the caller's may be the legacy default stream, which cannot be
    // captured), replayed on the caller's
    cudaGraph_t graph = nullptr;
    if (!g->cap_stream) D4B_CUDA(cudaStreamCreateWithFlags(&g->cap_stream, cudaStreamNonBlocking));
    D4B_CUDA(cudaStreamSynchronize(s));
    for (int l = 0; l <= top; ++l) g->mesh[l]->stream = g->cap_stream;
    d4b_comm_allreduce_reset_order();
    D4B_CUDA(cudaStreamBeginCapture(g->cap_stream, cudaStreamCaptureModeThreadLocal));
    enqueue(g->cap_stream);
    D4B_CUDA(cudaStreamEndCapture(g->cap_stream, &graph));
    d4b_comm_allreduce_reset_order();
    for (int l = 0; l <= top; ++l) g->mesh[l]->stream = s;
    cg->advance.resize(g->n_levels);   // the capture ran the host-side bookkeeping of one cycle
    for (int l = 0; l <= top; ++l) cg->advance[l] = d4b_comm_peer_epoch(g->mesh[l]) - e0[l];
    D4B_CUDA(cudaGraphInstantiate(&cg->exec, graph, 0));
    D4B_CUDA(cudaGraphDestroy(graph));
    D4B_CUDA(cudaGraphLaunch(cg->exec, s));
  } else {
    enqueue(s);
  }
  if (cg) cg->seen++;
  double r2 = 0.;
  D4B_CUDA(cudaMemcpyAsync(g->h_scal + 32, g->d_r2, sizeof(double), cudaMemcpyDeviceToHost, s));
  D4B_CUDA(cudaStreamSynchronize(s));
  r2 = g->h_scal[32];
  for (int l = 0; l <= top; ++l) g->mesh[l]->stream = saved[l];
  return r2;
}

struct TopVectors {
  double *u = nullptr;
  const double *rhs = nullptr;
  bool u_dev = true;
};

TopVectors stage_in(d4b_mg *g, double *u, const double *rhs) {
  d4b_mesh *mt = g->mesh[g->n_levels - 1];
  const size_t bytes = (size_t)mt->local_nodes * sizeof(double);
  TopVectors t;
  t.u_dev = is_dev(u);
  if (t.u_dev) t.u = u;
  else {
    if (!g->d_top_u) g->d_top_u = d4b_dev_alloc<double>(mt, (size_t)mt->local_nodes);
    D4B_CUDA(cudaMemcpyAsync(g->d_top_u, u, bytes, cudaMemcpyHostToDevice, mt->stream));
    t.u = g->d_top_u;
  }
  if (is_dev(rhs)) t.rhs = rhs;
  else {
    if (!g->d_top_rhs) g->d_top_rhs = d4b_dev_alloc<double>(mt, (size_t)mt->local_nodes);
    D4B_CUDA(cudaMemcpyAsync(g->d_top_rhs, rhs, bytes, cudaMemcpyHostToDevice, mt->stream));
    t.rhs = g->d_top_rhs;
  }
  return t;
}

void stage_out(d4b_mg *g, const TopVectors &t, double *u) {
  d4b_mesh *mt = g->mesh[g->n_levels - 1];
  if (!t.u_dev)
    D4B_CUDA(cudaMemcpyAsync(u, t.u, (size_t)mt->local_nodes * sizeof(double), cudaMemcpyDeviceToHost, mt->stream));
  D4B_CUDA(cudaStreamSynchronize(mt->stream));
}

}  // namespace

extern "C" d4b_mg_t *d4b_mg_create(const d4b_mg_desc_t *desc) {
  D4B_CHECK(desc && desc->n_levels >= 2, "d4b_mg_create: needs at least two levels (d4est_solver_multigrid.c:1434)");
  d4b_mg *g = new d4b_mg();
  g->n_levels = desc->n_levels;
  g->prm = *desc;
  {
    d4b_mesh *mt = desc->meshes[desc->n_levels - 1];
    D4B_CHECK(mt, "d4b_mg_create: the top level has no mesh");
    g->d_eigs = d4b_dev_alloc<double>(mt, (size_t)desc->n_levels + 1);
    g->d_scal = d4b_dev_alloc<double>(mt, 16);
    D4B_CUDA(cudaMemset(g->d_eigs, 0, ((size_t)desc->n_levels + 1) * sizeof(double)));
    D4B_CUDA(cudaMemset(g->d_scal, 0, 16 * sizeof(double)));
    D4B_CUDA(cudaMallocHost(&g->h_scal, 64 * sizeof(double)));
  }
  for (int l = 0; l < desc->n_levels; ++l) {
    D4B_CHECK(desc->meshes[l], "d4b_mg_create: level %d has no mesh", l);
    g->mesh.push_back(desc->meshes[l]);
  }
  g->prm.meshes = nullptr; g->prm.child_first = nullptr; g->prm.child_count = nullptr;
  const int top = desc->n_levels - 1;
  // transfer maps, validated: every fine element has exactly one parent, in order
  for (int l = 0; l < top; ++l) {
    d4b_mesh *mc = g->mesh[l], *mf = g->mesh[l + 1];
    int next = 0;
    for (int e = 0; e < mc->n_local; ++e) {
      const int cnt = desc->child_count[l][e];
      D4B_CHECK(cnt == 1 || cnt == 8, "level %d element %d: child count %d", l, e, cnt);
      D4B_CHECK(desc->child_first[l][e] == next, "level %d element %d: children are not consecutive", l, e);
      next += cnt;
    }
    D4B_CHECK(next == mf->n_local, "level %d: children cover %d of %d fine elements", l, next, mf->n_local);
    g->d_child_first.push_back(d4b_dev_upload<int32_t>(mc, desc->child_first[l], mc->n_local));
    g->d_child_count.push_back(d4b_dev_upload<int32_t>(mc, desc->child_count[l], mc->n_local));
  }
  for (int l = 0; l <= top; ++l) {
    d4b_mesh *m = g->mesh[l];
    g->d_r.push_back(d4b_dev_alloc<double>(m, (size_t)m->local_nodes));
    g->d_u.push_back(l < top ? d4b_dev_alloc<double>(m, (size_t)m->local_nodes) : nullptr);
    g->d_rhs.push_back(l < top ? d4b_dev_alloc<double>(m, (size_t)m->local_nodes) : nullptr);
  }
  // 1-D prolongation tables for degH <= degh: Lagrange basis on LGL(degH) evaluated at the
  // LGL(degh) nodes of the whole interval (half 0, build_p_prolong_1d d4est_operators.c:995-1012)
  // or of the lower / upper half (half 1, 2, build_hp_prolong_1d :944-993)
  std::vector<double> P;
  std::vector<int32_t> off(7 * 7 * 3, -1);
  for (int dH = 1; dH <= D4B_MAX_DEG; ++dH)
    for (int dh = dH; dh <= D4B_MAX_DEG; ++dh) {
      const int nH = dH + 1, nh = dh + 1;
      double xH[kMaxN], wH[kMaxN], xh[kMaxN], wh[kMaxN], y[kMaxN], E[kMaxN * kMaxN];
      d4b_tab_lobatto(nH, xH, wH);
      d4b_tab_lobatto(nh, xh, wh);
      for (int half = 0; half < 3; ++half) {
        for (int i = 0; i < nh; ++i) y[i] = half == 0 ? xh[i] : (half == 1 ? 0.5 * (xh[i] - 1.) : 0.5 * (xh[i] + 1.));
        d4b_tab_eval(nH, xH, nh, y, E);
        if (half == 0 && dH == dh)   // the reference copies instead of multiplying (:1115-1118)
          for (int i = 0; i < nh; ++i)
            for (int j = 0; j < nH; ++j) E[i * nH + j] = (i == j) ? 1. : 0.;
        off[((dH - 1) * 7 + (dh - 1)) * 3 + half] = (int32_t)P.size();
        P.insert(P.end(), E, E + nh * nH);
      }
    }
  g->d_P = d4b_dev_upload<double>(g->mesh[top], P.data(), P.size());
  g->d_P_off = d4b_dev_upload<int32_t>(g->mesh[top], off.data(), off.size());
  // a coarse element is never of higher degree than its children
  return g;
}

d4b_mesh *d4b_mg_top_mesh(d4b_mg *g) { return g->mesh[g->n_levels - 1]; }

extern "C" void d4b_mg_set_coarse_mass(d4b_mg_t *g, int32_t mode) {
  D4B_CHECK(g && mode >= 0 && mode <= 2, "d4b_mg_set_coarse_mass: bad argument");
  const int top = g->n_levels - 1;
  for (d4b_mesh *m : g->mesh) D4B_CUDA(cudaStreamSynchronize(m->stream));
  g->coarse_mass = mode;
  g->mc_mode = 0;   // dense coarse matrices (build_coarse_matrices) are rebuilt before the next V-cycle
  if (mode == 1) {
    // The reference builds the fine matrices of its matrix operator with the arguments of
    // d4est_quadrature_apply_fofufofvlilj shifted by one (d4est_solver_multigrid_matrix_operator.c:215-240 passes
    // u0 in the `vec` slot and NULL in the `u` slot), so f_u is evaluated at u = 0 whatever the linearisation
    // point is (src/Quadrature/d4est_quadrature.c:663-671): the coarse mass term never changes during a Newton solve.
    d4b_mesh *mt = g->mesh[top];
    D4B_CHECK(mt->d_nl_wja, "d4b_mg_set_coarse_mass(REFERENCE): the finest mesh has no nonlinear term "
                            "(d4b_mesh_set_nonlinear_term); use D4B_COARSE_MASS_GALERKIN with d4b_mesh_set_mass_diag");
    if (!g->d_W0) g->d_W0 = d4b_dev_alloc<double>(mt, (size_t)mt->local_nodes_quad);
    double *zero = nullptr;
    D4B_CUDA(cudaMalloc(&zero, std::max<size_t>((size_t)mt->local_nodes, 1) * sizeof(double)));
    D4B_CUDA(cudaMemsetAsync(zero, 0, (size_t)mt->local_nodes * sizeof(double), mt->stream));
    d4b_launch_quad_nl(mt, 6, zero, g->d_W0);
    D4B_CUDA(cudaStreamSynchronize(mt->stream));
    cudaFree(zero);
    // table form of the slices for levels with uniform families
    const bool tables = !(getenv("D4B_MG_QTABLE") && atoi(getenv("D4B_MG_QTABLE")) == 0);   // 0: general path everywhere (tests)
    if (g->d_Qt.empty() && tables) {
      g->d_Qt.assign(top, nullptr); g->q_nH3.assign(top, 0); g->q_nh3.assign(top, 0);
      for (int l = 0; l < top; ++l) {
        d4b_mesh *mc = g->mesh[l], *mf = g->mesh[l + 1];
        std::vector<int32_t> dc(mc->n_local), df(mf->n_local), cnt(mc->n_local);
        D4B_CUDA(cudaMemcpy(dc.data(), mc->d_deg, dc.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
        D4B_CUDA(cudaMemcpy(df.data(), mf->d_deg, df.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
        D4B_CUDA(cudaMemcpy(cnt.data(), g->d_child_count[l], cnt.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
        bool uniform = !dc.empty();
        for (int32_t c : cnt) uniform = uniform && c == 8;
        for (int32_t d : dc) uniform = uniform && d == dc[0];
        for (int32_t d : df) uniform = uniform && d == df[0];
        if (!uniform || (size_t)kQEpb * 8 * (df[0] + 1) * (df[0] + 1) * (df[0] + 1) * sizeof(double) > 200 * 1024) continue;
        const int nH = dc[0] + 1, nh = df[0] + 1, nH3 = nH * nH * nH, nh3 = nh * nh * nh, total = 8 * nh3;
        // 1-D half-interval prolongations [nh x nH] (the same tables d_P holds)
        double xH[kMaxN], wH[kMaxN], xh[kMaxN], wh[kMaxN], y[kMaxN];
        std::vector<double> P1[2];
        d4b_tab_lobatto(nH, xH, wH);
        d4b_tab_lobatto(nh, xh, wh);
        for (int half = 0; half < 2; ++half) {
          for (int i = 0; i < nh; ++i) y[i] = half == 0 ? 0.5 * (xh[i] - 1.) : 0.5 * (xh[i] + 1.);
          P1[half].resize((size_t)nh * nH);
          d4b_tab_eval(nH, xH, nh, y, P1[half].data());
        }
        auto Pstack = [&](int k, int j) {   // stacked prolongation entry: row k = (child, fine node), column j
          const int c = k / nh3, m = k % nh3;
          const int mx = m % nh, my = (m / nh) % nh, mz = m / (nh * nh), jx = j % nH, jy = (j / nH) % nH, jz = j / (nH * nH);
          return P1[c & 1][mx * nH + jx] * P1[(c >> 1) & 1][my * nH + jy] * P1[(c >> 2) & 1][mz * nH + jz];
        };
        std::vector<double> Qt((size_t)total * nH3);
        for (int i = 0; i < 8; ++i)
          for (int a = 0; a < nH3; ++a)
            for (int b = 0; b < nh3; ++b) {
              const int64_t f = (int64_t)i * nh3 * nH3 + (int64_t)a * nh3 + b;
              Qt[((size_t)i * nh3 + b) * nH3 + a] = Pstack((int)(f % total), (int)(f / total));
            }
        g->d_Qt[l] = d4b_dev_upload<double>(mc, Qt.data(), Qt.size());
        g->q_nH3[l] = nH3; g->q_nh3[l] = nh3;
      }
    }
  }
  if (mode && g->d_mw.empty()) {
    g->d_mw.assign(g->n_levels, nullptr);
    g->d_mt.assign(g->n_levels, nullptr);
    for (int l = 1; l <= top; ++l) {
      g->d_mw[l] = d4b_dev_alloc<double>(g->mesh[l], (size_t)g->mesh[l]->local_nodes);
      g->d_mt[l] = d4b_dev_alloc<double>(g->mesh[l], (size_t)g->mesh[l]->local_nodes);
    }
  }
  for (int l = 0; l < top; ++l) {
    d4b_mesh *m = g->mesh[l];
    m->config_epoch++;
    if (mode) {
      if (!m->d_addterm) m->d_addterm = d4b_dev_alloc<double>(m, (size_t)m->local_nodes);
      m->pre_apply = [g, l](d4b_mesh *, const double *u) { return coarse_mass_term(g, l, u); };
    } else {
      m->pre_apply = nullptr;
    }
  }
}

extern "C" void d4b_mg_destroy(d4b_mg_t *g) {
  if (!g) return;
  for (d4b_mesh *m : g->mesh) cudaStreamSynchronize(m->stream);
  for (auto p : g->d_child_first) cudaFree(p);
  for (auto p : g->d_child_count) cudaFree(p);
  for (auto p : g->d_u) if (p) cudaFree(p);
  for (auto p : g->d_rhs) if (p) cudaFree(p);
  for (auto p : g->d_r) if (p) cudaFree(p);
  for (auto p : g->d_mw) if (p) cudaFree(p);
  for (auto p : g->d_mt) if (p) cudaFree(p);
  if (g->d_W0) cudaFree(g->d_W0);
  for (auto p : g->d_Mc) if (p) cudaFree(p);
  if (g->d_unit) cudaFree(g->d_unit);
  for (auto p : g->d_Qt) if (p) cudaFree(p);
  for (int l = 0; l + 1 < g->n_levels; ++l) g->mesh[l]->pre_apply = nullptr;
  if (g->d_top_u) cudaFree(g->d_top_u);
  if (g->d_top_rhs) cudaFree(g->d_top_rhs);
  cudaFree(g->d_P); cudaFree(g->d_P_off);
  cudaFree(g->d_eigs); cudaFree(g->d_scal); cudaFreeHost(g->h_scal);
  for (auto &c : g->graphs) if (c.exec) cudaGraphExecDestroy(c.exec);
  if (g->d_r2) cudaFree(g->d_r2);
  if (g->cap_stream) cudaStreamDestroy(g->cap_stream);
  delete g;
}

extern "C" double d4b_mg_vcycle(d4b_mg_t *g, double *u, const double *rhs) {
  TopVectors t = stage_in(g, u, rhs);
  double r2 = vcycle_device(g, t.u, t.rhs);
  g->vcycle_num_finished++;
  stage_out(g, t, u);
  return r2;
}

extern "C" int32_t d4b_mg_solve(d4b_mg_t *g, double *u, const double *rhs, double *r2_history) {
  TopVectors t = stage_in(g, u, rhs);
  d4b_mesh *mt = g->mesh[g->n_levels - 1];
  const int64_t n = mt->local_nodes;
  // initial residual, d4est_solver_multigrid_compute_residual (:1351-1418)
  double *Au = mt->d_Au_stage;
  d4b_apply(mt, t.u, Au);
  resid_k<<<grid_for(n), 256, 0, mt->stream>>>(t.rhs, Au, g->d_r[g->n_levels - 1], n);
  double r2 = d4b_vec_dot(mt, g->d_r[g->n_levels - 1], g->d_r[g->n_levels - 1], n);
  double r2_last = r2;
  const double stoptol = g->prm.vcycle_rtol * g->prm.vcycle_rtol * r2 + g->prm.vcycle_atol * g->prm.vcycle_atol;
  if (r2_history) r2_history[0] = r2;
  g->vcycle_num_finished = 0;
  while (g->vcycle_num_finished < g->prm.vcycle_imax && r2 > stoptol) {
    // the smoother's final residual IS rhs - A u of the new iterate: no extra apply
    r2 = vcycle_device(g, t.u, t.rhs);
    g->vcycle_num_finished++;
    if (r2_history) r2_history[g->vcycle_num_finished] = r2;
    if (sqrt(r2 / r2_last) >= .99) break;
    r2_last = r2;
  }
  stage_out(g, t, u);
  return g->vcycle_num_finished;
}

extern "C" int64_t d4b_mg_graph_replays(const d4b_mg_t *g) { return g->n_replays; }
extern "C" void d4b_mg_get_eigs(const d4b_mg_t *g, double *eigs_out) {
  D4B_CUDA(cudaMemcpy(eigs_out, g->d_eigs, (size_t)g->n_levels * sizeof(double), cudaMemcpyDeviceToHost));
}

namespace {
void transfer_any(d4b_mg *g, int dir, int lc, const double *in, double *out) {
  D4B_CHECK(lc >= 0 && lc < g->n_levels - 1, "transfer: coarse level %d out of range", lc);
  d4b_mesh *m_in = dir == 0 ? g->mesh[lc] : g->mesh[lc + 1];
  d4b_mesh *m_out = dir == 0 ? g->mesh[lc + 1] : g->mesh[lc];
  cudaStream_t s = g->mesh[lc + 1]->stream;
  const double *din = in;
  double *dout = out, *tmp_in = nullptr, *tmp_out = nullptr;
  if (!is_dev(in)) {
    D4B_CUDA(cudaMalloc(&tmp_in, (size_t)m_in->local_nodes * sizeof(double)));
    D4B_CUDA(cudaMemcpyAsync(tmp_in, in, (size_t)m_in->local_nodes * sizeof(double), cudaMemcpyHostToDevice, s));
    din = tmp_in;
  }
  if (!is_dev(out)) {
    D4B_CUDA(cudaMalloc(&tmp_out, (size_t)m_out->local_nodes * sizeof(double)));
    dout = tmp_out;
  }
  transfer(g, dir, 0, lc, din, dout);
  if (tmp_out)
    D4B_CUDA(cudaMemcpyAsync(out, tmp_out, (size_t)m_out->local_nodes * sizeof(double), cudaMemcpyDeviceToHost, s));
  D4B_CUDA(cudaStreamSynchronize(s));
  if (tmp_in) cudaFree(tmp_in);
  if (tmp_out) cudaFree(tmp_out);
}
}  // namespace

extern "C" void d4b_mg_restrict(d4b_mg_t *g, int32_t lc, const double *fine, double *coarse) {
  transfer_any(g, 1, lc, fine, coarse);
}
extern "C" void d4b_mg_prolong(d4b_mg_t *g, int32_t lc, const double *coarse, double *fine) {
  transfer_any(g, 0, lc, coarse, fine);
}
