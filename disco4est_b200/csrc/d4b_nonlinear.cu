// d4b_nonlinear.cu -- mass-type nonlinear terms f(x, u) = a(x) (b(x) + u)^k as device functors,
// the build_residual callback, the Newton linearisation and the in-house Newton loop.
//
// Reference being replaced (all host code there):
//   d4est_quadrature_apply_fofufofvlj / _fofufofvlilj      src/Quadrature/d4est_quadrature.c:776-936, :593-774
//   multi_punctures_build_residual / _apply_jac            src/Problems/TwoPunctures/multi_punctures_fcns_with_opt.h:546-750
//   Aij_fcn, AijAij_fcn, psi_fcn (the puncture functor)    same file :292-384
//   neg_2pi_rho_up1_neg5 / neg_10pi_rho_up1_neg4           src/Problems/ConstantDensityStar/constant_density_star_fcns.h:334-360
//   d4est_solver_newton_solve                              src/Solver/d4est_solver_newton.c:136-367
#include <cmath>
#include <cstring>
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

struct NlParams {
  int kind, n_punctures;
  double bh[D4B_MAX_PUNCTURES][10];
  double rho0, R, c[3];
};

// One thread per quadrature point: a(x), b(x) of the functor; in: wja = w J, out: wja = w J a, nb = b
__global__ void nl_coeff_k(NlParams P, const double *__restrict__ xyz, int64_t LQ, double *__restrict__ wja,
                           double *__restrict__ nb) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < LQ; i += (int64_t)gridDim.x * blockDim.x) {
    const double x = xyz[i], y = xyz[LQ + i], z = xyz[2 * LQ + i];
    double a = 0., b = 0.;
    if (P.kind == D4B_NL_PUNCTURES) {
      // conformal factor 1 + sum m_n / (2 r_n) (+ u), extrinsic curvature of Bowen-York punctures
      double A[3][3] = {{0., 0., 0.}, {0., 0., 0.}, {0., 0., 0.}};
      double s = 0.;
      for (int n = 0; n < P.n_punctures; ++n) {
        const double *h = P.bh[n];
        const double d[3] = {x - h[1], y - h[2], z - h[3]};
        const double r = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        const double nv[3] = {d[0] / r, d[1] / r, d[2] / r};
        const double *Pm = h + 4, *S = h + 7;
        const double sxn[3] = {-nv[2] * S[1] + nv[1] * S[2], nv[2] * S[0] - nv[0] * S[2], -nv[1] * S[0] + nv[0] * S[1]};
        const double pdn = Pm[0] * nv[0] + Pm[1] * nv[1] + Pm[2] * nv[2];
        for (int ia = 0; ia < 3; ++ia)
          for (int ib = 0; ib < 3; ++ib) {
            const double g = (ia == ib) ? 1. : 0.;
            const double t1 = (3. / (2. * r * r)) * (Pm[ia] * nv[ib] + Pm[ib] * nv[ia] - (g - nv[ia] * nv[ib]) * pdn);
            const double t2 = (3. / (r * r * r)) * (sxn[ia] * nv[ib] + sxn[ib] * nv[ia]);
            A[ia][ib] += t1 + t2;
          }
        s += h[0] / (2. * r);
      }
      double K2 = 0.;
      for (int ia = 0; ia < 3; ++ia)
        for (int ib = 0; ib < 3; ++ib) K2 += A[ia][ib] * A[ia][ib];
      a = (-1. / 8.) * K2;
      b = 1. + s;
    } else {   // D4B_NL_CONSTANT_DENSITY_STAR
      const double dx = x - P.c[0], dy = y - P.c[1], dz = z - P.c[2];
      const double rho = (dx * dx + dy * dy + dz * dz > P.R * P.R) ? 0. : P.rho0;
      a = (-2. * M_PI) * rho;
      b = 0.;
    }
    wja[i] *= a;
    nb[i] = b;
  }
}

__global__ void mul_k(const double *__restrict__ a, double *__restrict__ y, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] *= a[i];
}
// F = Au + N - rhs (N, rhs optional)
__global__ void resid_sum_k(const double *__restrict__ N, const double *__restrict__ rhs, double *__restrict__ F,
                            int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    double v = F[i];
    if (N) v += N[i];
    if (rhs) v -= rhs[i];
    F[i] = v;
  }
}
__global__ void neg_k(double *__restrict__ x, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    x[i] = -1. * x[i];
}
// z = y ; z += a x   (d4est_linalg_vec_axpyeqz, d4est_linalg.c:217-221)
__global__ void axpyeqz_k(double a, const double *__restrict__ x, const double *__restrict__ y, double *__restrict__ z,
                          int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    z[i] = y[i] + a * x[i];
}

struct DevTmp {   // device view of a host-or-device input
  const double *p = nullptr;
  double *own = nullptr;
  DevTmp(d4b_mesh *m, const double *x, size_t n) {
    if (!x) return;
    if (is_dev(x)) { p = x; return; }
    D4B_CUDA(cudaMalloc(&own, std::max<size_t>(n, 1) * sizeof(double)));
    D4B_CUDA(cudaMemcpyAsync(own, x, n * sizeof(double), cudaMemcpyHostToDevice, m->stream));
    p = own;
  }
  ~DevTmp() { if (own) cudaFree(own); }
};

void drop_nonlinear(d4b_mesh *m) {
  const int64_t bytes = m->local_nodes_quad * (int64_t)sizeof(double);
  if (m->d_nl_wja) { cudaFree(m->d_nl_wja); m->d_nl_wja = nullptr; m->device_bytes -= bytes; }
  if (m->d_nl_b) { cudaFree(m->d_nl_b); m->d_nl_b = nullptr; m->device_bytes -= bytes; }
}

// F = A(u) + N(u) - rhs on device pointers, stream-ordered; the mesh's linearisation is masked
void residual_device(d4b_mesh *m, const double *du, const double *drhs, double *dF, double *dtmp) {
  const int64_t n = m->local_nodes;
  // every element in a Gauss-basis kernel: the functor is evaluated at the quadrature points INSIDE the apply (uh is
  // already there), like the linear term of the Jacobian -- one pass, no quadrature kernel.  D4B_FUSED_RESIDUAL=0 or
  // elements in the general kernel: apply + quadrature kernel.
  static const bool fuse = !(getenv("D4B_FUSED_RESIDUAL") && atoi(getenv("D4B_FUSED_RESIDUAL")) == 0);
  if (fuse && m->d_nl_wja && m->n_generic == 0 && !m->fast.empty() && !m->pre_apply) {
    m->epi = D4bEpilogue();
    m->nl_fused = true;
    d4b_apply_device(m, du, dF);
    m->nl_fused = false;
    if (drhs) resid_sum_k<<<grid_for(n), 256, 0, m->stream>>>(nullptr, drhs, dF, n);
    return;
  }
  double *W = m->d_W;
  m->d_W = nullptr;
  m->epi = D4bEpilogue();
  d4b_apply_device(m, du, dF);
  m->d_W = W;
  if (m->d_nl_wja) d4b_launch_quad_nl(m, 5, du, dtmp);
  if (m->d_nl_wja || drhs) resid_sum_k<<<grid_for(n), 256, 0, m->stream>>>(m->d_nl_wja ? dtmp : nullptr, drhs, dF, n);
}

void linearize_device(d4b_mesh *m, const double *du0) {
  D4B_CHECK(m->d_nl_wja, "d4b_mesh_linearize: no nonlinear term attached to this mesh");
  // recorded V-cycle graphs carry the POINTER of W and the kernel variants chosen for a mesh with a mass term: a new
  // linearisation point changes the contents only, the graphs stay valid (a Newton solve re-linearises every step)
  if (!m->d_W) { m->d_W = d4b_dev_alloc<double>(m, (size_t)m->local_nodes_quad); m->config_epoch++; }
  d4b_launch_quad_nl(m, 6, du0, m->d_W);
  m->w_version++;
}

}  // namespace

extern "C" void d4b_mesh_set_nonlinear_term(d4b_mesh_t *m, const d4b_nonlinear_desc_t *d) {
  D4B_CHECK(m, "d4b_mesh_set_nonlinear_term: NULL mesh");
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  drop_nonlinear(m);
  if (!d) return;
  const int64_t LQ = m->local_nodes_quad;
  m->d_nl_wja = d4b_dev_alloc<double>(m, (size_t)LQ);
  m->d_nl_b = d4b_dev_alloc<double>(m, (size_t)LQ);
  if (d->kind == D4B_NL_TABLE) {
    D4B_CHECK(d->a_quad && d->b_quad && d->jac_quad, "d4b_mesh_set_nonlinear_term: D4B_NL_TABLE needs a_quad, b_quad, jac_quad");
    DevTmp a(m, d->a_quad, LQ), b(m, d->b_quad, LQ), jac(m, d->jac_quad, LQ);
    d4b_launch_quad_op(m, 3, nullptr, jac.p, m->d_nl_wja);
    mul_k<<<grid_for(LQ), 256, 0, m->stream>>>(a.p, m->d_nl_wja, LQ);
    D4B_CUDA(cudaMemcpyAsync(m->d_nl_b, b.p, (size_t)LQ * sizeof(double), cudaMemcpyDeviceToDevice, m->stream));
    m->nl_k = d->exponent;
    D4B_CUDA(cudaStreamSynchronize(m->stream));
    return;
  }
  D4B_CHECK(d->kind == D4B_NL_PUNCTURES || d->kind == D4B_NL_CONSTANT_DENSITY_STAR,
            "d4b_mesh_set_nonlinear_term: unknown kind %d", d->kind);
  NlParams P;
  memset(&P, 0, sizeof(P));
  P.kind = d->kind;
  if (d->kind == D4B_NL_PUNCTURES) {
    D4B_CHECK(d->n_punctures >= 1 && d->n_punctures <= D4B_MAX_PUNCTURES, "d4b_mesh_set_nonlinear_term: %d punctures",
              d->n_punctures);
    P.n_punctures = d->n_punctures;
    memcpy(P.bh, d->punctures, sizeof(P.bh));
    m->nl_k = -7;
  } else {
    P.rho0 = d->star_rho0; P.R = d->star_R;
    for (int i = 0; i < 3; ++i) P.c[i] = d->star_c[i];
    m->nl_k = 5;
  }
  // positions and Jacobian at the quadrature points: the caller's (reference) arrays, or regenerated from the
  // map for meshes created with device geometry
  double *own_xyz = nullptr, *own_jac = nullptr;
  const double *xyz = nullptr, *jac = nullptr;
  DevTmp cx(m, d->xyz_quad, (size_t)3 * LQ), cj(m, d->jac_quad, LQ);
  if (d->xyz_quad && d->jac_quad) { xyz = cx.p; jac = cj.p; }
  else {
    D4B_CHECK(m->has_geom, "d4b_mesh_set_nonlinear_term: xyz_quad / jac_quad missing and the mesh has no device geometry");
    D4B_CUDA(cudaMalloc(&own_xyz, std::max<size_t>((size_t)3 * LQ, 1) * sizeof(double)));
    D4B_CUDA(cudaMalloc(&own_jac, std::max<size_t>((size_t)LQ, 1) * sizeof(double)));
    d4b_mesh_desc_t t;
    memset(&t, 0, sizeof(t));
    t.abi_version = D4B_ABI_VERSION;
    t.n_local = m->n_local;
    t.local_nodes = m->local_nodes; t.local_nodes_quad = LQ;
    t.deg = m->h_deg.data(); t.deg_quad = m->h_deg_quad.data();
    t.nodal_stride = m->h_nodal_stride.data(); t.quad_stride = m->h_quad_stride.data();
    d4b_geom_desc_t g = m->gdesc;
    g.elem_tree = m->g_tree.data(); g.elem_q = m->g_q.data(); g.elem_dq = m->g_dq.data();
    g.ghost_tree = g.ghost_q = g.ghost_dq = nullptr;
    d4b_geom_arrays_t out;
    memset(&out, 0, sizeof(out));
    out.xyz_quad = own_xyz; out.J_quad = own_jac;
    d4b_geom_compute(&t, &g, &out);
    xyz = own_xyz; jac = own_jac;
  }
  d4b_launch_quad_op(m, 3, nullptr, jac, m->d_nl_wja);
  nl_coeff_k<<<grid_for(LQ), 256, 0, m->stream>>>(P, xyz, LQ, m->d_nl_wja, m->d_nl_b);
  D4B_CUDA(cudaGetLastError());
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  if (own_xyz) cudaFree(own_xyz);
  if (own_jac) cudaFree(own_jac);
}

extern "C" void d4b_mesh_linearize(d4b_mesh_t *m, const double *u0) {
  D4B_CHECK(m, "d4b_mesh_linearize: NULL mesh");
  if (!u0) { d4b_mesh_set_mass_diag(m, nullptr); return; }
  DevTmp u(m, u0, (size_t)m->local_nodes);
  linearize_device(m, u.p);
  if (u.own) D4B_CUDA(cudaStreamSynchronize(m->stream));
}

extern "C" void d4b_build_residual(d4b_mesh_t *m, const double *u, const double *rhs, double *F) {
  D4B_CHECK(m && u && F, "d4b_build_residual: NULL argument");
  const size_t n = (size_t)m->local_nodes;
  DevTmp du(m, u, n), dr(m, rhs, n);
  double *dF = F, *own = nullptr, *tmp = nullptr;
  const bool F_dev = is_dev(F);
  if (!F_dev) { D4B_CUDA(cudaMalloc(&own, std::max<size_t>(n, 1) * sizeof(double))); dF = own; }
  if (m->d_nl_wja) {   // scratch of the mesh (shared with the smoother, never live at the same time)
    if (!m->d_work[2]) m->d_work[2] = d4b_dev_alloc<double>(m, n);
    tmp = m->d_work[2];
  }
  if (m->mpisize > 1) D4B_CHECK(m->d_ughost || m->n_ghost == 0, "d4b_build_residual: ghost buffer missing");
  residual_device(m, du.p, dr.p, dF, tmp);
  if (own) D4B_CUDA(cudaMemcpyAsync(F, own, n * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
  // stream-ordered for device pointers (d4b_sync to wait), synchronous otherwise
  if (own || du.own || dr.own) D4B_CUDA(cudaStreamSynchronize(m->stream));
  if (own) cudaFree(own);
}

extern "C" int32_t d4b_newton_solve(d4b_mesh_t *mr, d4b_mesh_t *mj, d4b_mg_t *pc, double *u,
                                    const d4b_newton_params_t *prm, d4b_newton_stats_t *st) {
  D4B_CHECK(mr && mj && u && prm, "d4b_newton_solve: NULL argument");
  D4B_CHECK(mr->local_nodes == mj->local_nodes, "d4b_newton_solve: residual and Jacobian meshes differ");
  D4B_CHECK(mj->d_nl_wja, "d4b_newton_solve: the Jacobian mesh has no nonlinear term (nothing to linearise)");
  const int64_t n = mr->local_nodes;
  const size_t bytes = (size_t)n * sizeof(double);
  const int gr = grid_for(n);
  cudaStream_t s = mj->stream;
  if (mr->stream != s) D4B_CUDA(cudaStreamSynchronize(mr->stream));
  std::vector<double *> own;
  auto vec = [&]() { double *p = nullptr; D4B_CUDA(cudaMalloc(&p, std::max<size_t>(bytes, 8))); own.push_back(p); return p; };
  const bool u_dev = is_dev(u);
  double *x = u_dev ? u : vec();
  if (!u_dev) D4B_CUDA(cudaMemcpyAsync(x, u, bytes, cudaMemcpyHostToDevice, s));
  double *const x_user = x;
  double *xt = vec(), *f0 = vec(), *ft = vec(), *step = vec(), *tmp = vec();
  cudaStream_t saved = mr->stream;
  mr->stream = s;   // one stream for both meshes during the solve
  d4b_newton_stats_t local;
  if (!st) st = &local;
  memset(st, 0, sizeof(*st));

  residual_device(mr, x, nullptr, f0, tmp);
  double fnrm = sqrt(d4b_vec_dot(mr, f0, f0, n));
  st->residual_norm[0] = fnrm;
  const double stop_tol = prm->atol + prm->rtol * fnrm;
  int itc = 0;
  while ((fnrm > stop_tol || itc < prm->imin) && itc < prm->imax) {
    D4B_CHECK(itc < 64, "d4b_newton_solve: more than 64 Newton steps");
    neg_k<<<gr, 256, 0, s>>>(f0, n);
    linearize_device(mj, x);                                         // u0 = x: the only use of x by the linear solve
    D4B_CUDA(cudaMemsetAsync(step, 0, bytes, s));
    const int its = d4b_fcg_solve(mj, pc, step, f0, prm->fcg_imax, prm->fcg_rtol, prm->fcg_atol, prm->fcg_flags, nullptr);
    st->fcg_iterations[itc] = its;
    st->jacobian_applies += its + 1;
    axpyeqz_k<<<gr, 256, 0, s>>>(1.0, step, x, xt, n);               // xt = x + step
    residual_device(mr, xt, nullptr, ft, tmp);
    std::swap(x, xt);
    std::swap(f0, ft);
    fnrm = sqrt(d4b_vec_dot(mr, f0, f0, n));
    ++itc;
    st->residual_norm[itc] = fnrm;
  }
  st->steps = itc;
  if (x != x_user) D4B_CUDA(cudaMemcpyAsync(x_user, x, bytes, cudaMemcpyDeviceToDevice, s));
  if (!u_dev) D4B_CUDA(cudaMemcpyAsync(u, x_user, bytes, cudaMemcpyDeviceToHost, s));
  D4B_CUDA(cudaStreamSynchronize(s));
  mr->stream = saved;
  for (double *p : own) cudaFree(p);
  return fnrm > stop_tol ? 1 : 0;
}
