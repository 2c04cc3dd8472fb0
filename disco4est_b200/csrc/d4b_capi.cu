// d4b_capi.cu -- extern "C" entry points of libd4est_b200 (include/d4est_b200.h).
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <vector>
#include "d4b_internal.h"

namespace {
d4b_abort_hook_t g_abort_hook = nullptr;
}

void d4b_abort_impl(const char *file, int line, const char *fmt, ...) {
  char msg[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(msg, sizeof(msg), fmt, ap);
  va_end(ap);
  if (g_abort_hook) g_abort_hook(file, line, msg);
  // reference convention: D4EST_ABORT -> SC_ABORT, never returns
  // (src/Utilities/d4est_util.h:171)
  fprintf(stderr, "[D4B_ABORT] %s:%d %s\n", file, line, msg);
  fflush(stderr);
  abort();
}

extern "C" void d4b_set_abort_hook(d4b_abort_hook_t hook) { g_abort_hook = hook; }
extern "C" int d4b_abi_version(void) { return D4B_ABI_VERSION; }

extern "C" int d4b_device_count(void) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

extern "C" void d4b_init(int device) {
  D4B_CHECK(d4b_device_count() > 0, "d4b_init: no CUDA device (there is no CPU fallback)");
  D4B_CUDA(cudaSetDevice(device));
  D4B_CUDA(cudaFree(0));
}

namespace {
bool is_device_ptr(const void *p) {
  cudaPointerAttributes at;
  cudaError_t e = cudaPointerGetAttributes(&at, p);
  if (e != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

}  // namespace

namespace {
__global__ void add_k(const double *__restrict__ x, double *__restrict__ y, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] += x[i];
}
}  // namespace

// du == nullptr: the Gauss-collocated input is already in m->d_uhat (the Gauss-resident smoother, d4b_solver.cu); with
// the peer halo the caller has also begun this apply's epoch (d4b_comm_peer_begin)
void d4b_apply_device(d4b_mesh *m, const double *du, double *dAu) {
  D4B_CHECK(m->epi.mode == 0 || m->n_generic == 0, "fused epilogue on a mesh with general-kernel elements");
  const bool have_uh = du == nullptr;
  D4B_CHECK(!have_uh || (m->n_generic == 0 && !m->pre_apply), "apply from uh on a mesh that needs the nodal input");
  // coarse multigrid level of a linearised problem: the Galerkin mass term of this input first
  const bool add = m->pre_apply && m->pre_apply(m, du);
  const bool add_after = add && (m->n_generic > 0 || m->fast.empty());
  if (add && !add_after) m->epi.add = m->d_addterm;
  if (m->mpisize == 1) {
    if (!have_uh) d4b_launch_fast(m, du, dAu, 0);
    d4b_launch_fast(m, du, dAu, 1);
  } else if (d4b_comm_peer_ready(m)) {
    // halo over peer memory (d4b_comm.cu): nothing is packed or sent; the boundary elements read the
    // neighbour ranks' uh over NVLink once those ranks have signalled that it is complete
    // compute stream:  uh(local) | interior elements ........................... | join
    // halo stream   :        wait uh(local) | signal neighbours | wait for theirs | boundary elements
    if (!have_uh) d4b_comm_peer_begin(m);
    // small meshes (coarse multigrid levels): an interior launch of a few microseconds hides nothing, and the two
    // cross-stream dependencies cost more than it runs -- everything in order on the compute stream
    static const int64_t split_min = getenv("D4B_PEER_SPLIT_MIN_NODES") ? atoll(getenv("D4B_PEER_SPLIT_MIN_NODES")) : (int64_t)1 << 20;
    if (m->local_nodes < split_min && d4b_fast_can_merge_halo(m)) {
      // every class in a kernel that waits by itself: ONE launch per class on the compute stream; the element groups
      // with a ghost neighbour come last in the class order and find the neighbours' flags set when they run
      // (no second stream, no event, no wait kernel, no extra launch)
      if (!have_uh) d4b_launch_fast(m, du, dAu, 0);
      d4b_comm_peer_signal(m, m->stream);
      d4b_launch_fast(m, du, dAu, 3);
    } else if (m->local_nodes < split_min) {
      if (!have_uh) d4b_launch_fast(m, du, dAu, 0);
      d4b_comm_peer_signal(m, m->stream);
      d4b_launch_fast(m, du, dAu, 1);
      d4b_comm_peer_wait(m, m->stream);
      d4b_launch_fast(m, du, dAu, 2);
    } else {
      // D4B_HALO_TIMING=1: event time stamps of the phases of every apply, averages printed by d4b_mesh_destroy
      static const bool timing = getenv("D4B_HALO_TIMING") && atoi(getenv("D4B_HALO_TIMING")) != 0;
      if (timing && m->halo_ev.empty()) {
        m->halo_ev.resize(6);
        for (cudaEvent_t &e : m->halo_ev) D4B_CUDA(cudaEventCreate(&e));
      } else if (timing) {   // the previous apply's stamps
        D4B_CUDA(cudaEventSynchronize(m->halo_ev[5]));
        for (int i = 0; i < 5; ++i) {
          float ms = 0.f;
          D4B_CUDA(cudaEventElapsedTime(&ms, m->halo_ev[0], m->halo_ev[i + 1]));
          m->halo_t[i] += ms;
        }
        ++m->halo_n;
      }
      if (timing) D4B_CUDA(cudaEventRecord(m->halo_ev[0], m->stream));      // 0: apply begins
      if (!have_uh) d4b_launch_fast(m, du, dAu, 0);
      D4B_CUDA(cudaEventRecord(m->ev_interp, m->stream));
      if (timing) D4B_CUDA(cudaEventRecord(m->halo_ev[1], m->stream));      // 1: uh(local) done
      D4B_CUDA(cudaStreamWaitEvent(m->comm_stream, m->ev_interp, 0));
      d4b_comm_peer_signal(m, m->comm_stream);
      d4b_launch_fast(m, du, dAu, 1);
      if (timing) D4B_CUDA(cudaEventRecord(m->halo_ev[2], m->stream));      // 2: interior elements done
      d4b_comm_peer_wait(m, m->comm_stream);
      if (timing) D4B_CUDA(cudaEventRecord(m->halo_ev[3], m->comm_stream)); // 3: neighbours have signalled
      cudaStream_t compute = m->stream;
      m->stream = m->comm_stream;
      d4b_launch_fast(m, du, dAu, 2);
      D4B_CUDA(cudaEventRecord(m->ev_boundary, m->comm_stream));
      if (timing) D4B_CUDA(cudaEventRecord(m->halo_ev[4], m->comm_stream)); // 4: boundary elements done
      m->stream = compute;
      D4B_CUDA(cudaStreamWaitEvent(m->stream, m->ev_boundary, 0));
      if (timing) D4B_CUDA(cudaEventRecord(m->halo_ev[5], m->stream));      // 5: joined
    }
  } else {
    // compute stream:  pack | uh(local) | interior elements ............ | join
    // halo stream   :       wait pack | send/recv | wait uh(local) | uh(ghosts) | boundary elements
    D4B_CHECK(!have_uh, "apply from uh needs the peer-memory halo");
    d4b_comm_exchange_begin(m, du);
    d4b_launch_fast(m, du, dAu, 0);
    D4B_CUDA(cudaEventRecord(m->ev_interp, m->stream));
    d4b_launch_fast(m, du, dAu, 1);
    cudaStream_t compute = m->stream;
    m->stream = m->comm_stream;
    D4B_CUDA(cudaStreamWaitEvent(m->comm_stream, m->ev_interp, 0));
    d4b_launch_fast(m, du, dAu, 2);
    D4B_CUDA(cudaEventRecord(m->ev_boundary, m->comm_stream));
    m->stream = compute;
    D4B_CUDA(cudaStreamWaitEvent(m->stream, m->ev_boundary, 0));
  }
  d4b_launch_generic(m, du, dAu, 0);
  if (add_after) {
    const int64_t n = m->local_nodes;
    add_k<<<(int)std::min<int64_t>((n + 255) / 256, 148 * 8), 256, 0, m->stream>>>(m->d_addterm, dAu, n);
  }
  m->epi.add = nullptr;
}

namespace {
// Host-pointer apply, pipelined: u goes up in chunks of consecutive elements, a chunk is applied as
// soon as the chunks holding its face neighbours have arrived (their uh is computed on arrival), and
// its part of Au goes down while later chunks are still going up -- the two copy engines run
// concurrently instead of one after the other.
void apply_host_pipelined(d4b_mesh *m, const double *u, double *Au) {
  const int K = (int)m->chunk_dep.size();
  if (!m->h2d_stream) {
    D4B_CUDA(cudaStreamCreateWithFlags(&m->h2d_stream, cudaStreamNonBlocking));
    D4B_CUDA(cudaStreamCreateWithFlags(&m->d2h_stream, cudaStreamNonBlocking));
    m->ev_h2d.resize(K); m->ev_done.resize(K);
    for (int k = 0; k < K; ++k) {
      D4B_CUDA(cudaEventCreateWithFlags(&m->ev_h2d[k], cudaEventDisableTiming));
      D4B_CUDA(cudaEventCreateWithFlags(&m->ev_done[k], cudaEventDisableTiming));
    }
  }
  D4B_CUDA(cudaStreamSynchronize(m->stream));   // the staging vectors are free
  m->epi = D4bEpilogue();
  double *du = m->d_u_stage, *dAu = m->d_Au_stage;
  // D4B_PIPE_TIMING=1: when the upload ended, when the first / last part of Au went down (ms after the first copy began)
  static const bool timing = getenv("D4B_PIPE_TIMING") && atoi(getenv("D4B_PIPE_TIMING")) != 0;
  static cudaEvent_t tev[4] = {nullptr, nullptr, nullptr, nullptr};
  if (timing && !tev[0]) for (cudaEvent_t &e : tev) D4B_CUDA(cudaEventCreate(&e));
  if (timing) D4B_CUDA(cudaEventRecord(tev[0], m->h2d_stream));
  bool first_down = true;
  // (one stream per direction: alternating the chunks between two streams per direction, so that the next copy is
  //  already queued at the copy engine when the current one ends, is slower -- 4.25 vs 3.96 ms)
  for (int i = 0; i < K; ++i) {
    const int k = m->chunk_order[i];
    const int64_t n0 = m->chunk_node[k], n1 = m->chunk_node[k + 1];
    if (n1 > n0)
      D4B_CUDA(cudaMemcpyAsync(du + n0, u + n0, (size_t)(n1 - n0) * sizeof(double), cudaMemcpyHostToDevice, m->h2d_stream));
    D4B_CUDA(cudaEventRecord(m->ev_h2d[i], m->h2d_stream));
  }
  if (timing) D4B_CUDA(cudaEventRecord(tev[1], m->h2d_stream));
  std::vector<char> launched(K, 0);
  for (int i = 0; i < K; ++i) {
    const int k = m->chunk_order[i];
    D4B_CUDA(cudaStreamWaitEvent(m->stream, m->ev_h2d[i], 0));
    d4b_launch_fast_range(m, du, dAu, 0, m->chunk_elem[k], m->chunk_elem[k + 1]);
    for (int j = 0; j <= i; ++j) {
      const int c = m->chunk_order[j];
      if (launched[c] || m->chunk_dep[c] > i) continue;
      launched[c] = 1;
      d4b_launch_fast_range(m, du, dAu, 1, m->chunk_elem[c], m->chunk_elem[c + 1]);
      D4B_CUDA(cudaEventRecord(m->ev_done[c], m->stream));
      cudaStream_t ds = m->d2h_stream;
      D4B_CUDA(cudaStreamWaitEvent(ds, m->ev_done[c], 0));
      const int64_t n0 = m->chunk_node[c], n1 = m->chunk_node[c + 1];
      if (timing && first_down) { D4B_CUDA(cudaEventRecord(tev[2], ds)); first_down = false; }
      if (n1 > n0)
        D4B_CUDA(cudaMemcpyAsync(Au + n0, dAu + n0, (size_t)(n1 - n0) * sizeof(double), cudaMemcpyDeviceToHost, ds));
    }
  }
  if (timing) D4B_CUDA(cudaEventRecord(tev[3], m->d2h_stream));
  D4B_CUDA(cudaStreamSynchronize(m->d2h_stream));
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  if (timing) {
    float t1 = 0.f, t2 = 0.f, t3 = 0.f;
    cudaEventElapsedTime(&t1, tev[0], tev[1]); cudaEventElapsedTime(&t2, tev[0], tev[2]); cudaEventElapsedTime(&t3, tev[0], tev[3]);
    fprintf(stderr, "[d4b] pipelined host apply, %d chunks: upload done %.3f ms | first download starts %.3f | last download done %.3f\n",
            K, t1, t2, t3);
  }
}

void apply_impl(d4b_mesh *m, const double *u, double *Au, int stiffness_only) {
  D4B_CHECK(m && u && Au, "d4b_apply: NULL argument");
  d4b_check_device(m);
  D4B_CHECK(m->n_ghost == 0 || m->d_ughost, "d4b_apply: ghost buffer missing");
  bool u_dev = is_device_ptr(u), Au_dev = is_device_ptr(Au);
  if (!u_dev && !Au_dev && !stiffness_only && !m->chunk_dep.empty() && u != Au && !m->pre_apply) {
    apply_host_pipelined(m, u, Au);
    return;
  }
  const double *du = u;
  double *dAu = Au;
  size_t bytes = (size_t)m->local_nodes * sizeof(double);
  if (!u_dev) {
    D4B_CUDA(cudaMemcpyAsync(m->d_u_stage, u, bytes, cudaMemcpyHostToDevice, m->stream));
    du = m->d_u_stage;
  }
  if (!Au_dev) dAu = m->d_Au_stage;
  if (stiffness_only) {
    d4b_launch_generic(m, du, dAu, 1);
  } else {
    m->epi = D4bEpilogue();
    d4b_apply_device(m, du, dAu);
  }
  if (!Au_dev) D4B_CUDA(cudaMemcpyAsync(Au, dAu, bytes, cudaMemcpyDeviceToHost, m->stream));
  if (!u_dev || !Au_dev) D4B_CUDA(cudaStreamSynchronize(m->stream));
}
}  // namespace

extern "C" void d4b_apply(d4b_mesh_t *m, const double *u, double *Au) { apply_impl(m, u, Au, 0); }
extern "C" void d4b_apply_stiffness(d4b_mesh_t *m, const double *u, double *Au) { apply_impl(m, u, Au, 1); }
// ---- volume quadrature operators ------------------------------------------------------
namespace {
// device view of a host-or-device array of n doubles (temporary device copy for host data)
struct DevIn {
  const double *p = nullptr;
  double *tmp = nullptr;
  DevIn(d4b_mesh *m, const double *x, int64_t n) {
    if (!x) return;
    if (is_device_ptr(x)) { p = x; return; }
    D4B_CUDA(cudaMalloc(&tmp, (size_t)n * sizeof(double)));
    D4B_CUDA(cudaMemcpyAsync(tmp, x, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, m->stream));
    p = tmp;
  }
  ~DevIn() { if (tmp) cudaFree(tmp); }
};
struct DevOut {
  double *p = nullptr, *tmp = nullptr, *host = nullptr;
  int64_t n = 0;
  DevOut(d4b_mesh *, double *x, int64_t n_) : n(n_) {
    if (is_device_ptr(x)) { p = x; return; }
    D4B_CUDA(cudaMalloc(&tmp, (size_t)n * sizeof(double)));
    p = tmp; host = x;
  }
  void finish(d4b_mesh *m) {
    if (tmp) D4B_CUDA(cudaMemcpyAsync(host, tmp, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    if (tmp) D4B_CUDA(cudaStreamSynchronize(m->stream));
  }
  ~DevOut() { if (tmp) cudaFree(tmp); }
};
}  // namespace

extern "C" void d4b_interp_to_quad(d4b_mesh_t *m, const double *u, double *u_quad) {
  D4B_CHECK(m && u && u_quad, "d4b_interp_to_quad: NULL argument");
  DevIn in(m, u, m->local_nodes);
  DevOut out(m, u_quad, m->local_nodes_quad);
  d4b_launch_quad_op(m, 0, in.p, nullptr, out.p);
  out.finish(m);
  if (in.tmp) D4B_CUDA(cudaStreamSynchronize(m->stream));
}

extern "C" void d4b_apply_galerkin(d4b_mesh_t *m, const double *in_quad, const double *jac_quad, double *out) {
  D4B_CHECK(m && in_quad && out, "d4b_apply_galerkin: NULL argument");
  DevIn in(m, in_quad, m->local_nodes_quad), jac(m, jac_quad, m->local_nodes_quad);
  DevOut o(m, out, m->local_nodes);
  d4b_launch_quad_op(m, 1, in.p, jac.p, o.p);
  o.finish(m);
  if (in.tmp || jac.tmp) D4B_CUDA(cudaStreamSynchronize(m->stream));
}

extern "C" void d4b_apply_mass(d4b_mesh_t *m, const double *x, const double *jac_quad, double *out) {
  D4B_CHECK(m && x && jac_quad && out, "d4b_apply_mass: NULL argument");
  DevIn in(m, x, m->local_nodes), jac(m, jac_quad, m->local_nodes_quad);
  DevOut o(m, out, m->local_nodes);
  d4b_launch_quad_op(m, 2, in.p, jac.p, o.p);
  o.finish(m);
  if (in.tmp || jac.tmp) D4B_CUDA(cudaStreamSynchronize(m->stream));
}

// rhs = M f - A(0), d4est_laplacian_build_rhs_with_strong_bc (src/dGMath/d4est_laplacian.c:15-140):
// the operator applied to u = 0 leaves the inhomogeneous Dirichlet terms of the boundary faces
extern "C" void d4b_build_rhs(d4b_mesh_t *m, const double *f, int32_t f_on_quad, const double *jac_quad, double *rhs) {
  D4B_CHECK(m && f && jac_quad && rhs, "d4b_build_rhs: NULL argument");
  DevIn in(m, f, f_on_quad ? m->local_nodes_quad : m->local_nodes), jac(m, jac_quad, m->local_nodes_quad);
  DevOut o(m, rhs, m->local_nodes);
  D4B_CUDA(cudaMemsetAsync(m->d_u_stage, 0, (size_t)m->local_nodes * sizeof(double), m->stream));
  m->epi = D4bEpilogue();
  d4b_apply_device(m, m->d_u_stage, m->d_Au_stage);
  d4b_launch_quad_op(m, f_on_quad ? 1 : 2, in.p, jac.p, o.p);
  d4b_vec_axpy(m, -1., m->d_Au_stage, o.p, m->local_nodes);
  o.finish(m);
  D4B_CUDA(cudaStreamSynchronize(m->stream));
}

extern "C" void d4b_estimator_bi(d4b_mesh_t *m, const d4b_estimator_desc_t *d, double *eta2, double *terms) {
  D4B_CHECK(m && d && eta2, "d4b_estimator_bi: NULL argument");
  D4B_CHECK(d->u && d->residual && d->jac_quad && d->diam_volume && d->sj_m_mortar_quad && d->hm_mortar_quad &&
            d->hp_mortar_quad, "d4b_estimator_bi: missing input array");
  D4B_CHECK(d->penalty_family >= D4B_EST_HOUSTON_MAXP_MINH && d->penalty_family <= D4B_EST_BI_MAX_P2_OVER_H,
            "d4b_estimator_bi: unknown penalty family %d", d->penalty_family);
  DevIn u(m, d->u, m->local_nodes), r(m, d->residual, m->local_nodes), jac(m, d->jac_quad, m->local_nodes_quad);
  DevIn diam(m, d->diam_volume, m->n_local), sj(m, d->sj_m_mortar_quad, m->mortar_nodes);
  DevIn hm(m, d->hm_mortar_quad, m->mortar_nodes), hp(m, d->hp_mortar_quad, m->mortar_nodes);
  DevIn g(m, d->dirichlet_g_mortar_quad, m->mortar_nodes);
  DevOut o(m, eta2, m->n_local);
  double *tp = nullptr, *ttmp = nullptr;
  if (terms) {
    if (is_device_ptr(terms)) tp = terms;
    else { D4B_CUDA(cudaMalloc(&ttmp, (size_t)4 * std::max(m->n_local, 1) * sizeof(double))); tp = ttmp; }
  }
  if (m->mpisize > 1) {   // ghost values of u (d4est_ghost_data_exchange, d4est_estimator_bi.c:454)
    d4b_comm_exchange_begin(m, u.p);
    d4b_comm_exchange_wait(m);
  }
  d4b_launch_estimator(m, u.p, r.p, jac.p, diam.p, sj.p, hm.p, hp.p, g.p, d->penalty_family, d->penalty_prefactor,
                       o.p, tp);
  o.finish(m);
  if (ttmp) {
    D4B_CUDA(cudaMemcpyAsync(terms, ttmp, (size_t)4 * m->n_local * sizeof(double), cudaMemcpyDeviceToHost, m->stream));
  }
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  if (ttmp) cudaFree(ttmp);
}

extern "C" void d4b_mesh_set_mass_diag(d4b_mesh_t *m, const double *jac_quad) {
  D4B_CHECK(m, "d4b_mesh_set_mass_diag: NULL mesh");
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  // recorded V-cycle graphs of this mesh go stale when the term appears or disappears (other kernel variants, another
  // pointer); new values in the same array leave them valid
  if (!jac_quad) {
    if (m->d_W) {
      cudaFree(m->d_W); m->d_W = nullptr; m->device_bytes -= m->local_nodes_quad * (int64_t)sizeof(double);
      m->config_epoch++;
      m->w_version++;
    }
    return;
  }
  if (!m->d_W) { m->d_W = d4b_dev_alloc<double>(m, (size_t)m->local_nodes_quad); m->config_epoch++; }
  m->w_version++;
  DevIn jac(m, jac_quad, m->local_nodes_quad);
  d4b_launch_quad_op(m, 3, nullptr, jac.p, m->d_W);
  D4B_CUDA(cudaStreamSynchronize(m->stream));
}

extern "C" int32_t d4b_host_pin(void *ptr, int64_t bytes) {
  if (!ptr || bytes <= 0) return (int32_t)cudaErrorInvalidValue;
  const cudaError_t e = cudaHostRegister(ptr, (size_t)bytes, cudaHostRegisterPortable);
  if (e != cudaSuccess) cudaGetLastError();
  return (int32_t)e;
}
extern "C" int32_t d4b_host_unpin(void *ptr) {
  const cudaError_t e = cudaHostUnregister(ptr);
  if (e != cudaSuccess) cudaGetLastError();
  return (int32_t)e;
}

extern "C" void d4b_sync(d4b_mesh_t *m) { D4B_CUDA(cudaStreamSynchronize(m->stream)); }
extern "C" void d4b_set_stream(d4b_mesh_t *m, void *cuda_stream) {
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  if (m->own_stream) { cudaStreamDestroy(m->stream); m->own_stream = false; }
  m->stream = (cudaStream_t)cuda_stream;
}
