// d4b_internal.h -- internal declarations of libd4est_b200 (not installed).
//
// Device-resident mesh for the matrix-free SIPG apply.  See DESIGN.md for the
// layout; include/d4est_b200.h is the public C ABI.
#ifndef D4B_INTERNAL_H
#define D4B_INTERNAL_H

#include <cuda_runtime.h>
#include <stdint.h>
#include <functional>
#include <vector>
#include "d4est_b200.h"

#define D4B_MAXN 8     // max LGL nodes per direction (deg <= 7)
#define D4B_MAXNQ 12   // max Gauss nodes per direction handled by the generic kernel
constexpr int kD4bMaxDevices = 64;   // per-device kernel attributes (dynamic shared memory) are tracked up to here

void d4b_abort_impl(const char *file, int line, const char *fmt, ...);
#define D4B_ABORT(...) d4b_abort_impl(__FILE__, __LINE__, __VA_ARGS__)
#define D4B_CHECK(cond, ...) do { if (!(cond)) D4B_ABORT(__VA_ARGS__); } while (0)
#define D4B_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    D4B_ABORT("CUDA error %s: %s", #call, cudaGetErrorString(e_)); } while (0)

// ---- 1-D operator tables (host builders, d4b_tables.cpp) --------------------
// Replace the lazily built tables of d4est_operators_t
// (src/dGMath/d4est_operators.c:196-304, :411-438, :855-872, :944-1012).
void d4b_tab_lobatto(int n, double *x, double *w);
void d4b_tab_gauss(int n, double *x, double *w);
// D[n x n] on LGL nodes, row-major
void d4b_tab_deriv(int n, const double *x, double *D);
// E[nq x n]: Lagrange basis on x[0..n) evaluated at y[0..nq)
void d4b_tab_eval(int n, const double *x, int nq, const double *y, double *E);

// Table block on the device: for every (deg, deg_q, half) the evaluation
// matrix E and for every deg the derivative matrix; offsets in doubles.
struct D4bTables {
  // E index: ((deg-1)*D4B_MAXNQ + (deg_q-1))*3 + half -> offset of [nq x n]
  int32_t e_off[D4B_MAXN * D4B_MAXNQ * 3];
  int32_t d_off[D4B_MAXN];      // deg-1 -> offset of D [n x n]
  int32_t w_off[D4B_MAXNQ];     // deg_q-1 -> offset of Gauss weights [nq]
};

// ---- fast-path element classes ----------------------------------------------
// A "regular" element: deg_quad == deg, all six faces conforming against a
// neighbour of the same degree (or a physical boundary), everything local.
struct D4bFastClass {
  int deg = 0;
  int n_elem = 0;                // elements taken by the fast kernel
  int n_interior = 0;            // the first n_interior of them have no ghost neighbour
  int n_ghost_hat = 0;           // ghost elements of this degree whose uh the class needs
  int32_t *d_ghost_hat = nullptr;
  int64_t ns0 = 0;               // ns_linear: the class is a run of consecutive elements, nodal offset of position i
  int ns_linear = 0;             //   = ns0 + i N^3
  int64_t hat_ns0 = 0;           // the same for d_hat_elem
  int hat_linear = 0;
  int32_t *d_elem = nullptr;     // [n_elem] element ids
  int64_t *d_elem_ns = nullptr, *d_hat_ns = nullptr, *d_ghost_hat_ns = nullptr;   // nodal offsets of d_elem / d_hat_elem / d_ghost_hat
  int n_hat = 0;                 // elements of this degree with deg_quad == deg (uh is computed for all)
  int32_t *d_hat_elem = nullptr;
  double *d_FG = nullptr;        // [n_elem][6][N^3] volume metric
  double *d_FGc = nullptr;       // kind 1, every element affine: [n_elem][6] metric without the quadrature weights
                                 // (G = w_a w_b w_c Gc at every point); d_FG is then not stored
  int64_t *d_face_rec = nullptr; // [6*n_elem] face record offset
  int64_t *d_face_nbr = nullptr; // [6*n_elem] neighbour nodal offset / -1
  int32_t *d_face_code = nullptr;
  std::vector<int32_t> h_elem, h_hat;   // host copies of d_elem / d_hat_elem (chunked host-pointer applies)
  int kind = 0;                  // 0: general conforming kernel, 1: tangential-free faces (d4b_flat_kernel.cuh),
                                 // 2: hp mortar kernel (d4b_hp_kernel.cuh), 3: its FACE_ONLY form (face split)
  void *d_face_meta = nullptr;   // kind 1: [6*n_elem] packed face descriptors
  // face split: kinds 0/1: per class position, 1 = the element has irregular faces whose contributions
  // arrive in m->d_extra; kind 3 (face class): per class position the 6-bit mask of those faces
  unsigned char *d_has_extra = nullptr;
  unsigned char *d_face_mask = nullptr;
};

// what the fast kernels do with a finished value s = (A u)[i] (fused smoother,
// d4est_solver_multigrid_smoother_cheby.c:130-152):
//   mode 0: Au[i] = s
//   mode 1: r = alpha (rhs[i] - s); p[i] = r + beta p[i]; u[i] += p[i]   (u is updated in place:
//           the kernels read the Gauss-collocated copy uhat, never u)
//   mode 2: Au[i] = rhs[i] - s
//   mode 3 (tangential-free kernel, HAT instantiation): the same recurrence as mode 1 carried in the Gauss basis --
//           s = (B x B x B) rh with B = I I^T, rhs / p are Gauss-basis vectors, uh_out[i] = uh[i] + p[i] -- so that
//           the next sweep needs no interpolation kernel (d4b_solver.cu, cheby_enqueue)
struct D4bEpilogue {
  int mode = 0;
  const double *rhs = nullptr;
  double *p = nullptr, *u = nullptr;
  double alpha = 0., beta = 0.;
  const double *coef = nullptr;   // mode 1: {alpha, beta} in device memory instead of the two arguments
  const double *add = nullptr;    // every mode: s += add[i] first (the coarse-level Galerkin mass term, d4b_mg.cu)
  double *uh_out = nullptr;       // mode 3
  // u . A u of this apply in its energy form, sum over quadrature points of grad . F + face terms (tangential-free
  // kernel): one partial sum per warp, launch after launch from dot_part[*dot_count] (d4b_launch_fast advances it)
  double *dot_part = nullptr;
  // interpolation kernel of this apply: the input vector is first updated, d = xpby_r + (*xpby_num / *xpby_den) d
  const double *xpby_r = nullptr, *xpby_num = nullptr, *xpby_den = nullptr;
};

// halo exchange with one neighbour rank; offsets / counts in doubles
struct D4bPeer {
  int rank = 0;
  int64_t send_off = 0, send_cnt = 0;   // into d_sendbuf
  int64_t recv_off = 0, recv_cnt = 0;   // into d_ughost
  int32_t send_elem_off = 0, send_elem_cnt = 0;   // its mirrors in d_send_src
  int32_t ghost_first = 0, ghost_cnt = 0;         // its ghosts (ghost ids are contiguous per owner)
};

// Halo over peer memory (d4b_comm.cu): instead of exchanging the mirrors' nodal values, every rank maps
// the Gauss-collocated vectors of its neighbour ranks (CUDA IPC over NVLink) and its boundary elements
// read the neighbours' lines where they lie.
struct D4bPeerHalo {
  int state = 0;                     // 0 not tried yet, 1 active, -1 unavailable (NCCL exchange is used)
  double *uh2 = nullptr;             // two copies of uhat, H doubles apart (IPC-exported); apply k writes copy k & 1
  int64_t H = 0;                     // the same on every rank, so that peer - local offsets hold for both copies
  unsigned long long *flags = nullptr;        // [mpisize]: flags[q] = last apply whose uhat rank q has finished
  std::vector<void *> mapped;                 // IPC mappings to close
  unsigned long long **d_signal = nullptr;    // per neighbour rank: address of its flags[my rank]
  int32_t *d_wait = nullptr;                  // neighbour ranks
  int n_nbr = 0;
  unsigned long long epoch = 0;             // host count of applies: which copy is current (parity)
  unsigned long long *d_epoch = nullptr;    // device count, advanced by the signal kernel, read by the waits
  int64_t *d_ghost_uh_off = nullptr; // per ghost: offset of its uhat lines relative to (uhat + local_nodes)
};

// Peer-memory halo inside the apply kernels (merged launch): element groups that reach class positions >= from
// hold elements with a ghost neighbour; before they read neighbour lines they wait until every neighbour rank has
// signalled this apply (flags[ranks[i]] >= epoch).  n == 0: no wait (one rank, NCCL halo, interior-only launches).
struct D4bHaloWait {
  const unsigned long long *flags = nullptr;
  const int32_t *ranks = nullptr;
  int n = 0, from = 0;
  const unsigned long long *epoch = nullptr;   // device counter of the mesh (advanced by this apply's signal kernel)
};

struct d4b_mesh {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = true;
  int32_t n_local = 0, n_ghost = 0, n_sides = 0;
  int64_t local_nodes = 0, local_nodes_quad = 0, mortar_nodes = 0, ghost_nodes = 0;
  int max_deg = 0, max_deg_quad = 0;
  int64_t device_bytes = 0;
  int64_t config_epoch = 0;      // bumped whenever a later launch sequence would differ (d4b_mesh_set_mass_diag)
  int64_t w_version = 0;         // bumped whenever the CONTENTS of d_W change (new linearisation point)

  // tables
  D4bTables tab;                 // host copy of offsets
  D4bTables *d_tab = nullptr;
  double *d_tabdata = nullptr;

  // per element
  int32_t *d_deg = nullptr, *d_deg_quad = nullptr;
  int64_t *d_nodal_stride = nullptr, *d_quad_stride = nullptr;
  int32_t *d_ghost_deg = nullptr, *d_ghost_deg_quad = nullptr;
  int64_t *d_ghost_nodal_stride = nullptr;
  // volume metric: G[c][quad point], c = 0..5 <-> (00,01,02,11,12,22),
  // G_{l'l} = w_r w_s w_t J sum_k rst[l][k] rst[l'][k]
  double *d_G = nullptr;
  // optional mass-type diagonal W = w_r w_s w_t jac at the quadrature points (d4b_mesh_set_mass_diag):
  // every apply adds V^T diag(W) V u
  double *d_W = nullptr;
  // per element-face: flux side record and position inside it
  int32_t *d_elem_side = nullptr;   // [6*n_local]
  int32_t *d_elem_sub = nullptr;    // [6*n_local]
  // side records
  int32_t *d_side_faces_m = nullptr, *d_side_faces_p = nullptr, *d_side_f_m = nullptr,
          *d_side_f_p = nullptr, *d_side_e_m = nullptr, *d_side_e_p = nullptr,
          *d_side_perm = nullptr, *d_side_sub_order = nullptr;
  int64_t *d_side_mortar_stride = nullptr;
  // folded mortar data, 8 planes of mortar_nodes doubles, (-)-side ordering:
  // 0..2 A_m[l] = sj sum_d n_d drst_m[l][d]   (x 1/2 on a big (-) side)
  // 3..5 A_p[l] = sj sum_d n_d drst_p[l][d]   (x 1/2 on a big (+) side)
  // 6    S     = sj sigma
  // 7    g     Dirichlet data (boundary sides), 0 elsewhere
  double *d_mortar = nullptr;

  // ghost values (filled by the exchange)
  double *d_ughost = nullptr;
  // exchange plan (d4b_comm.cu)
  int mpirank = 0, mpisize = 1;
  std::vector<D4bPeer> peers;
  int32_t n_send_elems = 0;
  int64_t *d_send_src = nullptr, *d_send_dst = nullptr;
  int32_t *d_send_cnt = nullptr;
  double *d_sendbuf = nullptr;
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ev_packed = nullptr, ev_recv = nullptr, ev_interp = nullptr, ev_boundary = nullptr;
  D4bPeerHalo peer;
  std::vector<cudaEvent_t> halo_ev;   // D4B_HALO_TIMING (d4b_capi.cu)
  double halo_t[5] = {0., 0., 0., 0., 0.};
  int64_t halo_n = 0;

  // staging for host-pointer calls
  double *d_u_stage = nullptr, *d_Au_stage = nullptr;
  // pipelined host-pointer apply (d4b_capi.cu): contiguous element chunks [chunk_elem[k], chunk_elem[k+1]),
  // sent in chunk_order; chunk_dep[k] = position in that order of the last chunk holding a face neighbour of k
  std::vector<int32_t> chunk_elem, chunk_dep, chunk_order;
  std::vector<int64_t> chunk_node;
  cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;
  std::vector<cudaEvent_t> ev_h2d, ev_done;
  // scratch for reductions / smoother
  double *d_red = nullptr;         // partial sums
  double *h_red = nullptr;         // pinned
  double *d_work[4] = {nullptr, nullptr, nullptr, nullptr};

  // nonlinear mass-type term f(x,u) = a(x) (b(x) + u)^k (d4b_nonlinear.cu): w J a and b at the quadrature points
  double *d_nl_wja = nullptr, *d_nl_b = nullptr;
  int nl_k = 0;
  bool nl_fused = false;   // the next apply adds the functor itself (residual) instead of the linear term W u
  // coarse multigrid level under a linearised finest level: pre_apply(u) leaves the Galerkin mass term
  // P^T M P u in d_addterm before every apply, the kernels' epilogue adds it (d4b_mg_set_coarse_mass)
  std::function<bool(d4b_mesh *, const double *)> pre_apply;
  double *d_addterm = nullptr;
  // meshes created with device geometry keep what is needed to regenerate positions / J on demand
  bool has_geom = false;
  d4b_geom_desc_t gdesc;
  std::vector<int32_t> h_deg, h_deg_quad, g_tree, g_q, g_dq;
  std::vector<int64_t> h_nodal_stride, h_quad_stride;

  // fast path
  std::vector<D4bFastClass> fast;
  int n_generic = 0;
  int32_t *d_generic_elem = nullptr;   // elements handled by the generic kernel
  // fast-path face data, see d4b_fast.cu
  double *d_fast_FF = nullptr;         // face records [7][N^2], one per face (d4b_fast.cu)
  double *d_uhat = nullptr;            // Gauss-collocated copy of the input vector
  double *d_flat_FF = nullptr;         // compact face records [3][N^2] of the tangential-free class
  int64_t n_flat_elem = 0;
  int64_t n_hp_elem = 0;               // elements of the mortar kernel (d4b_hp_kernel.cuh)
  double *d_extra = nullptr;           // face split: Gauss-basis contributions of irregular faces, nodal layout
  void *d_hp_tabs = nullptr;           // its table offsets / data
  double *d_hp_tabdata = nullptr;
  int64_t n_face_shared = 0, n_face_split = 0;
  int generic_smem = 0;
  D4bEpilogue epi;                     // epilogue of the next fast-kernel launches
  int64_t dot_count = 0;               // partial sums the launches of the current apply have claimed in epi.dot_part
  double *d_dot_part = nullptr;        // [dot_capacity]
  int64_t dot_capacity = 0;
  double *d_uhat_alt = nullptr, *d_rhs_hat = nullptr;   // Gauss-resident smoother: second copy of uhat, (I x I x I) rhs
  bool all_flat = false;               // every element runs in the tangential-free kernel
};

// the entry points launch on the mesh's stream: the mesh's device must be the caller's current one
inline void d4b_check_device(const d4b_mesh *m) {
  int cur = -1;
  cudaGetDevice(&cur);
  D4B_CHECK(cur == m->device, "mesh lives on device %d but the current device is %d (cudaSetDevice before the call)",
            m->device, cur);
}

template <typename T>
T *d4b_dev_alloc(d4b_mesh *m, size_t n);
template <typename T>
T *d4b_dev_upload(d4b_mesh *m, const T *h, size_t n);

// kernels / launchers
void d4b_launch_generic(d4b_mesh *m, const double *d_u, double *d_Au, int stiffness_only);
// volume quadrature operators on device pointers (d4b_generic.cu): mode 0 interpolate,
// 1 Galerkin integral, 2 mass matrix, 3 out_quad = w jac, 4 out = V^T diag(jac) V in (weights already in jac),
// 5 out = V^T (wja (b + V in)^k), 6 out_quad = k wja (b + V in)^(k-1)   (5, 6: d4b_launch_quad_nl)
void d4b_launch_quad_op(d4b_mesh *m, int mode, const double *d_in, const double *d_jac, double *d_out);
void d4b_launch_quad_nl(d4b_mesh *m, int mode, const double *d_in, double *d_out);
// d4est_estimator_bi_compute on device pointers (d4b_generic.cu); ghost values of u must be in m->d_ughost
void d4b_launch_estimator(d4b_mesh *m, const double *d_u, const double *d_r, const double *d_jac, const double *d_diam,
                          const double *d_sj, const double *d_hm, const double *d_hp, const double *d_g, int family,
                          double prefactor, double *d_eta2, double *d_terms);
// the whole apply on device pointers (halo exchange, fast kernels with m->epi, generic kernel)
void d4b_apply_device(d4b_mesh *m, const double *d_u, double *d_Au);
// d4b_mesh_create with the geometric factors given as DEVICE arrays (dg != nullptr; d4b_geom.cu) or taken from `d`
d4b_mesh *d4b_mesh_create_impl(const d4b_mesh_desc_t *d, const d4b_geom_arrays_t *dg);
// reads m->d_G, m->d_mortar and the side arrays on the device: they must be in place
void d4b_fast_setup(d4b_mesh *m, const d4b_mesh_desc_t *desc, const std::vector<int32_t> &elem_side,
                    std::vector<char> &is_fast);
// part 0: uh of local elements; 1: interior elements; 2 (after the halo exchange): uh of
// ghost elements + elements with a ghost neighbour
// part 3 (peer halo, every class in a kernel that can wait by itself): all elements in one launch per class
void d4b_launch_fast(d4b_mesh *m, const double *d_u, double *d_Au, int part);
bool d4b_fast_can_merge_halo(const d4b_mesh *m);
// one-rank meshes: what = 0: uh of the elements [e0, e1); what = 1: apply on the elements [e0, e1)
void d4b_launch_fast_range(d4b_mesh *m, const double *d_u, double *d_Au, int what, int e0, int e1);
// out = (I x I x I) in (to_gauss) or its inverse, over every collocated element
void d4b_launch_basis_change(d4b_mesh *m, const double *in, double *out, bool to_gauss);

// One smoothing of a multigrid level on device vectors (cg_eigs estimate when asked,
// Chebyshev sweeps, r = rhs - A u), enqueued without any host synchronisation; the estimate
// stays on the device and is re-used by later calls with compute_eigs == false (d4b_solver.cu)
// d_eig: one double of device memory owned by the caller (the level's estimate)
void d4b_smooth_async(d4b_mesh *m, double *du, const double *drhs, double *dr, const d4b_cheby_params_t &c,
                      bool compute_eigs, double *d_eig);
void d4b_dot_async(d4b_mesh *m, const double *x, const double *y, int64_t n, double *d_out);

// CG on device vectors with the stopping test on the device (d4b_krylov.cu); returns iterations
int d4b_cg_run(d4b_mesh *m, double *u, const double *rhs, double *r, double *d, int imax, double rtol, double atol,
               double *d_scal, double *h_scal);
d4b_mesh *d4b_mg_top_mesh(d4b_mg_t *g);

// halo exchange (d4b_comm.cu)
void d4b_comm_setup(d4b_mesh *m, const d4b_mesh_desc_t *d);
void d4b_comm_teardown(d4b_mesh *m);
void d4b_comm_exchange_begin(d4b_mesh *m, const double *d_u);
void d4b_comm_exchange_wait(d4b_mesh *m);
void d4b_comm_allreduce_sum(d4b_mesh *m, double *d_x, int n);
// peer-memory halo: collective set-up on first use (true when active), per-apply signal / wait
bool d4b_comm_peer_ready(d4b_mesh *m);
void d4b_comm_peer_begin(d4b_mesh *m);    // next epoch: points m->d_uhat at this apply's copy
void d4b_comm_peer_signal(d4b_mesh *m, cudaStream_t st);   // tell the neighbours this rank's uhat is complete
double *d4b_comm_peer_next_copy(d4b_mesh *m);
bool d4b_comm_graph_safe(d4b_mesh *m);
void d4b_comm_allreduce_reset_order();
int d4b_comm_peer_parity(const d4b_mesh *m);
void d4b_comm_peer_advance(d4b_mesh *m, unsigned long long epochs);
unsigned long long d4b_comm_peer_epoch(const d4b_mesh *m);
void d4b_comm_peer_teardown(d4b_mesh *m); // collective; before anything of the mesh is freed
void d4b_comm_peer_wait(d4b_mesh *m, cudaStream_t st);     // until every neighbour has signalled this epoch
// d4b_fast.cu: ghost neighbours of the conforming classes addressed through d_ghost_off (offsets relative to uhat)
void d4b_fast_patch_ghost_offsets(d4b_mesh *m, const int64_t *d_ghost_off);

#endif
