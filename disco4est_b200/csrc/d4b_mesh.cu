// d4b_mesh.cu -- builds the device-resident mesh from the flattened reference
// description (d4b_mesh_desc_t).  Runs once per mesh (after d4est_mesh_update,
// src/Mesh/d4est_mesh.c:2790-2890); nothing here is on the per-apply path.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "d4b_internal.h"

int d4b_generic_smem_bytes(int maxN, int maxNq);
void d4b_generic_configure(int smem_bytes);

template <typename T>
T *d4b_dev_alloc(d4b_mesh *m, size_t n) {
  T *p = nullptr;
  size_t bytes = std::max<size_t>(n, 1) * sizeof(T);
  D4B_CUDA(cudaMalloc(&p, bytes));
  m->device_bytes += (int64_t)bytes;
  return p;
}
template <typename T>
T *d4b_dev_upload(d4b_mesh *m, const T *h, size_t n) {
  T *p = d4b_dev_alloc<T>(m, n);
  if (n) D4B_CUDA(cudaMemcpy(p, h, n * sizeof(T), cudaMemcpyHostToDevice));
  return p;
}
template int32_t *d4b_dev_alloc<int32_t>(d4b_mesh *, size_t);
template int64_t *d4b_dev_alloc<int64_t>(d4b_mesh *, size_t);
template double *d4b_dev_alloc<double>(d4b_mesh *, size_t);
template int32_t *d4b_dev_upload<int32_t>(d4b_mesh *, const int32_t *, size_t);
template int64_t *d4b_dev_upload<int64_t>(d4b_mesh *, const int64_t *, size_t);
template double *d4b_dev_upload<double>(d4b_mesh *, const double *, size_t);

namespace {

// penalty functions, src/dGMath/d4est_laplacian_flux_sipg.c:945-1005
double penalty(int fcn, int deg_m, double h_m, int deg_p, double h_p, double pre) {
  double maxd = std::max(deg_m, deg_p), minh = std::min(h_m, h_p);
  switch (fcn) {
    case D4B_PENALTY_MAXP_SQR_OVER_MINH: return pre * maxd * maxd / minh;
    case D4B_PENALTY_MAXPP1_SQR_OVER_MINH: return pre * (maxd + 1) * (maxd + 1) / minh;
    case D4B_PENALTY_MEANP_SQR_OVER_MEANH: {
      double mp = .5 * (deg_m + deg_p), mh = .5 * (h_m + h_p);
      return pre * mp * mp / mh;
    }
    case D4B_PENALTY_MEAN_P_SQR_OVER_H:
      return pre * .5 * (deg_m * deg_m / h_m + deg_p * deg_p / h_p);
  }
  D4B_ABORT("unknown sipg_penalty_fcn %d", fcn);
  return 0.;
}

inline int perm_index(int n, int perm, int a, int b) {
  int row = (perm & 4) ? a : b, col = (perm & 4) ? b : a;
  if (perm & 2) row = n - 1 - row;
  if (perm & 1) col = n - 1 - col;
  return row * n + col;
}

void build_tables(d4b_mesh *m) {
  std::vector<double> data;
  D4bTables &T = m->tab;
  memset(&T, 0xff, sizeof(T));
  std::vector<double> xl(D4B_MAXN), wl(D4B_MAXN), xg(D4B_MAXNQ), wg(D4B_MAXNQ), y(D4B_MAXNQ);
  for (int dq = 1; dq <= m->max_deg_quad; ++dq) {
    d4b_tab_gauss(dq + 1, xg.data(), wg.data());
    T.w_off[dq - 1] = (int32_t)data.size();
    data.insert(data.end(), wg.begin(), wg.begin() + dq + 1);
  }
  for (int deg = 1; deg <= m->max_deg; ++deg) {
    int n = deg + 1;
    d4b_tab_lobatto(n, xl.data(), wl.data());
    std::vector<double> D(n * n);
    d4b_tab_deriv(n, xl.data(), D.data());
    T.d_off[deg - 1] = (int32_t)data.size();
    data.insert(data.end(), D.begin(), D.end());
    for (int dq = deg; dq <= m->max_deg_quad; ++dq) {
      int nq = dq + 1;
      d4b_tab_gauss(nq, xg.data(), wg.data());
      for (int half = 0; half < 3; ++half) {
        for (int k = 0; k < nq; ++k)
          y[k] = (half == 0) ? xg[k] : (half == 1) ? .5 * (xg[k] - 1.) : .5 * (xg[k] + 1.);
        std::vector<double> E(nq * n);
        d4b_tab_eval(n, xl.data(), nq, y.data(), E.data());
        T.e_off[((deg - 1) * D4B_MAXNQ + (dq - 1)) * 3 + half] = (int32_t)data.size();
        data.insert(data.end(), E.begin(), E.end());
      }
    }
  }
  m->d_tabdata = d4b_dev_upload<double>(m, data.data(), data.size());
  m->d_tab = (D4bTables *)d4b_dev_upload<int32_t>(m, (const int32_t *)&T, sizeof(T) / sizeof(int32_t));
}

}  // namespace

namespace {
struct PhaseTimer {   // D4B_TIMING=1: wall time of the set-up phases on stderr
  bool on = getenv("D4B_TIMING") && atoi(getenv("D4B_TIMING"));
  std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
  void lap(const char *what) {
    if (!on) return;
    auto n = std::chrono::steady_clock::now();
    fprintf(stderr, "[d4b_mesh_create] %-28s %8.3f s\n", what, std::chrono::duration<double>(n - t).count());
    t = n;
  }
};
}  // namespace

extern "C" d4b_mesh_t *d4b_mesh_create(const d4b_mesh_desc_t *d) {
  PhaseTimer pt;
  D4B_CHECK(d != nullptr, "d4b_mesh_create: NULL description");
  D4B_CHECK(d->abi_version == D4B_ABI_VERSION, "d4b_mesh_create: ABI version %d, library has %d",
            d->abi_version, D4B_ABI_VERSION);
  D4B_CHECK(d4b_device_count() > 0, "d4b_mesh_create: no CUDA device (there is no CPU fallback)");
  D4B_CHECK(d->bc_type == D4B_BC_DIRICHLET || d->bc_type == D4B_BC_ROBIN, "d4b_mesh_create: unknown bc_type %d", d->bc_type);
  D4B_CHECK(d->bc_type != D4B_BC_ROBIN || d->robin_coeff_mortar_quad, "d4b_mesh_create: Robin boundary without coefficients");
  d4b_mesh *m = new d4b_mesh();
  D4B_CUDA(cudaGetDevice(&m->device));
  D4B_CUDA(cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking));
  m->n_local = d->n_local; m->n_ghost = d->n_ghost; m->n_sides = d->n_sides;
  m->local_nodes = d->local_nodes; m->local_nodes_quad = d->local_nodes_quad;
  m->mortar_nodes = d->local_mortar_nodes_quad; m->ghost_nodes = d->ghost_nodes;
  const int E = d->n_local;
  for (int e = 0; e < E; ++e) {
    D4B_CHECK(d->deg[e] >= 1 && d->deg[e] <= D4B_MAX_DEG, "element %d: degree %d outside 1..%d", e,
              d->deg[e], D4B_MAX_DEG);
    D4B_CHECK(d->deg_quad[e] >= d->deg[e] && d->deg_quad[e] < D4B_MAXNQ,
              "element %d: deg_quad %d outside deg..%d", e, d->deg_quad[e], D4B_MAXNQ - 1);
    m->max_deg = std::max(m->max_deg, d->deg[e]);
    m->max_deg_quad = std::max(m->max_deg_quad, d->deg_quad[e]);
  }
  for (int g = 0; g < d->n_ghost; ++g) {
    m->max_deg = std::max(m->max_deg, d->ghost_deg[g]);
    m->max_deg_quad = std::max(m->max_deg_quad, d->ghost_deg_quad[g]);
  }
  D4B_CHECK(m->max_deg <= D4B_MAX_DEG && m->max_deg_quad < D4B_MAXNQ, "ghost degree out of range");
  build_tables(m);

  m->d_deg = d4b_dev_upload<int32_t>(m, d->deg, E);
  m->d_deg_quad = d4b_dev_upload<int32_t>(m, d->deg_quad, E);
  m->d_nodal_stride = d4b_dev_upload<int64_t>(m, d->nodal_stride, E);
  m->d_quad_stride = d4b_dev_upload<int64_t>(m, d->quad_stride, E);
  m->d_ghost_deg = d4b_dev_upload<int32_t>(m, d->ghost_deg, d->n_ghost);
  m->d_ghost_deg_quad = d4b_dev_upload<int32_t>(m, d->ghost_deg_quad, d->n_ghost);
  m->d_ghost_nodal_stride = d4b_dev_upload<int64_t>(m, d->ghost_nodal_stride, d->n_ghost);
  m->d_ughost = d4b_dev_alloc<double>(m, (size_t)d->ghost_nodes);
  d4b_comm_setup(m, d);

  // ---- volume metric G = w J (rst_xyz)(rst_xyz)^T, six symmetric planes
  std::vector<double> G((size_t)6 * d->local_nodes_quad);
  {
    const int64_t LQ = d->local_nodes_quad;
    // Gauss weights per quadrature degree, then one element per loop iteration (host threads)
    std::vector<std::vector<double>> wgs(D4B_MAXNQ + 1);
    for (int q = 1; q <= m->max_deg_quad; ++q) {
      std::vector<double> xg(q + 1);
      wgs[q].resize(q + 1);
      d4b_tab_gauss(q + 1, xg.data(), wgs[q].data());
    }
#pragma omp parallel for schedule(static)
    for (int e = 0; e < E; ++e) {
      int q = d->deg_quad[e], nq = q + 1;
      const std::vector<double> &wg = wgs[q];
      int64_t qs = d->quad_stride[e];
      for (int k = 0; k < nq * nq * nq; ++k) {
        int a = k % nq, b = (k / nq) % nq, c = k / (nq * nq);
        double wj = wg[a] * wg[b] * wg[c] * d->J_quad[qs + k];
        double r[3][3];
        for (int i = 0; i < 3; ++i)
          for (int j = 0; j < 3; ++j) r[i][j] = d->rst_xyz_quad[(i * 3 + j) * LQ + qs + k];
        int cc = 0;
        for (int i = 0; i < 3; ++i)
          for (int j = i; j < 3; ++j) {
            G[(size_t)cc * LQ + qs + k] = wj * (r[i][0] * r[j][0] + r[i][1] * r[j][1] + r[i][2] * r[j][2]);
            ++cc;
          }
      }
    }
    m->d_G = d4b_dev_upload<double>(m, G.data(), G.size());
  }

  pt.lap("tables, volume metric fold");
  // ---- element-face -> side record
  std::vector<int32_t> elem_side((size_t)6 * E, -1), elem_sub((size_t)6 * E, 0);
  for (int s = 0; s < d->n_sides; ++s) {
    int fm = d->side_f_m[s];
    for (int i = 0; i < d->side_faces_m[s]; ++i) {
      int32_t e = d->side_e_m[4 * s + i];
      if (e < 0) continue;
      D4B_CHECK(e < E && elem_side[6 * e + fm] == -1, "side %d: element %d face %d listed twice", s, e, fm);
      elem_side[6 * e + fm] = s;
      elem_sub[6 * e + fm] = i;
    }
  }
  for (size_t i = 0; i < elem_side.size(); ++i)
    D4B_CHECK(elem_side[i] >= 0, "element %d face %d has no flux side record", (int)(i / 6), (int)(i % 6));
  m->d_elem_side = d4b_dev_upload<int32_t>(m, elem_side.data(), elem_side.size());
  m->d_elem_sub = d4b_dev_upload<int32_t>(m, elem_sub.data(), elem_sub.size());

  // ---- folded mortar data (8 planes)
  const int64_t MP = d->local_mortar_nodes_quad;
  std::vector<double> mortar((size_t)8 * MP, 0.);
  auto degq_of = [&](int32_t r) { return r >= 0 ? d->deg_quad[r] : d->ghost_deg_quad[-r - 1]; };
  auto deg_of = [&](int32_t r) { return r >= 0 ? d->deg[r] : d->ghost_deg[-r - 1]; };
#pragma omp parallel for schedule(static)
  for (int s = 0; s < d->n_sides; ++s) {
    int faces_m = d->side_faces_m[s], faces_p = d->side_faces_p[s], perm = d->side_perm[s];
    int M = std::max(faces_m, faces_p);
    const int32_t *e_m = d->side_e_m + 4 * s, *e_p = d->side_e_p + 4 * s, *sub = d->side_sub_order + 4 * s;
    int64_t S = d->side_mortar_stride[s];
    int nq[4], off[4], offp[4], nqp[4], T = 0;
    int32_t em[4], ep[4];
    for (int i = 0; i < M; ++i) {
      em[i] = e_m[faces_m == 4 ? i : 0];
      ep[i] = faces_p ? e_p[faces_p == 4 ? sub[i] : 0] : 0;
      int q = degq_of(em[i]);
      if (faces_p) q = std::max(q, degq_of(ep[i]));
      nq[i] = q + 1;
      off[i] = T;
      T += nq[i] * nq[i];
    }
    D4B_CHECK(S >= 0 && S + T <= MP, "side %d: mortar stride out of range", s);
    for (int i = 0; i < M; ++i) nqp[M == 4 ? sub[i] : i] = nq[i] * nq[i];
    for (int j = 0, acc = 0; j < M; ++j) { offp[j] = acc; acc += nqp[j]; }
    double sm = (faces_m == 1 && M == 4) ? .5 : 1., sp = (faces_p == 1 && M == 4) ? .5 : 1.;
    for (int i = 0; i < M; ++i) {
      int deg_m = deg_of(em[i]), deg_p = faces_p ? deg_of(ep[i]) : deg_m;
      int j = (M == 4) ? sub[i] : 0;
      for (int k = 0; k < nq[i] * nq[i]; ++k) {
        int64_t ks = S + off[i] + k;
        double sj = d->sj_m_mortar_quad[ks];
        double n[3];
        for (int a = 0; a < 3; ++a) n[a] = d->n_m_mortar_quad[3 * S + (int64_t)a * T + off[i] + k];
        for (int l = 0; l < 3; ++l) {
          double am = 0.;
          for (int a = 0; a < 3; ++a)
            am += n[a] * d->drst_dxyz_m_mortar_quad[9 * S + (int64_t)(l + 3 * a) * T + off[i] + k];
          mortar[(size_t)l * MP + ks] = sm * sj * am;
        }
        double hm = d->hm_mortar_quad[ks];
        if (faces_p) {
          int kp = perm_index(nq[i], perm, k % nq[i], k / nq[i]);
          for (int l = 0; l < 3; ++l) {
            double ap = 0.;
            for (int a = 0; a < 3; ++a)
              ap += n[a] * d->drst_dxyz_p_mortar_quad_porder[9 * S + (int64_t)(l + 3 * a) * T + offp[j] + kp];
            mortar[(size_t)(3 + l) * MP + ks] = sp * sj * ap;
          }
          double hp = d->hp_mortar_quad[ks];
          mortar[(size_t)6 * MP + ks] = sj * d->side_penalty_mult[s] *
              penalty(d->sipg_penalty_fcn, deg_m, hm, deg_p, hp, d->sipg_penalty_prefactor);
        } else if (d->bc_type == D4B_BC_ROBIN) {
          // sj (c u_m - q) = S (u_m - g) with S = sj c, g = q / c; no gradient term, no term2:
          // the boundary branch of the kernels computes -sum A_m du_m + S (u_m - g) and -A_m (u_m - g)
          const double c = d->robin_coeff_mortar_quad[ks];
          const double q = d->robin_rhs_mortar_quad ? d->robin_rhs_mortar_quad[ks] : 0.;
          D4B_CHECK(c != 0. || q == 0., "side %d: Robin coefficient 0 with a non-zero right-hand side", s);
          for (int l = 0; l < 3; ++l) mortar[(size_t)l * MP + ks] = 0.;
          mortar[(size_t)6 * MP + ks] = sj * c;
          mortar[(size_t)7 * MP + ks] = (c != 0.) ? q / c : 0.;
        } else {
          mortar[(size_t)6 * MP + ks] = sj * d->side_penalty_mult[s] *
              penalty(d->sipg_penalty_fcn, deg_m, hm, deg_m, hm, d->sipg_penalty_prefactor);
          if (d->dirichlet_g_mortar_quad) mortar[(size_t)7 * MP + ks] = d->dirichlet_g_mortar_quad[ks];
        }
      }
    }
  }

  pt.lap("mortar fold");
  m->d_side_faces_m = d4b_dev_upload<int32_t>(m, d->side_faces_m, d->n_sides);
  m->d_side_faces_p = d4b_dev_upload<int32_t>(m, d->side_faces_p, d->n_sides);
  m->d_side_f_m = d4b_dev_upload<int32_t>(m, d->side_f_m, d->n_sides);
  m->d_side_f_p = d4b_dev_upload<int32_t>(m, d->side_f_p, d->n_sides);
  m->d_side_e_m = d4b_dev_upload<int32_t>(m, d->side_e_m, (size_t)4 * d->n_sides);
  m->d_side_e_p = d4b_dev_upload<int32_t>(m, d->side_e_p, (size_t)4 * d->n_sides);
  m->d_side_perm = d4b_dev_upload<int32_t>(m, d->side_perm, d->n_sides);
  m->d_side_sub_order = d4b_dev_upload<int32_t>(m, d->side_sub_order, (size_t)4 * d->n_sides);
  m->d_side_mortar_stride = d4b_dev_upload<int64_t>(m, d->side_mortar_stride, d->n_sides);

  // ---- fast-path classification; elements it does not take go to the
  // generic kernel.  D4B_FORCE_GENERIC=1 sends everything to the generic kernel.
  std::vector<char> is_fast(E, 0);
  const char *force = getenv("D4B_FORCE_GENERIC");
  if (!(force && force[0] == '1')) d4b_fast_setup(m, d, mortar, G, elem_side, is_fast);
  pt.lap("kernel classes, face records");
  std::vector<int32_t> generic;
  for (int e = 0; e < E; ++e)
    if (!is_fast[e]) generic.push_back(e);
  // elements of equal degree next to each other: the warps of a CTA then carry equal work
  std::stable_sort(generic.begin(), generic.end(), [&](int x, int y) {
    return d->deg_quad[x] != d->deg_quad[y] ? d->deg_quad[x] < d->deg_quad[y] : d->deg[x] < d->deg[y];
  });
  m->n_generic = (int)generic.size();
  m->d_generic_elem = d4b_dev_upload<int32_t>(m, generic.data(), generic.size());
  if (m->n_generic > 0 || true) {
    // the generic kernel also serves d4b_apply_stiffness, keep its data
    m->d_mortar = d4b_dev_upload<double>(m, mortar.data(), mortar.size());
  }
  m->generic_smem = d4b_generic_smem_bytes(m->max_deg + 1, m->max_deg_quad + 1);
  D4B_CHECK(m->generic_smem <= 227 * 1024, "generic kernel needs %d bytes of shared memory", m->generic_smem);
  d4b_generic_configure(m->generic_smem);

  // chunks of the pipelined host-pointer apply: one rank, every element in a Gauss-basis kernel, class
  // lists in element order (no interior / halo split)
  const int pipe_min = getenv("D4B_PIPELINE_MIN_ELEMS") ? atoi(getenv("D4B_PIPELINE_MIN_ELEMS")) : 4096;   // 0 = never
  if (d->mpisize <= 1 && m->n_generic == 0 && pipe_min > 0 && E >= pipe_min) {
    // 16 chunks: per-chunk cost (~7 us of DMA / event latency) against pipeline fill; profiles/r01_chunk_sweep.txt
    int K = E >= 32 * 64 ? 16 : (E >= 64 ? 8 : 1);
    if (const char *kc = getenv("D4B_PIPELINE_CHUNKS")) K = std::max(1, std::min(std::min(atoi(kc), 1024), E));
    m->chunk_elem.assign(K + 1, E);
    m->chunk_elem[0] = 0;
    for (int k = 1, e = 0; k < K; ++k) {
      const int64_t target = d->local_nodes * k / K;
      while (e < E && d->nodal_stride[e] < target) ++e;
      m->chunk_elem[k] = e;
    }
    auto chunk_of = [&](int e) {
      return (int)(std::upper_bound(m->chunk_elem.begin(), m->chunk_elem.end(), e) - m->chunk_elem.begin()) - 1;
    };
    // chunk adjacency through the flux sides
    std::vector<std::vector<char>> adj(K, std::vector<char>(K, 0));
    for (int s = 0; s < d->n_sides; ++s)
      for (int i = 0; i < d->side_faces_m[s]; ++i) {
        const int32_t em = d->side_e_m[4 * s + i];
        if (em < 0) continue;
        const int km = chunk_of(em);
        for (int j = 0; j < d->side_faces_p[s]; ++j) {
          const int32_t ep = d->side_e_p[4 * s + j];
          if (ep >= 0) { const int kp = chunk_of(ep); adj[km][kp] = adj[kp][km] = 1; }
        }
      }
    // transfer order: Cuthill-McKee on the chunk graph, so that a chunk's neighbours arrive soon after
    // it (Morton order would make half the mesh wait for the far octant)
    std::vector<int> degc(K, 0), order;
    for (int a = 0; a < K; ++a)
      for (int b = 0; b < K; ++b) degc[a] += (a != b && adj[a][b]);
    std::vector<char> seen(K, 0);
    while ((int)order.size() < K) {
      int start = -1;
      for (int a = 0; a < K; ++a)
        if (!seen[a] && (start < 0 || degc[a] < degc[start])) start = a;
      seen[start] = 1;
      size_t head = order.size();
      order.push_back(start);
      while (head < order.size()) {
        const int a = order[head++];
        std::vector<int> nb;
        for (int b = 0; b < K; ++b)
          if (!seen[b] && adj[a][b]) nb.push_back(b);
        std::stable_sort(nb.begin(), nb.end(), [&](int x, int y) { return degc[x] < degc[y]; });
        for (int b : nb) { seen[b] = 1; order.push_back(b); }
      }
    }
    std::vector<int> pos(K);
    for (int i = 0; i < K; ++i) pos[order[i]] = i;
    m->chunk_order.assign(order.begin(), order.end());
    m->chunk_dep.resize(K);   // position in the transfer order after which chunk k can be applied
    for (int k = 0; k < K; ++k) {
      int dp = pos[k];
      for (int b = 0; b < K; ++b)
        if (adj[k][b]) dp = std::max(dp, pos[b]);
      m->chunk_dep[k] = dp;
    }
    m->chunk_node.resize(K + 1);
    for (int k = 0; k <= K; ++k)
      m->chunk_node[k] = m->chunk_elem[k] < E ? d->nodal_stride[m->chunk_elem[k]] : d->local_nodes;
  }
  pt.lap("uploads, transfer chunks");
  m->d_u_stage = d4b_dev_alloc<double>(m, (size_t)d->local_nodes);
  m->d_Au_stage = d4b_dev_alloc<double>(m, (size_t)d->local_nodes);
  m->d_red = d4b_dev_alloc<double>(m, 4096);
  D4B_CUDA(cudaMallocHost(&m->h_red, 4096 * sizeof(double)));
  return m;
}

extern "C" void d4b_mesh_destroy(d4b_mesh_t *m) {
  if (!m) return;
  cudaStreamSynchronize(m->stream);
  void *ptrs[] = {m->d_tab, m->d_tabdata, m->d_deg, m->d_deg_quad, m->d_nodal_stride, m->d_quad_stride,
                  m->d_ghost_deg, m->d_ghost_deg_quad, m->d_ghost_nodal_stride, m->d_G, m->d_elem_side,
                  m->d_elem_sub, m->d_side_faces_m, m->d_side_faces_p, m->d_side_f_m, m->d_side_f_p,
                  m->d_side_e_m, m->d_side_e_p, m->d_side_perm, m->d_side_sub_order,
                  m->d_side_mortar_stride, m->d_mortar, m->d_ughost, m->d_u_stage, m->d_Au_stage,
                  m->d_red, m->d_work[0], m->d_work[1], m->d_work[2], m->d_work[3], m->d_generic_elem,
                  m->d_fast_FF, m->d_uhat, m->d_flat_FF, m->d_hp_tabs, m->d_hp_tabdata, m->d_extra, m->d_send_src, m->d_send_dst, m->d_send_cnt,
                  m->d_sendbuf, m->d_W, m->d_nl_wja, m->d_nl_b, m->d_addterm};
  for (void *p : ptrs)
    if (p) cudaFree(p);
  for (auto &c : m->fast) {
    void *cp[] = {c.d_elem, c.d_hat_elem, c.d_FG, c.d_FGc, c.d_face_rec, c.d_face_nbr, c.d_face_code, c.d_face_meta,
                  c.d_ghost_hat, c.d_has_extra, c.d_face_mask};
    for (void *p : cp)
      if (p) cudaFree(p);
  }
  if (m->h_red) cudaFreeHost(m->h_red);
  for (cudaEvent_t ev : m->ev_h2d) cudaEventDestroy(ev);
  for (cudaEvent_t ev : m->ev_done) cudaEventDestroy(ev);
  if (m->h2d_stream) cudaStreamDestroy(m->h2d_stream);
  if (m->d2h_stream) cudaStreamDestroy(m->d2h_stream);
  d4b_comm_teardown(m);
  if (m->own_stream) cudaStreamDestroy(m->stream);
  delete m;
}

extern "C" int64_t d4b_mesh_local_nodes(const d4b_mesh_t *m) { return m->local_nodes; }
extern "C" int64_t d4b_mesh_device_bytes(const d4b_mesh_t *m) { return m->device_bytes; }
extern "C" int64_t d4b_mesh_fast_elements(const d4b_mesh_t *m) { return m->n_local - m->n_generic; }
extern "C" int64_t d4b_mesh_flat_elements(const d4b_mesh_t *m) { return m->n_flat_elem; }
extern "C" int64_t d4b_mesh_hp_elements(const d4b_mesh_t *m) { return m->n_hp_elem; }
extern "C" int32_t d4b_mesh_launches_per_apply(const d4b_mesh_t *m) {
  int n = (m->n_generic > 0 || m->fast.empty()) ? 1 : 0;
  for (const auto &c : m->fast)
    n += (c.n_hat > 0) + (c.n_interior > 0) + (c.n_ghost_hat > 0) + (c.n_elem > c.n_interior);
  if (m->n_send_elems > 0) n += 1;   // pack kernel
  return n;
}
