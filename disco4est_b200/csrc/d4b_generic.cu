// d4b_generic.cu -- general hp element-centric SIPG apply kernel.
//
// One CTA per element; handles every case the reference's face iteration
// produces (d4est_mortars.c:601-802): variable degree, deg_quad > deg, hanging
// faces (1<->4), all face orientations, Dirichlet boundaries, ghost
// neighbours.  Each element's Au is written exactly once, without atomics:
// the element evaluates its own volume term
// (d4est_quadrature_apply_stiffness_matrix, d4est_quadrature.c:263-382) and,
// for each of its six faces, its own side of the SIPG flux
// (d4est_laplacian_flux_sipg_interface_aux, d4est_laplacian_flux_sipg.c:494-814;
// dirichlet_aux :15-237) reading the neighbour's nodal values directly.
//
// Uniform-degree conforming elements take the faster kernel in d4b_fast.cu;
// this kernel is the fallback for everything else and the in-library
// cross-check of the fast path (D4B_FORCE_GENERIC=1).
#include "d4b_internal.h"

namespace {

struct GenericArgs {
  const D4bTables *tab;
  const double *tabdata;
  const int32_t *deg, *deg_quad;
  const int64_t *nodal_stride, *quad_stride;
  const int32_t *ghost_deg, *ghost_deg_quad;
  const int64_t *ghost_nodal_stride;
  const double *G;
  const double *W;              // optional weighted diagonal at the quadrature points, or NULL
  int64_t local_nodes_quad;
  const int32_t *elem_side, *elem_sub;
  const int32_t *side_faces_m, *side_faces_p, *side_f_m, *side_f_p, *side_e_m, *side_e_p,
      *side_perm, *side_sub_order;
  const int64_t *side_mortar_stride;
  const double *mortar;
  int64_t mortar_nodes;
  const double *u, *ughost;
  double *Au;
  const int32_t *elem_list;
  int n_elem;
  int stiffness_only;
  int maxN, maxNq;
  int smem_doubles;             // shared memory of one warp (element), in doubles
};

__device__ __forceinline__ const double *tabE(const GenericArgs &a, int deg, int deg_q, int half) {
  return a.tabdata + a.tab->e_off[((deg - 1) * D4B_MAXNQ + (deg_q - 1)) * 3 + half];
}

// out[(c*nb + b)*nra + ar] = sum_l A(ar,l) in[(c*nb + b)*nca + l]   (dir 0)
// generic single-direction sweep on a 3-index array; A(row,col) = A[row*sr + col*sc]
// TPB = cooperating threads: 32 (one warp per element, apply kernel) or the whole CTA (quadrature operators)
template <bool WARP>
__device__ __forceinline__ void group_sync() {
  if (WARP) __syncwarp(); else __syncthreads();
}

template <bool WARP>
__device__ void sweep3(double *out, const double *in, const double *A, int sr, int sc,
                       int nrow, int ncol, int dir, int d0, int d1, int d2) {
  const int tid = WARP ? (threadIdx.x & 31) : threadIdx.x, nth = WARP ? 32 : blockDim.x;
  // input dims (d0,d1,d2) with d[dir] == ncol; output same with d[dir] = nrow
  int o0 = (dir == 0) ? nrow : d0, o1 = (dir == 1) ? nrow : d1, o2 = (dir == 2) ? nrow : d2;
  int total = o0 * o1 * o2;
  int istride = (dir == 0) ? 1 : (dir == 1) ? d0 : d0 * d1;
  for (int idx = tid; idx < total; idx += nth) {
    int i0 = idx % o0, i1 = (idx / o0) % o1, i2 = idx / (o0 * o1);
    int row = (dir == 0) ? i0 : (dir == 1) ? i1 : i2;
    int base = ((dir == 0) ? 0 : i0) + d0 * (((dir == 1) ? 0 : i1) + d1 * ((dir == 2) ? 0 : i2));
    double s = 0.;
    for (int l = 0; l < ncol; ++l) s += A[row * sr + l * sc] * in[base + l * istride];
    out[idx] = s;
  }
  group_sync<WARP>();
}

// 2-D: out[b*nra + a] = sum_l A(a,l) in[b*nca + l] (dir 0) or along b (dir 1)
__device__ void sweep2(double *out, const double *in, const double *A, int sr, int sc,
                       int nrow, int ncol, int dir, int d0, int d1) {
  const int tid = threadIdx.x & 31, nth = 32;
  int o0 = (dir == 0) ? nrow : d0, o1 = (dir == 1) ? nrow : d1;
  int total = o0 * o1;
  int istride = (dir == 0) ? 1 : d0;
  for (int idx = tid; idx < total; idx += nth) {
    int i0 = idx % o0, i1 = idx / o0;
    int row = (dir == 0) ? i0 : i1;
    int base = ((dir == 0) ? 0 : i0) + d0 * ((dir == 1) ? 0 : i1);
    double s = 0.;
    for (int l = 0; l < ncol; ++l) s += A[row * sr + l * sc] * in[base + l * istride];
    out[idx] = s;
  }
  __syncwarp();
}

__device__ __forceinline__ int vol_index(int N, int f, int a, int b, int n) {
  int dir = f >> 1;
  if (dir == 0) return n + N * (a + N * b);
  if (dir == 1) return a + N * (n + N * b);
  return a + N * (b + N * n);
}

// (-)-ordered face point (a,b) on an n x n grid -> (+)-ordered linear index
// (inverse reading of d4est_operators_reorient_face_data: flip0, flip1, transpose)
__device__ __forceinline__ int perm_index(int n, int perm, int a, int b) {
  int row = (perm & 4) ? a : b, col = (perm & 4) ? b : a;
  if (perm & 2) row = n - 1 - row;
  if (perm & 1) col = n - 1 - col;
  return row * n + col;
}

// traces of v (element nodal values, global or shared) on face f at the Gauss
// points of a mortar: out[0] = u, out[1..3] = du/d(rst), each nq*nq, own order.
// tr, t2d: scratch of N*N and N*nq doubles.
__device__ void traces_on_mortar(const GenericArgs &a, const double *v, int deg, int f, int dq,
                                 int hx, int hy, double *out, int ostride, double *tr, double *t2d) {
  const int lane = threadIdx.x & 31;
  int N = deg + 1, Nq = dq + 1;
  const double *D = a.tabdata + a.tab->d_off[deg - 1];
  const double *Ex = tabE(a, deg, dq, hx), *Ey = tabE(a, deg, dq, hy);
  int n = (f & 1) ? N - 1 : 0;
  for (int q = 0; q < 4; ++q) {
    for (int idx = lane; idx < N * N; idx += 32) {
      int fa = idx % N, fb = idx / N;
      int vi = vol_index(N, f, fa, fb, n);
      double s;
      if (q == 0) s = v[vi];
      else {
        int dir = q - 1;
        int stride = (dir == 0) ? 1 : (dir == 1) ? N : N * N;
        int i_dir = (dir == 0) ? vi % N : (dir == 1) ? (vi / N) % N : vi / (N * N);
        int base = vi - i_dir * stride;
        s = 0.;
        for (int l = 0; l < N; ++l) s += D[i_dir * N + l] * v[base + l * stride];
      }
      tr[idx] = s;
    }
    __syncwarp();
    sweep2(t2d, tr, Ex, N, 1, Nq, N, 0, N, N);
    sweep2(out + q * ostride, t2d, Ey, N, 1, Nq, N, 1, Nq, N);
  }
}

// one warp per element, kWarps elements per CTA; warps never synchronise with each other
constexpr int kMaxWarps = 4;
__global__ void __launch_bounds__(32 * kMaxWarps) generic_apply_kernel(GenericArgs a) {
  extern __shared__ double smem_all[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slot = blockIdx.x * (blockDim.x >> 5) + warp;
  if (slot >= a.n_elem) return;
  double *smem = smem_all + (size_t)warp * a.smem_doubles;
  int e = a.elem_list ? a.elem_list[slot] : slot;
  int p = a.deg[e], q = a.deg_quad[e];
  int N = p + 1, Nq = q + 1, n3 = N * N * N, q3 = Nq * Nq * Nq;
  int mN = a.maxN, mQ = a.maxNq;
  // shared layout
  double *s_u = smem;                         // mN^3
  double *s_out = s_u + mN * mN * mN;         // mN^3
  double *s_a = s_out + mN * mN * mN;         // mN^3
  double *s_t1 = s_a + mN * mN * mN;          // mN*mN*mQ
  double *s_t2 = s_t1 + mN * mN * mQ;         // mN*mQ*mQ
  double *s_g = s_t2 + mN * mQ * mQ;          // 3*mQ^3
  const double *ue = a.u + a.nodal_stride[e];
  const double *D = a.tabdata + a.tab->d_off[p - 1];
  const double *E = tabE(a, p, q, 0);
  for (int i = lane; i < n3; i += 32) { s_u[i] = ue[i]; s_out[i] = 0.; }
  __syncwarp();

  // ---- volume: sum_{l',l} D_l'^T V^T G_{l'l} V D_l u
  for (int l = 0; l < 3; ++l) {
    sweep3<true>(s_a, s_u, D, N, 1, N, N, l, N, N, N);
    sweep3<true>(s_t1, s_a, E, N, 1, Nq, N, 0, N, N, N);
    sweep3<true>(s_t2, s_t1, E, N, 1, Nq, N, 1, Nq, N, N);
    sweep3<true>(s_g + l * q3, s_t2, E, N, 1, Nq, N, 2, Nq, Nq, N);
  }
  {
    const double *G = a.G + a.quad_stride[e];
    int64_t LQ = a.local_nodes_quad;
    for (int k = lane; k < q3; k += 32) {
      double g0 = s_g[k], g1 = s_g[q3 + k], g2 = s_g[2 * q3 + k];
      double G00 = G[k], G01 = G[LQ + k], G02 = G[2 * LQ + k], G11 = G[3 * LQ + k],
             G12 = G[4 * LQ + k], G22 = G[5 * LQ + k];
      s_g[k] = G00 * g0 + G01 * g1 + G02 * g2;
      s_g[q3 + k] = G01 * g0 + G11 * g1 + G12 * g2;
      s_g[2 * q3 + k] = G02 * g0 + G12 * g1 + G22 * g2;
    }
    __syncwarp();
  }
  for (int lp = 0; lp < 3; ++lp) {
    // V^T: E transposed -> A(row=a, col=l) = E[l*N + a]
    sweep3<true>(s_t2, s_g + lp * q3, E, 1, N, N, Nq, 2, Nq, Nq, Nq);
    sweep3<true>(s_t1, s_t2, E, 1, N, N, Nq, 1, Nq, Nq, N);
    sweep3<true>(s_a, s_t1, E, 1, N, N, Nq, 0, Nq, N, N);
    // D_l'^T
    sweep3<true>(s_t1, s_a, D, 1, N, N, N, lp, N, N, N);
    for (int i = lane; i < n3; i += 32) s_out[i] += s_t1[i];
    __syncwarp();
  }

  // ---- mass-type diagonal: + V^T diag(W) V u (d4est_quadrature_apply_mass_matrix with
  // jac := J f, d4est_quadrature.c:385-477, as apply_fofufofvlilj calls it :593-774)
  if (a.W && !a.stiffness_only) {
    const double *W = a.W + a.quad_stride[e];
    sweep3<true>(s_t1, s_u, E, N, 1, Nq, N, 0, N, N, N);
    sweep3<true>(s_t2, s_t1, E, N, 1, Nq, N, 1, Nq, N, N);
    sweep3<true>(s_g, s_t2, E, N, 1, Nq, N, 2, Nq, Nq, N);
    for (int k = lane; k < q3; k += 32) s_g[k] *= W[k];
    __syncwarp();
    sweep3<true>(s_t2, s_g, E, 1, N, N, Nq, 2, Nq, Nq, Nq);
    sweep3<true>(s_t1, s_t2, E, 1, N, N, Nq, 1, Nq, Nq, N);
    sweep3<true>(s_a, s_t1, E, 1, N, N, Nq, 0, Nq, N, N);
    for (int i = lane; i < n3; i += 32) s_out[i] += s_a[i];
    __syncwarp();
  }

  // ---- faces
  if (!a.stiffness_only) {
    // scratch carved from the volume buffers (all free now)
    double *s_m = s_g;                       // 4 * mQ^2 : u_m, dm[3]
    double *s_p = s_m + 4 * mQ * mQ;         // 4 * mQ^2 : u_p, dp[3] in (+) order
    double *s_t = s_p + 4 * mQ * mQ;         // 4 * mQ^2 : t13, t2[3]
    double *s_tr = s_t1;                     // N^2
    double *s_2d = s_t2;                     // N*Nq
    double *s_r = s_a;                       // N^2
    for (int f = 0; f < 6; ++f) {
      int s = a.elem_side[6 * e + f];
      int my = a.elem_sub[6 * e + f];
      int faces_m = a.side_faces_m[s], faces_p = a.side_faces_p[s];
      int f_p = a.side_f_p[s], perm = a.side_perm[s];
      int M = faces_m > faces_p ? faces_m : faces_p;
      const int32_t *e_p = a.side_e_p + 4 * s, *sub = a.side_sub_order + 4 * s;
      const int32_t *e_m = a.side_e_m + 4 * s;
      int64_t S = a.side_mortar_stride[s];
      int m_big = (faces_m == 1 && M == 4), p_big = (faces_p == 1 && M == 4);
      // offsets of the mortars of this side
      int off = 0;
      int i_begin = (faces_m == 4) ? my : 0, i_end = (faces_m == 4) ? my + 1 : M;
      for (int i = 0; i < M; ++i) {
        int32_t em_i = e_m[(faces_m == 4) ? i : 0];
        int dq = (em_i >= 0) ? a.deg_quad[em_i] : a.ghost_deg_quad[-em_i - 1];
        int32_t ep_i = 0;
        if (faces_p) {
          ep_i = e_p[(faces_p == 4) ? sub[i] : 0];
          int qp = (ep_i >= 0) ? a.deg_quad[ep_i] : a.ghost_deg_quad[-ep_i - 1];
          if (qp > dq) dq = qp;
        }
        int nq = dq + 1, nq2 = nq * nq;
        if (i >= i_begin && i < i_end) {
          int hx = m_big ? 1 + (i & 1) : 0, hy = m_big ? 1 + (i >> 1) : 0;
          traces_on_mortar(a, s_u, p, f, dq, hx, hy, s_m, mQ * mQ, s_tr, s_2d);
          const double *mo = a.mortar + S + off;
          int64_t MP = a.mortar_nodes;
          if (faces_p == 0) {
            for (int k = lane; k < nq2; k += 32) {
              double du = s_m[k] - mo[7 * MP + k];
              double Am0 = mo[k], Am1 = mo[MP + k], Am2 = mo[2 * MP + k];
              double t1 = -(Am0 * s_m[mQ * mQ + k] + Am1 * s_m[2 * mQ * mQ + k] + Am2 * s_m[3 * mQ * mQ + k]);
              s_t[k] = t1 + mo[6 * MP + k] * du;
              s_t[mQ * mQ + k] = -Am0 * du;
              s_t[2 * mQ * mQ + k] = -Am1 * du;
              s_t[3 * mQ * mQ + k] = -Am2 * du;
            }
            __syncwarp();
          } else {
            int j = (M == 4) ? sub[i] : 0;
            int hxp = p_big ? 1 + (j & 1) : 0, hyp = p_big ? 1 + (j >> 1) : 0;
            int deg_p = (ep_i >= 0) ? a.deg[ep_i] : a.ghost_deg[-ep_i - 1];
            const double *vp = (ep_i >= 0) ? a.u + a.nodal_stride[ep_i]
                                           : a.ughost + a.ghost_nodal_stride[-ep_i - 1];
            traces_on_mortar(a, vp, deg_p, f_p, dq, hxp, hyp, s_p, mQ * mQ, s_tr, s_2d);
            for (int k = lane; k < nq2; k += 32) {
              int kp = perm_index(nq, perm, k % nq, k / nq);
              double du = s_m[k] - s_p[kp];
              double Am0 = mo[k], Am1 = mo[MP + k], Am2 = mo[2 * MP + k];
              double fm = Am0 * s_m[mQ * mQ + k] + Am1 * s_m[2 * mQ * mQ + k] + Am2 * s_m[3 * mQ * mQ + k];
              double fp = mo[3 * MP + k] * s_p[mQ * mQ + kp] + mo[4 * MP + k] * s_p[2 * mQ * mQ + kp] +
                          mo[5 * MP + k] * s_p[3 * mQ * mQ + kp];
              s_t[k] = -.5 * (fp + fm) + mo[6 * MP + k] * du;
              s_t[mQ * mQ + k] = -.5 * Am0 * du;
              s_t[2 * mQ * mQ + k] = -.5 * Am1 * du;
              s_t[3 * mQ * mQ + k] = -.5 * Am2 * du;
            }
            __syncwarp();
          }
          // test: E^T (w (x) w . t), lift, D_l^T for the term2 components
          const double *w = a.tabdata + a.tab->w_off[dq - 1];
          const double *Ex = tabE(a, p, dq, hx), *Ey = tabE(a, p, dq, hy);
          int n = (f & 1) ? N - 1 : 0;
          for (int c = 0; c < 4; ++c) {
            double *t = s_t + c * mQ * mQ;
            for (int k = lane; k < nq2; k += 32) t[k] *= w[k % nq] * w[k / nq];
            __syncwarp();
            sweep2(s_2d, t, Ey, 1, N, N, nq, 1, nq, nq);      // [N x nq] rows b
            sweep2(s_r, s_2d, Ex, 1, N, N, nq, 0, nq, N);     // [N x N]
            if (c == 0) {
              for (int idx = lane; idx < N * N; idx += 32)
                s_out[vol_index(N, f, idx % N, idx / N, n)] += s_r[idx];
            } else {
              int l = c - 1;
              int stride = (l == 0) ? 1 : (l == 1) ? N : N * N;
              if (l == (f >> 1)) {
                // normal direction: every volume node receives one contribution
                for (int v = lane; v < n3; v += 32) {
                  int i_l = (l == 0) ? v % N : (l == 1) ? (v / N) % N : v / (N * N);
                  int vface = v + (n - i_l) * stride;   // node on the face in the same line
                  int i0 = vface % N, i1 = (vface / N) % N, i2 = vface / (N * N);
                  int fa = (l == 0) ? i1 : i0, fb = (l == 2) ? i1 : i2;
                  s_out[v] += D[n * N + i_l] * s_r[fb * N + fa];
                }
              } else {
                // tangential: stays on the face
                for (int idx = lane; idx < N * N; idx += 32) {
                  int fa = idx % N, fb = idx / N;
                  int v = vol_index(N, f, fa, fb, n);
                  int i_l = (l == 0) ? v % N : (l == 1) ? (v / N) % N : v / (N * N);
                  // which face index runs along l
                  int along_a = ((f >> 1) == 0) ? (l == 1) : (l == 0);
                  double acc = 0.;
                  for (int jn = 0; jn < N; ++jn) {
                    int ra = along_a ? jn : fa, rb = along_a ? fb : jn;
                    acc += D[jn * N + i_l] * s_r[rb * N + ra];
                  }
                  s_out[v] += acc;
                }
              }
            }
            __syncwarp();
          }
        }
        off += nq2;
      }
    }
  }
  double *oe = a.Au + a.nodal_stride[e];
  for (int i = lane; i < n3; i += 32) oe[i] = s_out[i];
}

// ---- residual-based error estimator, d4est_estimator_bi_compute (src/Estimators/d4est_estimator_bi.c:349-602)
// One warp per element:  eta2_e = h_e^2 / p_e^2 * r_e^T M_e r_e
//   + sum over the mortars of the element's six flux sides of  w sj [ (c1 n.(grad u_m - grad u_p))^2 + (c2 (u_m - u_p))^2 ]
//   (boundary: w sj (c_d (u_m - g))^2), c1 / c2 / c_d = the Houston et al. or "bi" prefactors of
// d4est_estimator_bi.h:25-197.  n.grad u on either side comes from the folded mortar planes:
// sj n.grad u_m = sum_l A_m[l] du_m/d(rst_l) (and the same with A_p for the (+) side).
struct EstArgs {
  GenericArgs g;
  const double *r, *jac, *diam, *sj, *hm, *hp, *gbc;
  int family;
  double prefactor;
  double *eta2, *terms;   // [n_local], [4*n_local] or NULL
};

__device__ __forceinline__ void est_prefactors(int family, int deg_m, double h_m, int deg_p, double h_p, double pre,
                                               double &c_gradu, double &c_u, double &c_dir) {
  const double max_p = (double)(deg_m > deg_p ? deg_m : deg_p), min_h = h_m < h_p ? h_m : h_p;
  const double hop_m = h_m / (double)deg_m, hop_p = h_p / (double)deg_p, max_hop = hop_m > hop_p ? hop_m : hop_p;
  const double p2h_m = (double)deg_m * (double)deg_m / h_m, p2h_p = (double)deg_p * (double)deg_p / h_p;
  const double max_p2h = p2h_m > p2h_p ? p2h_m : p2h_p;
  switch (family) {
    case D4B_EST_HOUSTON_MAXP_MINH:
      c_gradu = sqrt(.5 * min_h / max_p); c_u = sqrt(.5 * pre * max_p * max_p / min_h);
      c_dir = sqrt(pre * max_p * max_p / min_h); break;
    case D4B_EST_HOUSTON_MAX_P2_OVER_H:
      c_gradu = sqrt(.5 * max_hop); c_u = sqrt(.5 * pre * max_p2h); c_dir = sqrt(pre * max_p2h); break;
    case D4B_EST_BI_MAXP_MINH:
      c_gradu = sqrt(min_h / max_p); c_u = sqrt(pre * max_p * max_p / min_h); c_dir = c_u; break;
    default:   // D4B_EST_BI_MAX_P2_OVER_H
      c_gradu = sqrt(max_hop); c_u = sqrt(pre * max_p2h); c_dir = c_u; break;
  }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(32 * kMaxWarps) estimator_kernel(EstArgs ea) {
  extern __shared__ double smem_all[];
  const GenericArgs &a = ea.g;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int e = blockIdx.x * (blockDim.x >> 5) + warp;
  if (e >= a.n_elem) return;
  double *smem = smem_all + (size_t)warp * a.smem_doubles;
  const int p = a.deg[e], q = a.deg_quad[e];
  const int N = p + 1, Nq = q + 1, n3 = N * N * N, q3 = Nq * Nq * Nq;
  const int mN = a.maxN, mQ = a.maxNq;
  double *s_u = smem, *s_out = s_u + mN * mN * mN, *s_a = s_out + mN * mN * mN;
  double *s_t1 = s_a + mN * mN * mN, *s_t2 = s_t1 + mN * mN * mQ, *s_g = s_t2 + mN * mQ * mQ;
  const double *E = tabE(a, p, q, 0);
  // ---- residual term: h^2 / p^2 * sum w J (V r)^2
  for (int i = lane; i < n3; i += 32) s_u[i] = ea.r[a.nodal_stride[e] + i];
  __syncwarp();
  sweep3<true>(s_t1, s_u, E, N, 1, Nq, N, 0, N, N, N);
  sweep3<true>(s_t2, s_t1, E, N, 1, Nq, N, 1, Nq, N, N);
  sweep3<true>(s_g, s_t2, E, N, 1, Nq, N, 2, Nq, Nq, N);
  double res = 0.;
  {
    const double *w = a.tabdata + a.tab->w_off[q - 1];
    const double *J = ea.jac + a.quad_stride[e];
    for (int k = lane; k < q3; k += 32)
      res += w[k % Nq] * w[(k / Nq) % Nq] * w[k / (Nq * Nq)] * J[k] * s_g[k] * s_g[k];
  }
  res = warp_sum(res) * ea.diam[e] * ea.diam[e] / ((double)p * (double)p);
  __syncwarp();
  // ---- face terms
  for (int i = lane; i < n3; i += 32) s_u[i] = a.u[a.nodal_stride[e] + i];
  __syncwarp();
  double je1 = 0., je2 = 0., jeb = 0.;
  double *s_m = s_g, *s_p = s_m + 4 * mQ * mQ, *s_tr = s_t1, *s_2d = s_t2;
  for (int f = 0; f < 6; ++f) {
    const int s = a.elem_side[6 * e + f], my = a.elem_sub[6 * e + f];
    const int faces_m = a.side_faces_m[s], faces_p = a.side_faces_p[s];
    const int f_p = a.side_f_p[s], perm = a.side_perm[s];
    const int M = faces_m > faces_p ? faces_m : faces_p;
    const int32_t *e_p = a.side_e_p + 4 * s, *sub = a.side_sub_order + 4 * s, *e_m = a.side_e_m + 4 * s;
    const int64_t S = a.side_mortar_stride[s];
    const int m_big = (faces_m == 1 && M == 4), p_big = (faces_p == 1 && M == 4);
    // the folded planes already carry the 1/2 of a hanging side (the reference halves dudx there,
    // d4est_laplacian_flux.c:907-915): sum_l A[l] du/d(rst_l) = sj n.grad u on either side
    int off = 0;
    const int i_begin = (faces_m == 4) ? my : 0, i_end = (faces_m == 4) ? my + 1 : M;
    for (int i = 0; i < M; ++i) {
      const int32_t em_i = e_m[(faces_m == 4) ? i : 0];
      int dq = (em_i >= 0) ? a.deg_quad[em_i] : a.ghost_deg_quad[-em_i - 1];
      int32_t ep_i = 0;
      if (faces_p) {
        ep_i = e_p[(faces_p == 4) ? sub[i] : 0];
        const int qp = (ep_i >= 0) ? a.deg_quad[ep_i] : a.ghost_deg_quad[-ep_i - 1];
        if (qp > dq) dq = qp;
      }
      const int nq = dq + 1, nq2 = nq * nq;
      if (i >= i_begin && i < i_end) {
        const int hx = m_big ? 1 + (i & 1) : 0, hy = m_big ? 1 + (i >> 1) : 0;
        traces_on_mortar(a, s_u, p, f, dq, hx, hy, s_m, mQ * mQ, s_tr, s_2d);
        const double *mo = a.mortar + S + off;
        const int64_t MP = a.mortar_nodes, ks0 = S + off;
        const double *w = a.tabdata + a.tab->w_off[dq - 1];
        int deg_p = p;
        if (faces_p) {
          const int j = (M == 4) ? sub[i] : 0;
          const int hxp = p_big ? 1 + (j & 1) : 0, hyp = p_big ? 1 + (j >> 1) : 0;
          deg_p = (ep_i >= 0) ? a.deg[ep_i] : a.ghost_deg[-ep_i - 1];
          const double *vp = (ep_i >= 0) ? a.u + a.nodal_stride[ep_i] : a.ughost + a.ghost_nodal_stride[-ep_i - 1];
          traces_on_mortar(a, vp, deg_p, f_p, dq, hxp, hyp, s_p, mQ * mQ, s_tr, s_2d);
        }
        for (int k = lane; k < nq2; k += 32) {
          const double sj = ea.sj[ks0 + k], wk = w[k % nq] * w[k / nq];
          const double h_m = ea.hm[ks0 + k];
          double c1, c2, cd;
          if (faces_p) {
            est_prefactors(ea.family, p, h_m, deg_p, ea.hp[ks0 + k], ea.prefactor, c1, c2, cd);
            const int kp = perm_index(nq, perm, k % nq, k / nq);
            const double gm = (mo[k] * s_m[mQ * mQ + k] + mo[MP + k] * s_m[2 * mQ * mQ + k] +
                               mo[2 * MP + k] * s_m[3 * mQ * mQ + k]) / sj;
            const double gp = (mo[3 * MP + k] * s_p[mQ * mQ + kp] + mo[4 * MP + k] * s_p[2 * mQ * mQ + kp] +
                               mo[5 * MP + k] * s_p[3 * mQ * mQ + kp]) / sj;
            const double du = s_m[k] - s_p[kp];
            je1 += wk * sj * c1 * c1 * (gm - gp) * (gm - gp);
            je2 += wk * sj * c2 * c2 * du * du;
          } else {
            est_prefactors(ea.family, p, h_m, p, h_m, ea.prefactor, c1, c2, cd);
            const double du = s_m[k] - (ea.gbc ? ea.gbc[ks0 + k] : 0.);
            jeb += wk * sj * cd * cd * du * du;
          }
        }
        __syncwarp();
      }
      off += nq2;
    }
  }
  je1 = warp_sum(je1); je2 = warp_sum(je2); jeb = warp_sum(jeb);
  if (lane == 0) {
    ea.eta2[e] = res + je1 + je2 + jeb;
    if (ea.terms) {
      ea.terms[e] = res; ea.terms[a.n_elem + e] = je1;
      ea.terms[2 * (int64_t)a.n_elem + e] = je2; ea.terms[3 * (int64_t)a.n_elem + e] = jeb;
    }
  }
}

// ---- volume quadrature operators, one CTA per element, any (deg, deg_quad) ----------
//   mode 0: out_quad = V in                      d4est_quadrature_interpolate            d4est_quadrature.c:966-1016
//   mode 1: out = V^T (w jac in_quad)            d4est_quadrature_apply_galerkin_integral :142-213
//   mode 2: out = V^T diag(w jac) V in           d4est_quadrature_apply_mass_matrix       :385-477
//   mode 3: out_quad = w jac                     (set-up of the fused diagonal)
struct QuadArgs {
  const D4bTables *tab;
  const double *tabdata;
  const int32_t *deg, *deg_quad;
  const int64_t *nodal_stride, *quad_stride;
  int mode, mQ;
  const double *in, *jac;
  double *out;
  const double *wja, *nb;   // modes 5, 6: w J a(x) and b(x) of f = a (b + u)^k
  int kexp;
};

// x^k for an integer exponent of either sign
__device__ __forceinline__ double ipow(double x, int k) {
  double r = 1., p = x;
  for (int n = k < 0 ? -k : k; n; n >>= 1) {
    if (n & 1) r *= p;
    p *= p;
  }
  return k < 0 ? 1. / r : r;
}

__global__ void __launch_bounds__(128) quad_ops_kernel(QuadArgs a) {
  extern __shared__ double smem[];
  const int e = blockIdx.x;
  const int p = a.deg[e], q = a.deg_quad[e];
  const int N = p + 1, Nq = q + 1, n3 = N * N * N, q3 = Nq * Nq * Nq;
  const int64_t ns = a.nodal_stride[e], qs = a.quad_stride[e];
  const int m3 = a.mQ * a.mQ * a.mQ;
  double *s0 = smem, *s1 = s0 + m3, *s2 = s1 + m3;
  const double *E = a.tabdata + a.tab->e_off[((p - 1) * D4B_MAXNQ + (q - 1)) * 3];
  const double *w = a.tabdata + a.tab->w_off[q - 1];
  if (a.mode == 3) {
    for (int k = threadIdx.x; k < q3; k += blockDim.x) {
      const int ia = k % Nq, ib = (k / Nq) % Nq, ic = k / (Nq * Nq);
      a.out[qs + k] = w[ic] * w[ib] * w[ia] * a.jac[qs + k];
    }
    return;
  }
  if (a.mode == 0 || a.mode == 2 || a.mode >= 4) {
    for (int i = threadIdx.x; i < n3; i += blockDim.x) s0[i] = a.in[ns + i];
    __syncthreads();
    sweep3<false>(s1, s0, E, N, 1, Nq, N, 0, N, N, N);
    sweep3<false>(s2, s1, E, N, 1, Nq, N, 1, Nq, N, N);
    sweep3<false>(s0, s2, E, N, 1, Nq, N, 2, Nq, Nq, N);
    if (a.mode == 0) {
      for (int k = threadIdx.x; k < q3; k += blockDim.x) a.out[qs + k] = s0[k];
      return;
    }
    if (a.mode == 6) {   // Newton linearisation: W = w J f_u(x, V u0) = k (w J a) (b + V u0)^(k-1)
      for (int k = threadIdx.x; k < q3; k += blockDim.x)
        a.out[qs + k] = (double)a.kexp * a.wja[qs + k] * ipow(a.nb[qs + k] + s0[k], a.kexp - 1);
      return;
    }
  } else {
    for (int k = threadIdx.x; k < q3; k += blockDim.x) s0[k] = a.in[qs + k];
    __syncthreads();
  }
  __syncthreads();
  for (int k = threadIdx.x; k < q3; k += blockDim.x) {
    const int ia = k % Nq, ib = (k / Nq) % Nq, ic = k / (Nq * Nq);
    if (a.mode == 4) s0[k] *= a.jac[qs + k];                                                   // weights folded in
    else if (a.mode == 5) s0[k] = a.wja[qs + k] * ipow(a.nb[qs + k] + s0[k], a.kexp);          // w J f(x, V u)
    else s0[k] *= w[ic] * w[ib] * w[ia] * (a.jac ? a.jac[qs + k] : 1.);
  }
  __syncthreads();
  sweep3<false>(s1, s0, E, 1, N, N, Nq, 2, Nq, Nq, Nq);
  sweep3<false>(s2, s1, E, 1, N, N, Nq, 1, Nq, Nq, N);
  sweep3<false>(s0, s2, E, 1, N, N, Nq, 0, Nq, N, N);
  for (int i = threadIdx.x; i < n3; i += blockDim.x) a.out[ns + i] = s0[i];
}

}  // namespace

void d4b_launch_quad_nl(d4b_mesh *m, int mode, const double *d_in, double *d_out) {
  D4B_CHECK(m->d_nl_wja && m->d_nl_b, "no nonlinear term attached to this mesh (d4b_mesh_set_nonlinear_term)");
  if (m->n_local == 0) return;
  QuadArgs a;
  a.tab = m->d_tab; a.tabdata = m->d_tabdata;
  a.deg = m->d_deg; a.deg_quad = m->d_deg_quad;
  a.nodal_stride = m->d_nodal_stride; a.quad_stride = m->d_quad_stride;
  a.mode = mode;
  a.mQ = m->max_deg_quad > m->max_deg ? m->max_deg_quad + 1 : m->max_deg + 1;
  a.in = d_in; a.jac = nullptr; a.out = d_out;
  a.wja = m->d_nl_wja; a.nb = m->d_nl_b; a.kexp = m->nl_k;
  const size_t smem = (size_t)3 * a.mQ * a.mQ * a.mQ * sizeof(double);
  D4B_CHECK(smem <= 48 * 1024, "quadrature operators: %zu bytes of shared memory", smem);
  quad_ops_kernel<<<m->n_local, 128, smem, m->stream>>>(a);
  D4B_CUDA(cudaGetLastError());
}

void d4b_launch_quad_op(d4b_mesh *m, int mode, const double *d_in, const double *d_jac, double *d_out) {
  if (m->n_local == 0) return;
  QuadArgs a;
  a.wja = nullptr; a.nb = nullptr; a.kexp = 0;
  a.tab = m->d_tab; a.tabdata = m->d_tabdata;
  a.deg = m->d_deg; a.deg_quad = m->d_deg_quad;
  a.nodal_stride = m->d_nodal_stride; a.quad_stride = m->d_quad_stride;
  a.mode = mode;
  a.mQ = m->max_deg_quad > m->max_deg ? m->max_deg_quad + 1 : m->max_deg + 1;
  a.in = d_in; a.jac = d_jac; a.out = d_out;
  const size_t smem = (size_t)3 * a.mQ * a.mQ * a.mQ * sizeof(double);
  D4B_CHECK(smem <= 48 * 1024, "quadrature operators: %zu bytes of shared memory", smem);
  quad_ops_kernel<<<m->n_local, 128, smem, m->stream>>>(a);
  D4B_CUDA(cudaGetLastError());
}

void d4b_generic_configure(int device, int smem_bytes);

void d4b_launch_generic(d4b_mesh *m, const double *d_u, double *d_Au, int stiffness_only) {
  bool force_all = (stiffness_only != 0) || (m->fast.empty());
  int n_elem = force_all ? m->n_local : m->n_generic;
  if (n_elem == 0) return;
  GenericArgs a;
  a.tab = m->d_tab; a.tabdata = m->d_tabdata;
  a.deg = m->d_deg; a.deg_quad = m->d_deg_quad;
  a.nodal_stride = m->d_nodal_stride; a.quad_stride = m->d_quad_stride;
  a.ghost_deg = m->d_ghost_deg; a.ghost_deg_quad = m->d_ghost_deg_quad;
  a.ghost_nodal_stride = m->d_ghost_nodal_stride;
  a.G = m->d_G; a.W = m->d_W; a.local_nodes_quad = m->local_nodes_quad;
  a.elem_side = m->d_elem_side; a.elem_sub = m->d_elem_sub;
  a.side_faces_m = m->d_side_faces_m; a.side_faces_p = m->d_side_faces_p;
  a.side_f_m = m->d_side_f_m; a.side_f_p = m->d_side_f_p;
  a.side_e_m = m->d_side_e_m; a.side_e_p = m->d_side_e_p;
  a.side_perm = m->d_side_perm; a.side_sub_order = m->d_side_sub_order;
  a.side_mortar_stride = m->d_side_mortar_stride;
  a.mortar = m->d_mortar; a.mortar_nodes = m->mortar_nodes;
  a.u = d_u; a.ughost = m->d_ughost; a.Au = d_Au;
  a.elem_list = force_all ? nullptr : m->d_generic_elem;
  a.n_elem = n_elem;
  a.stiffness_only = stiffness_only;
  a.maxN = m->max_deg + 1; a.maxNq = m->max_deg_quad + 1;
  // as many elements (warps) per CTA as the shared memory of an SM allows, at most kMaxWarps
  const int per_warp = m->generic_smem;
  int warps = (200 * 1024) / per_warp;
  warps = warps < 1 ? 1 : warps > kMaxWarps ? kMaxWarps : warps;
  a.smem_doubles = per_warp / (int)sizeof(double);
  d4b_generic_configure(m->device, per_warp * warps);
  generic_apply_kernel<<<(n_elem + warps - 1) / warps, 32 * warps, (size_t)per_warp * warps, m->stream>>>(a);
  D4B_CUDA(cudaGetLastError());
}

void d4b_launch_estimator(d4b_mesh *m, const double *d_u, const double *d_r, const double *d_jac, const double *d_diam,
                          const double *d_sj, const double *d_hm, const double *d_hp, const double *d_g, int family,
                          double prefactor, double *d_eta2, double *d_terms) {
  if (m->n_local == 0) return;
  EstArgs ea;
  GenericArgs &a = ea.g;
  a.tab = m->d_tab; a.tabdata = m->d_tabdata;
  a.deg = m->d_deg; a.deg_quad = m->d_deg_quad;
  a.nodal_stride = m->d_nodal_stride; a.quad_stride = m->d_quad_stride;
  a.ghost_deg = m->d_ghost_deg; a.ghost_deg_quad = m->d_ghost_deg_quad;
  a.ghost_nodal_stride = m->d_ghost_nodal_stride;
  a.G = m->d_G; a.W = nullptr; a.local_nodes_quad = m->local_nodes_quad;
  a.elem_side = m->d_elem_side; a.elem_sub = m->d_elem_sub;
  a.side_faces_m = m->d_side_faces_m; a.side_faces_p = m->d_side_faces_p;
  a.side_f_m = m->d_side_f_m; a.side_f_p = m->d_side_f_p;
  a.side_e_m = m->d_side_e_m; a.side_e_p = m->d_side_e_p;
  a.side_perm = m->d_side_perm; a.side_sub_order = m->d_side_sub_order;
  a.side_mortar_stride = m->d_side_mortar_stride;
  a.mortar = m->d_mortar; a.mortar_nodes = m->mortar_nodes;
  a.u = d_u; a.ughost = m->d_ughost; a.Au = nullptr;
  a.elem_list = nullptr; a.n_elem = m->n_local; a.stiffness_only = 0;
  a.maxN = m->max_deg + 1; a.maxNq = m->max_deg_quad + 1;
  const int per_warp = m->generic_smem;
  int warps = (200 * 1024) / per_warp;
  warps = warps < 1 ? 1 : warps > kMaxWarps ? kMaxWarps : warps;
  a.smem_doubles = per_warp / (int)sizeof(double);
  static int configured_dev[kD4bMaxDevices] = {};   // per device; 0 = the 48 KB default
  int &configured = configured_dev[m->device];
  if (per_warp * warps > 48 * 1024 && per_warp * warps > configured) {
    D4B_CUDA(cudaFuncSetAttribute(estimator_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, per_warp * warps));
    configured = per_warp * warps;
  }
  ea.r = d_r; ea.jac = d_jac; ea.diam = d_diam; ea.sj = d_sj; ea.hm = d_hm; ea.hp = d_hp; ea.gbc = d_g;
  ea.family = family; ea.prefactor = prefactor; ea.eta2 = d_eta2; ea.terms = d_terms;
  estimator_kernel<<<(m->n_local + warps - 1) / warps, 32 * warps, (size_t)per_warp * warps, m->stream>>>(ea);
  D4B_CUDA(cudaGetLastError());
}

int d4b_generic_smem_bytes(int maxN, int maxNq) {
  size_t d = (size_t)3 * maxN * maxN * maxN + (size_t)maxN * maxN * maxNq +
             (size_t)maxN * maxNq * maxNq + (size_t)3 * maxNq * maxNq * maxNq;
  size_t face = (size_t)12 * maxNq * maxNq;     // s_m, s_p, s_t inside the s_g region
  size_t g = (size_t)3 * maxNq * maxNq * maxNq;
  if (face > g) d += face - g;
  return (int)(d * sizeof(double));   // per warp (= per element in flight)
}

void d4b_generic_configure(int device, int smem_bytes) {
  // the attribute is per function and device, not per mesh: only ever raise it
  static int configured_dev[kD4bMaxDevices] = {};
  int &configured = configured_dev[device];
  if (smem_bytes <= 48 * 1024 || smem_bytes <= configured) return;
  D4B_CUDA(cudaFuncSetAttribute(generic_apply_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                smem_bytes));
  configured = smem_bytes;
}
