// d4b_fast.cu -- uniform-degree conforming SIPG apply: the hot kernel.
//
// Takes every element whose deg_quad == deg and whose six faces are physical
// boundaries or conforming faces against a LOCAL neighbour of the SAME degree
// (any face orientation); everything else goes to d4b_generic.cu.
//
// Algorithm (exact in exact arithmetic, so equal to the reference to
// round-off; cf. d4est_quadrature_apply_stiffness_matrix d4est_quadrature.c:263-382
// and d4est_laplacian_flux_sipg_interface_aux d4est_laplacian_flux_sipg.c:494-814):
//   * With Nq == N the element polynomial is carried in the Lagrange basis of
//     the N^3 tensor Gauss points ("Gauss-collocated"): uh = (I x I x I) u with
//     I = LGL->Gauss interpolation.  In that basis the reference-space gradient
//     is three 1-D sweeps with the Gauss-basis derivative matrix Dg, the
//     stiffness metric G acts point-wise, and a face trace / normal derivative
//     at the face Gauss points is a 1-D contraction of each line normal to the
//     face with the end-point vectors eL/eR (values) and dL/dR (derivatives).
//     No 2-D face interpolation is needed on the element's own side.
//   * Residuals are accumulated in the same basis (face terms are rank-1
//     updates along the normal lines) and mapped back once: Au = (I^T x I^T x I^T) rh.
//   * N^2 threads per element, each holding one line of N values in registers;
//     the three line orientations are exchanged through shared memory.
//   * The neighbour's side is evaluated from its nodal values read straight
//     from global memory (L2 hits): normal line -> trace + normal derivative,
//     2-D interpolation to the Gauss points, tangential derivatives there.
// Each element's Au is written exactly once, no atomics (deterministic).
#include <cstring>
#include "d4b_internal.h"

namespace {

struct FastTab {
  double I[8][8];    // LGL -> Gauss interpolation, [gauss][lgl]
  double Dg[8][8];   // derivative in the Gauss-point Lagrange basis, [point][basis]
  double e[2][8];    // Gauss basis evaluated at -1 / +1
  double d[2][8];    // Gauss basis derivative at -1 / +1
  double DL[2][8];   // rows 0 and N-1 of the LGL derivative matrix
  double w[8];       // Gauss weights
};

__constant__ FastTab c_tab[9];   // indexed by N

struct FastArgs {
  const int32_t *elem;          // [n_elem] element ids of this class
  int n_elem;
  const int64_t *nodal_stride, *quad_stride;
  const double *G;
  int64_t LQ;
  const double *mortar;
  int64_t MP;
  const int64_t *face_S;        // [6*n_local] mortar stride of (e,f)
  const int32_t *face_nbr;      // [6*n_local] neighbour element or -1 (boundary)
  const int32_t *face_code;     // [6*n_local] f_p | perm << 3
  const double *u;
  double *Au;
};

template <int N>
struct Cfg {
  static constexpr int NN = N * N;
  static constexpr int N3 = N * N * N;
  static constexpr int THREADS = (N == 7 || N == 6) ? 256 : 128;
  static constexpr int EPB = THREADS / NN;           // elements per CTA
  // shared doubles per element: V0,V1,V2 + FT (12*NN) + NB, NB2 (4*NN each)
  static constexpr int SH_PER_ELEM = 3 * N3 + 20 * NN;
};

__device__ __forceinline__ int perm_xy(int n, int perm, int a, int b, int &xp, int &yp) {
  int row = (perm & 4) ? a : b, col = (perm & 4) ? b : a;
  if (perm & 2) row = n - 1 - row;
  if (perm & 1) col = n - 1 - col;
  xp = col; yp = row;
  return row * n + col;
}

template <int N>
__global__ void __launch_bounds__(Cfg<N>::THREADS) fast_apply_kernel(FastArgs a) {
  using C = Cfg<N>;
  constexpr int NN = C::NN, N3 = C::N3;
  extern __shared__ double smem[];
  const FastTab &T = c_tab[N];
  const int slot = threadIdx.x / NN;
  const int lt = threadIdx.x % NN;
  const int lx = lt % N, ly = lt / N;
  const int ei = blockIdx.x * C::EPB + slot;
  const bool active = (slot < C::EPB) && (ei < a.n_elem);
  double *V0 = smem + (size_t)(slot < C::EPB ? slot : 0) * C::SH_PER_ELEM;
  double *V1 = V0 + N3, *V2 = V1 + N3;
  double *FT = V2 + N3;            // [6][2][NN] own trace / normal derivative
  double *NB = FT + 12 * NN;       // [2][2][NN] neighbour stage (one face pair)
  double *NB2 = NB + 4 * NN;       // [2][2][NN] neighbour stage 2 / weighted tangential term2
  const int e = active ? a.elem[ei] : 0;
  const int64_t ns = active ? a.nodal_stride[e] : 0;

  // ---- forward: uh = (I x I x I) u ------------------------------------------
  if (active) {
    double ul[N], v[N];
#pragma unroll
    for (int c = 0; c < N; ++c) ul[c] = a.u[ns + lx + N * ly + NN * c];
#pragma unroll
    for (int c2 = 0; c2 < N; ++c2) {
      double s = 0.;
#pragma unroll
      for (int c = 0; c < N; ++c) s += T.I[c2][c] * ul[c];
      v[c2] = s;
    }
#pragma unroll
    for (int c = 0; c < N; ++c) V0[lx + N * ly + NN * c] = v[c];
  }
  __syncthreads();
  if (active) {   // R-orientation: (b, c) = (lx, ly), line along a
    double l[N], v[N];
    double *p = V0 + N * lx + NN * ly;
#pragma unroll
    for (int i = 0; i < N; ++i) l[i] = p[i];
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.I[i2][i] * l[i];
      v[i2] = s;
    }
#pragma unroll
    for (int i = 0; i < N; ++i) p[i] = v[i];
  }
  __syncthreads();
  if (active) {   // S-orientation: (a, c) = (lx, ly), line along b
    double l[N], uh[N], g[N];
    double *p = V0 + lx + NN * ly;
#pragma unroll
    for (int i = 0; i < N; ++i) l[i] = p[N * i];
    double t2 = 0., n2 = 0., t3 = 0., n3 = 0.;
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.I[i2][i] * l[i];
      uh[i2] = s;
      t2 += T.e[0][i2] * s; n2 += T.d[0][i2] * s;
      t3 += T.e[1][i2] * s; n3 += T.d[1][i2] * s;
    }
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i2][i] * uh[i];
      g[i2] = s;
    }
#pragma unroll
    for (int i = 0; i < N; ++i) { p[N * i] = uh[i]; V2[lx + N * i + NN * ly] = g[i]; }
    FT[(2 * 2 + 0) * NN + lt] = t2; FT[(2 * 2 + 1) * NN + lt] = n2;
    FT[(3 * 2 + 0) * NN + lt] = t3; FT[(3 * 2 + 1) * NN + lt] = n3;
  }
  __syncthreads();
  if (active) {   // R-orientation: gradient along a, faces 0/1
    double uh[N];
    const double *p = V0 + N * lx + NN * ly;
#pragma unroll
    for (int i = 0; i < N; ++i) uh[i] = p[i];
    double t0 = 0., n0 = 0., t1 = 0., n1 = 0.;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      t0 += T.e[0][i] * uh[i]; n0 += T.d[0][i] * uh[i];
      t1 += T.e[1][i] * uh[i]; n1 += T.d[1][i] * uh[i];
    }
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i2][i] * uh[i];
      V1[i2 + N * lx + NN * ly] = s;
    }
    FT[(0 * 2 + 0) * NN + lt] = t0; FT[(0 * 2 + 1) * NN + lt] = n0;
    FT[(1 * 2 + 0) * NN + lt] = t1; FT[(1 * 2 + 1) * NN + lt] = n1;
  }
  __syncthreads();
  if (active) {   // T-orientation: (a, b) = (lx, ly): gradient along c, faces 4/5, metric
    double uh[N];
#pragma unroll
    for (int i = 0; i < N; ++i) uh[i] = V0[lx + N * ly + NN * i];
    double t4 = 0., n4 = 0., t5 = 0., n5 = 0.;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      t4 += T.e[0][i] * uh[i]; n4 += T.d[0][i] * uh[i];
      t5 += T.e[1][i] * uh[i]; n5 += T.d[1][i] * uh[i];
    }
    FT[(4 * 2 + 0) * NN + lt] = t4; FT[(4 * 2 + 1) * NN + lt] = n4;
    FT[(5 * 2 + 0) * NN + lt] = t5; FT[(5 * 2 + 1) * NN + lt] = n5;
    double ft[N];
    const double *G = a.G + a.quad_stride[e] + lx + N * ly;
#pragma unroll
    for (int c = 0; c < N; ++c) {
      double gt = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) gt += T.Dg[c][i] * uh[i];
      int idx = lx + N * ly + NN * c;
      double gr = V1[idx], gs = V2[idx];
      double G00 = G[NN * c], G01 = G[a.LQ + NN * c], G02 = G[2 * a.LQ + NN * c];
      double G11 = G[3 * a.LQ + NN * c], G12 = G[4 * a.LQ + NN * c], G22 = G[5 * a.LQ + NN * c];
      V1[idx] = G00 * gr + G01 * gs + G02 * gt;
      V2[idx] = G01 * gr + G11 * gs + G12 * gt;
      ft[c] = G02 * gr + G12 * gs + G22 * gt;
    }
    // V0 becomes the residual accumulator (each thread owns its own c-line here)
#pragma unroll
    for (int c = 0; c < N; ++c) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i][c] * ft[i];
      V0[lx + N * ly + NN * c] = s;
    }
  }

  // ---- faces, one opposite pair (same normal direction) at a time -------------
  // thread (lx, ly) is face Gauss point (lx, ly) of both faces of the pair
#pragma unroll 1
  for (int nd = 0; nd < 3; ++nd) {
    int nbr[2] = {-1, -1}, code[2] = {0, 0};
    int64_t fS[2] = {0, 0};
    if (active) {   // stage A: neighbour normal lines from global memory
#pragma unroll
      for (int sd = 0; sd < 2; ++sd) {
        const int f = 2 * nd + sd;
        nbr[sd] = a.face_nbr[6 * e + f];
        code[sd] = a.face_code[6 * e + f];
        fS[sd] = a.face_S[6 * e + f];
        if (nbr[sd] >= 0) {
          int fp = code[sd] & 7, ndp = fp >> 1, sp = fp & 1;
          const double *up = a.u + a.nodal_stride[nbr[sd]];
          int base, stride;
          if (ndp == 0) { base = N * lx + NN * ly; stride = 1; }
          else if (ndp == 1) { base = lx + NN * ly; stride = N; }
          else { base = lx + N * ly; stride = NN; }
          double dn = 0., tr = 0.;
#pragma unroll
          for (int l = 0; l < N; ++l) {
            double val = up[base + l * stride];
            dn += T.DL[sp][l] * val;
            if (l == (sp ? N - 1 : 0)) tr = val;
          }
          NB[(sd * 2 + 0) * NN + lt] = tr;
          NB[(sd * 2 + 1) * NN + lt] = dn;
        }
      }
    }
    __syncthreads();
    if (active) {   // stage B: interpolate along face x
#pragma unroll
      for (int sd = 0; sd < 2; ++sd) {
        if (nbr[sd] < 0) continue;
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const double *row = NB + (sd * 2 + q) * NN + N * ly;
          double s = 0.;
#pragma unroll
          for (int i = 0; i < N; ++i) s += T.I[lx][i] * row[i];
          NB2[(sd * 2 + q) * NN + lt] = s;
        }
      }
    }
    __syncthreads();
    if (active) {   // stage C: interpolate along face y
#pragma unroll
      for (int sd = 0; sd < 2; ++sd) {
        if (nbr[sd] < 0) continue;
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const double *col = NB2 + (sd * 2 + q) * NN + lx;
          double s = 0.;
#pragma unroll
          for (int i = 0; i < N; ++i) s += T.I[ly][i] * col[N * i];
          NB[(sd * 2 + q) * NN + lt] = s;
        }
      }
    }
    __syncthreads();
    double ve[2], vd[2];
    if (active) {   // stage D: fluxes at my face Gauss point
      const double wxy = T.w[lx] * T.w[ly];
      const int t1 = (nd == 0) ? 1 : 0, t2 = (nd == 2) ? 1 : 2;
#pragma unroll
      for (int sd = 0; sd < 2; ++sd) {
        const int f = 2 * nd + sd;
        const double *um_a = FT + (f * 2 + 0) * NN;
        double um = um_a[lt], dmn = FT[(f * 2 + 1) * NN + lt];
        double dm1 = 0., dm2 = 0.;
#pragma unroll
        for (int i = 0; i < N; ++i) {
          dm1 += T.Dg[lx][i] * um_a[N * ly + i];
          dm2 += T.Dg[ly][i] * um_a[N * i + lx];
        }
        const double *mo = a.mortar + fS[sd] + lt;
        double Am0 = mo[0], Am1 = mo[a.MP], Am2 = mo[2 * a.MP];
        double Amn = (nd == 0) ? Am0 : (nd == 1) ? Am1 : Am2;
        double Amt1 = (t1 == 0) ? Am0 : Am1, Amt2 = (t2 == 1) ? Am1 : Am2;
        double Fm = Amn * dmn + Amt1 * dm1 + Amt2 * dm2;
        double Ssig = mo[6 * a.MP];
        double du, t13, c2;
        if (nbr[sd] >= 0) {
          int fp = code[sd] & 7, perm = code[sd] >> 3;
          int ndp = fp >> 1;
          int t1p = (ndp == 0) ? 1 : 0, t2p = (ndp == 2) ? 1 : 2;
          int xp, yp;
          int kp = perm_xy(N, perm, lx, ly, xp, yp);
          const double *up_a = NB + (sd * 2 + 0) * NN;
          double up = up_a[kp], dpn = NB[(sd * 2 + 1) * NN + kp];
          double dp1 = 0., dp2 = 0.;
#pragma unroll
          for (int i = 0; i < N; ++i) {
            dp1 += T.Dg[xp][i] * up_a[N * yp + i];
            dp2 += T.Dg[yp][i] * up_a[N * i + xp];
          }
          double Ap0 = mo[3 * a.MP], Ap1 = mo[4 * a.MP], Ap2 = mo[5 * a.MP];
          double Apn = (ndp == 0) ? Ap0 : (ndp == 1) ? Ap1 : Ap2;
          double Apt1 = (t1p == 0) ? Ap0 : Ap1, Apt2 = (t2p == 1) ? Ap1 : Ap2;
          double Fp = Apn * dpn + Apt1 * dp1 + Apt2 * dp2;
          du = um - up;
          t13 = -.5 * (Fp + Fm) + Ssig * du;
          c2 = -.5;
        } else {
          du = um - mo[7 * a.MP];
          t13 = -Fm + Ssig * du;
          c2 = -1.;
        }
        double wdu = wxy * c2 * du;
        ve[sd] = wxy * t13;
        vd[sd] = Amn * wdu;
        NB2[(sd * 2 + 0) * NN + lt] = Amt1 * wdu;
        NB2[(sd * 2 + 1) * NN + lt] = Amt2 * wdu;
      }
    }
    __syncthreads();
    if (active) {   // stage E: transposed tangential derivatives, rank-1 update of the normal lines
#pragma unroll
      for (int sd = 0; sd < 2; ++sd) {
        const double *w1 = NB2 + (sd * 2 + 0) * NN, *w2 = NB2 + (sd * 2 + 1) * NN;
        double s = 0.;
#pragma unroll
        for (int i = 0; i < N; ++i) s += T.Dg[i][lx] * w1[N * ly + i] + T.Dg[i][ly] * w2[N * i + lx];
        ve[sd] += s;
      }
      int base, stride;
      if (nd == 0) { base = N * lx + NN * ly; stride = 1; }
      else if (nd == 1) { base = lx + NN * ly; stride = N; }
      else { base = lx + N * ly; stride = NN; }
#pragma unroll
      for (int l = 0; l < N; ++l)
        V0[base + l * stride] += T.e[0][l] * ve[0] + T.d[0][l] * vd[0] + T.e[1][l] * ve[1] + T.d[1][l] * vd[1];
    }
    // the next pair's stage-A barrier orders these V0 updates and the NB2 reads
  }
  __syncthreads();

  // ---- backward --------------------------------------------------------------
  if (active) {   // S-orientation: + Dg^T F_s
    double fs[N];
#pragma unroll
    for (int i = 0; i < N; ++i) fs[i] = V2[lx + N * i + NN * ly];
#pragma unroll
    for (int b = 0; b < N; ++b) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i][b] * fs[i];
      V0[lx + N * b + NN * ly] += s;
    }
  }
  __syncthreads();
  if (active) {   // R-orientation: + Dg^T F_r, then I^T along a
    double fr[N], tot[N];
    double *p = V0 + N * lx + NN * ly;
#pragma unroll
    for (int i = 0; i < N; ++i) fr[i] = V1[i + N * lx + NN * ly];
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i][i2] * fr[i];
      tot[i2] = p[i2] + s;
    }
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.I[i][i2] * tot[i];
      p[i2] = s;
    }
  }
  __syncthreads();
  if (active) {   // S-orientation: I^T along b
    double l[N];
    double *p = V0 + lx + NN * ly;
#pragma unroll
    for (int i = 0; i < N; ++i) l[i] = p[N * i];
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.I[i][i2] * l[i];
      p[N * i2] = s;
    }
  }
  __syncthreads();
  if (active) {   // T-orientation: I^T along c, write Au
    double l[N];
#pragma unroll
    for (int i = 0; i < N; ++i) l[i] = V0[lx + N * ly + NN * i];
#pragma unroll
    for (int c = 0; c < N; ++c) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.I[i][c] * l[i];
      a.Au[ns + lx + N * ly + NN * c] = s;
    }
  }
}

template <int N>
void launch_fast(d4b_mesh *m, const D4bFastClass &cls, const double *d_u, double *d_Au,
                 const int64_t *face_S, const int32_t *face_nbr, const int32_t *face_code) {
  using C = Cfg<N>;
  FastArgs a;
  a.elem = cls.d_elem; a.n_elem = cls.n_elem;
  a.nodal_stride = m->d_nodal_stride; a.quad_stride = m->d_quad_stride;
  a.G = m->d_G; a.LQ = m->local_nodes_quad;
  a.mortar = m->d_mortar; a.MP = m->mortar_nodes;
  a.face_S = face_S; a.face_nbr = face_nbr; a.face_code = face_code;
  a.u = d_u; a.Au = d_Au;
  size_t smem = (size_t)C::EPB * C::SH_PER_ELEM * sizeof(double);
  static bool configured = false;
  if (!configured) {
    D4B_CUDA(cudaFuncSetAttribute(fast_apply_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = true;
  }
  int grid = (cls.n_elem + C::EPB - 1) / C::EPB;
  fast_apply_kernel<N><<<grid, C::THREADS, smem, m->stream>>>(a);
  D4B_CUDA(cudaGetLastError());
}

void build_fast_tab(int N, FastTab &T) {
  memset(&T, 0, sizeof(T));
  double xl[8], wl[8], xg[8], wg[8], D[64], E[64], ends[2] = {-1., 1.};
  d4b_tab_lobatto(N, xl, wl);
  d4b_tab_gauss(N, xg, wg);
  d4b_tab_deriv(N, xl, D);
  d4b_tab_eval(N, xl, N, xg, E);
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < N; ++j) T.I[i][j] = E[i * N + j];
  double Dg[64];
  d4b_tab_deriv(N, xg, Dg);
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < N; ++j) T.Dg[i][j] = Dg[i * N + j];
  double Ee[16];
  d4b_tab_eval(N, xg, 2, ends, Ee);
  for (int s = 0; s < 2; ++s)
    for (int j = 0; j < N; ++j) T.e[s][j] = Ee[s * N + j];
  // derivative of the Gauss basis at +-1: d/dx of the interpolant = sum_k e[k] * Dg[k][j]
  // holds only inside the basis; use the exact relation psi_j'(x) = sum_k psi_k(x) Dg[k][j]
  for (int s = 0; s < 2; ++s)
    for (int j = 0; j < N; ++j) {
      double v = 0.;
      for (int k = 0; k < N; ++k) v += T.e[s][k] * Dg[k * N + j];
      T.d[s][j] = v;
    }
  for (int j = 0; j < N; ++j) { T.DL[0][j] = D[j]; T.DL[1][j] = D[(N - 1) * N + j]; T.w[j] = wg[j]; }
}

}  // namespace

void d4b_fast_setup(d4b_mesh *m, const d4b_mesh_desc_t *d, const std::vector<double> &,
                    const std::vector<int32_t> &elem_side, std::vector<char> &is_fast) {
  const int E = d->n_local;
  std::vector<int64_t> face_S((size_t)6 * E, 0);
  std::vector<int32_t> face_nbr((size_t)6 * E, -1), face_code((size_t)6 * E, 0);
  std::vector<std::vector<int32_t>> by_deg(D4B_MAX_DEG + 1);
  for (int e = 0; e < E; ++e) {
    bool ok = (d->deg_quad[e] == d->deg[e]);
    for (int f = 0; f < 6 && ok; ++f) {
      int s = elem_side[6 * e + f];
      if (d->side_faces_m[s] != 1 || d->side_faces_p[s] > 1) { ok = false; break; }
      face_S[6 * e + f] = d->side_mortar_stride[s];
      if (d->side_faces_p[s] == 1) {
        int32_t ep = d->side_e_p[4 * s];
        if (ep < 0 || d->deg[ep] != d->deg[e] || d->deg_quad[ep] != d->deg_quad[e]) { ok = false; break; }
        face_nbr[6 * e + f] = ep;
        face_code[6 * e + f] = d->side_f_p[s] | (d->side_perm[s] << 3);
      }
    }
    if (ok) { is_fast[e] = 1; by_deg[d->deg[e]].push_back(e); }
  }
  FastTab tabs[9];
  memset(tabs, 0, sizeof(tabs));
  bool any = false;
  for (int p = 1; p <= D4B_MAX_DEG; ++p) {
    if (by_deg[p].empty()) continue;
    any = true;
    build_fast_tab(p + 1, tabs[p + 1]);
    D4bFastClass c;
    c.deg = p;
    c.n_elem = (int)by_deg[p].size();
    c.d_elem = d4b_dev_upload<int32_t>(m, by_deg[p].data(), by_deg[p].size());
    m->fast.push_back(c);
  }
  if (!any) return;
  for (int p = 1; p <= D4B_MAX_DEG; ++p)
    if (!by_deg[p].empty())
      D4B_CUDA(cudaMemcpyToSymbol(c_tab, &tabs[p + 1], sizeof(FastTab), (size_t)(p + 1) * sizeof(FastTab)));
  m->d_fast_S = d4b_dev_upload<int64_t>(m, face_S.data(), face_S.size());
  m->d_fast_nbr = d4b_dev_upload<int32_t>(m, face_nbr.data(), face_nbr.size());
  m->d_fast_code = d4b_dev_upload<int32_t>(m, face_code.data(), face_code.size());
}

void d4b_launch_fast(d4b_mesh *m, const double *d_u, double *d_Au) {
  for (const D4bFastClass &c : m->fast) {
    switch (c.deg + 1) {
      case 2: launch_fast<2>(m, c, d_u, d_Au, m->d_fast_S, m->d_fast_nbr, m->d_fast_code); break;
      case 3: launch_fast<3>(m, c, d_u, d_Au, m->d_fast_S, m->d_fast_nbr, m->d_fast_code); break;
      case 4: launch_fast<4>(m, c, d_u, d_Au, m->d_fast_S, m->d_fast_nbr, m->d_fast_code); break;
      case 5: launch_fast<5>(m, c, d_u, d_Au, m->d_fast_S, m->d_fast_nbr, m->d_fast_code); break;
      case 6: launch_fast<6>(m, c, d_u, d_Au, m->d_fast_S, m->d_fast_nbr, m->d_fast_code); break;
      case 7: launch_fast<7>(m, c, d_u, d_Au, m->d_fast_S, m->d_fast_nbr, m->d_fast_code); break;
      case 8: launch_fast<8>(m, c, d_u, d_Au, m->d_fast_S, m->d_fast_nbr, m->d_fast_code); break;
      default: D4B_ABORT("fast path: unsupported degree %d", c.deg);
    }
  }
}
