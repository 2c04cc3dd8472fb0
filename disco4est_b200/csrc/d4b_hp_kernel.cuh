// d4b_hp_kernel.cuh -- the apply for hp-irregular elements whose own quadrature is collocated
// (deg_quad == deg) and whose face neighbours are collocated too: conforming faces between
// unequal degrees, hanging faces (1 <-> 4) in either direction, any orientation, boundaries,
// ghost neighbours.  It is the Gauss-basis formulation of d4b_fast.cu carried to the mortar
// space of the reference (d4est_laplacian_flux.c:232-1014, d4est_mortars.c:510-598):
//
//   * volume term and own traces exactly as in fast_apply_kernel (lines of the element's uh);
//   * a mortar (sub-face i of the flux side) has M = max(N, N') Gauss points per direction.
//     Own trace, normal and tangential derivatives at the element's N x N face Gauss points are
//     carried to the M x M mortar points by T (x) T, T = the Gauss(N) Lagrange basis evaluated at
//     the Gauss(M) points of the whole face or of the half the sub-face covers -- the same
//     polynomials the reference obtains from its p-/hp-prolongation followed by the LGL->Gauss
//     interpolation (d4est_operators.c:944-1012, :411-438);
//   * the neighbour's four face arrays come from its uh lines (global memory), are carried to the
//     same mortar points with its own T and read through the face permutation;
//   * SIPG fluxes at the mortar points with the folded mortar planes of d4b_mesh.cu, weights,
//     then T^T (x) T^T back to the element's face Gauss basis (the transposes the reference applies
//     in d4est_mortars_project_mass_mortar_onto_side), tangential Dg^T on the face and the rank-1
//     lifting along the normal lines.
//
// One warp per element for N >= 4, half a warp for N = 3, a quarter for N = 2 (groups never synchronise with each
// other): the number and size of an element's mortars vary, so the stages are ordered by __syncwarp only.  Own-degree loops
// are compile-time (template N, tables in constant memory), neighbour / mortar sizes are
// run-time (tables in global memory).  Included by d4b_fast.cu.

struct HpTabs {
  int32_t dg_off[9];           // n -> Dg in the Gauss(n) basis, [n][n]
  int32_t t_off[9][9][3];      // [n_from][m_to][half] -> T [m_to][n_from]
};

struct HpArgs {
  const int32_t *elem;
  int first, n_elem;           // this launch covers class positions [first, n_elem)
  const int32_t *deg, *ghost_deg;
  const int64_t *nodal_stride, *quad_stride, *ghost_nodal_stride;
  const double *G;             // 6 planes of LQ
  int64_t LQ;
  const double *W;
  const double *nlB;           // residual: the functor's b (mass_point, d4b_fast.cu)
  int nlK;
  const int32_t *elem_side, *elem_sub;
  const int32_t *side_faces_m, *side_faces_p, *side_f_p, *side_e_m, *side_e_p, *side_perm, *side_sub_order;
  const int64_t *side_mortar_stride;
  const double *mortar;        // 8 planes of MP
  int64_t MP;
  const double *uh;            // Gauss-collocated input, ghosts behind the local block
  int64_t local_nodes;
  double *Au;
  D4bEpilogue epi;
  const HpTabs *tabs;
  const double *tabdata;
  int mM;                      // largest Gauss count per direction on any mortar of the mesh
  int smem_doubles;            // shared doubles per element (lane group)
  // FACE_ONLY launches (face split): bit f of face_mask[class position] = face f is irregular;
  // their Gauss-basis contributions go to extra[nodal_stride + k] instead of Au
  const unsigned char *face_mask;
  double *extra;
};

// lanes per element when the shared memory allows it (see hp_apply_kernel and launch_hp)
#ifdef D4B_HP_FULL_WARP   // A/B switch: one warp per element at every degree
template <int N> struct HpGroup { static constexpr int size = 32; };
#else
template <int N> struct HpGroup { static constexpr int size = N == 2 ? 8 : N == 3 ? 16 : 32; };
#endif

template <int N>
__host__ __device__ constexpr int hp_smem_doubles(int mM) {
  // V0..V3, FT[6][2][NN], OF[4][NN], RF[4][NN], NF/QM/QP/TT/TMP [4][mM^2] each
  return 4 * N * N * N + 20 * N * N + 20 * mM * mM;
}

// idx / M for idx < 64, M <= 8 without an integer division
__device__ __forceinline__ int hp_div(int idx, float inv) { return (int)(((float)idx + .5f) * inv); }

// Four face arrays at once (u, d/dn, d/dt1, d/dt2): out[q][M x M] = (Ty (x) Tx) in[q][n x n].
// T* = [M][n] row-major in global memory; tmp: 4*n*M doubles.  NC = n when it is known at
// compile time (the element's own side), 0 otherwise (the neighbour's).
template <int NC>
__device__ __forceinline__ void hp_interp4(double *out, int ostride, const double *in, int istride, const double *Tx,
                                           const double *Ty, int n_rt, int M, double *tmp, int lane, int gs,
                                           unsigned gmask) {
  const int n = NC ? NC : n_rt, nM = n * M;
  const float invM = 1.f / (float)M;
  for (int idx = lane; idx < nM; idx += gs) {     // tmp[q][b*M + x] = sum_a Tx[x][a] in[q][b*n + a]
    const int b = hp_div(idx, invM), x = idx - b * M;
    double s0 = 0., s1 = 0., s2 = 0., s3 = 0.;
    const double *row = in + b * n;
#pragma unroll
    for (int a = 0; a < (NC ? NC : 8); ++a) {
      if (!NC && a >= n) break;
      const double t = __ldg(Tx + x * n + a);
      s0 += t * row[a]; s1 += t * row[istride + a]; s2 += t * row[2 * istride + a]; s3 += t * row[3 * istride + a];
    }
    tmp[idx] = s0; tmp[nM + idx] = s1; tmp[2 * nM + idx] = s2; tmp[3 * nM + idx] = s3;
  }
  __syncwarp(gmask);
  for (int idx = lane; idx < M * M; idx += gs) {  // out[q][y*M + x] = sum_b Ty[y][b] tmp[q][b*M + x]
    const int y = hp_div(idx, invM), x = idx - y * M;
    double s0 = 0., s1 = 0., s2 = 0., s3 = 0.;
#pragma unroll
    for (int b = 0; b < (NC ? NC : 8); ++b) {
      if (!NC && b >= n) break;
      const double t = __ldg(Ty + y * n + b);
      const double *c = tmp + b * M + x;
      s0 += t * c[0]; s1 += t * c[nM]; s2 += t * c[2 * nM]; s3 += t * c[3 * nM];
    }
    out[idx] = s0; out[ostride + idx] = s1; out[2 * ostride + idx] = s2; out[3 * ostride + idx] = s3;
  }
  __syncwarp(gmask);
}

// out[q][n x n] += (Ty (x) Tx)^T in[q][M x M], q = 0..3; tmp: 4*M*n doubles
template <int NC>
__device__ __forceinline__ void hp_project4(double *out, int ostride, const double *in, int istride, const double *Tx,
                                            const double *Ty, int M, double *tmp, int lane, int gs, unsigned gmask) {
  constexpr int n = NC;
  const int Mn = M * n;
  for (int idx = lane; idx < Mn; idx += gs) {     // tmp[q][y*n + a] = sum_x Tx[x][a] in[q][y*M + x]
    const int y = idx / n, a = idx - y * n;
    double s0 = 0., s1 = 0., s2 = 0., s3 = 0.;
    const double *row = in + y * M;
    for (int x = 0; x < M; ++x) {
      const double t = __ldg(Tx + x * n + a);
      s0 += t * row[x]; s1 += t * row[istride + x]; s2 += t * row[2 * istride + x]; s3 += t * row[3 * istride + x];
    }
    tmp[idx] = s0; tmp[Mn + idx] = s1; tmp[2 * Mn + idx] = s2; tmp[3 * Mn + idx] = s3;
  }
  __syncwarp(gmask);
  for (int idx = lane; idx < n * n; idx += gs) {  // out[q][b*n + a] += sum_y Ty[y][b] tmp[q][y*n + a]
    const int b = idx / n, a = idx - b * n;
    double s0 = 0., s1 = 0., s2 = 0., s3 = 0.;
    for (int y = 0; y < M; ++y) {
      const double t = __ldg(Ty + y * n + b);
      const double *c = tmp + y * n + a;
      s0 += t * c[0]; s1 += t * c[Mn]; s2 += t * c[2 * Mn]; s3 += t * c[3 * Mn];
    }
    out[idx] += s0; out[ostride + idx] += s1; out[2 * ostride + idx] += s2; out[3 * ostride + idx] += s3;
  }
  __syncwarp(gmask);
}

// FACE_ONLY: the face split of DESIGN.md 4b -- only the faces in face_mask, no volume term, no
// back-transform: the conforming kernels add the result to their accumulator.
template <int N, bool MASS, bool FACE_ONLY = false, int GS = 32>
__global__ void __launch_bounds__(128) hp_apply_kernel(HpArgs a) {
  constexpr int NN = N * N, N3 = N * N * N;
  extern __shared__ double smem[];
  // GS lanes per element: a p=1 element has 4 face points and 8 nodes, a whole warp for it leaves 3/4 of the lanes
  // idle in every loop.  Groups of one warp run different elements (different mortar loops): they synchronise with
  // __syncwarp(gmask) among their own lanes only.
  constexpr int EPW = 32 / GS;
  const int wlane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int grp = wlane / GS, lane = wlane % GS;
  const unsigned gmask = GS == 32 ? 0xffffffffu : (((1u << GS) - 1u) << (grp * GS));
  const int slot = a.first + (blockIdx.x * (blockDim.x >> 5) + warp) * EPW + grp;
  if (slot >= a.n_elem) return;
  const FastTab &T = c_tab[N];
  const int e = a.elem[slot];
  const int64_t ns = a.nodal_stride[e];
  const int mM2 = a.mM * a.mM;
  double *V0 = smem + ((size_t)warp * EPW + grp) * a.smem_doubles;
  double *V1 = V0 + N3, *V2 = V1 + N3, *V3 = V2 + N3;
  double *FT = V3 + N3;          // [6][2][NN] own trace / normal derivative at the face Gauss points
  double *OF = FT + 12 * NN;     // [4][NN]   own face: u, d/dn, d/dt1, d/dt2
  double *RF = OF + 4 * NN;      // [4][NN]   projected-back test integrals
  double *NF = RF + 4 * NN;      // [4][mM2]  neighbour face arrays at its own Gauss points
  double *QM = NF + 4 * mM2;     // [4][mM2]  own side at the mortar points
  double *QP = QM + 4 * mM2;     // [4][mM2]  neighbour at the mortar points, its order
  double *TT = QP + 4 * mM2;     // [4][mM2]  weighted fluxes
  double *TMP = TT + 4 * mM2;    // [4][mM2]

  // ---- lines along c: load uh, d/dt, faces 4/5
  for (int lt = lane; lt < NN; lt += GS) {
    double l[N];
#pragma unroll
    for (int c = 0; c < N; ++c) l[c] = a.uh[ns + lt + NN * c];
    double t4 = 0., n4 = 0., t5 = 0., n5 = 0.;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      V0[lt + NN * i] = l[i];
      t4 += T.e[0][i] * l[i]; n4 += T.d[0][i] * l[i];
      t5 += T.e[1][i] * l[i]; n5 += T.d[1][i] * l[i];
    }
#pragma unroll
    for (int c = 0; c < N; ++c) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[c][i] * l[i];
      V3[lt + NN * c] = s;
    }
    FT[(4 * 2 + 0) * NN + lt] = t4; FT[(4 * 2 + 1) * NN + lt] = n4;
    FT[(5 * 2 + 0) * NN + lt] = t5; FT[(5 * 2 + 1) * NN + lt] = n5;
  }
  __syncwarp(gmask);
  // ---- lines along b (faces 2/3) and along a (faces 0/1)
  for (int lt = lane; lt < NN; lt += GS) {
    const int lx = lt % N, ly = lt / N;
    double l[N];
#pragma unroll
    for (int i = 0; i < N; ++i) l[i] = V0[lx + N * i + NN * ly];
    double t2 = 0., n2 = 0., t3 = 0., n3 = 0.;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      t2 += T.e[0][i] * l[i]; n2 += T.d[0][i] * l[i];
      t3 += T.e[1][i] * l[i]; n3 += T.d[1][i] * l[i];
    }
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i2][i] * l[i];
      V2[lx + N * i2 + NN * ly] = s;
    }
    FT[(2 * 2 + 0) * NN + lt] = t2; FT[(2 * 2 + 1) * NN + lt] = n2;
    FT[(3 * 2 + 0) * NN + lt] = t3; FT[(3 * 2 + 1) * NN + lt] = n3;
#pragma unroll
    for (int i = 0; i < N; ++i) l[i] = V0[i + N * lx + NN * ly];
    double t0 = 0., n0 = 0., t1 = 0., n1 = 0.;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      t0 += T.e[0][i] * l[i]; n0 += T.d[0][i] * l[i];
      t1 += T.e[1][i] * l[i]; n1 += T.d[1][i] * l[i];
    }
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i2][i] * l[i];
      V1[i2 + N * lx + NN * ly] = s;
    }
    FT[(0 * 2 + 0) * NN + lt] = t0; FT[(0 * 2 + 1) * NN + lt] = n0;
    FT[(1 * 2 + 0) * NN + lt] = t1; FT[(1 * 2 + 1) * NN + lt] = n1;
  }
  __syncwarp(gmask);
  if (FACE_ONLY) {
    for (int k = lane; k < N3; k += GS) V0[k] = 0.;
    __syncwarp(gmask);
  }
  // ---- stiffness metric, point-wise; V0 becomes the residual accumulator
  if (!FACE_ONLY) {
    const double *G = a.G + a.quad_stride[e];
    const double *Wd = MASS ? a.W + a.quad_stride[e] : nullptr;
    const double *Bd = (MASS && a.nlB) ? a.nlB + a.quad_stride[e] : nullptr;
    for (int k = lane; k < N3; k += GS) {
      const double gr = V1[k], gs = V2[k], gt = V3[k];
      const double G00 = G[k], G01 = G[a.LQ + k], G02 = G[2 * a.LQ + k], G11 = G[3 * a.LQ + k],
                   G12 = G[4 * a.LQ + k], G22 = G[5 * a.LQ + k];
      V1[k] = G00 * gr + G01 * gs + G02 * gt;
      V2[k] = G01 * gr + G11 * gs + G12 * gt;
      V3[k] = G02 * gr + G12 * gs + G22 * gt;
      V0[k] = MASS ? mass_point(Wd, Bd, a.nlK, k, V0[k]) : 0.;
    }
  }
  __syncwarp(gmask);
  if (!FACE_ONLY) {
  // ---- r += Dg^T F along c, b, a
  for (int lt = lane; lt < NN; lt += GS) {
    double f[N];
#pragma unroll
    for (int i = 0; i < N; ++i) f[i] = V3[lt + NN * i];
#pragma unroll
    for (int c = 0; c < N; ++c) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i][c] * f[i];
      V0[lt + NN * c] += s;
    }
  }
  __syncwarp(gmask);
  for (int lt = lane; lt < NN; lt += GS) {
    const int lx = lt % N, ly = lt / N;
    double f[N];
#pragma unroll
    for (int i = 0; i < N; ++i) f[i] = V2[lx + N * i + NN * ly];
#pragma unroll
    for (int b = 0; b < N; ++b) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i][b] * f[i];
      V0[lx + N * b + NN * ly] += s;
    }
  }
  __syncwarp(gmask);
  for (int lt = lane; lt < NN; lt += GS) {
    const int lx = lt % N, ly = lt / N;
    double f[N];
#pragma unroll
    for (int i = 0; i < N; ++i) f[i] = V1[i + N * lx + NN * ly];
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i][i2] * f[i];
      V0[i2 + N * lx + NN * ly] += s;
    }
  }
  __syncwarp(gmask);

  }

  // ---- the six faces' descriptors, one lane per face: the chains elem_side -> side_* -> e_m / e_p -> deg, nodal_stride
  //      are in flight together instead of one after the other in the face loop (they were a fifth of the kernel's
  //      stall samples, profiles/r02_hp_kernel.md).  Sub-face 0 only -- the other three of a hanging side are read in
  //      the loop.  pack: my | faces_m << 2 | faces_p << 5 | f_p << 8 | perm << 11 | Mq0 << 14 | Np0 << 18
#ifndef D4B_HP_NO_DESC
  int fd_s = 0, fd_pack = 0;
  long long fd_S = 0, fd_up = 0;
  if (lane < 6) {
    fd_s = a.elem_side[6 * e + lane];
    const int fm = a.side_faces_m[fd_s], fp = a.side_faces_p[fd_s];
    fd_S = a.side_mortar_stride[fd_s];
    const int32_t em0 = a.side_e_m[4 * fd_s];
    const int dm = ((em0 >= 0) ? a.deg[em0] : a.ghost_deg[-em0 - 1]) + 1;
    int dp = 0;
    if (fp) {
      const int32_t ep0 = a.side_e_p[4 * fd_s + ((fp == 4) ? a.side_sub_order[4 * fd_s] : 0)];
      dp = ((ep0 >= 0) ? a.deg[ep0] : a.ghost_deg[-ep0 - 1]) + 1;
      fd_up = (ep0 >= 0) ? a.nodal_stride[ep0] : a.local_nodes + a.ghost_nodal_stride[-ep0 - 1];
    }
    fd_pack = (a.elem_sub[6 * e + lane] & 3) | fm << 2 | fp << 5 | (a.side_f_p[fd_s] & 7) << 8 |
              (a.side_perm[fd_s] & 7) << 11 | dm << 14 | dp << 18;
  }
#endif

  // ---- faces
#pragma unroll 1
  for (int f = 0; f < 6; ++f) {
    if (FACE_ONLY && !((a.face_mask[slot] >> f) & 1)) continue;
    const int nd = f >> 1, sd = f & 1;
#ifndef D4B_HP_NO_DESC
    const int s = __shfl_sync(gmask, fd_s, f, GS), pk = __shfl_sync(gmask, fd_pack, f, GS);
    const int64_t S = __shfl_sync(gmask, fd_S, f, GS), up0 = __shfl_sync(gmask, fd_up, f, GS);
    const int my = pk & 3, faces_m = (pk >> 2) & 7, faces_p = (pk >> 5) & 7, f_p = (pk >> 8) & 7, perm = (pk >> 11) & 7;
#else
    const int s = a.elem_side[6 * e + f], my = a.elem_sub[6 * e + f];
    const int faces_m = a.side_faces_m[s], faces_p = a.side_faces_p[s];
    const int f_p = a.side_f_p[s], perm = a.side_perm[s];
    const int64_t S = a.side_mortar_stride[s];
#endif
    const int M = faces_m > faces_p ? faces_m : faces_p;
    const int32_t *e_p = a.side_e_p + 4 * s, *e_m = a.side_e_m + 4 * s, *sub = a.side_sub_order + 4 * s;
    const bool m_big = (faces_m == 1 && M == 4), p_big = (faces_p == 1 && M == 4);
    // own face arrays in the order (u, d/dn, d/dt1, d/dt2), t1 / t2 = the face's a / b direction
    {
      const double *tr = FT + (f * 2 + 0) * NN;
      for (int lt = lane; lt < NN; lt += GS) {
        const int lx = lt % N, ly = lt / N;
        double d1 = 0., d2 = 0.;
#pragma unroll
        for (int i = 0; i < N; ++i) {
          d1 += T.Dg[lx][i] * tr[N * ly + i];
          d2 += T.Dg[ly][i] * tr[N * i + lx];
        }
        OF[lt] = tr[lt];
        OF[NN + lt] = FT[(f * 2 + 1) * NN + lt];
        OF[2 * NN + lt] = d1;
        OF[3 * NN + lt] = d2;
        RF[lt] = 0.; RF[NN + lt] = 0.; RF[2 * NN + lt] = 0.; RF[3 * NN + lt] = 0.;
      }
    }
    __syncwarp(gmask);
    const int t1 = (nd == 0) ? 1 : 0, t2 = (nd == 2) ? 1 : 2;
    const int i_begin = (faces_m == 4) ? my : 0, i_end = (faces_m == 4) ? my + 1 : M;
    int off = 0;
#pragma unroll 1
    for (int i = 0; i < M; ++i) {
      int Mq, Np = 0;
      int64_t up_off = 0;
#ifndef D4B_HP_NO_DESC
      if (i == 0) {
        Mq = (pk >> 14) & 15; Np = (pk >> 18) & 15; up_off = up0;
      } else
#endif
      {
        const int32_t em_i = e_m[(faces_m == 4) ? i : 0];
        Mq = ((em_i >= 0) ? a.deg[em_i] : a.ghost_deg[-em_i - 1]) + 1;
        if (faces_p) {
          const int32_t ep_i = e_p[(faces_p == 4) ? sub[i] : 0];
          Np = ((ep_i >= 0) ? a.deg[ep_i] : a.ghost_deg[-ep_i - 1]) + 1;
          up_off = (ep_i >= 0) ? a.nodal_stride[ep_i] : a.local_nodes + a.ghost_nodal_stride[-ep_i - 1];
        }
      }
      if (Np > Mq) Mq = Np;
      const int Mq2 = Mq * Mq;
      if (i >= i_begin && i < i_end) {
#ifndef D4B_HP_NO_PREFETCH
        {   // the sub-face's mortar metric is read at the end of the stage: ask for its sectors now
          const double *mo = a.mortar + S + off;
          for (int k = 4 * lane; k < Mq2 + 3; k += 4 * GS) {
            const double *pf = mo + (k < Mq2 ? k : Mq2 - 1);
#pragma unroll
            for (int q = 0; q < 8; ++q)
              if (q < 7 ? (faces_p || q < 3 || q == 6) : !faces_p)
                asm volatile("prefetch.global.L1 [%0];" ::"l"(pf + q * a.MP));
          }
        }
#endif
        const int hx = m_big ? 1 + (i & 1) : 0, hy = m_big ? 1 + (i >> 1) : 0;
        const double *Tx = a.tabdata + a.tabs->t_off[N][Mq][hx], *Ty = a.tabdata + a.tabs->t_off[N][Mq][hy];
        const bool own_id = (Mq == N) && !m_big;
        const double *qm = OF;
        int qms = NN;
        if (!own_id) {
          hp_interp4<N>(QM, mM2, OF, NN, Tx, Ty, N, Mq, TMP, lane, GS, gmask);
          qm = QM; qms = mM2;
        }
        const double *qp = nullptr;
        int qps = mM2, ndp = 0;
        if (faces_p) {
          const int sp = f_p & 1;
          ndp = f_p >> 1;
          const int Np2 = Np * Np;
          const double *up = a.uh + up_off;
          const FastTab &Tp = c_tab[Np];
          const float invNp = 1.f / (float)Np;
          for (int kp = lane; kp < Np2; kp += GS) {
            const int yp = hp_div(kp, invNp), xp = kp - yp * Np;
            int base, stride;
            if (ndp == 0) { base = Np * xp + Np2 * yp; stride = 1; }
            else if (ndp == 1) { base = xp + Np2 * yp; stride = Np; }
            else { base = xp + Np * yp; stride = Np2; }
            double tr = 0., dn = 0.;
            for (int l = 0; l < Np; ++l) {
              const double v = up[base + l * stride];
              tr += Tp.e[sp][l] * v;
              dn += Tp.d[sp][l] * v;
            }
            NF[kp] = tr;
            NF[mM2 + kp] = dn;
          }
          __syncwarp(gmask);
          const double *Dp = a.tabdata + a.tabs->dg_off[Np];
          for (int kp = lane; kp < Np2; kp += GS) {
            const int yp = hp_div(kp, invNp), xp = kp - yp * Np;
            double d1 = 0., d2 = 0.;
            for (int l = 0; l < Np; ++l) {
              d1 += __ldg(Dp + xp * Np + l) * NF[Np * yp + l];
              d2 += __ldg(Dp + yp * Np + l) * NF[Np * l + xp];
            }
            NF[2 * mM2 + kp] = d1;
            NF[3 * mM2 + kp] = d2;
          }
          __syncwarp(gmask);
          const int j = (M == 4) ? sub[i] : 0;
          const int hxp = p_big ? 1 + (j & 1) : 0, hyp = p_big ? 1 + (j >> 1) : 0;
          if (Np == Mq && !p_big) {
            qp = NF;
          } else {
            const double *Txp = a.tabdata + a.tabs->t_off[Np][Mq][hxp], *Typ = a.tabdata + a.tabs->t_off[Np][Mq][hyp];
            hp_interp4<0>(QP, mM2, NF, mM2, Txp, Typ, Np, Mq, TMP, lane, GS, gmask);
            qp = QP;
          }
        }
        // fluxes at the mortar points, (-)-side order
        {
          const double *mo = a.mortar + S + off;
          const FastTab &Tq = c_tab[Mq];
          const int t1p = (ndp == 0) ? 1 : 0, t2p = (ndp == 2) ? 1 : 2;
          const float invMq = 1.f / (float)Mq;
          for (int k = lane; k < Mq2; k += GS) {
            const int ky = hp_div(k, invMq), kx = k - ky * Mq;
            const double Am[3] = {mo[k], mo[a.MP + k], mo[2 * a.MP + k]};
            const double Sg = mo[6 * a.MP + k];
            const double um = qm[k];
            const double Fm = Am[nd] * qm[qms + k] + Am[t1] * qm[2 * qms + k] + Am[t2] * qm[3 * qms + k];
            double du, t13, c2;
            if (faces_p) {
              const int kp = perm_xy(Mq, perm, kx, ky);
              const double Ap[3] = {mo[3 * a.MP + k], mo[4 * a.MP + k], mo[5 * a.MP + k]};
              const double Fp = Ap[ndp] * qp[qps + kp] + Ap[t1p] * qp[2 * qps + kp] + Ap[t2p] * qp[3 * qps + kp];
              du = um - qp[kp];
              t13 = -.5 * (Fp + Fm) + Sg * du;
              c2 = -.5;
            } else {
              du = um - mo[7 * a.MP + k];
              t13 = -Fm + Sg * du;
              c2 = -1.;
            }
            const double wxy = Tq.w[kx] * Tq.w[ky];
            const double wdu = wxy * c2 * du;
            TT[k] = wxy * t13;
            TT[mM2 + k] = Am[nd] * wdu;
            TT[2 * mM2 + k] = Am[t1] * wdu;
            TT[3 * mM2 + k] = Am[t2] * wdu;
          }
        }
        __syncwarp(gmask);
        if (own_id) {
          for (int q = 0; q < 4; ++q)
            for (int lt = lane; lt < NN; lt += GS) RF[q * NN + lt] += TT[q * mM2 + lt];
          __syncwarp(gmask);
        } else {
          hp_project4<N>(RF, NN, TT, mM2, Tx, Ty, Mq, TMP, lane, GS, gmask);
        }
      }
      off += Mq2;
    }
    // tangential Dg^T on the face, rank-1 lifting along the normal lines
    for (int lt = lane; lt < NN; lt += GS) {
      const int lx = lt % N, ly = lt / N;
      double ve = RF[lt];
      const double vd = RF[NN + lt];
      const double *w1 = RF + 2 * NN, *w2 = RF + 3 * NN;
#pragma unroll
      for (int i = 0; i < N; ++i) ve += T.Dg[i][lx] * w1[N * ly + i] + T.Dg[i][ly] * w2[N * i + lx];
      int base, stride;
      if (nd == 0) { base = N * lx + NN * ly; stride = 1; }
      else if (nd == 1) { base = lx + NN * ly; stride = N; }
      else { base = lx + N * ly; stride = NN; }
#pragma unroll
      for (int l = 0; l < N; ++l) V0[base + l * stride] += T.e[sd][l] * ve + T.d[sd][l] * vd;
    }
    __syncwarp(gmask);
  }

  if (FACE_ONLY) {
    for (int k = lane; k < N3; k += GS) a.extra[ns + k] = V0[k];
  }
  if (!FACE_ONLY) {
  // ---- back to the LGL basis: I^T along a, b, c
  for (int lt = lane; lt < NN; lt += GS) {
    double l[N];
    double *p = V0 + N * lt;           // (b,c) = lt, line along a
#pragma unroll
    for (int i = 0; i < N; ++i) l[i] = p[i];
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.I[i][i2] * l[i];
      p[i2] = s;
    }
  }
  __syncwarp(gmask);
  for (int lt = lane; lt < NN; lt += GS) {
    const int lx = lt % N, ly = lt / N;
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
  __syncwarp(gmask);
  {
    const EpiCoef ec = epi_coef(a.epi);
    for (int lt = lane; lt < NN; lt += GS) {
      double l[N];
#pragma unroll
      for (int i = 0; i < N; ++i) l[i] = V0[lt + NN * i];
#pragma unroll
      for (int c = 0; c < N; ++c) {
        double s = 0.;
#pragma unroll
        for (int i = 0; i < N; ++i) s += T.I[i][c] * l[i];
        store_result(a.epi, ec, a.Au, ns + lt + NN * c, s);
      }
    }
  }
  }
}
