// d4b_fast.cu -- uniform-degree conforming SIPG apply: the hot kernels.
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
//     I = LGL->Gauss interpolation.  Kernel 1 (interp_kernel) writes uh for
//     every element once per apply (a pure streaming pass, 16 B/DOF).
//   * Kernel 2 (fast_apply_kernel) works entirely in that basis: the
//     reference-space gradient is three 1-D sweeps with the Gauss-basis
//     derivative matrix Dg, the stiffness metric G acts point-wise, and a face
//     trace / normal derivative AT THE FACE GAUSS POINTS is a 1-D contraction
//     of each line normal to the face with the end-point vectors e (values)
//     and d (derivatives) -- for the element itself from registers, for the
//     neighbour from its uh lines read straight from global memory (L2 hits).
//     No 2-D face interpolation is needed on either side.
//   * Residuals are accumulated in the same basis (face terms are rank-1
//     updates along the normal lines) and mapped back once:
//     Au = (I^T x I^T x I^T) rh.
//   * N^2 threads per element, each holding one line of N values in registers;
//     the three line orientations are exchanged through shared memory.
//   * Face geometric data (7 doubles per face Gauss point: A_m[3], A_p[3],
//     sj*sigma) is stored ONCE per face and shared by its two elements: the
//     non-owning side reads it through the face permutation with the sign
//     flipped (n_p sj_p = -n_m sj_m).  The set-up verifies that identity on the
//     reference's own arrays and falls back to per-side records if it fails.
// Each element's Au is written exactly once, no atomics (deterministic).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <map>
#include "d4b_internal.h"

namespace {

struct FastTab {
  double I[8][8];    // LGL -> Gauss interpolation, [gauss][lgl]
  double Dg[8][8];   // derivative in the Gauss-point Lagrange basis, [point][basis]
  double e[2][8];    // Gauss basis evaluated at -1 / +1
  double d[2][8];    // Gauss basis derivative at -1 / +1
  double w[8];       // Gauss weights
  double B[8][8];    // I I^T (Gauss x Gauss): the back-transform of the Gauss-resident smoother (HAT kernels)
  double Iinv[8][8]; // Gauss -> LGL, [lgl][gauss]: u = (Iinv x Iinv x Iinv) uh
};

__constant__ FastTab c_tab[9];   // indexed by N

// code bits of an element face
constexpr int kOwn = 1 << 6;     // this element owns the face record
constexpr int64_t kNoNbr = INT64_MIN;   // neighbour offset of a boundary face (offsets into a peer's uhat may be negative)

// what to do with a finished entry of A u (D4bEpilogue, d4b_internal.h)
struct EpiCoef { double alpha, beta; };
// the recurrence coefficients of this launch: kernel arguments, or -- when the spectrum
// estimate never left the device (d4b_solver.cu) -- a pair in device memory
__device__ __forceinline__ EpiCoef epi_coef(const D4bEpilogue &e) {
  EpiCoef c;
  if ((e.mode == 1 || e.mode == 3) && e.coef) { c.alpha = e.coef[0]; c.beta = e.coef[1]; }
  else { c.alpha = e.alpha; c.beta = e.beta; }
  return c;
}
__device__ __forceinline__ void store_result(const D4bEpilogue &e, const EpiCoef &c, double *Au, int64_t i, double s) {
  if (e.add) s += e.add[i];
  if (e.mode == 0) { Au[i] = s; return; }
  if (e.mode == 2) { Au[i] = e.rhs[i] - s; return; }
  const double r = (e.rhs[i] - s) * c.alpha;
  const double pn = r + c.beta * e.p[i];
  e.p[i] = pn;
  e.u[i] += pn;
}

// MASS kernels: the mass-type term of one quadrature point in the Gauss basis (point-wise there).  Linear, W uh -- the
// Jacobian of a nonlinear problem about u0, W = w J f_u(x, V u0) (d4b_mesh_linearize / d4b_mesh_set_mass_diag) -- or,
// with B set, the nonlinear functor itself, W (B + uh)^k with W = w J a (d4b_build_residual: the term
// d4est_quadrature_apply_fofufofvlj adds, src/Quadrature/d4est_quadrature.c:776-936, for f = a (b + u)^k)
__device__ __forceinline__ double mass_point(const double *__restrict__ W, const double *__restrict__ B, int k, int64_t q,
                                             double uh) {
  const double w = W[q];
  if (!B) return w * uh;
  const double x = B[q] + uh;
  double r = 1., p = x;
  for (int n = k < 0 ? -k : k; n; n >>= 1) {
    if (n & 1) r *= p;
    p *= p;
  }
  return w * (k < 0 ? 1. / r : r);
}

struct FastArgs {
  const int32_t *elem;          // [n_elem] element ids of this class
  int first;                    // this launch covers class positions [first, n_elem)
  int n_elem;
  const int64_t *elem_ns;       // [n_elem] nodal offset of each class position (one load instead of elem -> stride)
  const double *FG;             // [n_elem][6][N^3] volume metric
  const double *FF;             // face records, [7][N^2] each
  const int64_t *face_rec;      // [6*n_elem] offset of the face record in FF
  const int64_t *face_nbr;      // [6*n_elem] nodal offset of the neighbour or kNoNbr
  const int32_t *face_code;     // [6*n_elem] f_p | perm << 3 | own << 6
  const double *uh;             // Gauss-collocated input
  double *Au;
  D4bEpilogue epi;
  const double *W;              // MASS kernels: weighted diagonal at the quadrature points (d4b_mesh_set_mass_diag)
  const double *nlB;            // MASS kernels, residual: the functor's b at the quadrature points (see mass_point)
  int nlK;
  const int64_t *quad_stride;
  int prefetch;                 // L2 prefetch of the block's volume metric and face records at kernel start
  const unsigned char *has_extra;   // face split, see FlatArgs
  const double *extra;
};

// Placement of the N^3 block in shared memory for the three line orientations: an XOR swizzle for
// N = 4, 8 (conflict-free), padded pitches for the other N (a search over pitches and the element
// stride for the fewest 16-lane wavefronts, profiles/pad_search.py: N = 6 228 -> 162, N = 3 78 -> 57; flat kernel +2..3 % at p = 2, 4, 5).
template <int N> struct VLay { static constexpr int P1 = N, P2 = N * N, PAD = 0; };
template <> struct VLay<3> { static constexpr int P1 = 3, P2 = 18, PAD = 10; };
template <> struct VLay<5> { static constexpr int P1 = 5, P2 = 25, PAD = 10; };
template <> struct VLay<6> { static constexpr int P1 = 9, P2 = 54, PAD = 5; };
// doubles one block occupies (largest index + 1)
template <int N>
struct VSize { static constexpr int value = (N == 4 || N == 8) ? N * N * N : (VLay<N>::P2 + VLay<N>::P1) * (N - 1) + N; };
template <int N>
__device__ __forceinline__ int vix(int a, int b, int c) {
  if (N == 4) return 16 * c + ((a + 4 * b) ^ (5 * c));
  if (N == 8) return 64 * c + 8 * (b ^ (c & 1)) + (a ^ b);
  return a + VLay<N>::P1 * b + VLay<N>::P2 * c;
}

template <int N>
struct Cfg {
  static constexpr int NN = N * N;
  static constexpr int THREADS = (N == 7 || N == 6) ? 256 : 128;
  static constexpr int EPB = THREADS / NN;           // elements per CTA
};

// launch geometry of fast_apply_kernel by N = 2..8 (D4B_FAST_THREADS / D4B_FAST_MINB: compile-time
// overrides for A/B sweeps, profiles/build_variants.sh); a one-warp block synchronises with __syncwarp
#ifndef D4B_FAST_THREADS
#define D4B_FAST_THREADS {0, 0, 32, 128, 32, 128, 128, 256, 64}   // N = 7: 245 of 256 lanes instead of 49 of 64 (+6 %, profiles/r02_curved_256thread_ab.txt)
#endif
#ifndef D4B_FAST_MINB
#define D4B_FAST_MINB {0, 0, 16, 5, 20, 4, 4, 2, 5}
#endif
template <int N>
struct CurvedCfg {
  static constexpr int NN = N * N, N3 = N * N * N;
  static constexpr int kThreads[9] = D4B_FAST_THREADS;
  static constexpr int kMinBlocks[9] = D4B_FAST_MINB;
  static constexpr int THREADS = kThreads[N];
  static constexpr int MIN_BLOCKS = kMinBlocks[N];
  static constexpr int EPB = THREADS / NN;
  // shared doubles per element: V0,V1,V2 + FT (12*NN) + NB (8*NN) + NB2 (4*NN)
  static constexpr int VS = VSize<N>::value;              // doubles of one swizzled / padded N^3 block
  static constexpr int SH_PER_ELEM = 3 * VS + 24 * NN + VLay<N>::PAD;
};
template <int THREADS>
__device__ __forceinline__ void curved_sync() {
  if (THREADS == 32) { __syncwarp(); return; }
  __syncthreads();
}
__device__ __forceinline__ void l2_prefetch_bulk(const void *ptr, unsigned bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ptr), "r"(bytes) : "memory");
}

// my-order face point (a,b) -> neighbour-order (xp,yp); returns yp*n + xp
__device__ __forceinline__ int perm_xy(int n, int perm, int a, int b) {
  int row = (perm & 4) ? a : b, col = (perm & 4) ? b : a;
  if (perm & 2) row = n - 1 - row;
  if (perm & 1) col = n - 1 - col;
  return row * n + col;
}


// ---- kernel 1: uh = (I x I x I) u ---------------------------------------------
// MODE 0: uh = (I x I x I) u
// MODE 1: the direction update of CG fused in: d = r + (bnum / bden) d is written back and interpolated
//         (d4est_linalg_vec_xpby of d4est_solver_cg_eigs.c:257 followed by the next apply's first step)
// MODE 2: the way back, out = (Iinv x Iinv x Iinv) in
struct InterpArgs {
  int n_elem;
  const int64_t *elem_ns;          // [n_elem] nodal offset of each listed element
  int64_t ns0;                     // ns_linear: elem_ns[i] = ns0 + i N^3
  int ns_linear;
  const double *in;
  double *out;
  const double *r, *bnum, *bden;   // MODE 1
  double *d;                       // MODE 1: = in, written
};
template <int N, int MODE>
__global__ void __launch_bounds__(Cfg<N>::THREADS) interp_kernel(InterpArgs a) {
  using C = Cfg<N>;
  constexpr int NN = C::NN;
  __shared__ double sm[C::EPB * VSize<N>::value];
  const FastTab &T = c_tab[N];
  const int slot = threadIdx.x / NN, lt = threadIdx.x % NN, lx = lt % N, ly = lt / N;
  const int ei = blockIdx.x * C::EPB + slot;
  const bool active = (slot < C::EPB) && (ei < a.n_elem);
  double *V = sm + (slot < C::EPB ? slot : 0) * VSize<N>::value;
  const int64_t ns = !active ? 0 : a.ns_linear ? a.ns0 + (int64_t)ei * (N * NN) : a.elem_ns[ei];
  if (active) {
    double l[N];
#pragma unroll
    for (int c = 0; c < N; ++c) l[c] = a.in[ns + lx + N * ly + NN * c];
    if (MODE == 1) {
      const double beta = *a.bnum / *a.bden;
#pragma unroll
      for (int c = 0; c < N; ++c) {
        l[c] = a.r[ns + lx + N * ly + NN * c] + beta * l[c];
        a.d[ns + lx + N * ly + NN * c] = l[c];
      }
    }
#pragma unroll
    for (int c2 = 0; c2 < N; ++c2) {
      double s = 0.;
#pragma unroll
      for (int c = 0; c < N; ++c) s += (MODE == 2 ? T.Iinv[c2][c] : T.I[c2][c]) * l[c];
      V[vix<N>(lx, ly, c2)] = s;
    }
  }
  __syncthreads();
  if (active) {   // (b,c) = (lx,ly), line along a
    double l[N];
#pragma unroll
    for (int i = 0; i < N; ++i) l[i] = V[vix<N>(i, lx, ly)];
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += (MODE == 2 ? T.Iinv[i2][i] : T.I[i2][i]) * l[i];
      V[vix<N>(i2, lx, ly)] = s;
    }
  }
  __syncthreads();
  if (active) {   // (a,c) = (lx,ly), line along b; the result goes straight out
    double l[N];
#pragma unroll
    for (int i = 0; i < N; ++i) l[i] = V[vix<N>(lx, i, ly)];
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += (MODE == 2 ? T.Iinv[i2][i] : T.I[i2][i]) * l[i];
      a.out[ns + lx + N * i2 + NN * ly] = s;
    }
  }
}

// ---- kernel 2: the apply in the Gauss basis ------------------------------------
template <int N, bool MASS>
__global__ void __launch_bounds__(CurvedCfg<N>::THREADS, CurvedCfg<N>::MIN_BLOCKS) fast_apply_kernel(FastArgs a) {
  using C = CurvedCfg<N>;
  constexpr int NN = C::NN, N3 = C::N3;
  extern __shared__ double smem[];
  const FastTab &T = c_tab[N];
  const int slot = threadIdx.x / NN;
  const int lt = threadIdx.x % NN;
  const int lx = lt % N, ly = lt / N;
  const int ei = a.first + blockIdx.x * C::EPB + slot;
  const bool active = (slot < C::EPB) && (ei < a.n_elem);
  double *V0 = smem + (size_t)(slot < C::EPB ? slot : 0) * C::SH_PER_ELEM;
  double *V1 = V0 + C::VS, *V2 = V1 + C::VS;
  double *FT = V2 + C::VS;         // [6][2][NN] own trace / normal derivative at face Gauss points
  double *NB = FT + 12 * NN;       // [2][4][NN] neighbour: trace, normal, two tangential derivatives
  double *NB2 = NB + 8 * NN;       // [2][2][NN] weighted tangential term2
  const int64_t ns = active ? a.elem_ns[ei] : 0;
  double gt[N];
  // the volume metric and the face records are first needed after the forward sweeps: ask the L2 now
  if (a.prefetch) {
    if (threadIdx.x == 0) {
      constexpr unsigned kBlockBytes = C::EPB * 6 * N3 * sizeof(double);
      const int e0 = a.first + (int)blockIdx.x * C::EPB, left = a.n_elem - e0;
      l2_prefetch_bulk(a.FG + (size_t)e0 * (6 * N3), left >= C::EPB ? kBlockBytes : left * 6 * N3 * 8);
    }
    if (active) {
      constexpr int L = (7 * NN * 8 + 127) / 128;   // 128-byte lines of one face record
      for (int idx = lt; idx < 6 * L; idx += NN)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a.FF + a.face_rec[6 * ei + idx / L] + (idx % L) * 16));
    }
  }

  // ---- T-orientation (a,b) = (lx,ly): load uh, gradient along c, faces 4/5 ----
  if (active) {
    double uh[N];
#pragma unroll
    for (int c = 0; c < N; ++c) uh[c] = a.uh[ns + lx + N * ly + NN * c];
    double t4 = 0., n4 = 0., t5 = 0., n5 = 0.;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      V0[vix<N>(lx, ly, i)] = uh[i];
      t4 += T.e[0][i] * uh[i]; n4 += T.d[0][i] * uh[i];
      t5 += T.e[1][i] * uh[i]; n5 += T.d[1][i] * uh[i];
    }
#pragma unroll
    for (int c = 0; c < N; ++c) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[c][i] * uh[i];
      gt[c] = s;
    }
    FT[(4 * 2 + 0) * NN + lt] = t4; FT[(4 * 2 + 1) * NN + lt] = n4;
    FT[(5 * 2 + 0) * NN + lt] = t5; FT[(5 * 2 + 1) * NN + lt] = n5;
  }
  curved_sync<C::THREADS>();
  if (active) {
    {   // S-orientation: (a, c) = (lx, ly), line along b: gradient, faces 2/3
      double uh[N];
#pragma unroll
      for (int i = 0; i < N; ++i) uh[i] = V0[vix<N>(lx, i, ly)];
      double t2 = 0., n2 = 0., t3 = 0., n3 = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) {
        t2 += T.e[0][i] * uh[i]; n2 += T.d[0][i] * uh[i];
        t3 += T.e[1][i] * uh[i]; n3 += T.d[1][i] * uh[i];
      }
#pragma unroll
      for (int i2 = 0; i2 < N; ++i2) {
        double s = 0.;
#pragma unroll
        for (int i = 0; i < N; ++i) s += T.Dg[i2][i] * uh[i];
        V2[vix<N>(lx, i2, ly)] = s;
      }
      FT[(2 * 2 + 0) * NN + lt] = t2; FT[(2 * 2 + 1) * NN + lt] = n2;
      FT[(3 * 2 + 0) * NN + lt] = t3; FT[(3 * 2 + 1) * NN + lt] = n3;
    }
    {   // R-orientation: (b, c) = (lx, ly), line along a: gradient, faces 0/1
      double uh[N];
#pragma unroll
      for (int i = 0; i < N; ++i) uh[i] = V0[vix<N>(i, lx, ly)];
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
        V1[vix<N>(i2, lx, ly)] = s;
      }
      FT[(0 * 2 + 0) * NN + lt] = t0; FT[(0 * 2 + 1) * NN + lt] = n0;
      FT[(1 * 2 + 0) * NN + lt] = t1; FT[(1 * 2 + 1) * NN + lt] = n1;
    }
  }
  curved_sync<C::THREADS>();
  if (active) {   // T-orientation: metric, Dg^T along c; V0 becomes the residual accumulator
    double ft[N];
    const double *G = a.FG + (size_t)ei * (6 * N3) + lt;
#pragma unroll
    for (int c = 0; c < N; ++c) {
      const int idx = vix<N>(lx, ly, c);
      double gr = V1[idx], gs = V2[idx];
      double G00 = G[0 * N3 + NN * c], G01 = G[1 * N3 + NN * c], G02 = G[2 * N3 + NN * c];
      double G11 = G[3 * N3 + NN * c], G12 = G[4 * N3 + NN * c], G22 = G[5 * N3 + NN * c];
      V1[idx] = G00 * gr + G01 * gs + G02 * gt[c];
      V2[idx] = G01 * gr + G11 * gs + G12 * gt[c];
      ft[c] = G02 * gr + G12 * gs + G22 * gt[c];
    }
    // mass-type diagonal: in the Gauss basis V^T diag(W) V u is point-wise, r += W uh
    const double *Wd = MASS ? a.W + a.quad_stride[a.elem[ei]] + lt : nullptr;
    const double *Bd = (MASS && a.nlB) ? a.nlB + a.quad_stride[a.elem[ei]] + lt : nullptr;
    const double *xtr = (a.has_extra && a.has_extra[ei]) ? a.extra + ns + lt : nullptr;
#pragma unroll
    for (int c = 0; c < N; ++c) {
      double s = MASS ? mass_point(Wd, Bd, a.nlK, NN * c, V0[vix<N>(lx, ly, c)]) : 0.;
      if (xtr) s += xtr[NN * c];
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i][c] * ft[i];
      V0[vix<N>(lx, ly, c)] = s;
    }
  }

  // ---- faces, one opposite pair (same normal direction) at a time -------------
  // thread (lx, ly) is face Gauss point (lx, ly) of both faces of the pair
#pragma unroll
  for (int nd = 0; nd < 3; ++nd) {
    int code[2] = {0, 0};
    bool interior[2] = {false, false};
    double fd[2][7];   // face record values at my point: A_m[3], A_p[3], sj*sigma
    double um[2], fm[2];
    if (active) {
      const int t1 = (nd == 0) ? 1 : 0, t2 = (nd == 2) ? 1 : 2;
#pragma unroll
      for (int sd = 0; sd < 2; ++sd) {
        const int f = 2 * nd + sd;
        const int64_t nb = a.face_nbr[6 * ei + f];
        code[sd] = a.face_code[6 * ei + f];
        interior[sd] = nb != kNoNbr;
        const double *rec = a.FF + a.face_rec[6 * ei + f];
        const int perm = (code[sd] >> 3) & 7;
        const bool own = code[sd] & kOwn;
        // face record at my point; the non-owner reads through the permutation
        // and swaps / negates the two metric triples
        const int k = own ? lt : perm_xy(N, perm, lx, ly);
#pragma unroll
        for (int l = 0; l < 3; ++l) {
          double x = rec[l * NN + k], y = rec[(3 + l) * NN + k];
          fd[sd][l] = own ? x : -y;
          fd[sd][3 + l] = own ? y : -x;
        }
        fd[sd][6] = rec[6 * NN + k];
        if (interior[sd]) {   // neighbour normal line -> trace, normal derivative (its own order)
          const int fp = code[sd] & 7, ndp = fp >> 1, sp = fp & 1;
          const double *up = a.uh + nb;
          int base, stride;
          if (ndp == 0) { base = N * lx + NN * ly; stride = 1; }
          else if (ndp == 1) { base = lx + NN * ly; stride = N; }
          else { base = lx + N * ly; stride = NN; }
          double dn = 0., tr = 0.;
#pragma unroll
          for (int l = 0; l < N; ++l) {
            double val = up[base + l * stride];
            tr += T.e[sp][l] * val;
            dn += T.d[sp][l] * val;
          }
          NB[(sd * 4 + 0) * NN + lt] = tr;
          NB[(sd * 4 + 1) * NN + lt] = dn;
        }
        // own side
        const double *um_a = FT + (f * 2 + 0) * NN;
        double dmn = FT[(f * 2 + 1) * NN + lt];
        double dm1 = 0., dm2 = 0.;
#pragma unroll
        for (int i = 0; i < N; ++i) {
          dm1 += T.Dg[lx][i] * um_a[N * ly + i];
          dm2 += T.Dg[ly][i] * um_a[N * i + lx];
        }
        um[sd] = um_a[lt];
        fm[sd] = fd[sd][nd] * dmn + fd[sd][t1] * dm1 + fd[sd][t2] * dm2;
      }
    }
    curved_sync<C::THREADS>();
    if (active) {   // neighbour tangential derivatives at its own face point (lx, ly)
#pragma unroll
      for (int sd = 0; sd < 2; ++sd) {
        if (!interior[sd]) continue;
        const double *tr = NB + (sd * 4 + 0) * NN;
        double d1 = 0., d2 = 0.;
#pragma unroll
        for (int i = 0; i < N; ++i) {
          d1 += T.Dg[lx][i] * tr[N * ly + i];
          d2 += T.Dg[ly][i] * tr[N * i + lx];
        }
        NB[(sd * 4 + 2) * NN + lt] = d1;
        NB[(sd * 4 + 3) * NN + lt] = d2;
      }
    }
    curved_sync<C::THREADS>();
    double ve[2], vd[2];
    if (active) {   // fluxes at my face Gauss point
      const double wxy = T.w[lx] * T.w[ly];
      const int t1 = (nd == 0) ? 1 : 0, t2 = (nd == 2) ? 1 : 2;
#pragma unroll
      for (int sd = 0; sd < 2; ++sd) {
        double du, t13, c2;
        if (interior[sd]) {
          const int fp = code[sd] & 7, perm = (code[sd] >> 3) & 7;
          const int ndp = fp >> 1;
          const int kp = perm_xy(N, perm, lx, ly);
          const double *nb = NB + (sd * 4) * NN + kp;
          double up = nb[0], dpn = nb[NN], dp1 = nb[2 * NN], dp2 = nb[3 * NN];
          double Ap0 = fd[sd][3], Ap1 = fd[sd][4], Ap2 = fd[sd][5];
          double Apn = (ndp == 0) ? Ap0 : (ndp == 1) ? Ap1 : Ap2;
          double Apt1 = (ndp == 0) ? Ap1 : Ap0, Apt2 = (ndp == 2) ? Ap1 : Ap2;
          double Fp = Apn * dpn + Apt1 * dp1 + Apt2 * dp2;
          du = um[sd] - up;
          t13 = -.5 * (Fp + fm[sd]) + fd[sd][6] * du;
          c2 = -.5;
        } else {
          du = um[sd] - fd[sd][3];   // Dirichlet data g is stored in the A_p[0] slot
          t13 = -fm[sd] + fd[sd][6] * du;
          c2 = -1.;
        }
        double wdu = wxy * c2 * du;
        ve[sd] = wxy * t13;
        vd[sd] = fd[sd][nd] * wdu;
        NB2[(sd * 2 + 0) * NN + lt] = fd[sd][t1] * wdu;
        NB2[(sd * 2 + 1) * NN + lt] = fd[sd][t2] * wdu;
      }
    }
    curved_sync<C::THREADS>();
    if (active) {   // transposed tangential derivatives, rank-1 update of the normal lines
#pragma unroll
      for (int sd = 0; sd < 2; ++sd) {
        const double *w1 = NB2 + (sd * 2 + 0) * NN, *w2 = NB2 + (sd * 2 + 1) * NN;
        double s = 0.;
#pragma unroll
        for (int i = 0; i < N; ++i) s += T.Dg[i][lx] * w1[N * ly + i] + T.Dg[i][ly] * w2[N * i + lx];
        ve[sd] += s;
      }
#pragma unroll
      for (int l = 0; l < N; ++l) {
        const int v = (nd == 0) ? vix<N>(l, lx, ly) : (nd == 1) ? vix<N>(lx, l, ly) : vix<N>(lx, ly, l);
        V0[v] += T.e[0][l] * ve[0] + T.d[0][l] * vd[0] + T.e[1][l] * ve[1] + T.d[1][l] * vd[1];
      }
    }
    // the barrier after the next pair's first stage orders these V0 updates and NB2 reads
  }
  curved_sync<C::THREADS>();

  // ---- backward --------------------------------------------------------------
  if (active) {   // S-orientation: + Dg^T F_s
    double fs[N];
#pragma unroll
    for (int i = 0; i < N; ++i) fs[i] = V2[vix<N>(lx, i, ly)];
#pragma unroll
    for (int b = 0; b < N; ++b) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i][b] * fs[i];
      V0[vix<N>(lx, b, ly)] += s;
    }
  }
  curved_sync<C::THREADS>();
  if (active) {   // R-orientation: + Dg^T F_r, then I^T along a
    double fr[N], tot[N];
#pragma unroll
    for (int i = 0; i < N; ++i) fr[i] = V1[vix<N>(i, lx, ly)];
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i][i2] * fr[i];
      tot[i2] = V0[vix<N>(i2, lx, ly)] + s;
    }
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.I[i][i2] * tot[i];
      V0[vix<N>(i2, lx, ly)] = s;
    }
  }
  curved_sync<C::THREADS>();
  if (active) {   // S-orientation: I^T along b
    double l[N];
#pragma unroll
    for (int i = 0; i < N; ++i) l[i] = V0[vix<N>(lx, i, ly)];
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.I[i][i2] * l[i];
      V0[vix<N>(lx, i2, ly)] = s;
    }
  }
  curved_sync<C::THREADS>();
  if (active) {   // T-orientation: I^T along c, write Au
    double l[N];
    const EpiCoef ec = epi_coef(a.epi);
#pragma unroll
    for (int i = 0; i < N; ++i) l[i] = V0[vix<N>(lx, ly, i)];
#pragma unroll
    for (int c = 0; c < N; ++c) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.I[i][c] * l[i];
      store_result(a.epi, ec, a.Au, ns + lt + NN * c, s);
    }
  }
}

#include "d4b_flat_kernel.cuh"

template <int N>
void launch_flat(d4b_mesh *m, const D4bFastClass &cls, double *d_Au, int first, int last, const D4bHaloWait *halo = nullptr) {
  using C = FlatCfg<N>;
  if (last <= first) return;
  FlatArgs a;
  if (halo) a.halo = *halo;
  a.elem = cls.d_elem; a.first = first; a.n_elem = last;
  a.elem_ns = cls.d_elem_ns; a.ns0 = cls.ns0; a.ns_linear = cls.ns_linear;
  a.FG = cls.d_FG; a.FGc = cls.d_FGc; a.FF3 = m->d_flat_FF;
  a.meta = static_cast<const FaceMeta *>(cls.d_face_meta);
  a.uh = m->d_uhat; a.Au = d_Au; a.epi = m->epi;
  size_t smem = (size_t)C::EPB * C::SH_PER_ELEM * sizeof(double);
  const double *Wsrc = m->nl_fused ? m->d_nl_wja : m->d_W;
  a.W = Wsrc; a.nlB = m->nl_fused ? m->d_nl_b : nullptr; a.nlK = m->nl_k; a.quad_stride = m->d_quad_stride;
  a.has_extra = cls.d_has_extra; a.extra = m->d_extra;
  static bool configured_dev[kD4bMaxDevices] = {};   // the attribute is per device
  bool &configured = configured_dev[m->device];
  if (!configured) {
    D4B_CUDA(cudaFuncSetAttribute(flat_apply_kernel<N, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    D4B_CUDA(cudaFuncSetAttribute(flat_apply_kernel<N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    D4B_CUDA(cudaFuncSetAttribute(flat_apply_kernel<N, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    D4B_CUDA(cudaFuncSetAttribute(flat_apply_kernel<N, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    D4B_CUDA(cudaFuncSetAttribute(flat_apply_kernel<N, false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    D4B_CUDA(cudaFuncSetAttribute(flat_apply_kernel<N, false, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    D4B_CUDA(cudaFuncSetAttribute(flat_apply_kernel<N, false, false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    D4B_CUDA(cudaFuncSetAttribute(flat_apply_kernel<N, false, true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = true;
  }
  static const int prefetch = getenv("D4B_PREFETCH") ? atoi(getenv("D4B_PREFETCH")) : 2;
  a.prefetch = prefetch;
  int grid = (last - first + C::EPB - 1) / C::EPB;
  const int groups = grid;
  if (m->epi.dot_part) {   // this launch's partial sums follow those of the apply's earlier launches
    a.epi.dot_part = m->epi.dot_part + m->dot_count;
    m->dot_count += (int64_t)groups * (C::THREADS / 32);
    D4B_CHECK(m->dot_count <= m->dot_capacity, "energy dot: %lld partial sums, room for %lld", (long long)m->dot_count,
              (long long)m->dot_capacity);
  }
  if (m->epi.dot_part) {
    D4B_CHECK(!Wsrc && m->epi.mode != 3, "energy dot: not with a mass term or the Gauss-resident smoother");
    if (cls.d_FGc) flat_apply_kernel<N, false, true, false, true><<<grid, C::THREADS, smem, m->stream>>>(a);
    else flat_apply_kernel<N, false, false, false, true><<<grid, C::THREADS, smem, m->stream>>>(a);
  } else if (m->epi.mode == 3) {
    D4B_CHECK(!Wsrc && m->epi.uh_out && m->epi.rhs && m->epi.p, "Gauss-resident smoother: bad epilogue");
    if (cls.d_FGc) flat_apply_kernel<N, false, true, true><<<grid, C::THREADS, smem, m->stream>>>(a);
    else flat_apply_kernel<N, false, false, true><<<grid, C::THREADS, smem, m->stream>>>(a);
  } else if (cls.d_FGc) {
    if (Wsrc) flat_apply_kernel<N, true, true><<<grid, C::THREADS, smem, m->stream>>>(a);
    else flat_apply_kernel<N, false, true><<<grid, C::THREADS, smem, m->stream>>>(a);
  } else if (Wsrc) flat_apply_kernel<N, true><<<grid, C::THREADS, smem, m->stream>>>(a);
  else flat_apply_kernel<N, false><<<grid, C::THREADS, smem, m->stream>>>(a);
  D4B_CUDA(cudaGetLastError());
}

#include "d4b_curved_kernel.cuh"

// curved conforming elements, N <= 5: the shuffle kernel (D4B_CURVED_SHFL=0 selects the general conforming kernel)
inline bool use_curved_shfl(int N) {
  static const int on = getenv("D4B_CURVED_SHFL") ? atoi(getenv("D4B_CURVED_SHFL")) : 1;
  return on != 0 && N <= 5;
}

template <int N>
void launch_curved(d4b_mesh *m, const D4bFastClass &cls, double *d_Au, int first, int last, const D4bHaloWait *halo = nullptr) {
  using C = CurvedShflCfg<N>;
  if (last <= first) return;
  CurvedArgs a;
  if (halo) a.halo = *halo;
  a.elem = cls.d_elem; a.first = first; a.n_elem = last;
  a.elem_ns = cls.d_elem_ns; a.ns0 = cls.ns0; a.ns_linear = cls.ns_linear;
  a.FG = cls.d_FG; a.FF = m->d_fast_FF;
  a.meta = static_cast<const FaceMeta *>(cls.d_face_meta);
  a.uh = m->d_uhat; a.Au = d_Au; a.epi = m->epi;
  const double *Wsrc = m->nl_fused ? m->d_nl_wja : m->d_W;
  a.W = Wsrc; a.nlB = m->nl_fused ? m->d_nl_b : nullptr; a.nlK = m->nl_k; a.quad_stride = m->d_quad_stride;
  a.has_extra = cls.d_has_extra; a.extra = m->d_extra;
  static const int prefetch = getenv("D4B_PREFETCH") ? atoi(getenv("D4B_PREFETCH")) : 1;
  a.prefetch = prefetch;
  const size_t smem = (size_t)C::EPB * C::SH_PER_ELEM * sizeof(double);
  const int grid = (last - first + C::EPB - 1) / C::EPB;
  if (Wsrc) curved_apply_kernel<N, true><<<grid, C::THREADS, smem, m->stream>>>(a);
  else curved_apply_kernel<N, false><<<grid, C::THREADS, smem, m->stream>>>(a);
  D4B_CUDA(cudaGetLastError());
}
template <> void launch_curved<6>(d4b_mesh *, const D4bFastClass &, double *, int, int, const D4bHaloWait *) {}
template <> void launch_curved<7>(d4b_mesh *, const D4bFastClass &, double *, int, int, const D4bHaloWait *) {}
template <> void launch_curved<8>(d4b_mesh *, const D4bFastClass &, double *, int, int, const D4bHaloWait *) {}

#include "d4b_hp_kernel.cuh"

template <int N, int GS>
void launch_hp_gs(d4b_mesh *m, const HpArgs &a, int first, int last, bool mass, bool face_only) {
  constexpr int EPW = 32 / GS;   // elements per warp
  const size_t per_warp = (size_t)EPW * a.smem_doubles * sizeof(double);
  int warps = (int)((200 * 1024) / per_warp);
  warps = warps < 1 ? 1 : warps > 4 ? 4 : warps;
  const size_t smem = per_warp * warps;
  static size_t configured_dev[kD4bMaxDevices] = {};   // per device; 0 = the 48 KB default
  size_t &configured = configured_dev[m->device];
  if (smem > 48 * 1024 && smem > configured) {
    D4B_CUDA(cudaFuncSetAttribute(hp_apply_kernel<N, false, false, GS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    D4B_CUDA(cudaFuncSetAttribute(hp_apply_kernel<N, true, false, GS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    D4B_CUDA(cudaFuncSetAttribute(hp_apply_kernel<N, false, true, GS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = smem;
  }
  const int grid = (last - first + warps * EPW - 1) / (warps * EPW);
  if (face_only) hp_apply_kernel<N, false, true, GS><<<grid, 32 * warps, smem, m->stream>>>(a);
  else if (mass) hp_apply_kernel<N, true, false, GS><<<grid, 32 * warps, smem, m->stream>>>(a);
  else hp_apply_kernel<N, false, false, GS><<<grid, 32 * warps, smem, m->stream>>>(a);
  D4B_CUDA(cudaGetLastError());
}

template <int N>
void launch_hp(d4b_mesh *m, const D4bFastClass &cls, double *d_Au, int first, int last, bool face_only = false) {
  if (last <= first) return;
  HpArgs a;
  a.elem = cls.d_elem; a.first = first; a.n_elem = last;
  a.deg = m->d_deg; a.ghost_deg = m->d_ghost_deg;
  a.nodal_stride = m->d_nodal_stride; a.quad_stride = m->d_quad_stride;
  a.ghost_nodal_stride = m->peer.state == 1 ? m->peer.d_ghost_uh_off : m->d_ghost_nodal_stride;
  const double *Wsrc = m->nl_fused ? m->d_nl_wja : m->d_W;
  a.G = m->d_G; a.LQ = m->local_nodes_quad; a.W = Wsrc; a.nlB = m->nl_fused ? m->d_nl_b : nullptr; a.nlK = m->nl_k;
  a.elem_side = m->d_elem_side; a.elem_sub = m->d_elem_sub;
  a.side_faces_m = m->d_side_faces_m; a.side_faces_p = m->d_side_faces_p; a.side_f_p = m->d_side_f_p;
  a.side_e_m = m->d_side_e_m; a.side_e_p = m->d_side_e_p; a.side_perm = m->d_side_perm;
  a.side_sub_order = m->d_side_sub_order; a.side_mortar_stride = m->d_side_mortar_stride;
  a.mortar = m->d_mortar; a.MP = m->mortar_nodes;
  a.uh = m->d_uhat; a.local_nodes = m->local_nodes; a.Au = d_Au; a.epi = m->epi;
  a.tabs = static_cast<const HpTabs *>(m->d_hp_tabs); a.tabdata = m->d_hp_tabdata;
  a.mM = m->max_deg + 1;
  a.smem_doubles = hp_smem_doubles<N>(a.mM);
  a.face_mask = cls.d_face_mask; a.extra = m->d_extra;
  // Lanes per element: HpGroup<N> (8 at N = 2, 16 at N = 3) as long as four CTAs of an SM still fit their shared memory,
  // else a whole warp -- the lane groups of one warp run their mortar loops one after the other where they differ, and
  // with few resident warps nothing covers that (profiles/r02_hp_kernel.md: a gain up to max degree 3, a loss at 5).
  constexpr int GSmin = HpGroup<N>::size;
  const bool pack = GSmin < 32 && 4 * (4 * (size_t)(32 / GSmin) * a.smem_doubles * sizeof(double) + 1024) <= 228 * 1024;
  if (pack) launch_hp_gs<N, GSmin>(m, a, first, last, Wsrc != nullptr, face_only);
  else launch_hp_gs<N, 32>(m, a, first, last, Wsrc != nullptr, face_only);
}

void build_hp_tabs(d4b_mesh *m) {
  HpTabs T;
  memset(&T, 0xff, sizeof(T));
  std::vector<double> data;
  double xg[8][8], wg[8];
  for (int n = 2; n <= 8; ++n) d4b_tab_gauss(n, xg[n - 1], wg);
  for (int n = 2; n <= 8; ++n) {
    std::vector<double> D(n * n);
    d4b_tab_deriv(n, xg[n - 1], D.data());
    T.dg_off[n] = (int32_t)data.size();
    data.insert(data.end(), D.begin(), D.end());
    for (int mt = n; mt <= 8; ++mt)
      for (int half = 0; half < 3; ++half) {
        double y[8];
        for (int k = 0; k < mt; ++k)
          y[k] = (half == 0) ? xg[mt - 1][k] : (half == 1) ? .5 * (xg[mt - 1][k] - 1.) : .5 * (xg[mt - 1][k] + 1.);
        std::vector<double> E(mt * n);
        d4b_tab_eval(n, xg[n - 1], mt, y, E.data());
        T.t_off[n][mt][half] = (int32_t)data.size();
        data.insert(data.end(), E.begin(), E.end());
      }
  }
  m->d_hp_tabdata = d4b_dev_upload<double>(m, data.data(), data.size());
  m->d_hp_tabs = d4b_dev_upload<int32_t>(m, (const int32_t *)&T, sizeof(T) / sizeof(int32_t));
}

template <int N>
void launch_interp(d4b_mesh *m, const D4bFastClass &cls, const double *d_u, int phase) {
  using C = Cfg<N>;
  InterpArgs a;
  memset(&a, 0, sizeof(a));
  if (phase == 0) {   // local elements
    if (cls.n_hat == 0) return;
    int grid = (cls.n_hat + C::EPB - 1) / C::EPB;
    a.elem_ns = cls.d_hat_ns; a.n_elem = cls.n_hat; a.in = d_u; a.out = m->d_uhat;
    a.ns0 = cls.hat_ns0; a.ns_linear = cls.hat_linear;
    if (m->epi.xpby_r) {   // CG: d = r + beta d on the way (d4b_solver.cu)
      a.r = m->epi.xpby_r; a.bnum = m->epi.xpby_num; a.bden = m->epi.xpby_den; a.d = const_cast<double *>(d_u);
      interp_kernel<N, 1><<<grid, C::THREADS, 0, m->stream>>>(a);
    } else {
      interp_kernel<N, 0><<<grid, C::THREADS, 0, m->stream>>>(a);
    }
  } else {            // ghost elements, from the receive buffer to the tail of uhat
    if (cls.n_ghost_hat == 0) return;
    int grid = (cls.n_ghost_hat + C::EPB - 1) / C::EPB;
    a.elem_ns = cls.d_ghost_hat_ns; a.n_elem = cls.n_ghost_hat;
    a.in = m->d_ughost; a.out = m->d_uhat + m->local_nodes;
    interp_kernel<N, 0><<<grid, C::THREADS, 0, m->stream>>>(a);
  }
  D4B_CUDA(cudaGetLastError());
}

// out = (M x M x M) in over every collocated element of the class: M = I (to the Gauss basis) or its inverse
template <int N>
void launch_basis_change(d4b_mesh *m, const D4bFastClass &cls, const double *in, double *out, bool to_gauss) {
  using C = Cfg<N>;
  if (cls.n_hat == 0) return;
  InterpArgs a;
  memset(&a, 0, sizeof(a));
  a.elem_ns = cls.d_hat_ns; a.n_elem = cls.n_hat; a.in = in; a.out = out;
  a.ns0 = cls.hat_ns0; a.ns_linear = cls.hat_linear;
  const int grid = (cls.n_hat + C::EPB - 1) / C::EPB;
  if (to_gauss) interp_kernel<N, 0><<<grid, C::THREADS, 0, m->stream>>>(a);
  else interp_kernel<N, 2><<<grid, C::THREADS, 0, m->stream>>>(a);
  D4B_CUDA(cudaGetLastError());
}

template <int N>
void launch_fast(d4b_mesh *m, const D4bFastClass &cls, double *d_Au, int first, int last) {
  using C = CurvedCfg<N>;
  if (last <= first) return;
  FastArgs a;
  a.elem = cls.d_elem; a.first = first; a.n_elem = last;
  a.elem_ns = cls.d_elem_ns;
  a.FG = cls.d_FG; a.FF = m->d_fast_FF;
  a.face_rec = cls.d_face_rec; a.face_nbr = cls.d_face_nbr; a.face_code = cls.d_face_code;
  a.uh = m->d_uhat; a.Au = d_Au; a.epi = m->epi;
  size_t smem = (size_t)C::EPB * C::SH_PER_ELEM * sizeof(double);
  const double *Wsrc = m->nl_fused ? m->d_nl_wja : m->d_W;
  a.W = Wsrc; a.nlB = m->nl_fused ? m->d_nl_b : nullptr; a.nlK = m->nl_k; a.quad_stride = m->d_quad_stride;
  a.has_extra = cls.d_has_extra; a.extra = m->d_extra;
  static bool configured_dev[kD4bMaxDevices] = {};
  bool &configured = configured_dev[m->device];
  if (!configured) {
    D4B_CUDA(cudaFuncSetAttribute(fast_apply_kernel<N, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    D4B_CUDA(cudaFuncSetAttribute(fast_apply_kernel<N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = true;
  }
  static const int prefetch = getenv("D4B_PREFETCH") ? atoi(getenv("D4B_PREFETCH")) : 1;
  a.prefetch = prefetch;
  int grid = (last - first + C::EPB - 1) / C::EPB;
  if (Wsrc) fast_apply_kernel<N, true><<<grid, C::THREADS, smem, m->stream>>>(a);
  else fast_apply_kernel<N, false><<<grid, C::THREADS, smem, m->stream>>>(a);
  D4B_CUDA(cudaGetLastError());
}

template <int N>
void launch_interp_range(d4b_mesh *m, const D4bFastClass &cls, const double *d_u, int lo, int hi) {
  using C = Cfg<N>;
  if (hi <= lo) return;
  InterpArgs a;
  memset(&a, 0, sizeof(a));
  a.elem_ns = cls.d_hat_ns + lo; a.n_elem = hi - lo; a.in = d_u; a.out = m->d_uhat;
  a.ns0 = cls.hat_ns0 + (int64_t)lo * (N * N * N); a.ns_linear = cls.hat_linear;
  interp_kernel<N, 0><<<(hi - lo + C::EPB - 1) / C::EPB, C::THREADS, 0, m->stream>>>(a);
  D4B_CUDA(cudaGetLastError());
}

void build_fast_tab(int N, FastTab &T) {
  memset(&T, 0, sizeof(T));
  double xl[8], wl[8], xg[8], wg[8], E[64], ends[2] = {-1., 1.};
  d4b_tab_lobatto(N, xl, wl);
  d4b_tab_gauss(N, xg, wg);
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
  // psi_j'(x) = sum_k psi_k(x) Dg[k][j] (psi_j' has degree N-2, exact in the basis)
  for (int s = 0; s < 2; ++s)
    for (int j = 0; j < N; ++j) {
      double v = 0.;
      for (int k = 0; k < N; ++k) v += T.e[s][k] * Dg[k * N + j];
      T.d[s][j] = v;
    }
  for (int j = 0; j < N; ++j) T.w[j] = wg[j];
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < N; ++j) {
      long double v = 0.L;
      for (int k = 0; k < N; ++k) v += (long double)T.I[i][k] * (long double)T.I[j][k];
      T.B[i][j] = (double)v;
    }
  // Iinv = I^-1 (Gauss-Jordan with partial pivoting in long double; N <= 8, well conditioned)
  long double A[8][16];
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < N; ++j) { A[i][j] = T.I[i][j]; A[i][N + j] = (i == j) ? 1.L : 0.L; }
  for (int c = 0; c < N; ++c) {
    int piv = c;
    for (int r = c + 1; r < N; ++r) if (fabsl(A[r][c]) > fabsl(A[piv][c])) piv = r;
    for (int j = 0; j < 2 * N; ++j) std::swap(A[c][j], A[piv][j]);
    const long double inv = 1.L / A[c][c];
    for (int j = 0; j < 2 * N; ++j) A[c][j] *= inv;
    for (int r = 0; r < N; ++r) {
      if (r == c) continue;
      const long double f = A[r][c];
      for (int j = 0; j < 2 * N; ++j) A[r][j] -= f * A[c][j];
    }
  }
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < N; ++j) T.Iinv[i][j] = (double)A[i][N + j];
}


// ---- set-up on the device: everything that touches per-point data reads m->d_mortar / m->d_G where they already
// are (uploaded or, with device geometry, generated there); only flags and index lists cross the bus ----
struct RecSrc { int32_t side, boundary, NN, nd, ndp, pad; int64_t dst; };

// per conforming side: bit 0 = tangential-free (face_is_flat), bit 1 = consistent with the record of the other side
__global__ void side_flags_kernel(int n_sides, int64_t MP, const double *__restrict__ mortar,
                                  const int32_t *__restrict__ faces_m, const int32_t *__restrict__ faces_p,
                                  const int32_t *__restrict__ f_m, const int32_t *__restrict__ f_p,
                                  const int32_t *__restrict__ e_m, const int32_t *__restrict__ e_p,
                                  const int32_t *__restrict__ perm, const int64_t *__restrict__ mstride,
                                  const int32_t *__restrict__ deg_quad, const int32_t *__restrict__ elem_side,
                                  unsigned char *__restrict__ flags) {
  const int s = blockIdx.x;
  if (s >= n_sides) return;
  __shared__ int not_flat;
  __shared__ double sh_scale[64], sh_err[64];
  if (threadIdx.x == 0) not_flat = 0;
  __syncthreads();
  if (faces_m[s] != 1 || faces_p[s] > 1) { if (threadIdx.x == 0) flags[s] = 0; return; }
  const int32_t em = e_m[4 * s];
  if (em < 0) { if (threadIdx.x == 0) flags[s] = 0; return; }
  const int N = deg_quad[em] + 1, NN = N * N;
  const int64_t S = mstride[s];
  const int nd = f_m[s] >> 1;
  const bool has_p = faces_p[s] == 1;
  const int ndp = has_p ? (f_p[s] >> 1) : 0;
  for (int k = threadIdx.x; k < NN; k += blockDim.x) {
    for (int l = 0; l < 3; ++l) {
      if (l != nd && mortar[(size_t)l * MP + S + k] != 0.) not_flat = 1;
      if (has_p && l != ndp && mortar[(size_t)(3 + l) * MP + S + k] != 0.) not_flat = 1;
    }
  }
  // consistency with the other side's record (only sides with a local neighbour of the same size)
  double scale = 0., err = 0.;
  bool cons_valid = false;
  if (has_p && e_p[4 * s] >= 0) {
    const int32_t ep = e_p[4 * s];
    const int s2 = elem_side[6 * ep + f_p[s]];
    if (s2 >= 0 && deg_quad[ep] + 1 == N && faces_m[s2] == 1 && faces_p[s2] == 1) {
      cons_valid = true;
      const int64_t S2 = mstride[s2];
      const int pm = perm[s];
      for (int k = threadIdx.x; k < NN; k += blockDim.x) {
        const int kp = perm_xy(N, pm, k % N, k / N);
        for (int l = 0; l < 3; ++l) {
          const double am = mortar[(size_t)l * MP + S + k], ap = mortar[(size_t)(3 + l) * MP + S + k];
          const double oam = mortar[(size_t)l * MP + S2 + kp], oap = mortar[(size_t)(3 + l) * MP + S2 + kp];
          scale = fmax(scale, fmax(fabs(am), fabs(ap)));
          err = fmax(err, fmax(fabs(am + oap), fabs(ap + oam)));
        }
        const double sg = mortar[(size_t)6 * MP + S + k], osg = mortar[(size_t)6 * MP + S2 + kp];
        scale = fmax(scale, fabs(sg));
        err = fmax(err, fabs(sg - osg));
      }
    }
  }
  sh_scale[threadIdx.x] = scale; sh_err[threadIdx.x] = err;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int t = 1; t < blockDim.x; ++t) { scale = fmax(scale, sh_scale[t]); err = fmax(err, sh_err[t]); }
    flags[s] = (unsigned char)((not_flat ? 0 : 1) | ((cons_valid && err <= 1e-12 * scale) ? 2 : 0));
  }
}

// face records from the folded mortar planes: 7 planes (+ g in the A_p[0] slot of boundary faces), or the
// compact 3 planes (A_m[n], A_p[n'] or g, sj sigma)
__global__ void gather_records_kernel(int n, const RecSrc *__restrict__ src, int compact, int64_t MP,
                                      const double *__restrict__ mortar, const int64_t *__restrict__ mstride,
                                      double *__restrict__ out) {
  const int r = blockIdx.x;
  if (r >= n) return;
  const RecSrc q = src[r];
  const int64_t S = mstride[q.side];
  double *o = out + q.dst;
  for (int k = threadIdx.x; k < q.NN; k += blockDim.x) {
    if (compact) {
      o[k] = mortar[(size_t)q.nd * MP + S + k];
      o[q.NN + k] = q.boundary ? mortar[(size_t)7 * MP + S + k] : mortar[(size_t)(3 + q.ndp) * MP + S + k];
      o[2 * q.NN + k] = mortar[(size_t)6 * MP + S + k];
    } else {
      for (int pl = 0; pl < 7; ++pl) o[pl * q.NN + k] = mortar[(size_t)pl * MP + S + k];
      if (q.boundary) o[3 * q.NN + k] = mortar[(size_t)7 * MP + S + k];
    }
  }
}

// FG[i][cpl][k] = G[cpl][quad_stride[elem[i]] + k]
__global__ void gather_fg_kernel(int n_elem, int N3, int64_t LQ, const int32_t *__restrict__ elem,
                                 const int64_t *__restrict__ quad_stride, const double *__restrict__ G,
                                 double *__restrict__ FG) {
  const int i = blockIdx.x;
  if (i >= n_elem) return;
  const int64_t qs = quad_stride[elem[i]];
  for (int j = threadIdx.x; j < 6 * N3; j += blockDim.x)
    FG[(size_t)i * 6 * N3 + j] = G[(size_t)(j / N3) * LQ + qs + j % N3];
}

// affine test of a class: Gc[i][cpl] = G / (w_a w_b w_c) at point 0, *bad += 1 if some point deviates
__global__ void affine_kernel(int n_elem, int N, int64_t LQ, const int32_t *__restrict__ elem,
                              const int64_t *__restrict__ quad_stride, const double *__restrict__ G,
                              double *__restrict__ Gc, int *__restrict__ bad) {
  const int i = blockIdx.x;
  if (i >= n_elem) return;
  const FastTab &T = c_tab[N];
  const int NN = N * N, N3 = N * NN;
  const int64_t qs = quad_stride[elem[i]];
  __shared__ double gc[6];
  __shared__ double scale;
  if (threadIdx.x == 0) {
    double sc = 0.;
    for (int cpl = 0; cpl < 6; ++cpl) {
      gc[cpl] = G[(size_t)cpl * LQ + qs] / (T.w[0] * T.w[0] * T.w[0]);
      Gc[(size_t)i * 6 + cpl] = gc[cpl];
      sc = fmax(sc, fabs(gc[cpl]));
    }
    scale = sc;
  }
  __syncthreads();
  for (int j = threadIdx.x; j < 6 * N3; j += blockDim.x) {
    const int cpl = j / N3, k = j % N3;
    const double w3 = T.w[k % N] * T.w[(k / N) % N] * T.w[k / NN];
    if (fabs(G[(size_t)cpl * LQ + qs + k] - gc[cpl] * w3) > 1e-14 * scale * w3) atomicAdd(bad, 1);
  }
}

}  // namespace

void d4b_fast_setup(d4b_mesh *m, const d4b_mesh_desc_t *d, const std::vector<int32_t> &elem_side,
                    std::vector<char> &is_fast) {
  const int E = d->n_local;
  const int64_t MP = d->local_mortar_nodes_quad, LQ = d->local_nodes_quad;
  std::vector<std::vector<int32_t>> by_deg(D4B_MAX_DEG + 1), hat_by_deg(D4B_MAX_DEG + 1);
  for (int e = 0; e < E; ++e) {
    if (d->deg_quad[e] == d->deg[e]) hat_by_deg[d->deg[e]].push_back(e);
    bool ok = (d->deg_quad[e] == d->deg[e]);
    for (int f = 0; f < 6 && ok; ++f) {
      int s = elem_side[6 * e + f];
      if (d->side_faces_m[s] != 1 || d->side_faces_p[s] > 1) { ok = false; break; }
      if (d->side_faces_p[s] == 1) {
        int32_t ep = d->side_e_p[4 * s];
        const int dp = ep >= 0 ? d->deg[ep] : d->ghost_deg[-ep - 1];
        const int dqp = ep >= 0 ? d->deg_quad[ep] : d->ghost_deg_quad[-ep - 1];
        if (dp != d->deg[e] || dqp != d->deg_quad[e]) { ok = false; break; }
      }
    }
    if (ok) { is_fast[e] = 1; by_deg[d->deg[e]].push_back(e); }
  }
  // hp-irregular elements the mortar kernel (d4b_hp_kernel.cuh) takes: collocated quadrature on the
  // element and on every element of its six flux sides; D4B_NO_HP=1 leaves them to the general kernel
  std::vector<std::vector<int32_t>> hp_by_deg(D4B_MAX_DEG + 1);
  const bool allow_hp = !(getenv("D4B_NO_HP") && atoi(getenv("D4B_NO_HP")) != 0);
  auto collocated = [&](int32_t r) {
    return r >= 0 ? d->deg_quad[r] == d->deg[r] : d->ghost_deg_quad[-r - 1] == d->ghost_deg[-r - 1];
  };
  // Face split: such an element runs in the conforming kernels with its irregular faces
  // switched off (an all-zero boundary record) and a FACE_ONLY launch of the mortar kernel supplies those
  // faces' contributions through m->d_extra; split_mask[e] = the irregular faces.
  std::vector<unsigned char> split_mask(E, 0);
  std::vector<std::vector<int32_t>> split_by_deg(D4B_MAX_DEG + 1);
  std::vector<int32_t> eligible;
  int64_t n_irregular_faces = 0;
  for (int e = 0; e < E && allow_hp; ++e) {
    if (is_fast[e] || d->deg_quad[e] != d->deg[e]) continue;
    bool ok = true;
    unsigned mask = 0;
    for (int f = 0; f < 6 && ok; ++f) {
      const int s = elem_side[6 * e + f];
      for (int i = 0; i < d->side_faces_m[s] && ok; ++i) ok = collocated(d->side_e_m[4 * s + i]);
      for (int i = 0; i < d->side_faces_p[s] && ok; ++i) ok = collocated(d->side_e_p[4 * s + i]);
      bool regular = d->side_faces_m[s] == 1 && d->side_faces_p[s] <= 1;
      if (regular && d->side_faces_p[s] == 1) {
        const int32_t ep = d->side_e_p[4 * s];
        regular = (ep >= 0 ? d->deg[ep] : d->ghost_deg[-ep - 1]) == d->deg[e];
      }
      if (!regular) { mask |= 1u << f; ++n_irregular_faces; }
    }
    if (!ok) continue;
    is_fast[e] = 1;
    split_mask[e] = (unsigned char)mask;
    eligible.push_back(e);
  }
  const bool any_hp = !eligible.empty();
  // element-wise mortar kernel or face split?  The split pays when an irregular element has few irregular
  // faces (AMR interfaces: one or two); with most faces irregular it would process the element twice.
  // D4B_SPLIT=0/1 overrides.
  bool split = any_hp && n_irregular_faces <= (int64_t)(2.5 * (double)eligible.size());
  if (getenv("D4B_SPLIT")) split = any_hp && atoi(getenv("D4B_SPLIT")) != 0;
  for (int e : eligible) {
    if (split) split_by_deg[d->deg[e]].push_back(e);
    else { hp_by_deg[d->deg[e]].push_back(e); split_mask[e] = 0; }
  }
  if (split)
    for (int p = 1; p <= D4B_MAX_DEG; ++p) {
      std::vector<int32_t> merged(by_deg[p].size() + split_by_deg[p].size());
      std::merge(by_deg[p].begin(), by_deg[p].end(), split_by_deg[p].begin(), split_by_deg[p].end(), merged.begin());
      by_deg[p].swap(merged);
    }
  bool any = any_hp;
  for (int p = 1; p <= D4B_MAX_DEG; ++p) any = any || !by_deg[p].empty();
  if (!any) return;
  if (any_hp) build_hp_tabs(m);
  const bool allow_flat = !(getenv("D4B_NO_FLAT") && atoi(getenv("D4B_NO_FLAT")) != 0);

  // per-side flags from the folded mortar planes where they lie (device): tangential-free, and -- for a side
  // with a local neighbour -- consistent with the record written from the other side (n_p sj_p = -n_m sj_m)
  std::vector<unsigned char> side_flags((size_t)std::max(d->n_sides, 1), 0);
  if (d->n_sides > 0) {
    unsigned char *d_flags = nullptr;
    D4B_CUDA(cudaMalloc(&d_flags, (size_t)d->n_sides));
    side_flags_kernel<<<d->n_sides, 64, 0, m->stream>>>(d->n_sides, MP, m->d_mortar, m->d_side_faces_m, m->d_side_faces_p,
                                                        m->d_side_f_m, m->d_side_f_p, m->d_side_e_m, m->d_side_e_p,
                                                        m->d_side_perm, m->d_side_mortar_stride, m->d_deg_quad,
                                                        m->d_elem_side, d_flags);
    D4B_CUDA(cudaGetLastError());
    D4B_CUDA(cudaMemcpyAsync(side_flags.data(), d_flags, (size_t)d->n_sides, cudaMemcpyDeviceToHost, m->stream));
    D4B_CUDA(cudaStreamSynchronize(m->stream));
    cudaFree(d_flags);
  }
  auto face_is_flat = [&](int s, int) -> bool { return side_flags[s] & 1; };

  // ---- face records, one per face, shared by both sides when consistent.  Only their places are decided
  // here; gather_records_kernel fills them from m->d_mortar afterwards.
  int64_t ff_size = 0, ff3_size = 0;
  std::vector<RecSrc> src7, src3;
  std::vector<int64_t> rec_of((size_t)6 * E, -1), rec3_of((size_t)6 * E, -1);   // owning (element, face) -> record
  auto write_record = [&](int s, int NN, bool boundary) -> int64_t {
    const int64_t off = ff_size;
    ff_size += (int64_t)7 * NN;
    src7.push_back(RecSrc{s, boundary ? 1 : 0, NN, 0, 0, 0, off});
    return off;
  };
  auto write_record3 = [&](int s, int NN, bool boundary) -> int64_t {
    const int64_t idx = ff3_size / ((int64_t)3 * NN);
    const int nd = d->side_f_m[s] >> 1, ndp = boundary ? 0 : (d->side_f_p[s] >> 1);
    src3.push_back(RecSrc{s, boundary ? 1 : 0, NN, nd, ndp, 0, ff3_size});
    ff3_size += (int64_t)3 * NN;
    return idx;
  };
  // can side s use the record written from the other side?
  auto consistent = [&](int s, int, int, int) -> bool { return side_flags[s] & 2; };

  // nodal offsets of a list of elements (the kernels' first load: one array instead of list -> stride table)
  auto upload_ns = [&](const std::vector<int32_t> &els, const int64_t *stride) -> int64_t * {
    std::vector<int64_t> ns(els.size());
    for (size_t i = 0; i < els.size(); ++i) ns[i] = stride[els[i]];
    return d4b_dev_upload<int64_t>(m, ns.data(), ns.size());
  };
  auto set_hat_linear = [&](D4bFastClass &c, const std::vector<int32_t> &els, int n3) {
    c.hat_linear = !els.empty();
    c.hat_ns0 = els.empty() ? 0 : d->nodal_stride[els[0]];
    for (size_t i = 0; i < els.size() && c.hat_linear; ++i)
      c.hat_linear = d->nodal_stride[els[i]] == c.hat_ns0 + (int64_t)i * n3;
  };
  FastTab tabs[9];
  memset(tabs, 0, sizeof(tabs));
  int64_t n_shared = 0, n_split = 0;
  for (int p = 1; p <= D4B_MAX_DEG; ++p) {
    std::vector<int32_t> ghost_hat_p;
    for (int g = 0; g < d->n_ghost; ++g)
      if (d->ghost_deg[g] == p && d->ghost_deg_quad[g] == p) ghost_hat_p.push_back(g);
    // with hp elements around, every collocated element may be somebody's neighbour: its uh is needed
    const bool hat_needed = any_hp && (!hat_by_deg[p].empty() || !ghost_hat_p.empty());
    if (by_deg[p].empty() && hp_by_deg[p].empty() && !hat_needed) continue;
    const int N = p + 1, NN = N * N, N3 = N * NN;
    build_fast_tab(N, tabs[N]);
    D4B_CUDA(cudaMemcpyToSymbol(c_tab, &tabs[N], sizeof(FastTab), (size_t)N * sizeof(FastTab)));
    std::vector<int32_t> els_kind[2];
    for (int e : by_deg[p]) {
      bool flat = allow_flat;
      for (int f = 0; f < 6 && flat; ++f)
        if (!((split_mask[e] >> f) & 1)) flat = face_is_flat(elem_side[6 * e + f], NN);
      els_kind[flat ? 1 : 0].push_back(e);
    }
    // elements without a ghost neighbour first: they run while the halo is in flight
    auto touches_ghost = [&](int e) {
      for (int f = 0; f < 6; ++f) {
        const int s = elem_side[6 * e + f];
        for (int i = 0; i < d->side_faces_m[s]; ++i) if (d->side_e_m[4 * s + i] < 0) return true;
        for (int i = 0; i < d->side_faces_p[s]; ++i) if (d->side_e_p[4 * s + i] < 0) return true;
      }
      return false;
    };
    int n_interior_kind[2] = {0, 0};
    for (int kind = 0; kind < 2; ++kind) {
      auto mid = std::stable_partition(els_kind[kind].begin(), els_kind[kind].end(),
                                       [&](int e) { return !touches_ghost(e); });
      n_interior_kind[kind] = (int)(mid - els_kind[kind].begin());
    }
    const std::vector<int32_t> &ghost_hat = ghost_hat_p;
    bool hat_done = false;
    // compact records are addressed by index * 3 NN: keep the array a multiple of this degree's record
    while (ff3_size % ((int64_t)3 * NN)) ++ff3_size;
    int64_t zero7 = -1, zero3 = -1;   // all-zero records that switch a face off (the arrays start zeroed)
    if (!split_by_deg[p].empty()) {
      zero7 = ff_size;
      ff_size += (int64_t)7 * NN;
      zero3 = ff3_size / ((int64_t)3 * NN);
      ff3_size += (int64_t)3 * NN;
      // the FACE_ONLY class runs before this degree's conforming classes (same stream)
      std::vector<int32_t> els = split_by_deg[p];
      auto mid = std::stable_partition(els.begin(), els.end(), [&](int e) { return !touches_ghost(e); });
      std::vector<unsigned char> masks(els.size());
      for (size_t i = 0; i < els.size(); ++i) masks[i] = split_mask[els[i]];
      D4bFastClass c;
      c.deg = p;
      c.kind = 3;
      c.n_hat = (int)hat_by_deg[p].size();
      c.d_hat_elem = d4b_dev_upload<int32_t>(m, hat_by_deg[p].data(), hat_by_deg[p].size());
      c.d_hat_ns = upload_ns(hat_by_deg[p], d->nodal_stride); set_hat_linear(c, hat_by_deg[p], N3);
      c.h_hat = hat_by_deg[p];
      c.n_ghost_hat = (int)ghost_hat.size();
      c.d_ghost_hat = d4b_dev_upload<int32_t>(m, ghost_hat.data(), ghost_hat.size());
      c.d_ghost_hat_ns = upload_ns(ghost_hat, d->ghost_nodal_stride);
      hat_done = true;
      c.n_elem = (int)els.size();
      c.n_interior = (int)(mid - els.begin());
      c.d_elem = d4b_dev_upload<int32_t>(m, els.data(), els.size());
      c.d_elem_ns = upload_ns(els, d->nodal_stride);
      c.h_elem = els;
      D4B_CUDA(cudaMalloc(&c.d_face_mask, masks.size()));
      D4B_CUDA(cudaMemcpy(c.d_face_mask, masks.data(), masks.size(), cudaMemcpyHostToDevice));
      m->n_hp_elem += c.n_elem;
      if (!m->d_extra) m->d_extra = d4b_dev_alloc<double>(m, (size_t)d->local_nodes);
      m->fast.push_back(c);
    }
    for (int kind = 0; kind < 2; ++kind) {
      const std::vector<int32_t> &els = els_kind[kind];
      if (els.empty()) continue;
      D4bFastClass c;
      c.deg = p;
      c.kind = kind;
      if (!hat_done) {
        c.n_hat = (int)hat_by_deg[p].size();
        c.d_hat_elem = d4b_dev_upload<int32_t>(m, hat_by_deg[p].data(), hat_by_deg[p].size());
        c.d_hat_ns = upload_ns(hat_by_deg[p], d->nodal_stride); set_hat_linear(c, hat_by_deg[p], N3);
        c.h_hat = hat_by_deg[p];
        c.n_ghost_hat = (int)ghost_hat.size();
        c.d_ghost_hat = d4b_dev_upload<int32_t>(m, ghost_hat.data(), ghost_hat.size());
        c.d_ghost_hat_ns = upload_ns(ghost_hat, d->ghost_nodal_stride);
        hat_done = true;
      }
      c.n_elem = (int)els.size();
      c.n_interior = n_interior_kind[kind];
      c.d_elem = d4b_dev_upload<int32_t>(m, els.data(), els.size());
      c.d_elem_ns = upload_ns(els, d->nodal_stride);
      c.ns0 = d->nodal_stride[els[0]];
      c.ns_linear = 1;
      for (size_t i = 0; i < els.size() && c.ns_linear; ++i)
        c.ns_linear = d->nodal_stride[els[i]] == c.ns0 + (int64_t)i * N3;
      std::vector<int64_t> face_rec((size_t)6 * els.size()), face_nbr((size_t)6 * els.size(), kNoNbr);
      std::vector<int32_t> face_code((size_t)6 * els.size(), 0);
      std::vector<unsigned char> has_extra(els.size(), 0);
      bool any_extra = false;
      for (size_t i = 0; i < els.size(); ++i) {
        const int e = els[i];
        has_extra[i] = split_mask[e] != 0;
        any_extra = any_extra || has_extra[i];
        for (int f = 0; f < 6; ++f) {
          const int s = elem_side[6 * e + f];
          const size_t fi = 6 * i + f;
          if ((split_mask[e] >> f) & 1) {   // irregular face: off here, supplied by the FACE_ONLY launch
            face_rec[fi] = kind ? zero3 : zero7;
            face_code[fi] = kOwn;
            continue;
          }
          if (d->side_faces_p[s] == 0) {
            face_rec[fi] = kind ? write_record3(s, NN, true) : write_record(s, NN, true);
            face_code[fi] = kOwn;
            continue;
          }
          const int32_t ep = d->side_e_p[4 * s];
          const int fp = d->side_f_p[s], perm = d->side_perm[s];
          if (ep < 0) {   // ghost neighbour: its uh lives behind the local block; this side owns the record
            face_nbr[fi] = d->local_nodes + d->ghost_nodal_stride[-ep - 1];
            face_rec[fi] = kind ? write_record3(s, NN, false) : write_record(s, NN, false);
            face_code[fi] = fp | (perm << 3) | kOwn;
            continue;
          }
          face_nbr[fi] = d->nodal_stride[ep];
          const int64_t mykey = (int64_t)e * 6 + f, okey = (int64_t)ep * 6 + fp;
          const int s2 = elem_side[6 * ep + fp];
          std::vector<int64_t> &recs = kind ? rec3_of : rec_of;
          bool i_own = mykey < okey;
          int64_t off;
          if (i_own) {
            if (recs[mykey] < 0) recs[mykey] = kind ? write_record3(s, NN, false) : write_record(s, NN, false);
            off = recs[mykey];
          } else if (consistent(s, s2, N, perm)) {
            if (recs[okey] < 0) recs[okey] = kind ? write_record3(s2, NN, false) : write_record(s2, NN, false);
            off = recs[okey];
            ++n_shared;
          } else {
            off = kind ? write_record3(s, NN, false) : write_record(s, NN, false);
            i_own = true;
            ++n_split;
          }
          face_rec[fi] = off;
          face_code[fi] = fp | (perm << 3) | (i_own ? kOwn : 0);
        }
      }
      c.h_elem = els;
      if (any_extra) {
        D4B_CUDA(cudaMalloc(&c.d_has_extra, has_extra.size()));
        D4B_CUDA(cudaMemcpy(c.d_has_extra, has_extra.data(), has_extra.size(), cudaMemcpyHostToDevice));
      }
      // tangential-free class whose elements are all affine (J and d(rst)/d(xyz) constant over the element, so
      // G / (w_a w_b w_c) is one number per component): six doubles per element instead of 6 N^3.  Exact up to
      // 1e-14 of the largest component (the reference computes J numerically point by point); D4B_AFFINE=0 keeps
      // the per-point metric.
      bool affine = kind == 1 && !(getenv("D4B_AFFINE") && atoi(getenv("D4B_AFFINE")) == 0);
      if (affine) {
        double *d_Gc = d4b_dev_alloc<double>(m, (size_t)6 * els.size());
        int *d_bad = nullptr, bad = 0;
        D4B_CUDA(cudaMalloc(&d_bad, sizeof(int)));
        D4B_CUDA(cudaMemsetAsync(d_bad, 0, sizeof(int), m->stream));
        affine_kernel<<<(int)els.size(), 64, 0, m->stream>>>((int)els.size(), N, LQ, c.d_elem, m->d_quad_stride, m->d_G,
                                                             d_Gc, d_bad);
        D4B_CUDA(cudaGetLastError());
        D4B_CUDA(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, m->stream));
        D4B_CUDA(cudaStreamSynchronize(m->stream));
        cudaFree(d_bad);
        affine = bad == 0;
        if (affine) c.d_FGc = d_Gc;
        else { cudaFree(d_Gc); m->device_bytes -= (int64_t)(6 * els.size() * sizeof(double)); }
      }
      if (!affine) {
        c.d_FG = d4b_dev_alloc<double>(m, (size_t)6 * N3 * els.size());
        gather_fg_kernel<<<(int)els.size(), 128, 0, m->stream>>>((int)els.size(), N3, LQ, c.d_elem, m->d_quad_stride,
                                                                 m->d_G, c.d_FG);
        D4B_CUDA(cudaGetLastError());
      }
      if (kind == 0) {
        c.d_face_rec = d4b_dev_upload<int64_t>(m, face_rec.data(), face_rec.size());
        c.d_face_nbr = d4b_dev_upload<int64_t>(m, face_nbr.data(), face_nbr.size());
        c.d_face_code = d4b_dev_upload<int32_t>(m, face_code.data(), face_code.size());
        if (N <= 5) {   // packed descriptors of the shuffle kernel (d4b_curved_kernel.cuh): rec = offset in doubles
          std::vector<FaceMeta> meta(face_rec.size());
          for (size_t j = 0; j < meta.size(); ++j) {
            D4B_CHECK(face_rec[j] < INT32_MAX, "curved path: face records exceed 2^31 doubles");
            meta[j].nbr = face_nbr[j];
            meta[j].rec = (int32_t)face_rec[j];
            meta[j].code = face_code[j];
          }
          c.d_face_meta = d4b_dev_upload<int64_t>(m, reinterpret_cast<const int64_t *>(meta.data()), meta.size() * 2);
        }
      } else {
        std::vector<FaceMeta> meta(face_rec.size());
        for (size_t j = 0; j < meta.size(); ++j) {
          D4B_CHECK(face_rec[j] < INT32_MAX, "fast path: too many face records");
          meta[j].nbr = face_nbr[j];
          meta[j].rec = (int32_t)face_rec[j];
          meta[j].code = face_code[j];
        }
        static_assert(sizeof(FaceMeta) == 16, "FaceMeta layout");
        c.d_face_meta = d4b_dev_upload<int64_t>(m, reinterpret_cast<const int64_t *>(meta.data()), meta.size() * 2);
        m->n_flat_elem += c.n_elem;
      }
      m->fast.push_back(c);
    }
    if (!hp_by_deg[p].empty() || !hat_done) {   // the mortar-kernel class (possibly empty: it then only carries uh)
      std::vector<int32_t> els = hp_by_deg[p];
      auto side_touches_ghost = [&](int e) {
        for (int f = 0; f < 6; ++f) {
          const int s = elem_side[6 * e + f];
          for (int i = 0; i < d->side_faces_m[s]; ++i) if (d->side_e_m[4 * s + i] < 0) return true;
          for (int i = 0; i < d->side_faces_p[s]; ++i) if (d->side_e_p[4 * s + i] < 0) return true;
        }
        return false;
      };
      auto mid = std::stable_partition(els.begin(), els.end(), [&](int e) { return !side_touches_ghost(e); });
      D4bFastClass c;
      c.deg = p;
      c.kind = 2;
      if (!hat_done) {
        c.n_hat = (int)hat_by_deg[p].size();
        c.d_hat_elem = d4b_dev_upload<int32_t>(m, hat_by_deg[p].data(), hat_by_deg[p].size());
        c.d_hat_ns = upload_ns(hat_by_deg[p], d->nodal_stride); set_hat_linear(c, hat_by_deg[p], N3);
        c.h_hat = hat_by_deg[p];
        c.n_ghost_hat = (int)ghost_hat.size();
        c.d_ghost_hat = d4b_dev_upload<int32_t>(m, ghost_hat.data(), ghost_hat.size());
        c.d_ghost_hat_ns = upload_ns(ghost_hat, d->ghost_nodal_stride);
        hat_done = true;
      }
      c.n_elem = (int)els.size();
      c.n_interior = (int)(mid - els.begin());
      c.d_elem = d4b_dev_upload<int32_t>(m, els.data(), els.size());
      c.d_elem_ns = upload_ns(els, d->nodal_stride);
      c.h_elem = els;
      m->n_hp_elem += c.n_elem;
      m->fast.push_back(c);
    }
  }
  m->n_face_shared = n_shared;
  m->n_face_split = n_split;
  auto fill = [&](int64_t size, const std::vector<RecSrc> &src, int compact) -> double * {
    if (size == 0) return nullptr;
    double *out = d4b_dev_alloc<double>(m, (size_t)size);
    D4B_CUDA(cudaMemsetAsync(out, 0, (size_t)size * sizeof(double), m->stream));
    if (!src.empty()) {
      RecSrc *d_src = nullptr;
      D4B_CUDA(cudaMalloc(&d_src, src.size() * sizeof(RecSrc)));
      D4B_CUDA(cudaMemcpyAsync(d_src, src.data(), src.size() * sizeof(RecSrc), cudaMemcpyHostToDevice, m->stream));
      gather_records_kernel<<<(int)src.size(), 64, 0, m->stream>>>((int)src.size(), d_src, compact, MP, m->d_mortar,
                                                                   m->d_side_mortar_stride, out);
      D4B_CUDA(cudaGetLastError());
      D4B_CUDA(cudaStreamSynchronize(m->stream));
      cudaFree(d_src);
    }
    return out;
  };
  m->d_fast_FF = fill(ff_size, src7, 0);
  m->d_flat_FF = fill(ff3_size, src3, 1);
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  m->d_uhat = d4b_dev_alloc<double>(m, (size_t)(d->local_nodes + d->ghost_nodes));
  // every element in the tangential-free kernel: the solvers may use its energy dot and Gauss-resident smoother
  int64_t n_flat = 0;
  for (const D4bFastClass &c : m->fast) if (c.kind == 1) n_flat += c.n_elem;
  m->all_flat = m->n_generic == 0 && n_flat == d->n_local && d->n_local > 0;
  for (const D4bFastClass &c : m->fast) m->all_flat = m->all_flat && (c.kind == 1 || c.n_elem == 0);
}

// every class runs in a kernel with the in-kernel halo wait (tangential-free or the curved shuffle kernel)
bool d4b_fast_can_merge_halo(const d4b_mesh *m) {
  static const bool off = getenv("D4B_PEER_MERGED") && atoi(getenv("D4B_PEER_MERGED")) == 0;
  if (off || m->n_generic > 0 || m->fast.empty()) return false;
  for (const D4bFastClass &c : m->fast) {
    if (c.n_elem == 0) continue;
    if (c.kind == 1) continue;
    if (c.kind == 0 && use_curved_shfl(c.deg + 1) && c.d_face_meta) continue;
    return false;
  }
  return true;
}

void d4b_launch_basis_change(d4b_mesh *m, const double *in, double *out, bool to_gauss) {
  for (const D4bFastClass &c : m->fast) {
#define D4B_CASE(n) case n: launch_basis_change<n>(m, c, in, out, to_gauss); break;
    switch (c.deg + 1) {
      D4B_CASE(2) D4B_CASE(3) D4B_CASE(4) D4B_CASE(5) D4B_CASE(6) D4B_CASE(7) D4B_CASE(8)
      default: D4B_ABORT("fast path: unsupported degree %d", c.deg);
    }
#undef D4B_CASE
  }
}

namespace {
// neighbour offsets >= local_nodes address a ghost's lines behind the local block of uhat: replace them by the
// offset of the same lines in the owner's uhat (peer memory), relative to this rank's uhat
__global__ void patch_ghost_nbr_k(int64_t *__restrict__ nbr, int step, int64_t n, int64_t local_nodes,
                                  const int64_t *__restrict__ ghost_stride, int n_ghost,
                                  const int64_t *__restrict__ ghost_off) {
  for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    const int64_t v = nbr[j * step];
    if (v == kNoNbr || v < local_nodes) continue;
    const int64_t key = v - local_nodes;
    int lo = 0, hi = n_ghost - 1;   // the last ghost whose first node is <= key
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (ghost_stride[mid] <= key) lo = mid; else hi = mid - 1;
    }
    nbr[j * step] = ghost_off[lo] + (key - ghost_stride[lo]);
  }
}
}  // namespace

void d4b_fast_patch_ghost_offsets(d4b_mesh *m, const int64_t *d_ghost_off) {
  if (m->n_ghost == 0) return;
  for (const D4bFastClass &c : m->fast) {
    const int64_t n = (int64_t)6 * c.n_elem;
    if (n == 0) continue;
    const int grid = (int)std::min<int64_t>((n + 255) / 256, 1184);
    if (c.d_face_nbr)
      patch_ghost_nbr_k<<<grid, 256, 0, m->stream>>>(c.d_face_nbr, 1, n, m->local_nodes, m->d_ghost_nodal_stride,
                                                     m->n_ghost, d_ghost_off);
    if (c.d_face_meta)
      patch_ghost_nbr_k<<<grid, 256, 0, m->stream>>>(static_cast<int64_t *>(c.d_face_meta), 2, n, m->local_nodes,
                                                     m->d_ghost_nodal_stride, m->n_ghost, d_ghost_off);
    D4B_CUDA(cudaGetLastError());
  }
}

void d4b_launch_fast_range(d4b_mesh *m, const double *d_u, double *d_Au, int what, int e0, int e1) {
  for (const D4bFastClass &c : m->fast) {
    const std::vector<int32_t> &list = what == 0 ? c.h_hat : c.h_elem;
    const int lo = (int)(std::lower_bound(list.begin(), list.end(), e0) - list.begin());
    const int hi = (int)(std::lower_bound(list.begin(), list.end(), e1) - list.begin());
    if (hi <= lo) continue;
#define D4B_CASE(n) case n: \
      if (what == 0) launch_interp_range<n>(m, c, d_u, lo, hi); \
      else if (c.kind >= 2) launch_hp<n>(m, c, d_Au, lo, hi, c.kind == 3); \
      else if (c.kind) launch_flat<n>(m, c, d_Au, lo, hi); \
      else if (use_curved_shfl(n) && c.d_face_meta) launch_curved<n>(m, c, d_Au, lo, hi); \
      else launch_fast<n>(m, c, d_Au, lo, hi); \
      break;
    switch (c.deg + 1) {
      D4B_CASE(2) D4B_CASE(3) D4B_CASE(4) D4B_CASE(5) D4B_CASE(6) D4B_CASE(7) D4B_CASE(8)
      default: D4B_ABORT("fast path: unsupported degree %d", c.deg);
    }
#undef D4B_CASE
  }
}

// part 0: uh of the local elements; part 1: elements without a ghost neighbour;
// part 2 (after the halo arrived): uh of the ghost elements + elements with a ghost neighbour.
// Launches go to m->stream (d4b_apply_device points it at the high-priority halo stream for part 2).
void d4b_launch_fast(d4b_mesh *m, const double *d_u, double *d_Au, int part) {
  if (part == 0 || (part == 2 && m->peer.state != 1)) {   // (peer halo: the ghosts' uh is read where its owner wrote it)
    const int phase = part == 0 ? 0 : 1;
    for (const D4bFastClass &c : m->fast) {
#define D4B_CASE(n) case n: launch_interp<n>(m, c, d_u, phase); break;
      switch (c.deg + 1) {
        D4B_CASE(2) D4B_CASE(3) D4B_CASE(4) D4B_CASE(5) D4B_CASE(6) D4B_CASE(7) D4B_CASE(8)
        default: D4B_ABORT("fast path: unsupported degree %d", c.deg);
      }
#undef D4B_CASE
    }
  }
  if (part == 0) return;
  if (part == 3) {   // one launch per class; the groups from n_interior on wait for the neighbour ranks by themselves
    for (const D4bFastClass &c : m->fast) {
      D4bHaloWait h;
      h.flags = m->peer.flags; h.ranks = m->peer.d_wait; h.n = c.n_elem > c.n_interior ? m->peer.n_nbr : 0;
      h.from = c.n_interior; h.epoch = m->peer.d_epoch;
#define D4B_CASE(n) case n: if (c.kind == 1) launch_flat<n>(m, c, d_Au, 0, c.n_elem, &h); else launch_curved<n>(m, c, d_Au, 0, c.n_elem, &h); break;
      switch (c.deg + 1) {
        D4B_CASE(2) D4B_CASE(3) D4B_CASE(4) D4B_CASE(5) D4B_CASE(6) D4B_CASE(7) D4B_CASE(8)
        default: D4B_ABORT("fast path: unsupported degree %d", c.deg);
      }
#undef D4B_CASE
    }
    return;
  }
  for (const D4bFastClass &c : m->fast) {
    const int first = part == 1 ? 0 : c.n_interior, last = part == 1 ? c.n_interior : c.n_elem;
#define D4B_CASE(n) case n: if (c.kind >= 2) launch_hp<n>(m, c, d_Au, first, last, c.kind == 3); else if (c.kind) launch_flat<n>(m, c, d_Au, first, last); else if (use_curved_shfl(n) && c.d_face_meta) launch_curved<n>(m, c, d_Au, first, last); else launch_fast<n>(m, c, d_Au, first, last); break;
    switch (c.deg + 1) {
      D4B_CASE(2) D4B_CASE(3) D4B_CASE(4) D4B_CASE(5) D4B_CASE(6) D4B_CASE(7) D4B_CASE(8)
      default: D4B_ABORT("fast path: unsupported degree %d", c.deg);
    }
#undef D4B_CASE
  }
}
