// d4b_curved_kernel.cuh -- the apply for curved conforming elements (all seven planes of the folded
// face metric are live: A_m[3], A_p[3], sj*sigma) in the same line-fused form as d4b_flat_kernel.cuh.
//
// What a curved face adds to the tangential-free case (d4est_laplacian_flux_sipg.c:494-814):
//   * the trial side needs the two TANGENTIAL reference derivatives of both traces at the face Gauss
//     points:  F_m = A_m[n] dn u_m + A_m[t1] d1 u_m + A_m[t2] d2 u_m  and the same for the neighbour in
//     its own frame;
//   * the test side needs the transposed tangential derivatives of  A_m[t] w c (u_m - u_p).
// Both are 2-D operations ON THE FACE, whose N^2 Gauss points are exactly the N^2 threads of the element
// in the orientation whose lines are normal to that face: thread (x, y) holds the value at face point
// (x, y), its row and column mates are N lanes of the same warp.  All six 1-D contractions per face are
// therefore warp shuffles (an element's N^2 threads never straddle a warp: N <= 5), the face work fuses
// into the three line orientations of the volume sweep exactly as in the tangential-free kernel, and no
// face array ever goes through shared memory (the general conforming kernel stages four face arrays of
// both sides per face pair through shared memory with three barriers per pair).
//   * the neighbour's tangential derivatives are taken in MY face frame from its trace at my Gauss
//     points and mapped to its frame by the face permutation (swap / negate: the Gauss points are
//     symmetric, Dg[N-1-i][N-1-j] = -Dg[i][j]).
// Included by d4b_fast.cu after d4b_flat_kernel.cuh (shares FaceMeta, block_sync, the tables).

struct CurvedArgs {
  const int32_t *elem;
  int first, n_elem;
  const int64_t *elem_ns;       // [n_elem] nodal offset of each class position
  int64_t ns0;                  // ns_linear: elem_ns[i] = ns0 + i N^3, no load needed (see FlatArgs)
  int ns_linear;
  const double *FG;             // [n_elem][6][N^3]
  const double *FF;             // 7-plane face records [7][NN]; meta.rec = offset in doubles
  const FaceMeta *meta;         // [6*n_elem]
  const double *uh;
  double *Au;
  D4bEpilogue epi;
  int prefetch;
  const double *W;
  const double *nlB;            // residual: the functor's b (mass_point, d4b_fast.cu)
  int nlK;
  const int64_t *quad_stride;
  const unsigned char *has_extra;
  const double *extra;
  D4bHaloWait halo;
};

template <int N>
struct CurvedShflCfg {
  static constexpr int NN = N * N, N3 = N * N * N;
  static constexpr int THREADS = 32;                 // one warp per block: __syncwarp only
  static constexpr int EPB = 32 / NN;                // elements per warp (N = 2: 8, 3: 3, 4: 2, 5: 1)
#ifndef D4B_CURVED_MINB
#define D4B_CURVED_MINB {0, 0, 16, 16, 16, 8, 0, 0, 0}
#endif
  static constexpr int kMinBlocks[9] = D4B_CURVED_MINB;
  static constexpr int MIN_BLOCKS = kMinBlocks[N];
  static constexpr int VS = VSize<N>::value;
  static constexpr int SH_PER_ELEM = 3 * VS + 12 * NN + 12 + VLay<N>::PAD;
};

// sum_i D[x][i] v(i, y) and sum_i D[y][i] v(x, i) over the element's face points (lanes base + i + N j)
template <int N>
__device__ __forceinline__ void face_grad(const FastTab &T, double v, int base, int x, int y, double &dx, double &dy) {
  dx = 0.; dy = 0.;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const double vr = __shfl_sync(0xffffffffu, v, base + i + N * y);
    const double vc = __shfl_sync(0xffffffffu, v, base + x + N * i);
    dx += T.Dg[x][i] * vr;
    dy += T.Dg[y][i] * vc;
  }
}
// transposed: sum_i D[i][x] a(i, y) + sum_i D[i][y] b(x, i)
template <int N>
__device__ __forceinline__ double face_grad_t(const FastTab &T, double a, double b, int base, int x, int y) {
  double s = 0.;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const double ar = __shfl_sync(0xffffffffu, a, base + i + N * y);
    const double bc = __shfl_sync(0xffffffffu, b, base + x + N * i);
    s += T.Dg[i][x] * ar + T.Dg[i][y] * bc;
  }
  return s;
}

// What one face needs from memory at this thread's face Gauss point (x, y): its record (own view) and the neighbour's
// trace / normal derivative.  Split from the arithmetic so that the loads of BOTH faces of a pair are in flight
// before the first shuffle: the shuffles are convergence points, and with load and use in one function the six
// faces' L2 latencies were paid one after the other (a quarter of the kernel's stall samples, profiles/
// r02_ncu_curved_p3_sphere_stalls.txt).
struct CurvedFaceData {
  double Am[3], Ap[3], S, trp, dnp;
};
template <int N>
__device__ __forceinline__ void curved_face_load(const CurvedArgs &a, const FastTab &T, const FaceMeta &fm, bool active,
                                                 int64_t ns, int x, int y, CurvedFaceData &d) {
  constexpr int NN = N * N;
  const int code = fm.code;
  const bool own = code & kOwn, interior = fm.nbr != kNoNbr;
  const bool swp = code & (4 << 3), frow = code & (2 << 3), fcol = code & (1 << 3);
  int row = swp ? x : y, col = swp ? y : x;
  row = frow ? N - 1 - row : row;
  col = fcol ? N - 1 - col : col;
#pragma unroll
  for (int l = 0; l < 3; ++l) { d.Am[l] = 0.; d.Ap[l] = 0.; }
  d.S = 0.; d.trp = 0.; d.dnp = 0.;
  const int ndp = (code >> 1) & 3, sp = code & 1;
  if (active) {
    const int k = own ? (x + N * y) : (col + N * row);
    const double *r = a.FF + (size_t)fm.rec + k;
    double rp[3], rq[3];
#pragma unroll
    for (int l = 0; l < 3; ++l) { rp[l] = r[l * NN]; rq[l] = r[(3 + l) * NN]; }
    d.S = r[6 * NN];
    // neighbour line normal to its face through its point (col, row), read from the face inwards
    const int stride = (ndp == 0) ? 1 : ((ndp == 1) ? N : NN);
    const int b0 = (ndp == 0) ? (N * col + NN * row) : ((ndp == 1) ? (col + NN * row) : (col + N * row));
    const double *p = a.uh + (interior ? fm.nbr : ns) + b0 + (sp ? (N - 1) * stride : 0);
    const int step = sp ? -stride : stride;
    double v[N];
#pragma unroll
    for (int l = 0; l < N; ++l) v[l] = p[l * step];
#pragma unroll
    for (int l = 0; l < 3; ++l) {
      d.Am[l] = own ? rp[l] : -rq[l];
      d.Ap[l] = own ? rq[l] : -rp[l];
    }
#pragma unroll
    for (int l = 0; l < N; ++l) { d.trp += T.e[0][l] * v[l]; d.dnp += T.d[0][l] * v[l]; }
    d.dnp = sp ? -d.dnp : d.dnp;
  }
}

// The arithmetic of one face; ND = my normal direction (compile time).  Every lane of the warp must call it
// (shuffles); lanes without an element carry zeros.
// OWN_T: the caller supplies the own tangential derivatives (dmx, dmy in) and takes the two tangential test
// weights back (w1, w2 out) instead of their transposed face derivative being folded into ve.
template <int N, int ND, bool OWN_T = false>
__device__ __forceinline__ void curved_face(const FastTab &T, const FaceMeta &fm, const CurvedFaceData &d, int base, int x,
                                            int y, double trm, double dnm, double wxy, double &ve, double &vd,
                                            double dmx = 0., double dmy = 0., double *w1 = nullptr, double *w2 = nullptr) {
  constexpr int T1 = (ND == 0) ? 1 : 0, T2 = (ND == 2) ? 1 : 2;   // my tangential directions, x <-> T1, y <-> T2
  const int code = fm.code;
  const bool interior = fm.nbr != kNoNbr;
  const bool swp = code & (4 << 3), frow = code & (2 << 3), fcol = code & (1 << 3);
  const int ndp = (code >> 1) & 3;
  const double *Am = d.Am, *Ap = d.Ap;
  const double S = d.S, trp = d.trp, dnp = d.dnp;
  // tangential derivatives on the face, both traces, in my frame
  double dpx, dpy;
  if (!OWN_T) face_grad<N>(T, trm, base, x, y, dmx, dmy);
  face_grad<N>(T, trp, base, x, y, dpx, dpy);
  const double Fm = Am[ND] * dnm + Am[T1] * dmx + Am[T2] * dmy;
  // neighbour frame: its face point (col, row) <-> its tangential directions (t1', t2') in increasing order
  const double dp1 = (swp ? dpy : dpx) * (fcol ? -1. : 1.), dp2 = (swp ? dpx : dpy) * (frow ? -1. : 1.);
  const double Apn = (ndp == 0) ? Ap[0] : (ndp == 1) ? Ap[1] : Ap[2];
  const double Apt1 = (ndp == 0) ? Ap[1] : Ap[0], Apt2 = (ndp == 2) ? Ap[1] : Ap[2];
  const double Fp = Apn * dnp + Apt1 * dp1 + Apt2 * dp2;
  // interior: t13 = -1/2 (F_m + F_p) + S du, t2 = -1/2 A_m du; boundary: t13 = -F_m + S (u_m - g), t2 = -A_m (u_m - g),
  // Dirichlet / Robin data g in the A_p[0] slot
  const double du = trm - (interior ? trp : Ap[0]);
  const double t13 = interior ? (-.5 * (Fp + Fm) + S * du) : (-Fm + S * du);
  const double wdu = wxy * (interior ? -.5 : -1.) * du;
  vd = Am[ND] * wdu;
  if (OWN_T) { ve = wxy * t13; *w1 = Am[T1] * wdu; *w2 = Am[T2] * wdu; }
  else ve = wxy * t13 + face_grad_t<N>(T, Am[T1] * wdu, Am[T2] * wdu, base, x, y);
}

template <int N, bool MASS>
__global__ void __launch_bounds__(CurvedShflCfg<N>::THREADS, CurvedShflCfg<N>::MIN_BLOCKS) curved_apply_kernel(CurvedArgs a) {
  using C = CurvedShflCfg<N>;
  constexpr int NN = C::NN, N3 = C::N3;
  extern __shared__ double smem[];
  const FastTab &T = c_tab[N];
  const int slot = threadIdx.x / NN;
  const int lt = threadIdx.x % NN;
  const int lx = lt % N, ly = lt / N;
  const int base = (slot < C::EPB ? slot : 0) * NN;          // first lane of my element
  const int blk = blockIdx.x;
  const int ei = a.first + blk * C::EPB + slot;
  const bool active = (slot < C::EPB) && (ei < a.n_elem);
  double *V0 = smem + (size_t)(slot < C::EPB ? slot : 0) * C::SH_PER_ELEM;
  double *V1 = V0 + C::VS, *V2 = V1 + C::VS;
  double *FB = V2 + C::VS;                                // [6][2][NN], thread-private columns
  FaceMeta *META = reinterpret_cast<FaceMeta *>(FB + 12 * NN);
  const int64_t ns = !active ? 0 : a.ns_linear ? a.ns0 + (int64_t)ei * N3 : a.elem_ns[ei];
  const double wxy = T.w[lx] * T.w[ly];
  // both faces of a pair loaded before the first is used: +3 % at N = 2, 3, 5; at N = 4 the 18 extra live registers
  // spill under the 128-register cap of 16 blocks per SM and fewer blocks cost more than the latency
  // (profiles/r02_curved_hoist_ab.txt)
  constexpr bool kHoist = N != 4;
  if (threadIdx.x == 0 && a.prefetch) {
    constexpr unsigned kBlockBytes = C::EPB * 6 * N3 * sizeof(double);
    const int e0 = a.first + blk * C::EPB, left = a.n_elem - e0;
    l2_prefetch(a.FG + (size_t)e0 * (6 * N3), left >= C::EPB ? kBlockBytes : left * 6 * N3 * 8);
  }
  // ---- T-orientation (a,b) = (lx,ly), line along c: only stage uh; its c-derivative and the traces of faces 4/5
  // are taken in the metric phase from the staged copy (keeps them out of the registers during the S / R faces)
  if (active) {
    for (int j = lt; j < 6; j += NN) {
      const FaceMeta fm = a.meta[6 * (size_t)ei + j];
      META[j] = fm;
      if (a.prefetch) {   // the record's seven planes are first touched after the forward sweeps
        constexpr int L = (7 * NN * 8 + 127) / 128;
        for (int q = 0; q < L; ++q) l2_prefetch_line(a.FF + fm.rec + q * 16);
      }
    }
#pragma unroll
    for (int c = 0; c < N; ++c) V0[vix<N>(lx, ly, c)] = a.uh[ns + lt + NN * c];
  }
  if (a.halo.n > 0 && a.first + (blk + 1) * C::EPB > a.halo.from) halo_wait(a.halo);   // see d4b_flat_kernel.cuh
  __syncwarp();
  const FaceMeta zero_meta = {kNoNbr, 0, kOwn};
  {   // S-orientation: (a,c) = (lx,ly), line along b: gradient, faces 2/3
    double t2 = 0., n2 = 0., t3 = 0., n3 = 0.;
    if (active) {
      double uh[N];
#pragma unroll
      for (int i = 0; i < N; ++i) uh[i] = V0[vix<N>(lx, i, ly)];
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
    }
    double ve, vd;
    const FaceMeta fa = active ? META[2] : zero_meta, fb = active ? META[3] : zero_meta;
    CurvedFaceData da, db;
    curved_face_load<N>(a, T, fa, active, ns, lx, ly, da);
    if (kHoist) curved_face_load<N>(a, T, fb, active, ns, lx, ly, db);
    curved_face<N, 1>(T, fa, da, base, lx, ly, t2, n2, wxy, ve, vd);
    if (active) { FB[(2 * 2 + 0) * NN + lt] = ve; FB[(2 * 2 + 1) * NN + lt] = vd; }
    if (!kHoist) curved_face_load<N>(a, T, fb, active, ns, lx, ly, db);
    curved_face<N, 1>(T, fb, db, base, lx, ly, t3, n3, wxy, ve, vd);
    if (active) { FB[(3 * 2 + 0) * NN + lt] = ve; FB[(3 * 2 + 1) * NN + lt] = vd; }
  }
  {   // R-orientation: (b,c) = (lx,ly), line along a: gradient, faces 0/1
    double t0 = 0., n0 = 0., t1 = 0., n1 = 0.;
    if (active) {
      double uh[N];
#pragma unroll
      for (int i = 0; i < N; ++i) uh[i] = V0[vix<N>(i, lx, ly)];
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
    }
    double ve, vd;
    const FaceMeta fa = active ? META[0] : zero_meta, fb = active ? META[1] : zero_meta;
    CurvedFaceData da, db;
    curved_face_load<N>(a, T, fa, active, ns, lx, ly, da);
    if (kHoist) curved_face_load<N>(a, T, fb, active, ns, lx, ly, db);
    curved_face<N, 0>(T, fa, da, base, lx, ly, t0, n0, wxy, ve, vd);
    if (active) { FB[(0 * 2 + 0) * NN + lt] = ve; FB[(0 * 2 + 1) * NN + lt] = vd; }
    if (!kHoist) curved_face_load<N>(a, T, fb, active, ns, lx, ly, db);
    curved_face<N, 0>(T, fb, db, base, lx, ly, t1, n1, wxy, ve, vd);
    if (active) { FB[(1 * 2 + 0) * NN + lt] = ve; FB[(1 * 2 + 1) * NN + lt] = vd; }
  }
  __syncwarp();
  // ---- T-orientation: faces 4/5, metric, Dg^T along c; V0 becomes the residual.  The thread of face point
  // (a,b) holds the whole c-line of all three gradient components here, so the own tangential derivatives
  // at faces 4/5 are traces of d_r u, d_s u along that line and the tangential test weights go straight
  // into the flux lines F_r, F_s: no shuffles for the own side of this face pair.
  {
    double gr[N], gs[N], gt[N];
    double d14 = 0., d24 = 0., d15 = 0., d25 = 0., tr4 = 0., dn4 = 0., tr5 = 0., dn5 = 0.;
#pragma unroll
    for (int c = 0; c < N; ++c) { gr[c] = 0.; gs[c] = 0.; gt[c] = 0.; }
    if (active) {
      double uh[N];
#pragma unroll
      for (int c = 0; c < N; ++c) {
        const int idx = vix<N>(lx, ly, c);
        uh[c] = V0[idx]; gr[c] = V1[idx]; gs[c] = V2[idx];
        d14 += T.e[0][c] * gr[c]; d24 += T.e[0][c] * gs[c];
        d15 += T.e[1][c] * gr[c]; d25 += T.e[1][c] * gs[c];
        tr4 += T.e[0][c] * uh[c]; dn4 += T.d[0][c] * uh[c];
        tr5 += T.e[1][c] * uh[c]; dn5 += T.d[1][c] * uh[c];
      }
#pragma unroll
      for (int c = 0; c < N; ++c) {
        double sd = 0.;
#pragma unroll
        for (int i = 0; i < N; ++i) sd += T.Dg[c][i] * uh[i];
        gt[c] = sd;
      }
    }
    double ve0, vd0, ve1, vd1, w14, w24, w15, w25;
    const FaceMeta fa = active ? META[4] : zero_meta, fb = active ? META[5] : zero_meta;
    CurvedFaceData da, db;
    curved_face_load<N>(a, T, fa, active, ns, lx, ly, da);
    if (kHoist) curved_face_load<N>(a, T, fb, active, ns, lx, ly, db);
    curved_face<N, 2, true>(T, fa, da, base, lx, ly, tr4, dn4, wxy, ve0, vd0, d14, d24, &w14, &w24);
    if (!kHoist) curved_face_load<N>(a, T, fb, active, ns, lx, ly, db);
    curved_face<N, 2, true>(T, fb, db, base, lx, ly, tr5, dn5, wxy, ve1, vd1, d15, d25, &w15, &w25);
    if (active) {
      double ft[N];
      const double *G = a.FG + (size_t)ei * (6 * N3) + lt;
#pragma unroll
      for (int c = 0; c < N; ++c) {
        const int idx = vix<N>(lx, ly, c);
        double G00 = G[0 * N3 + NN * c], G01 = G[1 * N3 + NN * c], G02 = G[2 * N3 + NN * c];
        double G11 = G[3 * N3 + NN * c], G12 = G[4 * N3 + NN * c], G22 = G[5 * N3 + NN * c];
        V1[idx] = G00 * gr[c] + G01 * gs[c] + G02 * gt[c] + T.e[0][c] * w14 + T.e[1][c] * w15;
        V2[idx] = G01 * gr[c] + G11 * gs[c] + G12 * gt[c] + T.e[0][c] * w24 + T.e[1][c] * w25;
        ft[c] = G02 * gr[c] + G12 * gs[c] + G22 * gt[c];
      }
      const double *Wd = MASS ? a.W + a.quad_stride[a.elem[ei]] + lt : nullptr;
      const double *Bd = (MASS && a.nlB) ? a.nlB + a.quad_stride[a.elem[ei]] + lt : nullptr;
      const double *xtr = (a.has_extra && a.has_extra[ei]) ? a.extra + ns + lt : nullptr;
#pragma unroll
      for (int c = 0; c < N; ++c) {
        double s = T.e[0][c] * ve0 + T.d[0][c] * vd0 + T.e[1][c] * ve1 + T.d[1][c] * vd1;
        if (MASS) s += mass_point(Wd, Bd, a.nlK, NN * c, V0[vix<N>(lx, ly, c)]);
        if (xtr) s += xtr[NN * c];
#pragma unroll
        for (int i = 0; i < N; ++i) s += T.Dg[i][c] * ft[i];
        V0[vix<N>(lx, ly, c)] = s;
      }
    }
  }
  __syncwarp();
  if (active) {   // S-orientation: + Dg^T F_s + faces 2/3
    double fs[N];
#pragma unroll
    for (int i = 0; i < N; ++i) fs[i] = V2[vix<N>(lx, i, ly)];
    const double ve0 = FB[(2 * 2 + 0) * NN + lt], vd0 = FB[(2 * 2 + 1) * NN + lt];
    const double ve1 = FB[(3 * 2 + 0) * NN + lt], vd1 = FB[(3 * 2 + 1) * NN + lt];
#pragma unroll
    for (int b = 0; b < N; ++b) {
      double s = T.e[0][b] * ve0 + T.d[0][b] * vd0 + T.e[1][b] * ve1 + T.d[1][b] * vd1;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i][b] * fs[i];
      V0[vix<N>(lx, b, ly)] += s;
    }
  }
  __syncwarp();
  if (active) {   // R-orientation: + Dg^T F_r + faces 0/1, then I^T along a
    double fr[N], tot[N];
#pragma unroll
    for (int i = 0; i < N; ++i) fr[i] = V1[vix<N>(i, lx, ly)];
    const double ve0 = FB[(0 * 2 + 0) * NN + lt], vd0 = FB[(0 * 2 + 1) * NN + lt];
    const double ve1 = FB[(1 * 2 + 0) * NN + lt], vd1 = FB[(1 * 2 + 1) * NN + lt];
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = T.e[0][i2] * ve0 + T.d[0][i2] * vd0 + T.e[1][i2] * ve1 + T.d[1][i2] * vd1;
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
  __syncwarp();
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
  __syncwarp();
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
