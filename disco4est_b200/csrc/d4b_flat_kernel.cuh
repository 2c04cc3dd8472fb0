// d4b_flat_kernel.cuh -- the apply for elements whose six faces are
// "tangential-free": on every face Gauss point the folded face metric
// A[l] = sj sum_d n_d drst[l][d] vanishes for the two directions l tangential
// to the face, on both sides (exactly, tested at set-up).  Axis-aligned bricks
// and any mesh whose grid lines meet the faces orthogonally are of this kind.
// Then the SIPG face terms (d4est_laplacian_flux_sipg.c:494-814) need only
// traces and NORMAL derivatives, both of which are 1-D contractions of the
// line normal to the face, which the thread owning face Gauss point (x,y)
// already holds in registers (own side) or reads from the neighbour's uh
// (other side).  The face work therefore fuses into the three line
// orientations of the volume sweep and needs no shared-memory exchange:
//
//   forward  T/S/R: line -> gradient component, traces of the two faces normal
//                   to the line, neighbour traces, fluxes (kept per thread)
//   pointwise     : stiffness metric G
//   backward T/S/R: Dg^T of the flux line + rank-1 lifting of both faces
//   I^T x I^T x I^T back to the LGL basis, one store of Au.
//
// Included by d4b_fast.cu (shares its constant tables and helpers).

struct FaceMeta {     // 16 bytes per element face
  int64_t nbr;        // offset of the neighbour's uh block, kNoNbr on the boundary
  int32_t rec;        // record index into FF3 (records of 3 planes [NN])
  int32_t code;       // f_p | perm << 3 | own << 6
};

// the first lanes of the block poll the neighbour ranks' flags (see D4bHaloWait); the caller synchronises the block
__device__ __forceinline__ void halo_wait(const D4bHaloWait &h) {
  for (int i = threadIdx.x; i < h.n; i += blockDim.x) {
    const unsigned long long *p = h.flags + h.ranks[i];
    const unsigned long long epoch = *h.epoch;
    unsigned long long v, t0 = 0;
    for (int spin = 0;; ++spin) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
      if (v >= epoch) break;
      unsigned long long now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (spin == 0) t0 = now;
      if (now - t0 > 30000000000ull) __trap();   // 30 s: a neighbour never produced this apply (no printf: it would
                                                 // put a stack frame into every instantiation of the apply kernels)
      __nanosleep(64);
    }
  }
}

struct FlatArgs {
  const int32_t *elem;
  int first;                    // this launch covers class positions [first, n_elem)
  int n_elem;
  const int64_t *elem_ns;       // [n_elem] nodal offset of each class position
  int64_t ns0;                  // ns_linear: elem_ns[i] = ns0 + i N^3 (the class is a run of consecutive elements):
  int ns_linear;                //   the kernel's first dependent load is not needed at all
  const double *FG;             // [n_elem][6][N^3]
  const double *FGc;            // CG kernels: [n_elem][6], metric of an affine element without the weights
  const double *FF3;            // compact face records: A_m[n], A_p[n] (g on the boundary), sj*sigma
  const FaceMeta *meta;         // [6*n_elem]
  const double *uh;
  double *Au;
  D4bEpilogue epi;
  int prefetch;                 // 1: bulk-prefetch the block's volume metric into L2 at kernel start
  const double *W;              // MASS kernels: weighted diagonal at the quadrature points (d4b_mesh_set_mass_diag)
  const double *nlB;            // MASS kernels, residual: the functor's b at the quadrature points (mass_point, d4b_fast.cu)
  int nlK;
  const int64_t *quad_stride;
  // face split: elements with has_extra[class position] add extra[nodal_stride + k] (Gauss basis, the
  // contributions of their irregular faces from hp_apply_kernel<.., FACE_ONLY>) to the accumulator
  const unsigned char *has_extra;
  const double *extra;
  D4bHaloWait halo;
};

// ask the L2 for `bytes` (multiple of 16) at a 16-byte aligned global address
__device__ __forceinline__ void l2_prefetch(const void *ptr, unsigned bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ptr), "r"(bytes) : "memory");
}

// per-line L2 prefetch (no alignment requirement)
__device__ __forceinline__ void l2_prefetch_line(const void *ptr) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr));
}

template <int N>
struct FlatCfg {
  static constexpr int NN = N * N, N3 = N * N * N;
  // threads per block by N = 2..8, from A/B sweeps on B200 (profiles/): small
  // blocks keep barriers cheap, a one-warp block needs none; N = 6, 7 take 256-thread blocks because N^2 = 36, 49
  // threads per element fill 252 / 245 of 256 lanes but only 108 of 128 / 49 of 64 (N = 7: +23 %, N = 6: +2 %,
  // profiles/r02_flat_256thread_ab.txt)
  // (D4B_FLAT_THREADS / D4B_FLAT_MINB: compile-time overrides for those sweeps, profiles/build_variants.sh)
#ifndef D4B_FLAT_THREADS
#define D4B_FLAT_THREADS {0, 0, 32, 64, 32, 128, 256, 256, 64}
#endif
#ifndef D4B_FLAT_MINB
#define D4B_FLAT_MINB {0, 0, 1, 12, 24, 5, 2, 3, 8}
#endif
  static constexpr int kThreads[9] = D4B_FLAT_THREADS;
  static constexpr int kMinBlocks[9] = D4B_FLAT_MINB;   // resident blocks per SM the register budget is cut for
  static constexpr int THREADS = kThreads[N];
  static constexpr int EPB = THREADS / NN;
  // V0,V1,V2 + per-thread face results [6][2][NN] + 6 face metas (2 doubles each)
  // (the element stride is part of the tuned layouts of VLay: do not grow it)
  static constexpr int VS = VSize<N>::value;
  static constexpr int SH_PER_ELEM = 3 * VS + 12 * NN + 12 + VLay<N>::PAD;
  static constexpr int MIN_BLOCKS = kMinBlocks[N];
};

// a one-warp block only needs a warp-level barrier
template <int THREADS>
__device__ __forceinline__ void block_sync() {
#ifndef D4B_FLAT_NO_SYNCWARP
  if (THREADS == 32) { __syncwarp(); return; }
#endif
  __syncthreads();
}

// flux of one face at this thread's face Gauss point (x, y) (my face ordering).
// Branch-free so that the loads of several faces can be in flight together: a
// boundary face reads (and ignores) a line of the element itself.
template <int N>
__device__ __forceinline__ void flat_face(const FlatArgs &a, const FastTab &T, const FaceMeta &fm, int64_t ns, int x,
                                          int y, double trm, double dnm, double wxy, double &ve, double &vd) {
  constexpr int NN = N * N;
  const int code = fm.code;
  const bool own = code & kOwn, interior = fm.nbr != kNoNbr;
  int row = (code & (4 << 3)) ? x : y, col = (code & (4 << 3)) ? y : x;
  row = (code & (2 << 3)) ? N - 1 - row : row;
  col = (code & (1 << 3)) ? N - 1 - col : col;
  const int k = own ? (x + N * y) : (col + N * row);
  const double *r = a.FF3 + (size_t)fm.rec * (3 * NN) + k;
  const double r0 = r[0], r1 = r[NN], S = r[2 * NN];
  // neighbour line normal to its face f_p through its point (col,row); read it
  // from the face inwards so that the -1 end tables serve both ends
  // (e[1][l] = e[0][N-1-l], d[1][l] = -d[0][N-1-l])
  const int ndp = (code >> 1) & 3, sp = code & 1;
  const int stride = (ndp == 0) ? 1 : ((ndp == 1) ? N : NN);
  const int base = (ndp == 0) ? (N * col + NN * row) : ((ndp == 1) ? (col + NN * row) : (col + N * row));
  double v[N];
  double trp = 0., dnp = 0.;
  const double *p0 = a.uh + (interior ? fm.nbr : ns) + base;
  {
    const double *p = p0 + (sp ? (N - 1) * stride : 0);
    const int step = sp ? -stride : stride;
#pragma unroll
    for (int l = 0; l < N; ++l) v[l] = p[l * step];
#pragma unroll
    for (int l = 0; l < N; ++l) {
      trp += T.e[0][l] * v[l];
      dnp += T.d[0][l] * v[l];
    }
    dnp = sp ? -dnp : dnp;
  }
  const double Amn = own ? r0 : -r1;
  const double Apn = own ? r1 : -r0;
  const double du = trm - (interior ? trp : r1);   // boundary: Dirichlet data in the A_p slot
  const double cm = interior ? .5 : 1., cp = interior ? .5 : 0.;
  ve = wxy * (S * du - cm * Amn * dnm - cp * Apn * dnp);
  vd = -wxy * cm * Amn * du;
}

// CG: every element of the class is affine (J and d(rst)/d(xyz) constant over the element, exact test at
// set-up): the six metric components are one number each per element, times the Gauss weights
// HAT: the smoother recurrence carried in the Gauss basis (D4bEpilogue mode 3): back-transform with B = I I^T,
// rhs / p / the updated iterate are Gauss-basis vectors, the new uh goes to epi.uh_out (neighbours still read a.uh)
// DOT: u . A u of the launch in its energy form, one partial sum per warp in epi.dot_part (cg_eigs, d4b_solver.cu)
template <int N, bool MASS, bool CG = false, bool HAT = false, bool DOT = false>
__global__ void __launch_bounds__(FlatCfg<N>::THREADS, FlatCfg<N>::MIN_BLOCKS) flat_apply_kernel(FlatArgs a) {
  using C = FlatCfg<N>;
  constexpr int NN = C::NN, N3 = C::N3;
  extern __shared__ double smem[];
  const FastTab &T = c_tab[N];
  const int slot = threadIdx.x / NN;
  const int lt = threadIdx.x % NN;
  const int lx = lt % N, ly = lt / N;
  {
  const int blk = blockIdx.x;
  const int ei = a.first + blk * C::EPB + slot;
  const bool active = (slot < C::EPB) && (ei < a.n_elem);
  double *V0 = smem + (size_t)(slot < C::EPB ? slot : 0) * C::SH_PER_ELEM;
  double *V1 = V0 + C::VS, *V2 = V1 + C::VS;
  double *FB = V2 + C::VS;                                // [6][2][NN], thread-private columns
  FaceMeta *META = reinterpret_cast<FaceMeta *>(FB + 12 * NN);
  constexpr bool dot = DOT;
  double dsum = 0.;   // dot: this thread's share of u . A u in its energy form, grad . F (+ mass, irregular faces) + faces
  const int64_t ns = !active ? 0 : a.ns_linear ? a.ns0 + (int64_t)ei * N3 : a.elem_ns[ei];
  const double wxy = T.w[lx] * T.w[ly];
  double gt[N];
  double tr4 = 0., dn4 = 0., tr5 = 0., dn5 = 0.;

  // the volume metric is the largest stream and is first needed half-way
  // through the kernel: ask the L2 for this block's share now
  if (!CG && threadIdx.x == 0 && a.prefetch) {
    constexpr unsigned kBlockBytes = C::EPB * 6 * N3 * sizeof(double);
    const int e0 = a.first + blk * C::EPB, left = a.n_elem - e0;
    l2_prefetch(a.FG + (size_t)e0 * (6 * N3), left >= C::EPB ? kBlockBytes : left * 6 * N3 * 8);
  }
  // fused smoother: its three operand streams are only touched by the very last
  // phase; have them in L2 by then
  if (active && (a.epi.mode == 1 || HAT) && a.prefetch) {
    for (int off = lt * 16; off < N3; off += NN * 16) {
      l2_prefetch_line(a.epi.rhs + ns + off);
      l2_prefetch_line(a.epi.p + ns + off);
      if (!HAT) l2_prefetch_line(a.epi.u + ns + off);
    }
  }
  // ---- T-orientation (a,b) = (lx,ly), line along c ----
  if (active) {
    double uh[N];
#pragma unroll
    for (int c = 0; c < N; ++c) uh[c] = a.uh[ns + lt + NN * c];
    for (int j = lt; j < 6; j += NN) {
      const FaceMeta fm = a.meta[6 * (size_t)ei + j];
      META[j] = fm;
      if ((3 * NN * 8) % 16 == 0 && a.prefetch > 1) l2_prefetch(a.FF3 + (size_t)fm.rec * (3 * NN), 3 * NN * 8);
      if (a.prefetch > 2 && fm.nbr != kNoNbr)   // the neighbour's uh block: its lines are read two phases later
        for (int off = 0; off < N3; off += 16) l2_prefetch_line(a.uh + fm.nbr + off);
    }
#pragma unroll
    for (int i = 0; i < N; ++i) {
      V0[vix<N>(lx, ly, i)] = uh[i];
      tr4 += T.e[0][i] * uh[i]; dn4 += T.d[0][i] * uh[i];
      tr5 += T.e[1][i] * uh[i]; dn5 += T.d[1][i] * uh[i];
    }
#pragma unroll
    for (int c = 0; c < N; ++c) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[c][i] * uh[i];
      gt[c] = s;
    }
  }
  // merged launch with the peer-memory halo (small meshes): the groups holding elements with a ghost neighbour --
  // the last ones of the class order -- wait for the neighbours' uh here (their own loads are already in flight);
  // nothing before this point reads a neighbour.  On large meshes these groups are a launch of their own on the
  // halo stream instead (d4b_capi.cu): at the end of one big launch their NVLink latency is the tail of the kernel
  // (0.375 vs 0.368 ms at 2 GPUs), moved forward they spin for a neighbour that may be most of an apply behind and
  // hold SM slots meanwhile (0.418 ms) -- profiles/r02_peer_merged_launch_ab.txt
  if (a.halo.n > 0 && a.first + (blk + 1) * C::EPB > a.halo.from) halo_wait(a.halo);
  block_sync<C::THREADS>();
  if (active) {
    // the lifted face term tested with the element's own polynomial: sum_c uh_c (e_c ve + d_c vd) = tr ve + dn vd
    {   // faces 4/5 (normal c), face point (a,b)
      double ve, vd;
      flat_face<N>(a, T, META[4], ns, lx, ly, tr4, dn4, wxy, ve, vd);
      FB[(4 * 2 + 0) * NN + lt] = ve; FB[(4 * 2 + 1) * NN + lt] = vd;
      if (dot) dsum += tr4 * ve + dn4 * vd;
      flat_face<N>(a, T, META[5], ns, lx, ly, tr5, dn5, wxy, ve, vd);
      FB[(5 * 2 + 0) * NN + lt] = ve; FB[(5 * 2 + 1) * NN + lt] = vd;
      if (dot) dsum += tr5 * ve + dn5 * vd;
    }
    {   // S-orientation: (a,c) = (lx,ly), line along b: gradient, faces 2/3
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
      double ve, vd;
      flat_face<N>(a, T, META[2], ns, lx, ly, t2, n2, wxy, ve, vd);
      FB[(2 * 2 + 0) * NN + lt] = ve; FB[(2 * 2 + 1) * NN + lt] = vd;
      if (dot) dsum += t2 * ve + n2 * vd;
      flat_face<N>(a, T, META[3], ns, lx, ly, t3, n3, wxy, ve, vd);
      FB[(3 * 2 + 0) * NN + lt] = ve; FB[(3 * 2 + 1) * NN + lt] = vd;
      if (dot) dsum += t3 * ve + n3 * vd;
    }
    {   // R-orientation: (b,c) = (lx,ly), line along a: gradient, faces 0/1
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
      double ve, vd;
      flat_face<N>(a, T, META[0], ns, lx, ly, t0, n0, wxy, ve, vd);
      FB[(0 * 2 + 0) * NN + lt] = ve; FB[(0 * 2 + 1) * NN + lt] = vd;
      if (dot) dsum += t0 * ve + n0 * vd;
      flat_face<N>(a, T, META[1], ns, lx, ly, t1, n1, wxy, ve, vd);
      FB[(1 * 2 + 0) * NN + lt] = ve; FB[(1 * 2 + 1) * NN + lt] = vd;
      if (dot) dsum += t1 * ve + n1 * vd;
    }
  }
  block_sync<C::THREADS>();
  if (active) {   // T-orientation: metric, Dg^T along c + faces 4/5; V0 becomes the residual
    double ft[N];
    const double *G = CG ? nullptr : a.FG + (size_t)ei * (6 * N3) + lt;
    double gc[6];
    if (CG) {
#pragma unroll
      for (int q = 0; q < 6; ++q) gc[q] = a.FGc[(size_t)ei * 6 + q] * wxy;
    }
#pragma unroll
    for (int c = 0; c < N; ++c) {
      const int idx = vix<N>(lx, ly, c);
      double gr = V1[idx], gs = V2[idx];
      double G00, G01, G02, G11, G12, G22;
      if (CG) {
        const double wc = T.w[c];
        G00 = gc[0] * wc; G01 = gc[1] * wc; G02 = gc[2] * wc; G11 = gc[3] * wc; G12 = gc[4] * wc; G22 = gc[5] * wc;
      } else {
        G00 = G[0 * N3 + NN * c]; G01 = G[1 * N3 + NN * c]; G02 = G[2 * N3 + NN * c];
        G11 = G[3 * N3 + NN * c]; G12 = G[4 * N3 + NN * c]; G22 = G[5 * N3 + NN * c];
      }
      const double Fr = G00 * gr + G01 * gs + G02 * gt[c];
      const double Fs = G01 * gr + G11 * gs + G12 * gt[c];
      V1[idx] = Fr;
      V2[idx] = Fs;
      ft[c] = G02 * gr + G12 * gs + G22 * gt[c];
      if (dot) dsum += gr * Fr + gs * Fs + gt[c] * ft[c];
    }
    const double ve0 = FB[(4 * 2 + 0) * NN + lt], vd0 = FB[(4 * 2 + 1) * NN + lt];
    const double ve1 = FB[(5 * 2 + 0) * NN + lt], vd1 = FB[(5 * 2 + 1) * NN + lt];
    // mass-type diagonal: in the Gauss basis V^T diag(W) V u is point-wise, r += W uh
    const double *Wd = MASS ? a.W + a.quad_stride[a.elem[ei]] + lt : nullptr;
    const double *Bd = (MASS && a.nlB) ? a.nlB + a.quad_stride[a.elem[ei]] + lt : nullptr;
    const double *xtr = (a.has_extra && a.has_extra[ei]) ? a.extra + ns + lt : nullptr;
#pragma unroll
    for (int c = 0; c < N; ++c) {
      double s = T.e[0][c] * ve0 + T.d[0][c] * vd0 + T.e[1][c] * ve1 + T.d[1][c] * vd1;
      if (MASS) {
        const double uc = V0[vix<N>(lx, ly, c)], wu = mass_point(Wd, Bd, a.nlK, NN * c, uc);
        s += wu;
        if (dot) dsum += wu * uc;
      }
      if (xtr) {
        s += xtr[NN * c];
        if (dot) dsum += xtr[NN * c] * V0[vix<N>(lx, ly, c)];
      }
#pragma unroll
      for (int i = 0; i < N; ++i) s += T.Dg[i][c] * ft[i];
      V0[vix<N>(lx, ly, c)] = s;
    }
  }
  if (dot) {   // one partial sum per warp, fixed order (deterministic)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dsum += __shfl_xor_sync(0xffffffffu, dsum, o);
    if ((threadIdx.x & 31) == 0) a.epi.dot_part[(size_t)blk * (C::THREADS / 32) + (threadIdx.x >> 5)] = dsum;
  }
  block_sync<C::THREADS>();
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
  block_sync<C::THREADS>();
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
      for (int i = 0; i < N; ++i) s += (HAT ? T.B[i][i2] : T.I[i][i2]) * tot[i];
      V0[vix<N>(i2, lx, ly)] = s;
    }
  }
  block_sync<C::THREADS>();
  if (active) {   // S-orientation: I^T along b
    double l[N];
#pragma unroll
    for (int i = 0; i < N; ++i) l[i] = V0[vix<N>(lx, i, ly)];
#pragma unroll
    for (int i2 = 0; i2 < N; ++i2) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += (HAT ? T.B[i][i2] : T.I[i][i2]) * l[i];
      V0[vix<N>(lx, i2, ly)] = s;
    }
  }
  block_sync<C::THREADS>();
  if (active) {   // T-orientation: I^T along c, write Au
    double l[N];
    const EpiCoef ec = epi_coef(a.epi);
#pragma unroll
    for (int i = 0; i < N; ++i) l[i] = V0[vix<N>(lx, ly, i)];
#pragma unroll
    for (int c = 0; c < N; ++c) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += (HAT ? T.B[i][c] : T.I[i][c]) * l[i];
      if (HAT) {
        const int64_t i = ns + lt + NN * c;
        const double r = (a.epi.rhs[i] - s) * ec.alpha;
        const double pn = r + ec.beta * a.epi.p[i];
        a.epi.p[i] = pn;
        a.epi.uh_out[i] = a.uh[i] + pn;
      } else {
        store_result(a.epi, ec, a.Au, ns + lt + NN * c, s);
      }
    }
  }
  }   // element group
}
