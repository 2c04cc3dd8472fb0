/* d4est_b200.h -- C ABI of libd4est_b200.so
 *
 * B200-native (sm_100a) replacement for the one data-parallel hot path of
 * trevor-vincent/disco4est: the matrix-free SIPG operator application
 *   Au = K u + F(u)
 * (d4est_laplacian_apply_aij, src/dGMath/d4est_laplacian.c:317-399, and its
 * with_opt twin, src/dGMath/d4est_laplacian_with_opt.c:289-368), the BLAS-1
 * sweeps around it (src/LinearAlgebra/d4est_linalg.c:179-236) and the
 * Chebyshev / cg_eigs smoother kernels that wrap it
 * (src/Solver/d4est_solver_multigrid_smoother_cheby.c:81-176,
 *  src/Solver/d4est_solver_cg_eigs.c:115-276).
 *
 * Plain C types only: pointers, sizes, enums.  No torch, no C++ in signatures.
 * Error convention follows the reference (src/Utilities/d4est_util.h:171,
 * D4EST_ABORT): entry points do not return error codes; any CUDA failure or
 * invalid input calls the abort hook (d4b_set_abort_hook), which by default
 * prints "[D4B_ABORT] file:line msg" and calls abort().  There is NO CPU
 * fallback: without a CUDA device every compute entry point aborts.
 */
#ifndef D4EST_B200_H
#define D4EST_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define D4B_ABI_VERSION 1
#define D4B_MAX_DEG 7          /* p = 1..7 (N = 2..8 nodes per direction)   */
#define D4B_FACES 6
#define D4B_HALF 4             /* sub-faces of a hanging 3-D face            */

/* penalty functions, src/dGMath/d4est_laplacian_flux_sipg.c:945-1005 */
typedef enum {
  D4B_PENALTY_MAXP_SQR_OVER_MINH = 0,
  D4B_PENALTY_MAXPP1_SQR_OVER_MINH = 1,
  D4B_PENALTY_MEANP_SQR_OVER_MEANH = 2,
  D4B_PENALTY_MEAN_P_SQR_OVER_H = 3
} d4b_penalty_t;

/* boundary condition on physical boundary faces,
 * src/dGMath/d4est_laplacian_flux.h:171-186 */
typedef enum { D4B_BC_DIRICHLET = 0, D4B_BC_ROBIN = 1 } d4b_bc_t;

/* Flattened, read-only description of one rank's mesh in the REFERENCE'S OWN
 * array layouts (so that the adapter on the reference side only copies
 * pointers).  Everything is a host pointer; d4b_mesh_create() copies what it
 * needs and re-lays it out for the device, the caller may free afterwards.
 *
 * Element / node ordering = reference ordering: element id = position in the
 * local tree/quadrant iteration (d4est_mesh.c:2394-2490), node i of element e
 * lives at nodal_stride[e] + i with i = ir + N*(is + N*it), r fastest
 * (d4est_operators.c:1385-1410).                                           */
typedef struct d4b_mesh_desc
{
  int32_t abi_version;          /* = D4B_ABI_VERSION                        */
  int32_t n_local;              /* local elements                           */
  int32_t n_ghost;              /* ghost elements (0 on one rank)           */
  int32_t n_sides;              /* flux "side" records, see below           */
  int64_t local_nodes;          /* d4est_mesh_local_sizes_t.local_nodes     */
  int64_t local_nodes_quad;     /* .local_nodes_quad                        */
  int64_t local_mortar_nodes_quad; /* .local_mortar_nodes_quad              */
  int64_t ghost_nodes;          /* sum of ghost (deg+1)^3                   */

  /* ---- per local element [n_local]  (d4est_element_data.h:13-46) ---- */
  const int32_t *deg;
  const int32_t *deg_quad;
  const int64_t *nodal_stride;
  const int64_t *quad_stride;

  /* ---- per ghost element [n_ghost], in ghost-id order ---- */
  const int32_t *ghost_deg;
  const int32_t *ghost_deg_quad;
  const int64_t *ghost_nodal_stride;  /* offset into the ghost receive buffer
                                         (d4est_ghost_data.c:258-268)       */

  /* ---- volume geometric factors (d4est_mesh.h:117-162) ---- */
  const double *J_quad;         /* [local_nodes_quad]                       */
  const double *rst_xyz_quad;   /* [(i*3+j)*local_nodes_quad + quad_stride]
                                   = d(rst_i)/d(xyz_j), d4est_mesh.c:2757-2777 */

  /* ---- mortar geometric factors, all indexed by side_mortar_stride ----
   * scalar s[k]            at            stride + k
   * vector v[d][k]         at 3*stride + d*total + k
   * matrix m[i][j][k]      at 9*stride + (i + 3*j)*total + k
   * with total = mortar quadrature nodes of the side
   * (d4est_mesh.c:962-1000, d4est_laplacian_flux.c:425-449)              */
  const double *sj_m_mortar_quad;
  const double *n_m_mortar_quad;
  const double *drst_dxyz_m_mortar_quad;
  const double *drst_dxyz_p_mortar_quad_porder;  /* (+)-side ordering       */
  const double *hm_mortar_quad;
  const double *hp_mortar_quad;
  /* Dirichlet data g at the boundary-face quadrature points, same scalar
   * indexing as sj (only boundary sides are read); NULL = homogeneous, which
   * is what every apply_lhs of the reference uses.                         */
  const double *dirichlet_g_mortar_quad;

  /* ---- flux side records [n_sides] ----
   * One record per call the reference's face iteration makes to the flux
   * interface/boundary function (d4est_mortars.c:601-802): a (-) side of 1 or
   * 4 elements on face f_m against a (+) side of 1 or 4 elements (0 =
   * physical boundary).  Element references: >= 0 local element id,
   * < 0 ghost element -(gid+1).                                            */
  const int32_t *side_faces_m;     /* 1 or 4                                */
  const int32_t *side_faces_p;     /* 0 (boundary), 1 or 4                  */
  const int32_t *side_f_m;         /* face of the (-) elements, 0..5        */
  const int32_t *side_f_p;         /* face of the (+) elements              */
  const int32_t *side_e_m;         /* [4*n_sides]                           */
  const int32_t *side_e_p;         /* [4*n_sides], in (+)-side order        */
  /* face-node permutation (+)->(-) ordering, bit0 flip dir0, bit1 flip dir1,
   * bit2 transpose, applied in that order (what
   * d4est_operators_reorient_face_data does, d4est_operators.c:1993-2087)  */
  const int32_t *side_perm;
  /* sub-face reorder: (-)-ordered sub-face i is (+)-ordered sub-face
   * side_sub_order[4*s+i] (d4est_reference_reorient_face_order)            */
  const int32_t *side_sub_order;   /* [4*n_sides]                           */
  const int64_t *side_mortar_stride;
  const double  *side_penalty_mult; /* region multiplier,
                                       d4est_laplacian_flux_sipg.c:1165-1199 */

  /* ---- flux parameters (d4est_laplacian_flux_sipg.h:18-37) ---- */
  double  sipg_penalty_prefactor;
  int32_t sipg_penalty_fcn;        /* d4b_penalty_t                         */
  int32_t bc_type;                 /* d4b_bc_t                              */

  /* ---- ghost exchange plan (d4est_ghost_data.c:142-255) ---- */
  int32_t mpirank, mpisize;
  int32_t n_mirrors;
  const int32_t *mirror_elem;          /* [n_mirrors] local element ids     */
  const int32_t *mirror_proc_offsets;  /* [mpisize+1]                       */
  const int32_t *mirror_proc_mirrors;  /* indices into mirror_elem          */
  const int32_t *ghost_proc_offsets;   /* [mpisize+1] ghost ranges per rank */

  /* ---- Robin boundary (bc_type = D4B_BC_ROBIN), d4est_laplacian_flux_sipg_robin_aux
   * (src/dGMath/d4est_laplacian_flux_sipg.c:339-455): on a boundary face the operator adds
   * V^T w sj (coeff u_m - rhs) and nothing else.  Both arrays use the scalar indexing of sj and
   * hold the callbacks d4est_laplacian_robin_bc_t.robin_coeff / .robin_rhs evaluated at the
   * boundary quadrature points (only boundary sides are read).  rhs may be NULL (= 0, what the
   * Jacobian and residual of the two-puncture problems use, multi_puncture_fcns.h:790-804);
   * a point with coeff = 0 and rhs != 0 is rejected.                                        */
  const double *robin_coeff_mortar_quad;
  const double *robin_rhs_mortar_quad;
} d4b_mesh_desc_t;

typedef struct d4b_mesh d4b_mesh_t;    /* opaque device-resident mesh */

/* ---------------------------------------------------------------- runtime */
typedef void (*d4b_abort_hook_t)(const char *file, int line, const char *msg);
void d4b_set_abort_hook(d4b_abort_hook_t hook);
int  d4b_abi_version(void);
/* number of CUDA devices visible (0 if none / driver missing); never aborts */
int  d4b_device_count(void);
/* bind this process to a device (reference: one MPI rank per device)       */
void d4b_init(int device);

/* ------------------------------------------------------------------- ranks */
/* One process per GPU.  Replaces p4est->mpicomm for this path: the halo
 * exchange (d4est_ghost_data_exchange, src/Mesh/d4est_ghost_data.c:142-255)
 * and the scalar reductions (sc_allreduce, src/Solver/d4est_solver_cg_eigs.c:181-245)
 * run over NCCL.  Rank 0 obtains an id, the caller broadcasts its
 * D4B_COMM_ID_BYTES bytes with its own transport (MPI_Bcast in the reference,
 * torch.distributed in the harness), every rank calls d4b_comm_init, then
 * creates meshes whose desc carries mpirank/mpisize and the exchange plan.
 * Single-rank users call none of these and need no NCCL.                    */
#define D4B_COMM_ID_BYTES 128
void d4b_comm_get_unique_id(void *id_out);
void d4b_comm_init(const void *id, int rank, int size);
void d4b_comm_destroy(void);
int  d4b_comm_size(void);
int  d4b_comm_rank(void);

/* ------------------------------------------------------------------- mesh */
/* replaces the per-mesh setup the reference does lazily inside
 * d4est_mesh_update (src/Mesh/d4est_mesh.c:2790-2890) for the apply path   */
d4b_mesh_t *d4b_mesh_create(const d4b_mesh_desc_t *desc);
void        d4b_mesh_destroy(d4b_mesh_t *mesh);
int64_t     d4b_mesh_local_nodes(const d4b_mesh_t *mesh);
/* bytes of device memory held by the mesh (geometric factors + tables)     */
int64_t     d4b_mesh_device_bytes(const d4b_mesh_t *mesh);
/* elements taken by the uniform-degree conforming kernel (the rest go through
 * the general hp kernel); for reporting and tests                           */
int64_t     d4b_mesh_fast_elements(const d4b_mesh_t *mesh);
/* of those, elements whose six faces carry no tangential metric terms */
int64_t     d4b_mesh_flat_elements(const d4b_mesh_t *mesh);
/* elements taken by the hp mortar kernel: collocated quadrature (deg_quad == deg) on the element
 * and its face neighbours, faces between unequal degrees and / or hanging faces             */
int64_t     d4b_mesh_hp_elements(const d4b_mesh_t *mesh);
/* Page-lock `bytes` of host memory at `ptr` (cudaHostRegister) / release it again.  The reference's solver vectors are
 * P4EST_ALLOC'd, i.e. pageable (d4est_elliptic_data_t, src/EllipticSystem/d4est_elliptic_data.h): a host-pointer
 * d4b_apply then stages every copy through the driver's bounce buffer and the chunked host apply cannot overlap its
 * transfers.  Pinned once after allocation they go at link speed.  Returns 0 on success, a CUDA error code otherwise
 * (the vectors then simply stay pageable); d4b_host_unpin before the memory is freed.                             */
int32_t     d4b_host_pin(void *ptr, int64_t bytes);
int32_t     d4b_host_unpin(void *ptr);
/* CUDA kernels one d4b_apply launches on this mesh (for launch accounting)  */
int32_t     d4b_mesh_launches_per_apply(const d4b_mesh_t *mesh);
/* How this mesh's applies obtain ghost data (replaces d4est_ghost_data_exchange, src/Mesh/d4est_ghost_data.c:142-255):
 * 0 one rank, no exchange; 1 NCCL send/recv of the mirrors' nodal values; 2 peer memory -- the boundary elements read
 * the neighbour ranks' Gauss-collocated vectors over NVLink (CUDA IPC mappings, one node; chosen in the first apply
 * when every rank can map its neighbours and no element needs the general kernel; D4B_PEER_HALO=0 keeps mode 1);
 * -1 multi-rank mesh that has not been applied yet.  With mode 2 d4b_mesh_destroy is collective.              */
int32_t     d4b_mesh_halo_mode(const d4b_mesh_t *mesh);

/* --------------------------------------------- geometric factors on device */
/* Replaces the geometry half of d4est_mesh_data_compute (src/Mesh/d4est_mesh.c:2543-2755)
 * and d4est_mesh_compute_mortar_quadrature_quantities (:481-1140,
 * d4est_mortars_compute_geometric_data_on_mortar_aux, src/Mesh/d4est_mortars.c:10-277) for
 * the analytic maps ([geometry] DX_compute_method = analytic, JAC_compute_method =
 * numerical): every volume and mortar quadrature point evaluates dX/drst of the tree map
 * in a CUDA thread, inverts it and writes the reference's arrays.  The derivatives are our
 * own differentiation of the maps X (d4est_geometry_brick.c, d4est_geometry_cubed_sphere.c:
 * 127-167, :315-400), not the reference's generated expressions.                        */
typedef enum {
  D4B_GEOM_BRICK = 0,          /* [geometry] name = brick: one tree, [X0,X1]x[Y0,Y1]x[Z0,Z1]       */
  D4B_GEOM_CUBED_SPHERE = 1    /* name = cubed_sphere: 13 trees (p8est_connectivity_new_sphere):
                                  0-5 outer shell R1..R2, 6-11 inner shell R0..R1, 12 centre cube */
} d4b_geom_type_t;

/* d4est_mesh_face_h_t, same numeric values (src/Mesh/d4est_mesh.h:32-42) */
typedef enum {
  D4B_FACE_H_J_DIV_SJ_QUAD = 0,
  D4B_FACE_H_J_DIV_SJ_MIN_LOBATTO = 1,
  D4B_FACE_H_J_DIV_SJ_MEAN_LOBATTO = 2,
  D4B_FACE_H_J_DIV_SJ_MAX_LOBATTO = 3,
  D4B_FACE_H_TREE_H = 4,
  D4B_FACE_H_VOLUME_DIV_AREA = 5
} d4b_face_h_t;

#define D4B_ROOT_LEN_LOG2 30   /* P8EST_MAXLEVEL: quadrant coordinates live in [0, 2^30) */

typedef struct d4b_geom_desc
{
  int32_t geom_type;                  /* d4b_geom_type_t                                   */
  int32_t face_h_type;                /* d4b_face_h_t                                      */
  int32_t compactify_inner_shell, compactify_outer_shell;
  double  X0, X1, Y0, Y1, Z0, Z1;     /* brick extents (d4est_geometry_brick_attr_t)       */
  double  R0, R1, R2;                 /* cubed sphere radii (…cubed_sphere_attr_t)         */
  /* p4est position of every local / ghost element (d4est_element_data_t.tree, q, dq)      */
  const int32_t *elem_tree;           /* [n_local]                                         */
  const int32_t *elem_q;              /* [3*n_local]                                       */
  const int32_t *elem_dq;             /* [n_local]                                         */
  const int32_t *ghost_tree, *ghost_q, *ghost_dq;
} d4b_geom_desc_t;

/* outputs in the reference's layouts (see d4b_mesh_desc_t); each pointer may be a host
 * pointer, a device pointer, or NULL (not wanted)                                        */
typedef struct d4b_geom_arrays
{
  double *xyz;                        /* [3*local_nodes]: d*local_nodes + nodal_stride + i */
  double *J_quad, *rst_xyz_quad;
  double *sj_m_mortar_quad, *n_m_mortar_quad;
  double *drst_dxyz_m_mortar_quad, *drst_dxyz_p_mortar_quad_porder;
  double *hm_mortar_quad, *hp_mortar_quad;
  double *xyz_quad;                   /* [3*local_nodes_quad]: d*local_nodes_quad + quad_stride + k
                                         (d4est_mesh_data_t.xyz_quad) */
} d4b_geom_arrays_t;

/* `topo` supplies sizes, degrees, strides and side records; its geometric arrays are not read */
void d4b_geom_compute(const d4b_mesh_desc_t *topo, const d4b_geom_desc_t *geom,
                      const d4b_geom_arrays_t *out);
/* d4b_mesh_create with the geometric factors generated on the device instead of uploaded */
d4b_mesh_t *d4b_mesh_create_with_geometry(const d4b_mesh_desc_t *topo, const d4b_geom_desc_t *geom);

/* ------------------------------------------------------------------ apply */
/* Au = A u; replaces d4est_laplacian_apply_aij / _with_opt_apply_aij
 * (d4est_laplacian.c:317-399).  u, Au: local_nodes doubles in reference node
 * order; each may be a host or a device pointer (cudaPointerGetAttributes).
 * Au is fully overwritten (d4est_quadrature.c:337).  Synchronous when any
 * pointer is a host pointer; stream-ordered otherwise (d4b_sync to wait).  */
void d4b_apply(d4b_mesh_t *mesh, const double *u, double *Au);
/* volume stiffness only: d4est_laplacian_apply_stiffness_matrix
 * (d4est_laplacian.c:197-234)                                              */
void d4b_apply_stiffness(d4b_mesh_t *mesh, const double *u, double *Au);
void d4b_sync(d4b_mesh_t *mesh);

/* ------------------------------------------- volume quadrature operators */
/* On every local element, LGL(deg) <-> Gauss(deg_quad); V = the 3-D interpolation matrix,
 * w = the tensor Gauss weights.  Nodal arrays have local_nodes entries (reference node
 * order), *_quad arrays local_nodes_quad entries at quad_stride[e] + k.  Host or device
 * pointers; synchronous when any is a host pointer.
 *   d4b_interp_to_quad : u_quad = V u                   d4est_quadrature_interpolate
 *                                                        (src/Quadrature/d4est_quadrature.c:966-1016)
 *   d4b_apply_galerkin : out = V^T (w jac_quad in_quad)  d4est_quadrature_apply_galerkin_integral (:142-213);
 *                                                        jac_quad may be NULL (= 1)
 *   d4b_apply_mass     : out = V^T diag(w jac_quad) V x  d4est_quadrature_apply_mass_matrix (:385-477)
 * With jac_quad = J these are the reference's operators; with jac_quad = J f(x, V u0) --
 * f evaluated by the caller through its own d4est_xyzu_fcn_t -- d4b_apply_mass is the
 * Newton-linearised mass-type term of d4est_quadrature_apply_fofufofvlilj (QUAD_APPLY_MATRIX,
 * :593-774) and d4b_apply_galerkin(mesh, f_quad, J, out) the residual term of
 * d4est_quadrature_apply_fofufofvlj (:776-936).                                           */
void d4b_interp_to_quad(d4b_mesh_t *mesh, const double *u, double *u_quad);
void d4b_apply_galerkin(d4b_mesh_t *mesh, const double *in_quad, const double *jac_quad, double *out);
void d4b_apply_mass(d4b_mesh_t *mesh, const double *x, const double *jac_quad, double *out);
/* Right-hand side with strongly imposed boundary data, d4est_laplacian_build_rhs_with_strong_bc
 * (src/dGMath/d4est_laplacian.c:15-140):  rhs = M f - A(0), where A(0) is the operator of `mesh`
 * applied to u = 0 -- only the inhomogeneous Dirichlet terms of the boundary faces survive, so the
 * mesh must have been created with the problem's boundary data in dirichlet_g_mortar_quad.
 *   f_on_quad = 0 (INIT_FIELD_ON_LOBATTO): f at the LGL nodes, M f = V^T diag(w jac_quad) V f
 *   f_on_quad = 1 (INIT_FIELD_ON_QUAD):    f at the quadrature points, M f = V^T (w jac_quad f)
 * jac_quad = J_quad.  Host or device pointers.                                                   */
void d4b_build_rhs(d4b_mesh_t *mesh, const double *f, int32_t f_on_quad, const double *jac_quad, double *rhs);
/* Residual-based a-posteriori error estimator, d4est_estimator_bi_compute
 * (src/Estimators/d4est_estimator_bi.c:349-602): for every local element
 *   eta2[e] = h_e^2 / p_e^2 |r|^2_{L2(e)}
 *           + sum over its mortars of  int sj [ (c1 n.(grad u_m - grad u_p))^2 + (c2 (u_m - u_p))^2 ]
 *           + on boundary faces        int sj (c_d (u_m - g))^2
 * with the penalty prefactors of d4est_estimator_bi.h:25-197 (family below).  `residual` is what the
 * problem's build_residual callback left in Au (nodal), diam_volume the reference's per-element size
 * (d4est_mesh_data_t.diam_volume), sj / hm / hp its mortar arrays (the mesh does not keep them),
 * dirichlet_g the boundary data at the boundary quadrature points (NULL = zero).  terms (may be
 * NULL) receives the four parts like estimator_vtk: [4][n_local] = residual, Je1, Je2, Je2 boundary.
 * All pointers host or device.                                                                  */
typedef enum {
  D4B_EST_HOUSTON_MAXP_MINH = 0,       /* houston_{gradu,u,u_dirichlet}_prefactor_maxp_minh   */
  D4B_EST_HOUSTON_MAX_P2_OVER_H = 1,   /* houston_gradu_prefactor_max_h_over_p, houston_u[_dirichlet]_prefactor_max_p2_over_h */
  D4B_EST_BI_MAXP_MINH = 2,            /* bi_gradu_prefactor_maxp_minh, bi_u_prefactor_conforming_maxp_minh */
  D4B_EST_BI_MAX_P2_OVER_H = 3         /* bi_gradu_prefactor_max_h_over_p, bi_u_prefactor_conforming_max_p2_over_h */
} d4b_estimator_penalty_t;
typedef struct d4b_estimator_desc
{
  const double *u, *residual;                 /* [local_nodes] */
  const double *jac_quad;                     /* [local_nodes_quad] J_quad */
  const double *diam_volume;                  /* [n_local] */
  const double *sj_m_mortar_quad, *hm_mortar_quad, *hp_mortar_quad;   /* [local_mortar_nodes_quad] */
  const double *dirichlet_g_mortar_quad;      /* [local_mortar_nodes_quad] or NULL */
  int32_t penalty_family;                     /* d4b_estimator_penalty_t */
  double  penalty_prefactor;
} d4b_estimator_desc_t;
void d4b_estimator_bi(d4b_mesh_t *mesh, const d4b_estimator_desc_t *desc, double *eta2, double *terms);
/* Fuse a mass-type diagonal into the operator: after this call every d4b_apply (and the
 * smoother, cg_eigs and V-cycle built on it) computes
 *   Au = A u + V^T diag(w jac_quad) V u,
 * i.e. the Jacobian apply of the reference's nonlinear problems (SIPG Laplacian +
 * d4est_quadrature_apply_fofufofvlilj term, e.g. constant_density_star_apply_jac,
 * src/Problems/ConstantDensityStar/constant_density_star_fcns.h:528-600) in one pass: the
 * kernels add W * u at the Gauss points, 8 more bytes per quadrature point.  jac_quad is
 * copied (host or device); NULL removes the term.  d4b_apply_stiffness is unaffected.     */
void d4b_mesh_set_mass_diag(d4b_mesh_t *mesh, const double *jac_quad);
/* run all work of this mesh on the caller's CUDA stream (a cudaStream_t passed
 * as void*; NULL = the legacy default stream).  The library otherwise uses a
 * private non-blocking stream.                                             */
void d4b_set_stream(d4b_mesh_t *mesh, void *cuda_stream);

/* --------------------------------------------------------------- BLAS-1  */
/* device-resident vector ops, src/LinearAlgebra/d4est_linalg.c:179-236;
 * all pointers are DEVICE pointers                                          */
double d4b_vec_dot(d4b_mesh_t *mesh, const double *x, const double *y, int64_t n);
void   d4b_vec_axpy(d4b_mesh_t *mesh, double a, const double *x, double *y, int64_t n);
void   d4b_vec_xpby(d4b_mesh_t *mesh, const double *x, double b, double *y, int64_t n);
void   d4b_vec_scale(d4b_mesh_t *mesh, double a, double *x, int64_t n);

/* d4b_vec_dot_noalloc has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  It is d4b_vec_dot without a mesh
 * and without any allocation: the per-block partial sums of the two-pass reduction live in a
 * file-scope device array of the library (csrc/d4b_static.cu) instead of the mesh's
 * cudaMalloc'ed buffer, and the path calls no cudaMalloc, cudaFree or allocating routine.
 *
 *   out[v] = sum_i x[v*n + i] * y[v*n + i],  v = 0 .. nvec-1
 *
 * x, y (nvec*n doubles each) and out (nvec doubles) are DEVICE pointers on the current
 * device; the two kernels are enqueued on cuda_stream (a cudaStream_t passed as void*, NULL =
 * the legacy default stream) and the call does not synchronise.  For one vector pair the sum
 * is bit-identical to d4b_vec_dot on one rank (same grid, same reduction trees); there is no
 * allreduce over ranks.
 * Limit: nvec <= D4B_DOT_NOALLOC_MAX_VECS; more is rejected with D4B_NOALLOC_E_TOO_LARGE.
 * The scratch is D4B_DOT_NOALLOC_MAX_VECS x 592 doubles = 303 104 bytes, held by every GPU
 * that loads the library until the process exits.  n is not limited.
 * Sharing: every call on a device uses the same array, which keeps the last call's partial
 * sums; each call overwrites all it reads, so nothing needs clearing, but calls that overlap
 * on the GPU (other streams, concurrent graph replays) RACE -- serialise them.
 * Set-up (the device address of the array, once per device) happens on the first call with
 * that device current; nvec = 0 does the set-up and launches nothing -- call it once before
 * a stream capture.
 * Returns D4B_NOALLOC_OK or an error code; never aborts.  On error a message is written to
 * err (err_len bytes, NUL-terminated; err may be NULL); for E_ARG / E_TOO_LARGE / a failed
 * set-up nothing has been launched, E_CUDA reports a failed kernel launch (`out` is then
 * undefined).                                                                                */
#define D4B_DOT_NOALLOC_MAX_VECS 64
#define D4B_NOALLOC_OK           0
#define D4B_NOALLOC_E_ARG        1   /* negative size or NULL pointer             */
#define D4B_NOALLOC_E_TOO_LARGE  2   /* nvec > D4B_DOT_NOALLOC_MAX_VECS           */
#define D4B_NOALLOC_E_CUDA       3   /* a CUDA call failed; err has its message   */
int32_t d4b_vec_dot_noalloc(const double *x, const double *y, int64_t n, int32_t nvec, double *out,
                            void *cuda_stream, char *err, int64_t err_len);

/* ---------------------------------------------------- Chebyshev smoother  */
/* parameters of [mg_smoother_cheby], d4est_solver_multigrid_smoother_cheby.h:7-27 */
typedef struct d4b_cheby_params
{
  int32_t cheby_imax;
  int32_t cheby_eigs_cg_imax;
  double  cheby_eigs_lmax_lmin_ratio;
  double  cheby_eigs_max_multiplier;
  int32_t cheby_use_new_cg_eigs;
  int32_t cheby_use_zero_guess_for_eigs;
  int32_t cheby_eigs_reuse_fromdownvcycle;
  int32_t cheby_eigs_reuse_fromlastvcycle;
} d4b_cheby_params_t;

/* cg_eigs (src/Solver/d4est_solver_cg_eigs.c:115-276): runs imax CG steps on
 * A x = rhs starting from x0 = u and returns the Gershgorin bound of the
 * Lanczos tridiagonal.  As in the reference, u IS ADVANCED by those CG steps
 * (cg_eigs works on vecs->u in place, :206) unless use_zero_guess is set, in
 * which case a scratch zero vector is iterated and u is untouched
 * (d4est_solver_multigrid_smoother_cheby.c:281-305).  u, rhs host or device. */
double d4b_cg_eigs(d4b_mesh_t *mesh, double *u, const double *rhs,
                   int32_t imax, int32_t use_new, int32_t use_zero_guess);

/* Chebyshev recurrence (…smoother_cheby.c:81-176) with spectrum [lmin,lmax]:
 * iter sweeps on A u = rhs, then r = rhs - A u.  u is updated in place, r is
 * written.  Host or device pointers.  The apply and the recurrence are fused
 * on the device: each iteration streams u, p, rhs once.                    */
void d4b_cheby_smooth(d4b_mesh_t *mesh, double *u, const double *rhs, double *r,
                      int32_t iter, double lmin, double lmax);

/* ------------------------------------------------------------- multigrid  */
/* Device-resident V-cycle, replaces d4est_solver_multigrid_vcycle / _solve
 * (src/Solver/d4est_solver_multigrid.c:751-1349, :1420-1506) for a hierarchy
 * the caller describes level by level.  Level numbering follows the
 * reference: level 0 is the coarsest, n_levels-1 the finest ("toplevel").
 * The reference creates level l-1 from level l with p4est_coarsen_ext +
 * p4est_balance_ext and records, per coarse element, which fine elements it
 * came from (d4est_solver_multigrid_refine_data_t, d4est_solver_multigrid.h:12-18,
 * callbacks.h:23-108); a family is contiguous in the fine element order, so
 * that record is (first child, child count):
 *   count 8: h-coarsened family in z-order, coarse degree = min of the children
 *            -> restriction = apply_hp_prolong_transpose, prolongation = apply_hp_prolong
 *   count 1: same element (possibly lower degree: p-coarsening, or an element
 *            re-split by the balance step)
 *            -> apply_p_prolong_transpose / apply_p_prolong (d4est_operators.c:1091-1132, :1689-1749) */
typedef struct d4b_mg_desc
{
  int32_t n_levels;
  d4b_mesh_t *const *meshes;             /* [n_levels], coarsest first; not owned  */
  const int32_t *const *child_first;     /* [n_levels-1][n_local of level l]: first element of level l+1 */
  const int32_t *const *child_count;     /* [n_levels-1][n_local of level l]: 1 or 8 */
  /* [multigrid] section, d4est_solver_multigrid.c:392-405 */
  int32_t vcycle_imax;
  double  vcycle_rtol, vcycle_atol;
  /* [mg_smoother_cheby] */
  d4b_cheby_params_t smoother;
  /* bottom solver: 0 = mg_bottom_solver_cg (…bottom_solver_cg.c:48-200),
   *                1 = mg_bottom_solver_cheby (…bottom_solver_cheby.c:60-115)  */
  int32_t bottom_solver;
  int32_t bottom_iter;
  double  bottom_rtol, bottom_atol;
  d4b_cheby_params_t bottom_cheby;
} d4b_mg_desc_t;

typedef struct d4b_mg d4b_mg_t;
d4b_mg_t *d4b_mg_create(const d4b_mg_desc_t *desc);
void      d4b_mg_destroy(d4b_mg_t *mg);
/* one V-cycle on A u = rhs of the finest level (u updated in place; host or
 * device pointers); returns the global |rhs - A u|^2 after it                */
double    d4b_mg_vcycle(d4b_mg_t *mg, double *u, const double *rhs);
/* d4est_solver_multigrid_solve: V-cycles until vcycle_imax, the residual
 * tolerance, or stagnation (sqrt(r2/r2_last) >= 0.99).  Returns the number of
 * V-cycles; r2_history (may be NULL) receives the initial and every
 * subsequent global squared residual, [vcycle_imax + 1].                     */
int32_t   d4b_mg_solve(d4b_mg_t *mg, double *u, const double *rhs, double *r2_history);
/* largest-eigenvalue estimates the smoother used last, [n_levels]           */
void      d4b_mg_get_eigs(const d4b_mg_t *mg, double *eigs_out);
/* V-cycles that were replayed from a recorded CUDA graph instead of being enqueued kernel by kernel (a cycle is
 * recorded the second time it is asked for with the same vectors; device plumbing of d4est_solver_multigrid_vcycle,
 * src/Solver/d4est_solver_multigrid.c:751-1349; D4B_MG_GRAPH=0 / D4B_MG_GRAPH_MULTI=0 switch it off for all /
 * multi-rank hierarchies)                                                                                     */
int64_t   d4b_mg_graph_replays(const d4b_mg_t *mg);
/* inter-grid transfer alone (tests): coarse level l <- / -> fine level l+1;
 * device or host pointers                                                    */
void      d4b_mg_restrict(d4b_mg_t *mg, int32_t coarse_level, const double *fine, double *coarse);
void      d4b_mg_prolong(d4b_mg_t *mg, int32_t coarse_level, const double *coarse, double *fine);

/* ----------------------------------------------------------- outer Krylov */
/* Device-resident versions of the reference's in-house Krylov solvers (the PETSc KSP wrappers
 * stay on the host and reach the device through apply_lhs / pc_apply).  u: initial guess in,
 * solution out; host or device pointers; return value = iterations performed.
 *
 * d4b_cg_solve:  d4est_solver_cg_solve (src/Solver/d4est_solver_cg.c:75-197), unpreconditioned CG,
 *   runs while i < imax and |r|^2 > atol^2 + |r_0|^2 rtol^2.
 * d4b_fcg_solve: d4est_solver_fcg_solve (src/Solver/d4est_solver_fcg_improved.c:103-346), flexible
 *   CG with tol = atol + rtol |r_0|; pc = the V-cycle hierarchy whose finest mesh is `mesh`
 *   (each application = d4est_krylov_pc_multigrid_apply: zero guess, d4b_mg_solve), or NULL
 *   for the identity (do_not_use_preconditioner = 1).  r2_history (may be NULL) receives
 *   |r_k|^2 at the start of every iteration, [imax].
 *   flags = 0 reproduces what the reference executes: its direction updates are written
 *   d4est_linalg_vec_axpyeqz(c, d, v, d) (:260-262) and axpyeqz copies y into z before adding c x
 *   (src/LinearAlgebra/d4est_linalg.c:217-221), so with z aliasing x they yield d_k = (1 + c) v_k,
 *   q_k = (1 + c) w_k.  Iterates and iteration counts match the reference only with that
 *   arithmetic.  D4B_FCG_TEXTBOOK selects the recurrence the reference's comments state
 *   (d_k = v_k + c d_{k-1}, Notay & Napov).                                               */
#define D4B_FCG_TEXTBOOK 1
int32_t d4b_cg_solve(d4b_mesh_t *mesh, double *u, const double *rhs, int32_t imax, double rtol, double atol);
int32_t d4b_fcg_solve(d4b_mesh_t *mesh, d4b_mg_t *pc, double *u, const double *rhs, int32_t imax,
                      double rtol, double atol, int32_t flags, double *r2_history);

/* ------------------------------------------- nonlinear mass-type terms (Newton)  */
/* The shipped nonlinear problems are  -lap u + f(x, u) = 0  with a point-wise f; the reference
 * integrates f and f_u against the basis element by element on the host,
 *   residual:  d4est_quadrature_apply_fofufofvlj   (src/Quadrature/d4est_quadrature.c:776-936)
 *   Jacobian:  d4est_quadrature_apply_fofufofvlilj (:593-774, QUAD_APPLY_MATRIX)
 * with f given as a d4est_xyzu_fcn_t callback.  Every shipped f has the form
 *       f(x, u) = a(x) (b(x) + u)^k ,   f_u = k a(x) (b(x) + u)^(k-1) ,   k an integer:
 *   two punctures     a = -K^2(x) / 8,  b = 1 + sum_n m_n / (2 r_n),  k = -7
 *                     (multi_punctures_neg_1o8_K2_psi_neg7 / _plus_7o8_K2_psi_neg8,
 *                      src/Problems/TwoPunctures/multi_punctures_fcns_with_opt.h:292-423)
 *   constant density  a = -2 pi rho(x), b = 0, k = 5, rho = rho0 inside radius R of c, else 0
 *   star              (neg_2pi_rho_up1_neg5 / neg_10pi_rho_up1_neg4,
 *                      src/Problems/ConstantDensityStar/constant_density_star_fcns.h:334-360)
 * These are DEVICE FUNCTORS here: a(x), b(x) are evaluated once per mesh by a kernel at the
 * quadrature points, (b + V u)^k inside the residual / linearisation kernels; no callback is
 * evaluated on the host and only u crosses the bus.                                              */
typedef enum {
  D4B_NL_PUNCTURES = 1,              /* a, b from the puncture list                              */
  D4B_NL_CONSTANT_DENSITY_STAR = 2,  /* a from the density step                                  */
  D4B_NL_TABLE = 3                   /* a_quad, b_quad, exponent supplied by the caller           */
} d4b_nonlinear_kind_t;
#define D4B_MAX_PUNCTURES 10
typedef struct d4b_nonlinear_desc
{
  int32_t kind;
  int32_t n_punctures;
  double  punctures[D4B_MAX_PUNCTURES][10];  /* M, X, Y, Z, PX, PY, PZ, SX, SY, SZ (black_hole_data_t) */
  double  star_rho0, star_R, star_c[3];
  int32_t exponent;                  /* D4B_NL_TABLE only                                         */
  const double *a_quad, *b_quad;     /* D4B_NL_TABLE only, [local_nodes_quad]                     */
  const double *xyz_quad;            /* [3*local_nodes_quad] d4est_mesh_data_t.xyz_quad; NULL for a mesh
                                        created by d4b_mesh_create_with_geometry (positions kept)  */
  const double *jac_quad;            /* [local_nodes_quad] J_quad; NULL likewise                  */
} d4b_nonlinear_desc_t;
/* attach f to the mesh (desc == NULL removes it); host or device arrays                          */
void d4b_mesh_set_nonlinear_term(d4b_mesh_t *mesh, const d4b_nonlinear_desc_t *desc);
/* Newton linearisation about u0: from now on every apply of `mesh` (d4b_apply, smoother, Krylov)
 * is the Jacobian  A + V^T diag(w J f_u(x, V u0)) V  -- what multi_punctures_apply_jac computes
 * (multi_punctures_fcns_with_opt.h:546-624).  u0 host or device; u0 == NULL drops the term.       */
void d4b_mesh_linearize(d4b_mesh_t *mesh, const double *u0);
/* the build_residual callback of d4est_elliptic_eqns_t (src/EllipticSystem/d4est_elliptic_eqns.h:26-37)
 * on the device:  F = A(u) + V^T (w J f(x, V u)) - rhs   with A(u) the operator of `mesh` INCLUDING its
 * boundary data, the middle term only when a nonlinear term is attached, rhs == NULL meaning 0:
 *   Poisson  F = A u - rhs            (poisson_build_residual, src/Problems/Poisson/poisson_sinx_fcns.h:133-149)
 *   punctures F = A u + N(u)          (multi_punctures_build_residual, ...with_opt.h:706-750)
 * Any linearisation of `mesh` is ignored here.  Host or device pointers; stream-ordered when all are device
 * pointers (d4b_sync to wait), synchronous otherwise.                                               */
void d4b_build_residual(d4b_mesh_t *mesh, const double *u, const double *rhs, double *F);

/* The reference's multigrid "matrix operator" for such Jacobians
 * (src/Solver/d4est_solver_multigrid_matrix_operator.c:5-47, :170-250): below the finest level the mass-type
 * term is meant to be the Galerkin product  M_l = P^T M_{l+1} P  of the finest level's V^T diag(W) V, element
 * family by element family, stored as dense matrices.  Applied here matrix-free on every coarse level of `mg`
 * whenever the finest mesh is linearised (prolong to the finest level, weight, come back down):
 *   mode 0 (default): coarse levels are plain Laplacians
 *   mode 1: what the reference EXECUTES.  d4est_operators_compute_PT_mat_P (src/dGMath/d4est_operators.c:608-661)
 *           transposes the stacked prolongation of a family as a whole and then uses a contiguous slice of that
 *           transpose as "P_i^T" (:638, :650), which is the transpose only for a single child (p-coarsening);
 *           for an h-coarsened family of 8 the slice is another matrix Q_i, so M_l = sum_i Q_i M_i P_i.
 *           And its fine matrices are built with the arguments of d4est_quadrature_apply_fofufofvlilj shifted by
 *           one (d4est_solver_multigrid_matrix_operator.c:215-240: u0 lands in the unused `vec` slot, NULL in `u`),
 *           so f_u is evaluated at u = 0 for every linearisation point: W = w J f_u(x, 0) on the coarse levels.
 *           Needs the nonlinear term on the finest mesh.  Iterates and iteration counts of the reference's Newton-Krylov runs are reproduced only with this.
 *   mode 2: the Galerkin product the reference's comments describe.                                          */
#define D4B_COARSE_MASS_OFF 0
#define D4B_COARSE_MASS_REFERENCE 1
#define D4B_COARSE_MASS_GALERKIN 2
void d4b_mg_set_coarse_mass(d4b_mg_t *mg, int32_t mode);

/* d4est_solver_newton_solve (src/Solver/d4est_solver_newton.c:136-367) with d4est_solver_fcg_solve as the
 * Krylov callback and the V-cycle hierarchy `pc` (may be NULL) as preconditioner, device-resident:
 *   F_0 = F(u); while (|F| > atol + rtol |F_0| or it < imin) and it < imax:
 *     linearise about u; solve J step = -F with FCG from a zero guess; u += step; F = F(u).
 * mesh_res supplies the residual (its boundary data included), mesh_jac the Jacobian (boundary data zeroed,
 * as the reference asks of flux_data_for_jac); they may be the same handle.  Both need the nonlinear term.
 * u: initial guess in, solution out (host or device).  Returns 0 when converged, 1 otherwise (ierr).    */
typedef struct d4b_newton_params
{
  double  atol, rtol;                /* [d4est_solver_newton]                                     */
  int32_t imin, imax;
  double  fcg_atol, fcg_rtol;        /* [d4est_solver_fcg]                                        */
  int32_t fcg_imax, fcg_flags;
} d4b_newton_params_t;
typedef struct d4b_newton_stats
{
  int32_t steps;                     /* Newton steps taken                                        */
  int32_t fcg_iterations[64];        /* per step                                                  */
  double  residual_norm[65];         /* |F| before the first and after every step                 */
  int64_t jacobian_applies;          /* fine-level operator applies inside FCG (not the V-cycles)  */
} d4b_newton_stats_t;
int32_t d4b_newton_solve(d4b_mesh_t *mesh_res, d4b_mesh_t *mesh_jac, d4b_mg_t *pc, double *u,
                         const d4b_newton_params_t *params, d4b_newton_stats_t *stats);

#ifdef __cplusplus
}
#endif
#endif /* D4EST_B200_H */
