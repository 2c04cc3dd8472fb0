#!/bin/bash
# build_ref.sh -- compile the hot-path subset of the UNMODIFIED reference
# (/root/reference) serially (no MPI, no PETSc, no HDF5) into oracle/_ref/.
#
# TEST INFRASTRUCTURE.  Sources are compiled where they lie; nothing from
# /root/reference is copied into the repository.  Outputs (objects, the
# extracted third-party tarballs, binaries) go to oracle/_ref/ only, which is
# git-ignored but travels to the GPU box.
#
# The reference's own build system (CMake + bundled ExternalProjects) is NOT
# run: it requires MPI, PETSc, HDF5 and OpenBLAS, three of whose tarballs are
# absent.  Instead: libsc 2.8.3 / p4est 2.8 / zlog 1.2.15 (tarballs under
# /root/reference/third_party) are compiled with hand-written config headers
# (oracle/refbuild/*.h, no-MPI mode), BLAS/LAPACK symbols come from
# oracle/refbuild/miniblas.c, and the PETSc/HDF5-only files are left out
# (shim headers in oracle/refbuild/shim give their types only).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${D4B_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
B="$OUT/build"
JOBS="${JOBS:-8}"
if [ ! -d "$REF/src" ]; then
  echo "build_ref.sh: $REF not present; keeping prebuilt oracle/_ref as is"
  exit 0
fi
mkdir -p "$B/obj/sc" "$B/obj/p4" "$B/obj/zlog" "$B/obj/d8"
cd "$B"
[ -d libsc-2.8.3 ] || tar xzf "$REF/third_party/libsc-2.8.3.tar.gz"
[ -d p4est-2.8 ] || tar xzf "$REF/third_party/p4est-2.8.tar.gz"
[ -d zlog-1.2.15 ] || tar xzf "$REF/third_party/zlog-1.2.15.tar.gz"

OPT="${REF_CFLAGS:--O2}"
TPINC="-I$HERE/refbuild -Ilibsc-2.8.3/src -Ilibsc-2.8.3/iniparser -Ilibsc-2.8.3/libb64 -Ip4est-2.8/src -Izlog-1.2.15/src"
CMDS="$B/cmds.txt"
: > "$CMDS"
emit() { # src obj flags
  if [ ! -f "$2" ] || [ "$1" -nt "$2" ]; then echo "gcc $3 -c $1 -o $2" >> "$CMDS"; fi
}
for f in libsc-2.8.3/src/*.c libsc-2.8.3/iniparser/dictionary.c libsc-2.8.3/iniparser/iniparser.c \
         libsc-2.8.3/iniparser/inistring.c libsc-2.8.3/libb64/cdecode.c libsc-2.8.3/libb64/cencode.c; do
  case $f in *sc_v4l2*|*sc_getopt*) continue;; esac
  emit $f obj/sc/$(basename $f .c).o "$OPT -w -fPIC $TPINC"
done
for f in p4est-2.8/src/p4est*.c p4est-2.8/src/p8est*.c; do
  case $f in *empty*) continue;; esac
  emit $f obj/p4/$(basename $f .c).o "$OPT -w -fPIC $TPINC"
done
for f in zlog-1.2.15/src/*.c; do
  case $f in *zlog-chk-conf*) continue;; esac
  emit $f obj/zlog/$(basename $f .c).o "$OPT -w -fPIC -Izlog-1.2.15/src"
done

DINC=""
for d in EllipticSystem Geometry Solver Estimators dGMath LinearAlgebra Problems Utilities pXest hpAMR Mesh IO Quadrature Kron Databox Dictionary; do
  DINC="$DINC -I$REF/src/$d"
done
DFLAGS="-std=gnu99 $OPT -w -fPIC -DDIM=3 -DNDEBUG $DINC -I$REF/third_party/mpsort -I$REF/third_party/PPF -I$HERE/refbuild/shim $TPINC"
# the hot-path subset of CMakeLists.txt:77-170 (SRCS + SRCS_3D_ONLY) that does
# not need PETSc / HDF5 / raw MPI
D8SRCS="
Utilities/ini.c Utilities/d4est_util.c LinearAlgebra/d4est_linalg.c
EllipticSystem/d4est_elliptic_eqns.c EllipticSystem/d4est_elliptic_data.c
dGMath/d4est_operators.c dGMath/d4est_gradient.c dGMath/d4est_lgl.c dGMath/d4est_reference.c
dGMath/d4est_laplacian.c dGMath/d4est_laplacian_flux.c dGMath/d4est_laplacian_flux_sipg.c
dGMath/d4est_laplacian_flux_sipg_penalty_debugger.c
Mesh/d4est_mortars_with_opt.c dGMath/d4est_laplacian_with_opt.c dGMath/d4est_laplacian_with_opt_flux.c
dGMath/d4est_laplacian_with_opt_flux_sipg.c
Geometry/d4est_geometry_brick.c Geometry/d4est_geometry.c Geometry/d4est_geometry_cubed_sphere.c
Geometry/d4est_connectivity_cubed_sphere.c Geometry/d4est_geometry_general_wedge.c
Geometry/d4est_geometry_hole_in_a_box.c
hpAMR/d4est_amr.c hpAMR/d4est_amr_uniform.c hpAMR/d4est_amr_random.c
Mesh/d4est_element_data.c Mesh/d4est_mesh.c Mesh/d4est_xyz_functions.c Mesh/d4est_mortars.c
Mesh/d4est_mortars_aux.c Mesh/d4est_ghost_data.c Mesh/d4est_ghost.c
Quadrature/d4est_quadrature.c Quadrature/d4est_quadrature_legendre.c Quadrature/d4est_quadrature_lobatto.c
Solver/d4est_solver_multigrid.c Solver/d4est_solver_multigrid_bottom_solver_cg.c
Solver/d4est_solver_multigrid_bottom_solver_cheby.c Solver/d4est_solver_multigrid_element_data_updater.c
Solver/d4est_solver_multigrid_logger_residual.c Solver/d4est_solver_multigrid_profiler_basic.c
Solver/d4est_solver_multigrid_mesh_analyzer.c Solver/d4est_solver_multigrid_smoother_cheby.c
Solver/d4est_solver_multigrid_smoother_none.c Solver/d4est_solver_multigrid_bottom_solver_reuse_smoother.c
Solver/d4est_solver_cg_eigs.c Solver/d4est_krylov_pc_multigrid.c
Solver/d4est_solver_fcg_improved.c Solver/d4est_solver_cg.c
Solver/d4est_solver_newton.c Solver/d4est_solver_multigrid_matrix_operator.c
Estimators/d4est_estimator_bi.c
"
for s in $D8SRCS; do
  emit $REF/src/$s obj/d8/$(basename $s .c).o "$DFLAGS"
done
emit $HERE/refbuild/miniblas.c obj/d8/miniblas.o "$OPT -fPIC"
emit $HERE/refbuild/ref_stubs.c obj/d8/ref_stubs.o "$DFLAGS"
emit $HERE/../disco4est_b200/adapter/d4est_b200_adapter.c obj/d8/d4est_b200_adapter.o \
     "$DFLAGS -DD4EST_B200_ADAPTER_NO_DEVICE -I$HERE/../include -I$HERE/../disco4est_b200/adapter"
emit $HERE/ref_driver.c obj/d8/ref_driver.o "$DFLAGS -I$REF/src/Problems/TwoPunctures -I$HERE -I$HERE/../include -I$HERE/../disco4est_b200/adapter"
# device-linked variant: the same driver and adapter with the C ABI library behind them
emit $HERE/../disco4est_b200/adapter/d4est_b200_adapter.c obj/d8/d4est_b200_adapter_dev.o \
     "$DFLAGS -I$HERE/../include -I$HERE/../disco4est_b200/adapter"
emit $HERE/ref_driver.c obj/d8/ref_driver_dev.o "$DFLAGS -DD4B_WITH_DEVICE -I$REF/src/Problems/TwoPunctures -I$HERE -I$HERE/../include -I$HERE/../disco4est_b200/adapter"

if [ -s "$CMDS" ]; then
  echo "build_ref.sh: compiling $(wc -l < "$CMDS") files"
  xargs -P "$JOBS" -I{} sh -c '{}' < "$CMDS"
fi
D8OBJS=""
for s in $D8SRCS; do D8OBJS="$D8OBJS obj/d8/$(basename $s .c).o"; done
gcc -o "$OUT/ref_driver" obj/d8/ref_driver.o obj/d8/d4est_b200_adapter.o $D8OBJS \
    obj/d8/miniblas.o obj/d8/ref_stubs.o obj/p4/*.o obj/sc/*.o obj/zlog/*.o -lm -lpthread
echo "build_ref.sh: built $OUT/ref_driver"
if [ -f "$HERE/../disco4est_b200/libd4est_b200.so" ]; then
  gcc -o "$OUT/ref_driver_b200" obj/d8/ref_driver_dev.o obj/d8/d4est_b200_adapter_dev.o $D8OBJS \
      obj/d8/miniblas.o obj/d8/ref_stubs.o obj/p4/*.o obj/sc/*.o obj/zlog/*.o \
      -L"$HERE/../disco4est_b200" -ld4est_b200 -Wl,-rpath,'$ORIGIN/../../disco4est_b200' \
      -Wl,--allow-shlib-undefined -lm -lpthread
  echo "build_ref.sh: built $OUT/ref_driver_b200 (reference + libd4est_b200.so; runs on a GPU box only)"
fi
