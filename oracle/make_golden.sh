#!/bin/bash
# make_golden.sh -- regenerate tests/golden/*.d4b.gz from the UNMODIFIED
# reference (oracle/_ref/ref_driver, built by oracle/build_ref.sh).
# TEST INFRASTRUCTURE.  Each fixture holds the flattened mesh description, the
# seeded input vector u (srand(102321), as d4est_test_laplacian_speedup.c:429)
# and the reference's outputs Au_plain / Au_with_opt (+ its 1-D tables).
# --quad adds the reference's volume quadrature operators (interpolate, Galerkin
# integral, mass matrix, fofufofvlj / fofufofvlilj with a star-like f).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
DRV="$HERE/_ref/ref_driver"
G="$HERE/../tests/golden"
mkdir -p "$G"
run() { # name ini args...
  local name=$1 ini=$2; shift 2
  "$DRV" "$HERE/configs/$ini" "$G/$name.d4b" "$@" > "$G/$name.json"
  gzip -9nf "$G/$name.d4b"
  echo "$name: $(cat "$G/$name.json")"
}
geom() { # name ini args...: sidecar for the device geometry generator (element positions, map
  # parameters, hm/hp under every face_h_type); the main dump of the same run is discarded
  local name=$1 ini=$2; shift 2
  "$DRV" "$HERE/configs/$ini" /tmp/d4b_geom_$$.d4b "$@" --geom "$G/$name.geom.d4b" > /dev/null
  rm -f /tmp/d4b_geom_$$.d4b
  gzip -9nf "$G/$name.geom.d4b"
}
variant() { # new.ini base.ini sed-expression...: derived configuration under /tmp (penalty function / multipliers)
  local new=$1 base=$2; shift 2
  sed "$@" "$HERE/configs/$base" > "$HERE/configs/$new"
}
geom brick_l2_p3 brick_poisson.ini --uniform-deg 3
geom sphere_p2 sphere_p.ini --uniform-deg 2
geom sphere_hp_amr1 sphere_hp.ini --amr 1 --bc poly
run brick_l2_p1 brick_poisson.ini                       # the 4^3 p=1 mesh of d4est_test_poisson_sinx.sh
run brick_l2_p2_bc brick_poisson.ini --uniform-deg 2 --bc poly
run brick_l2_p3 brick_poisson.ini --uniform-deg 3 --cheby 15 15 --quad   # + volume quadrature operators; + cg_eigs / Chebyshev smoother outputs
run brick_l2_p5 brick_poisson.ini --uniform-deg 5
run brick_l2_p7 brick_poisson.ini --uniform-deg 7
run sphere_p2 sphere_p.ini --uniform-deg 2 --cheby 8 10 --quad # conforming, 13 trees, all face orientations; smoother outputs
run sphere_hp_amr0 sphere_hp.ini --bc poly               # per-region degrees + quadrature increments
run sphere_hp_amr1 sphere_hp.ini --amr 1 --bc poly       # random hp refinement: hanging faces
run sphere_p_amr1 sphere_p.ini --amr 1 --bc poly           # random hp refinement with deg_quad = deg: the hp mortar kernel's mesh
run brick_l2_p2_robin brick_poisson.ini --uniform-deg 2 --bc robin   # Robin boundary (coeff 1/|x|-like, non-zero rhs)
run sphere_p_amr1_robin sphere_p.ini --amr 1 --bc robin               # Robin on the outer sphere, random hp refinement (two-puncture-like)
run sphere_hp_amr1_zero sphere_hp.ini --amr 1 --cheby 8 10 --quad  # hanging faces, homogeneous BC, smoother outputs
run brick_mg_l2_p2 brick_poisson_mg.ini --uniform-deg 2 --mg 4 --fcg 30   # + in-house FCG (multigrid-preconditioned and not) and CG: iterates, iteration counts; + the reference multigrid: residual and iterate after each of 4 V-cycles
# the other three penalty functions (d4est_laplacian_flux_sipg.c:945-1005; meanp_sqr_over_meanh is what
# src/Problems/Okendon/options.input and BoyenYorkModel/options.input select) and region multipliers != 1
# (:1165-1199), on the 13-tree sphere with per-region degrees 1 / 2 / 1 (p-jumps and h-jumps on the region
# interfaces, so the four functions and the averaged multipliers all differ) and inhomogeneous Dirichlet data
variant _pen_maxpp1.ini sphere_p.ini -e 's/^sipg_penalty_fcn = .*/sipg_penalty_fcn = maxpp1_sqr_over_minh/'
variant _pen_meanp_mult.ini sphere_p.ini -e 's/^sipg_penalty_fcn = .*/sipg_penalty_fcn = meanp_sqr_over_meanh\nsipg_use_region_multiplier = 1\nsipg_use_region_multipliers = 1\nsipg_number_of_regions = 3\nregion0_multiplier = 0.7\nregion1_multiplier = 1.9\nregion2_multiplier = 3.1\nsipg_region0_multiplier = 0.7\nsipg_region1_multiplier = 1.9\nsipg_region2_multiplier = 3.1/'
variant _pen_meanph.ini sphere_p.ini -e 's/^sipg_penalty_fcn = .*/sipg_penalty_fcn = mean_p_sqr_over_h/'
run sphere_p_pen_maxpp1 _pen_maxpp1.ini --bc poly --no-opt
run sphere_p_pen_meanp_mult _pen_meanp_mult.ini --bc poly
run sphere_p_pen_meanph _pen_meanph.ini --bc poly
rm -f "$HERE"/configs/_pen_*.ini
# BASELINE configs[4] in small: two punctures on the 13-tree sphere (Robin outer boundary), the reference's in-house
# Newton -> FCG -> multigrid (matrix operator for the mass-type term): Newton steps, FCG iterations per step,
# residual norms, final u; residual F(u) and Jacobian J(u0) v samples with the unmodified puncture callbacks
"$DRV" "$HERE/configs/sphere_punctures.ini" "$G/sphere_punctures_newton.d4b" --newton | tail -1 > "$G/sphere_punctures_newton.json"
gzip -9nf "$G/sphere_punctures_newton.d4b"
echo "sphere_punctures_newton: $(cat "$G/sphere_punctures_newton.json")"
# the bench V-cycle's exact configuration: Chebyshev bottom solver (d4est_solver_multigrid_bottom_solver_cheby.c:60-115)
variant _mg_cheby.ini brick_poisson_mg.ini -e 's/^bottom_solver_name = .*/bottom_solver_name = mg_bottom_solver_cheby/'
cat >> "$HERE/configs/_mg_cheby.ini" <<'INI'

[mg_bottom_solver_cheby]
cheby_imax = 15
cheby_eigs_cg_imax = 15
cheby_eigs_lmax_lmin_ratio = 30.
cheby_eigs_max_multiplier = 1.
cheby_print_residual_norm = 0
cheby_print_spectral_bound = 0
cheby_print_spectral_bound_iterations = 0
INI
run brick_mg_l2_p2_cheby _mg_cheby.ini --uniform-deg 2 --mg 4 --fcg 30
rm -f "$HERE"/configs/_mg_cheby.ini
