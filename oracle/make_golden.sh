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
