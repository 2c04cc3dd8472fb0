#!/bin/bash
# dropin_check.sh -- run the UNMODIFIED reference with eqns->apply_lhs = d4est_b200_apply_lhs
# (oracle/_ref/ref_driver_b200, needs a GPU).  TEST INFRASTRUCTURE.
# 1. one apply through the reference's own call path, compared with d4est_laplacian_apply_aij
# 2. the reference's multigrid solver (its smoother, cg_eigs, bottom CG) driving our apply:
#    residual history and solution against the stock run.
HERE="$(cd "$(dirname "$0")" && pwd)"
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
DRV="$HERE/_ref/ref_driver_b200"
set -e
echo "== brick 4^3 p=3, multigrid 4 V-cycles"
"$DRV" "$HERE/configs/brick_poisson_mg.ini" - --uniform-deg 3 --mg 4 --fcg 30 --b200 | tail -5
echo "== brick 8^3 p=2, multigrid 3 V-cycles"
sed 's/min_level = 2/min_level = 3/' "$HERE/configs/brick_poisson_mg.ini" > /tmp/d4b_mg_l3.ini
"$DRV" /tmp/d4b_mg_l3.ini - --uniform-deg 2 --mg 3 --b200 | tail -3
echo "== cubed sphere, random hp AMR (hanging faces, all orientations): one apply"
"$DRV" "$HERE/configs/sphere_hp.ini" - --amr 1 --b200 | tail -2
echo "== cubed sphere uniform p=2 (curved conforming kernel): one apply"
"$DRV" "$HERE/configs/sphere_p.ini" - --uniform-deg 2 --b200 | tail -2
echo "== cubed sphere, random hp AMR with deg_quad = deg (mortar kernel / face split, fused smoother): preconditioner slot"
"$DRV" "$HERE/configs/sphere_p_mg.ini" - --amr 1 --mg 3 --b200 | grep pc_vcycles
echo "== Robin boundary (two-puncture-like), random hp AMR: one apply"
"$DRV" "$HERE/configs/sphere_p.ini" - --amr 1 --bc robin --b200 | grep b200_apply_rel_l2
echo "== inhomogeneous Dirichlet data, random hp AMR: right-hand side and estimator"
"$DRV" "$HERE/configs/sphere_p.ini" - --amr 1 --bc poly --b200 | grep "build_rhs\|estimator"
