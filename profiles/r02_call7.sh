cd $GRAFT_REPO_ROOT
for aff in 0 1; do for bulk in 0 1; do
  echo "AFFINE=$aff BULK=$bulk"; D4B_AFFINE=$aff D4B_FLAT_BULK=$bulk python profiles/flat_bench.py 6 3 2>&1 | tail -1
done; done
for bulk in 0 1; do echo "AFFINE=0 BULK=$bulk 32^3"; D4B_AFFINE=0 D4B_FLAT_BULK=$bulk python profiles/flat_bench.py 5 1 2 4 5 6 7 2>&1 | tail -6; done
python -m pytest tests/test_gpu_apply.py -q -x 2>&1 | tail -2
D4B_FLAT_BULK=1 D4B_AFFINE=0 python -m pytest tests/test_gpu_apply.py -q -x 2>&1 | tail -2
