cd $GRAFT_REPO_ROOT
(timeout 900 python -m pytest tests/test_gpu_apply.py tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -3)
