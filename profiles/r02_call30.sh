cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
D4B_MG_DEBUG=1 timeout 600 python -m pytest tests/test_multi_rank.py -m gpu -q -x -k vcycle_graph 2>&1 | grep -v "^$" | grep "d4b mg\|passed\|failed\|Error" | head -60
