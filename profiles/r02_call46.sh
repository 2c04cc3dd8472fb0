cd $GRAFT_REPO_ROOT
python profiles/newton_vcycle_breakdown.py 4 2>&1 | tail -6
