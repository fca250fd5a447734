cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_ppo_cuda.py tests/test_graph_cuda.py -m gpu -q 2>&1 | tail -8
