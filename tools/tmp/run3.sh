cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gemm_cuda.py tests/test_ppo_cuda.py tests/test_graph_cuda.py -m gpu -x -q 2>&1 | tail -5
for s in 1 2 4; do echo share=$s; URLG_WGRAD_SHARE=$s python tools/lstm_update_profile.py go2 4096 1 2>&1 | tail -2; done
URLG_WGRAD_SHARE=1 timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/p3_launches_update_share1.csv python tools/lstm_update_profile.py go2 4096 1 > gpurun_out/p3_ncu1.log 2>&1
URLG_WGRAD_SHARE=2 timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/p3_launches_update_share2.csv python tools/lstm_update_profile.py go2 4096 1 > gpurun_out/p3_ncu2.log 2>&1
python tools/lstm_update_profile.py g1 4096 1 2>&1 | tail -2
