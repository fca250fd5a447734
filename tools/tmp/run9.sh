cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gemm_cuda.py tests/test_ppo_cuda.py tests/test_graph_cuda.py tests/test_checkpoint_cuda.py -m gpu -x -q 2>&1 | tail -2
export URLG_WGRAD_SHARE=2
for a in 0 1; do echo adam_images=$a; URLG_ADAM_IMAGES=$a python tools/lstm_update_profile.py go2 4096 1 2>&1 | tail -2; done
timeout 600 ncu --profile-from-start off --cache-control none --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/p6_launches_update_warm.csv python tools/lstm_update_profile.py go2 4096 1 > gpurun_out/p6_ncu.log 2>&1
grep -E "finalize|adam" gpurun_out/p6_launches_update_warm.csv | head -4 | awk -F'","' '{print $5, $NF}'
