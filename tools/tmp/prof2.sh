cd $GRAFT_REPO_ROOT
python tools/lstm_update_profile.py go2 4096 1 2>&1 | tail -2
URLG_EXP_NOFX=1 python tools/lstm_update_profile.py go2 4096 1 2>&1 | tail -2
URLG_EXP_NOFX=1 timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/p2_launches_update_nofx.csv python tools/lstm_update_profile.py go2 4096 1 > gpurun_out/p2_ncu_upd.log 2>&1
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_bf16 -c 4 -f -o gpurun_out/r2_bwd_minibatch_go2 python tools/lstm_update_profile.py go2 4096 1 > gpurun_out/p2_ncu_upd_full.log 2>&1
ls -la gpurun_out | tail -4
