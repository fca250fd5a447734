cd $GRAFT_REPO_ROOT
URLG_WGRAD_SHARE=2 timeout 600 ncu --profile-from-start off --cache-control none --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/p5_launches_update_warm.csv python tools/lstm_update_profile.py go2 4096 1 > gpurun_out/p5_ncu.log 2>&1
echo done
