cd $GRAFT_REPO_ROOT
for s in 2; do echo share=$s; URLG_WGRAD_SHARE=$s python tools/lstm_update_profile.py go2 4096 1 2>&1 | tail -2; done
URLG_WGRAD_SHARE=2 timeout 600 ncu --profile-from-start off --cache-control none --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/p6_launches_update_warm.csv python tools/lstm_update_profile.py go2 4096 1 > gpurun_out/p6_ncu.log 2>&1
grep finalize gpurun_out/p6_launches_update_warm.csv | head -3 | awk -F'","' '{print $NF}'
