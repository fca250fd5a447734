cd $GRAFT_REPO_ROOT
URLG_WGRAD_SHARE=2 timeout 600 ncu --profile-from-start off --cache-control none --set full --import-source on --clock-control none -k regex:grads_finalize -c 2 -f -o gpurun_out/p7_finalize python tools/lstm_update_profile.py go2 4096 1 > gpurun_out/p7_ncu.log 2>&1
echo done
