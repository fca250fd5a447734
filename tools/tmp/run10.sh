cd $GRAFT_REPO_ROOT
export URLG_WGRAD_SHARE=2
BF="--steps 10 --warmup 3 --no-cpu --no-large --no-configs --no-eager-gpu"
for a in 0 1 0 1; do echo adam_images=$a; URLG_ADAM_IMAGES=$a python bench.py $BF 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['update_ms'], d['rollout_ms'], d['ppo_iter_ms'], d['roofline']['frac'])"; done
timeout 600 ncu --profile-from-start off --cache-control none --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/p6_launches_update_warm.csv python tools/lstm_update_profile.py go2 4096 1 > gpurun_out/p6_ncu.log 2>&1
grep -E "finalize" gpurun_out/p6_launches_update_warm.csv | head -2 | awk -F'","' '{print $5, $NF}'
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/p10_launches_update_g1.csv python tools/lstm_update_profile.py g1 4096 1 > gpurun_out/p10_ncu_g1.log 2>&1
tail -2 gpurun_out/p10_ncu_g1.log
