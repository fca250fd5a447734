set -x
cd $GRAFT_REPO_ROOT
BF="--steps 2 --warmup 3 --no-cpu --no-large --no-configs --no-eager-gpu"
python bench.py $BF > gpurun_out/p1_bench_plain.json 2> gpurun_out/p1_bench_plain.err || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py $BF > gpurun_out/p1_ncu_bench.log 2>&1
python tools/k2_bench.py --sizes 262144 --once > gpurun_out/p1_k2_once.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:post_physics -s 4 -c 2 -f -o gpurun_out/r2_k2_v7_262144 python tools/k2_bench.py --sizes 262144 --once > gpurun_out/p1_ncu_k2.log 2>&1
python tools/lstm_update_profile.py go2 4096 1 > gpurun_out/p1_upd_plain.log 2>&1 || exit 1
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_update_go2.csv python tools/lstm_update_profile.py go2 4096 1 > gpurun_out/p1_ncu_upd.log 2>&1
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -c 14 -f -o gpurun_out/r2_update_minibatch_go2 python tools/lstm_update_profile.py go2 4096 1 > gpurun_out/p1_ncu_upd_full.log 2>&1
tail -3 gpurun_out/p1_upd_plain.log
ls -la gpurun_out | tail -12
