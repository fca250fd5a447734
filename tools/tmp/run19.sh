cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_v2.txt 2>&1; tail -3 gpurun_out/r2_pytest_gpu_v2.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python bench.py > gpurun_out/r2_bench_go2_n1_v2.json 2> gpurun_out/p19_bench.err; tail -2 gpurun_out/p19_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_go2_reference_arm.json 2> gpurun_out/p19_ref.err; tail -c 600 gpurun_out/r2_bench_go2_reference_arm.json
BF="--steps 2 --warmup 3 --no-cpu --no-large --no-configs --no-eager-gpu"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r2_launches_bench_v2.csv python bench.py $BF > gpurun_out/p19_ncu_bench.log 2>&1
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_update_go2_v2.csv python tools/lstm_update_profile.py go2 4096 1 > gpurun_out/p19_ncu_upd.log 2>&1
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"gemm_bf16|mlp_fused|grads_finalize" -c 6 -f -o gpurun_out/r2_tensor_minibatch_v2 python tools/lstm_update_profile.py go2 4096 1 > gpurun_out/p19_ncu_full.log 2>&1
ls -la gpurun_out | tail -5
