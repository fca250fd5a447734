cd $GRAFT_REPO_ROOT
export MASTER_ADDR=127.0.0.1
timeout 300 python -m pytest tests/test_peer_exchange_cuda.py -m gpu -x -q 2>&1 | tail -4
for m in 1 0; do
URLG_PEER_MODE=$m URLG_PARITY_NLOC=4096 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2955$m tools/multi_gpu_parity.py > gpurun_out/p22_parity_m$m.json 2> gpurun_out/p22_parity_m$m.err; python -c "
import json; d=json.loads(open('gpurun_out/p22_parity_m$m.json').read().strip().splitlines()[-1]); print('mode=$m', [(g['vs_eager_max_abs']) for g in d['update_graph_with_nccl']], d['ranks_bit_identical'], d['pass'], d['update_param_max_abs'])"; grep -E "Error|error" gpurun_out/p22_parity_m$m.err | head -3
URLG_PEER_MODE=$m timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2956$m bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --no-eager-gpu --no-configs --no-large > gpurun_out/p22_bench2_m$m.json 2> gpurun_out/p22_bench2_m$m.err; python -c "
import json; d=json.loads(open('gpurun_out/p22_bench2_m$m.json').read().strip().splitlines()[-1]); print('mode=$m', d['update_ms'], d['ppo_iter_ms'])"
done
