cd $GRAFT_REPO_ROOT
export MASTER_ADDR=127.0.0.1
timeout 300 python -m pytest tests/test_peer_exchange_cuda.py -m gpu -x -q 2>&1 | tail -15
URLG_PARITY_NLOC=4096 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_parity.py > gpurun_out/r2_multi_gpu_parity_n2_4096_peer.json 2> gpurun_out/p14_parity.err; tail -c 900 gpurun_out/r2_multi_gpu_parity_n2_4096_peer.json; grep -E "Error|error" gpurun_out/p14_parity.err | head -5
for x in 0 1; do
URLG_PEER_FUSED=$x timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$x bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --no-eager-gpu --no-configs --no-large > gpurun_out/p14_bench2_f$x.json 2> gpurun_out/p14_bench2_f$x.err; python -c "
import json; d=json.loads(open('gpurun_out/p14_bench2_f$x.json').read().strip().splitlines()[-1]); print('fused=$x', d['run']['gradient_exchange'][:40], d['update_ms'], d['ppo_iter_ms'], d['value'])"; grep -E "Error|error" gpurun_out/p14_bench2_f$x.err | head -3
done
