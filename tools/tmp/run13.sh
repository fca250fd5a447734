cd $GRAFT_REPO_ROOT
export MASTER_ADDR=127.0.0.1
timeout 300 python -m pytest tests/test_peer_exchange_cuda.py tests/test_gemm_cuda.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_parity.py > gpurun_out/r2_multi_gpu_parity_n2_128_peer.json 2> gpurun_out/p13_parity128.err; tail -c 700 gpurun_out/r2_multi_gpu_parity_n2_128_peer.json; tail -4 gpurun_out/p13_parity128.err
for x in 0 1; do
URLG_PEER_XCHG=$x timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$x bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --no-eager-gpu --no-configs --no-large > gpurun_out/p13_bench2_x$x.json 2> gpurun_out/p13_bench2_x$x.err; python -c "
import json; d=json.loads(open('gpurun_out/p13_bench2_x$x.json').read().strip().splitlines()[-1]); print('xchg=$x', d['run']['gradient_exchange'], d['update_ms'], d['ppo_iter_ms'], d['value'])"; tail -2 gpurun_out/p13_bench2_x$x.err
done
