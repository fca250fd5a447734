cd $GRAFT_REPO_ROOT
export MASTER_ADDR=127.0.0.1
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu --no-eager-gpu --no-large > gpurun_out/r2_bench_go2_n8.json 2> gpurun_out/p21_bench8.err
python -c "
import json; d=json.loads(open('gpurun_out/r2_bench_go2_n8.json').read().strip().splitlines()[-1]); print(d['run']['gradient_exchange'][:60], d['value'], d['rollout_ms'], d['update_ms'], d['ppo_iter_ms'], d['e2e']['value']); [print(k, v['value'], v['ppo_iter_ms'], v['update_ms']) for k,v in d['configs'].items()]"
grep -E "Error|error" gpurun_out/p21_bench8.err | head -5
URLG_PEER_XCHG=0 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu --no-eager-gpu --no-large --no-configs > gpurun_out/r2_bench_go2_n8_nccl.json 2> gpurun_out/p21_bench8n.err
python -c "
import json; d=json.loads(open('gpurun_out/r2_bench_go2_n8_nccl.json').read().strip().splitlines()[-1]); print(d['run']['gradient_exchange'][:60], d['value'], d['rollout_ms'], d['update_ms'], d['ppo_iter_ms'])"
