cd $GRAFT_REPO_ROOT
export MASTER_ADDR=127.0.0.1
BF="--gpus 8 --steps 10 --warmup 3 --no-cpu --no-eager-gpu --no-large --no-configs"
for m in 1 0 nccl 1; do
if [ $m = nccl ]; then export URLG_PEER_XCHG=0; else export URLG_PEER_XCHG=1; export URLG_PEER_MODE=$m; fi
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2957${m:0:1} bench.py $BF > gpurun_out/p23_bench8_$m.json 2> gpurun_out/p23_bench8_$m.err
python -c "
import json; d=json.loads(open('gpurun_out/p23_bench8_$m.json').read().strip().splitlines()[-1]); print('$m', d['run']['gradient_exchange'][:50], d['value'], d['rollout_ms'], d['update_ms'], d['ppo_iter_ms'])"
done
