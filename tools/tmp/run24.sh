cd $GRAFT_REPO_ROOT
export MASTER_ADDR=127.0.0.1
BF="--gpus 8 --steps 20 --warmup 3 --no-cpu --no-eager-gpu --no-large --no-configs"
i=0
for m in nccl 1 nccl 0; do
i=$((i+1))
if [ $m = nccl ]; then export URLG_PEER_XCHG=0; else export URLG_PEER_XCHG=1; export URLG_PEER_MODE=$m; fi
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2958$i bench.py $BF > gpurun_out/p24_bench8_$i.json 2> gpurun_out/p24_bench8_$i.err
python -c "
import json; d=json.loads(open('gpurun_out/p24_bench8_$i.json').read().strip().splitlines()[-1]); print('$m', d['run']['gradient_exchange'][:30], d['value'], d['rollout_ms'], d['update_ms'], d['ppo_iter_ms'])"
done
