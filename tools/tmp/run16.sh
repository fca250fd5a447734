cd $GRAFT_REPO_ROOT
export MASTER_ADDR=127.0.0.1
for i in 1 2; do
URLG_PEER_XCHG=0 URLG_PARITY_NLOC=4096 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2953$i tools/multi_gpu_parity.py > gpurun_out/p16_parity_nccl$i.json 2> gpurun_out/p16_parity_nccl$i.err; python -c "
import json; d=json.loads(open('gpurun_out/p16_parity_nccl$i.json').read().strip().splitlines()[-1]); print('nccl', d['gradient_exchange'], [(g['vs_eager_max_abs']) for g in d['update_graph_with_nccl']], d['ranks_bit_identical'], d['pass'])"
done
