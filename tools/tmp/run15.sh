cd $GRAFT_REPO_ROOT
export MASTER_ADDR=127.0.0.1
for f in 1 0 1; do
URLG_PEER_FUSED=$f URLG_PARITY_NLOC=4096 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2952$f tools/multi_gpu_parity.py > gpurun_out/p15_parity_f$f.json 2> gpurun_out/p15_parity_f$f.err; python -c "
import json; d=json.loads(open('gpurun_out/p15_parity_f$f.json').read().strip().splitlines()[-1]); print('fused=$f', [(g['vs_eager_max_abs']) for g in d['update_graph_with_nccl']], d['ranks_bit_identical'], d['pass'])"
done
