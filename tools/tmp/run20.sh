cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_env_cuda.py tests/test_graph_cuda.py -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 --no-cpu --no-large --no-configs --no-eager-gpu > gpurun_out/p20_bench.json 2> gpurun_out/p20_bench.err; tail -3 gpurun_out/p20_bench.err
python -c "
import json; d=json.loads(open('gpurun_out/p20_bench.json').read().strip().splitlines()[-1]); e=d['e2e']; print(d['value'], d['rollout_ms'], 'e2e', e['value'], e['rollout_ms'], e.get('step_graphs_ms'), e.get('one_graph_ms'), e.get('h2d_link_gbs'), e.get('h2d_floor_ms'), e['h2d_bytes_per_step'], e['eager']['rollout_ms'])"
