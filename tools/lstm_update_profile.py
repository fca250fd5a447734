#!/usr/bin/env python
"""One PPO update of a recurrent task bracketed by cudaProfilerStart/Stop, for
  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file X python tools/lstm_update_profile.py g1 4096
Without ncu it prints the event-timed update (eager and, with --graph, graph replay)."""
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from test_graph_cuda import make_runner, eager_rollout

task = sys.argv[1] if len(sys.argv) > 1 else "g1"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
minibatches = int(sys.argv[3]) if len(sys.argv) > 3 else 0        # 0: the whole update
r = make_runner(task, n)
alg = r.alg
if minibatches:
    alg.num_learning_epochs, alg.num_mini_batches_run = 1, minibatches


def one(profile=False):
    c = eager_rollout(r)
    with torch.inference_mode():
        alg.compute_returns(c)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if profile:
        torch.cuda.profiler.start()
    e0.record()
    alg.update()
    e1.record()
    torch.cuda.synchronize()
    if profile:
        torch.cuda.profiler.stop()
    return e0.elapsed_time(e1)


for _ in range(2):
    one()
print("update ms (eager):", [round(one(), 3) for _ in range(3)])
print("update ms (profiled bracket):", round(one(True), 3))
