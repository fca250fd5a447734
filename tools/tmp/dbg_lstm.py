import sys, os, json
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from tests import test_ppo_cuda as T
for i in range(5):
    rep = T.one_update_report(True, 47, 50, 12, [32], 4096, 24, keep_vectors=True)
    pt = {k: float("%.2e" % v) for k, v in rep["per_tensor_delta_rel_l2"].items()}
    print(i, "max", rep["cuda_vs_oracle32"]["param_max_abs"], json.dumps(pt), flush=True)
    print(i, "kl", [round(a[0]/b[0]-1, 6) for a, b in zip(rep["lr_trace_cuda"], rep["lr_trace_oracle"])], flush=True)
