#!/usr/bin/env python
"""One PPO update of the CUDA path against the oracle at the BASELINE sizes (N = 4096 envs, T = 24), measured on the
UPDATE DELTA, with the fp32 oracle's own distance to an fp64 evaluation of the same update printed beside it (the floor
any fp32 evaluation order can reach), and the two crafted adaptive-KL scenarios.  Writes one JSON document:

    python tools/ppo_update_parity.py > profiles/r2_ppo_parity_4096.json
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import contextlib  # noqa: E402
import torch  # noqa: E402
from tests import test_ppo_cuda as T  # noqa: E402

N = int(os.environ.get("URLG_PARITY_N", "4096"))
out = {"what": "one PPO.update (5 epochs x 4 minibatches), CUDA path vs oracle/rsl_rl_port, distances on theta_after - theta_before",
       "threads": torch.get_num_threads(), "cases": []}
for name, args in (("go2_mlp", (False, 48, None, 12, [512, 256, 128], N, 24)),
                   ("g1_lstm", (True, 47, 50, 12, [32], N, 24))):
    for seed in range(5):              # every seed is recorded, also the ones with a clip-boundary row (see tests/test_ppo_cuda.py)
        with contextlib.redirect_stdout(sys.stderr):       # the modules print their layer summaries like rsl_rl does
            rep = T.one_update_report(*args, with_fp64=True, seed=seed)
        rep["case"] = name
        rep["seed"] = seed
        out["cases"].append(rep)
        print(name, seed, json.dumps({k: rep[k] for k in ("cuda_vs_oracle32", "oracle32_vs_oracle64", "delta_scale", "lr_final")}),
              file=sys.stderr, flush=True)
for kl in (0.0202, 0.00483):
    with contextlib.redirect_stdout(sys.stderr):
        rep = T.one_update_report(False, 48, None, 12, [512, 256, 128], 1024, 24, kl_target=kl, with_fp64=False)
    rep["case"] = "go2_mlp_kl_%g" % kl
    out["cases"].append(rep)
    print(rep["case"], rep["lr_trace_cuda"], file=sys.stderr, flush=True)
print(json.dumps(out, indent=1))
