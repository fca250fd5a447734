import sys, os, json
sys.path.insert(0, os.getcwd())
from tests import test_ppo_cuda as T
for rec, args in ((True, (47, 50, 12, [32])), (False, (48, None, 12, [512, 256, 128]))):
    for seed in range(5):
        rep = T.one_update_report(rec, *args, 4096, 24, seed=seed)
        d = rep["cuda_vs_oracle32"]
        print("rec" if rec else "mlp", "seed", seed, "max_abs %.3e delta_rel_l2 %.3e beyond %d lr %.3e" % (d["param_max_abs"], d["delta_rel_l2"], d["n_beyond_1e-3"], rep["lr_final"]["cuda"]), flush=True)
