#!/usr/bin/env python
"""Multi-GPU parity: W ranks, each owning N_loc envs, against the ORACLE's single-process PPO over the
union of W*N_loc envs — global advantage normalisation (all-reduced moments) to 1e-4 and parameters after one PPO
update (20 steps, NCCL gradient / KL all-reduce per step) to 1e-3.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/multi_gpu_parity.py

Rank 0 prints one JSON line (kept under profiles/).  Test infrastructure: imports oracle/.
"""
import copy
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def main():
    from tests import test_ppo_cuda as T
    from oracle import rsl_rl_port as O
    from unitree_rl_gym_b200.rsl_rl import modules as M, algorithms as A
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    n_loc = int(os.environ.get("URLG_PARITY_NLOC", "128"))          # 4096 = the BASELINE per-GPU size (oracle over the union: ~1 min on CPU)
    steps, n_obs, n_act, hidden = 24, 48, 12, [512, 256, 128]
    n_glob = n_loc * world

    # ---- the oracle over the union (every rank builds it: cheap, deterministic, avoids shipping tensors around)
    torch.manual_seed(0)
    oac = O.ActorCritic(n_obs, n_obs, n_act, hidden, hidden, "elu", init_noise_std=0.8)
    pac = M.ActorCritic(n_obs, n_obs, n_act, hidden, hidden, "elu", init_noise_std=0.8)
    pac.load_state_dict(copy.deepcopy(oac.state_dict()))
    oppo = O.PPO(oac, device="cpu", **T.ALG)
    oppo.init_storage(n_glob, steps, [n_obs], [None], [n_act])
    g = torch.Generator().manual_seed(3)
    for t in range(steps):
        obs = torch.randn(n_glob, n_obs, generator=g)
        oppo.sample_noise = torch.randn(n_glob, n_act, generator=g)
        with torch.no_grad():
            oppo.act(obs, obs)
        rew = torch.randn(n_glob, generator=g)
        dones = torch.rand(n_glob, generator=g) < 0.05
        tos = dones & (torch.rand(n_glob, generator=g) < 0.5)
        with torch.no_grad():
            oppo.process_env_step(rew, dones, {"time_outs": tos})
    last = torch.randn(n_glob, n_obs, generator=g)
    with torch.no_grad():
        oppo.compute_returns(last)

    # ---- this rank's shard of the same rollout, through the product classes (NCCL world)
    pppo = A.PPO(pac, device=f"cuda:{torch.cuda.current_device()}", **T.ALG)
    pppo.init_storage(n_loc, steps, [n_obs], [None], [n_act])
    lo, hi = rank * n_loc, (rank + 1) * n_loc
    so, sp = oppo.storage, pppo.storage
    for name in ("observations", "actions", "mu", "sigma", "rewards", "values", "actions_log_prob", "dones"):
        getattr(sp, name).copy_(getattr(so, name)[:, lo:hi])
    sp.step = steps
    pppo.compute_returns(last[lo:hi].cuda())            # local scan + all-reduced (sum, sumsq, n) + normalise
    torch.cuda.synchronize()
    adv_err = float((sp.advantages.cpu() - so.advantages[:, lo:hi]).abs().max())
    ret_err = float((sp.returns.cpu() - so.returns[:, lo:hi]).abs().max())

    # local minibatch permutations; the oracle's global permutation is their union, minibatch by minibatch
    mini_loc = (n_loc * steps) // T.ALG["num_mini_batches"]
    perms = [torch.randperm(T.ALG["num_mini_batches"] * mini_loc, generator=torch.Generator().manual_seed(10 + r))
             for r in range(world)]
    glob = []
    for i in range(T.ALG["num_mini_batches"]):
        for r in range(world):
            p = perms[r][i * mini_loc:(i + 1) * mini_loc]
            glob.append((p // n_loc) * n_glob + r * n_loc + p % n_loc)
    oppo.perm, pppo.perm = torch.cat(glob), perms[rank]
    before = {n: p.detach().clone() for n, p in oac.named_parameters()}
    ov, osur = oppo.update()
    eng = pppo.engine
    pppo.cuda_graph_update = True                  # update 1 runs eagerly (warm-up), 2 captures + replays, 3 replays
    snap = [t.clone() for t in (eng.flat, eng.exp_avg, eng.exp_avg_sq, eng.stats)]

    def against_oracle():
        worst_rel, worst_abs, n_in, n_tot = 0.0, 0.0, 0, 0
        for (n, p), (_, q) in zip(pac.named_parameters(), oac.named_parameters()):
            a, b = p.detach().cpu().numpy(), q.detach().numpy()
            worst_rel = max(worst_rel, T.rel_l2(a, b))
            d = np.abs(a - b)
            worst_abs = max(worst_abs, float(d.max()))
            n_in += int((d <= 1e-3).sum())
            n_tot += d.size
        return worst_rel, worst_abs, n_in, n_tot

    pv, psur = pppo.update()
    torch.cuda.synchronize()
    worst_rel, worst_abs, n_in, n_tot = against_oracle()
    eager_flat = eng.flat.clone()
    # the same update again from the same state, now with the 20 minibatch steps AND their NCCL all-reduces as one graph
    graph_rounds = []
    for mode in ("capture+replay", "replay"):
        for t, c in zip((eng.flat, eng.exp_avg, eng.exp_avg_sq, eng.stats), snap):
            t.copy_(c)
        eng.sync_weights()
        sp.step = steps
        gv, gsur = pppo.update()
        torch.cuda.synchronize()
        g_rel, g_abs, g_in, g_tot = against_oracle()
        graph_rounds.append({"mode": mode, "graph": pppo._update_graph is not None, "update_param_rel_l2_worst": g_rel,
                             "update_param_max_abs": g_abs, "update_frac_within_1e-3": g_in / g_tot,
                             "vs_eager_max_abs": float((eng.flat - eager_flat).abs().max()), "losses": [gv, gsur]})
    moved = max(float((q.detach() - before[n]).abs().max()) for n, q in oac.named_parameters())
    # every rank must hold identical parameters
    flat = pppo.engine.flat.clone()
    ref = flat.clone()
    dist.broadcast(ref, 0)
    same = bool(torch.equal(flat, ref))
    flags = torch.tensor([int(same)], device="cuda")
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    res = {"world": world, "n_loc": n_loc, "steps": steps, "returns_max_abs_err": ret_err, "advantages_max_abs_err": adv_err,
           "update_param_rel_l2_worst": worst_rel, "update_param_max_abs": worst_abs, "update_frac_within_1e-3": n_in / n_tot,
           "oracle_param_moved_max": moved, "lr": [oppo.learning_rate, pppo.learning_rate],
           "losses": [ov, pv, osur, psur], "ranks_bit_identical": bool(flags.item()),
           "gradient_exchange": pppo.exchange, "n_loc": n_loc, "update_graph_with_nccl": graph_rounds, "update_graph_launches": pppo.update_graph_launches}
    graph_ok = all(g["graph"] and g["update_param_rel_l2_worst"] < 1e-3 and g["update_frac_within_1e-3"] >= 0.9999 and
                   g["update_param_max_abs"] <= 3e-3 for g in graph_rounds) or not pppo._graph_nccl
    ok = (graph_ok and adv_err < 1e-4 and ret_err < 1e-4 and worst_rel < 1e-3 and n_in / n_tot >= 0.9999 and worst_abs <= 3e-3 and
          abs(oppo.learning_rate - pppo.learning_rate) < 1e-6 * oppo.learning_rate + 1e-9 and res["ranks_bit_identical"])
    res["pass"] = bool(ok)
    if rank == 0:
        print(json.dumps(res), flush=True)
    pppo.release_graphs()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
