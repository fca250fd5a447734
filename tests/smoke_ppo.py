"""PPO half of __graft_entry__.smoke(): GAE and one minibatch (gather, tcgen05 forward, PPO head, backward) through
the C ABI on cuda:0, checked against the oracle's rsl_rl restatement — the small cases of tests/test_ppo_cuda.py."""


def smoke_ppo():
    from tests import test_ppo_cuda as T
    T.test_gae_matches_oracle(64, 24)
    T.test_gae_hand_case()
    T.test_minibatch_gradients_match_oracle(False, 48, None, 12, [512, 256, 128], 130, 7)
    print("smoke: PPO half (GAE 1e-4, minibatch gradients rel-L2 < 1e-3) matches the oracle")
