#!/usr/bin/env python
"""Per-shape timing of the tcgen05 GEMM (urlg_tc_gemm) on the shapes of one Go2 PPO minibatch (M = 24576 rows) and of
the rollout forward (M = 4096), split-bf16 operands, the same descriptors mlp_forward / mlp_backward build
(csrc/policy_kernels.cu, csrc/ppo_kernels.cu).  Each shape: 20 launches in one CUDA graph, CUDA events around the replay.

    python tools/gemm_bench.py [--rows 24576]
"""
import argparse
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    from unitree_rl_gym_b200 import _lib
    import bench
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=24576)
    ap.add_argument("--only", default="", help="comma-separated shape names; with --profile: 3 eager launches each (for ncu)")
    ap.add_argument("--profile", action="store_true")
    ap.add_argument("--waves", type=int, default=1, help="split-K weight gradients: tiles x splits <= waves x 148")
    a = ap.parse_args()
    R = a.rows
    L = _lib.lib()
    hbm, _, tf_sus, _ = bench.measured_peaks()
    dims = [48, 512, 256, 128, 12]
    bf = lambda *s: (torch.randn(*s, device="cuda") * 0.1).bfloat16()
    rup = lambda x, m: (x + m - 1) // m * m
    shapes = []
    # name, M, N, K, a_mn, b_mn, flags, splits, block_n, epilogue
    for l in range(4):
        shapes.append((f"fwd{l}", R, dims[l + 1], dims[l], 0, 0, 1 if l < 3 else 0, 1, 0, "bias" + ("+elu+bf16" if l < 3 else "+f32")))
    for l in (3, 2, 1):
        nout, nin = (16 if l == 3 else dims[l + 1]), dims[l]
        shapes.append((f"dgrad{l}", R, nin, nout, 0, 1, 0, 1, min(256, rup(nin, 64)), "dact+colsum+bf16"))
    for l in (3, 2, 1, 0):
        nout, nin = dims[l + 1], dims[l]
        bn = min(256, rup(nin, 64))
        bn_eff = min(bn, 128)                                   # split-bf16 tiles are at most 128 wide
        tiles = ((nout + 127) // 128) * ((nin + bn_eff - 1) // bn_eff)
        splits = max(1, min((a.waves * 148) // tiles, ((R + 63) // 64) // 4))
        shapes.append((f"wgrad{l}", nout, nin, R, 1, 1, 2, splits, bn, "atomic f32"))
    # launch floor: one tile, one k-block
    shapes.append(("floor", 128, 64, 64, 0, 0, 1, 1, 0, "bias+elu+bf16"))
    only = set(a.only.split(",")) if a.only else None
    for name, M, N, K, a_mn, b_mn, flags, splits, bn, epi in shapes:
        if only and name not in only:
            continue
        if not a_mn:
            lda = rup(K, 8); A, Al = bf(M, lda), bf(M, lda)
        else:
            lda = rup(M, 8); A, Al = bf(K, lda), bf(K, lda)
        if not b_mn:
            ldb = rup(K, 8); B, Bl = bf(N, ldb), bf(N, ldb)
        else:
            ldb = rup(N, 8); B, Bl = bf(K, ldb), bf(K, ldb)
        d = _lib.GemmDesc()
        d.A, d.lda, d.B, d.ldb, d.A_lo, d.B_lo = A.data_ptr(), lda, B.data_ptr(), ldb, Al.data_ptr(), Bl.data_ptr()
        d.a_mn, d.b_mn, d.M, d.N, d.K, d.block_n, d.k_splits, d.flags = a_mn, b_mn, M, N, K, bn, splits, flags
        keep = []
        mn = M * N
        by = 2 * 2 * (M * K + N * K)
        if "bf16" in epi:
            ld = rup(N, 8)
            o, ol = torch.zeros(M, ld, device="cuda", dtype=torch.bfloat16), torch.zeros(M, ld, device="cuda", dtype=torch.bfloat16)
            d.out_bf16, d.out_bf16_lo, d.ld_bf16 = o.data_ptr(), ol.data_ptr(), ld
            keep += [o, ol]; by += 4 * mn
        else:
            o = torch.zeros(M, N, device="cuda")
            d.out_f32, d.ld_f32 = o.data_ptr(), N
            keep += [o]; by += 4 * mn
        if "bias" in epi:
            b = torch.randn(N, device="cuda"); d.bias = b.data_ptr(); keep.append(b)
        if "dact" in epi:
            y, yl = bf(M, rup(N, 8)), bf(M, rup(N, 8))
            d.dact_ref, d.dact_ref_lo, d.ld_ref = y.data_ptr(), yl.data_ptr(), rup(N, 8)
            cs = torch.zeros(N, device="cuda"); d.colsum = cs.data_ptr()
            keep += [y, yl, cs]; by += 4 * mn
        fn = lambda: _lib.check(L.urlg_tc_gemm(C.byref(d), _lib.cur_stream()))
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        if a.profile:
            continue
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(20):
                fn()
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); e1.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 20
        fl = 2.0 * M * N * K
        print(json.dumps({"gemm": name, "M": M, "N": N, "K": K, "splits": splits, "epilogue": epi, "us": round(us, 2),
                          "tflops": round(fl / us / 1e6, 1), "alg_gbs": round(by / us / 1e3, 1),
                          "frac_hbm": round(by / us / 1e3 / hbm, 3), "frac_tensor": round(fl / us / 1e6 / tf_sus, 3)}), flush=True)


if __name__ == "__main__":
    main()
