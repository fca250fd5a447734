"""Standalone GPU probe of urlg_tc_gemm (tcgen05 GEMM): each variant runs in its own subprocess under a
timeout so a protocol bug cannot take the other variants (or the box) with it.
    python tests/gpu_probe_gemm.py            # all variants
    python tests/gpu_probe_gemm.py <idx>      # one variant (child)"""
import ctypes as C
import subprocess
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

# (M, N, K, a_mn, b_mn, flags, k_splits, bias, dact, colsum)
VARIANTS = [
    (128, 64, 64, 0, 0, 0, 1, False, False, False),
    (4096, 512, 48, 0, 0, 1, 1, True, False, False),       # fwd layer 1 (K tail, ELU)
    (4096, 256, 512, 0, 0, 1, 1, True, False, False),      # fwd layer 2
    (1000, 12, 128, 0, 0, 0, 1, True, False, False),       # fwd head (N tail, M tail)
    (4096, 128, 256, 0, 1, 0, 1, False, True, True),       # dgrad (B MN-major) + ELU' + colsum
    (4096, 128, 16, 0, 1, 0, 1, False, False, False),      # dgrad from padded head grads (K=16)
    (512, 48, 4096, 1, 1, 2, 8, False, False, False),      # wgrad both MN-major, split-K atomic
    (256, 512, 3000, 1, 1, 2, 5, False, False, False),     # wgrad, ragged K
    (128, 12, 2048, 1, 1, 2, 4, False, False, False),      # wgrad head: A = h3 [B,128], B = dmu [B,16 padded]
]


def run(idx):
    import torch
    from unitree_rl_gym_b200 import _lib
    M, N, K, a_mn, b_mn, flags, splits, use_bias, use_dact, use_colsum = VARIANTS[idx]
    g = torch.Generator(device="cuda").manual_seed(idx)
    pad = lambda n, m: (n + m - 1) // m * m
    if not a_mn:
        lda = pad(K, 8); A = torch.zeros(M, lda, device="cuda", dtype=torch.bfloat16)
        A[:, :K] = torch.randn(M, K, device="cuda", generator=g).bfloat16(); Af = A[:, :K].float()
    else:
        lda = pad(M, 8); A = torch.zeros(K, lda, device="cuda", dtype=torch.bfloat16)
        A[:, :M] = torch.randn(K, M, device="cuda", generator=g).bfloat16(); Af = A[:, :M].float().t()
    if not b_mn:
        ldb = pad(K, 8); B = torch.zeros(N, ldb, device="cuda", dtype=torch.bfloat16)
        B[:, :K] = torch.randn(N, K, device="cuda", generator=g).bfloat16(); Bf = B[:, :K].float()
    else:
        ldb = pad(N, 8); B = torch.zeros(K, ldb, device="cuda", dtype=torch.bfloat16)
        B[:, :N] = torch.randn(K, N, device="cuda", generator=g).bfloat16(); Bf = B[:, :N].float().t()
    bias = torch.randn(N, device="cuda", generator=g) if use_bias else None
    ref_y = (torch.randn(M, N, device="cuda", generator=g)).bfloat16() if use_dact else None
    colsum = torch.zeros(N, device="cuda") if use_colsum else None
    out = torch.zeros(M, N, device="cuda")
    outb = torch.zeros(M, pad(N, 8), device="cuda", dtype=torch.bfloat16)
    d = _lib.GemmDesc()
    d.A, d.lda, d.B, d.ldb = A.data_ptr(), lda, B.data_ptr(), ldb
    d.a_mn, d.b_mn, d.M, d.N, d.K = a_mn, b_mn, M, N, K
    d.block_n, d.k_splits = 0, splits
    d.out_f32, d.ld_f32 = out.data_ptr(), N
    d.out_bf16, d.ld_bf16 = (outb.data_ptr(), outb.shape[1]) if not (flags & 2) else (None, 0)
    d.bias = bias.data_ptr() if bias is not None else None
    d.dact_ref, d.ld_ref = (ref_y.data_ptr(), N) if ref_y is not None else (None, 0)
    d.colsum = colsum.data_ptr() if colsum is not None else None
    d.flags = flags
    _lib.check(_lib.lib().urlg_tc_gemm(C.byref(d), _lib.cur_stream()))
    torch.cuda.synchronize()
    want = Af.double() @ Bf.double().t()
    if bias is not None:
        want = want + bias.double()
    if flags & 1:
        want = torch.nn.functional.elu(want)
    if ref_y is not None:
        y = ref_y.double()
        want = want * torch.where(y > 0, torch.ones_like(y), y + 1)
    err = (out.double() - want).abs().max().item()
    scale = want.abs().max().item()
    ok = err <= 2e-3 * max(scale, 1.0)
    msg = f"variant {idx} {VARIANTS[idx]}: max_err {err:.3e} (scale {scale:.3e})"
    if not (flags & 2):
        eb = (outb[:, :N].double() - want).abs().max().item()
        ok = ok and eb <= 1e-2 * max(scale, 1.0)
        msg += f" bf16_err {eb:.3e}"
    if colsum is not None:
        ec = (colsum.double() - want.sum(0)).abs().max().item()
        ok = ok and ec <= 2e-3 * max(want.sum(0).abs().max().item(), 1.0)
        msg += f" colsum_err {ec:.3e}"
    print(("OK   " if ok else "FAIL ") + msg, flush=True)
    return 0 if ok else 1


if __name__ == "__main__":
    if len(sys.argv) > 1:
        sys.exit(run(int(sys.argv[1])))
    bad = 0
    for i in range(len(VARIANTS)):
        try:
            r = subprocess.run([sys.executable, __file__, str(i)], timeout=120, capture_output=True, text=True)
            sys.stdout.write(r.stdout)
            if r.returncode != 0:
                bad += 1
                sys.stdout.write(f"variant {i} rc={r.returncode}\n" + r.stderr[-600:] + "\n")
        except subprocess.TimeoutExpired:
            bad += 1
            print(f"variant {i} TIMEOUT", flush=True)
    print("probe done, failures:", bad)
    sys.exit(1 if bad else 0)
