"""urlg_tc_gemm / urlg_tc_gemm_group through the C ABI: the four epilogue modes of the PPO half (forward bias+ELU with
bf16 hi/lo outputs, fp32 head, dgrad with ELU' reference + bias-gradient column sums, split-K weight gradient) against
an fp64 reference, each alone and all four as ONE grouped launch — non-atomic outputs of the group must be bit-identical
to the separate launches (same tiles, same arithmetic), atomic ones equal to summation-order tolerance."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

rup = lambda x, m: (x + m - 1) // m * m


def split(x):
    hi = x.bfloat16()
    return hi, (x - hi.float()).bfloat16()


class Problem:
    def __init__(self, kind, M, N, K, seed):
        from unitree_rl_gym_b200 import _lib
        g = torch.Generator(device="cuda").manual_seed(seed)
        rn = lambda *s: torch.randn(*s, device="cuda", generator=g)
        self.kind, self.M, self.N, self.K = kind, M, N, K
        d = self.d = _lib.GemmDesc()
        a_mn = b_mn = 0
        if kind == "wgrad":
            a_mn = b_mn = 1
        elif kind == "dgrad":
            b_mn = 1
        Af, Bf = rn(M, K) * 0.5, rn(N, K) * 0.5                       # logical A[m,k], B[n,k]
        def image(x, mn):                                             # K-major [rows, k] or MN-major [k, rows], padded pitch
            t = x.t().contiguous() if mn else x
            ld = rup(t.shape[1], 8)
            buf = torch.zeros(t.shape[0], ld, device="cuda")
            buf[:, :t.shape[1]] = t
            return split(buf) + (ld,)
        self.A, self.Al, lda = image(Af, a_mn)
        self.B, self.Bl, ldb = image(Bf, b_mn)
        self.Af = (self.A.float() + self.Al.float())[:, :(M if a_mn else K)]
        self.Bf = (self.B.float() + self.Bl.float())[:, :(N if b_mn else K)]
        if a_mn: self.Af = self.Af.t()
        if b_mn: self.Bf = self.Bf.t()
        d.A, d.A_lo, d.lda, d.B, d.B_lo, d.ldb = self.A.data_ptr(), self.Al.data_ptr(), lda, self.B.data_ptr(), self.Bl.data_ptr(), ldb
        d.a_mn, d.b_mn, d.M, d.N, d.K = a_mn, b_mn, M, N, K
        self.bias = self.ref = self.colsum = self.out = self.ob = self.obl = None
        if kind in ("fwd", "head"):
            self.bias = rn(N)
            d.bias = self.bias.data_ptr()
        if kind in ("fwd", "dgrad"):
            ld = rup(N, 8)
            self.ob = torch.zeros(M, ld, device="cuda", dtype=torch.bfloat16)
            self.obl = torch.zeros(M, ld, device="cuda", dtype=torch.bfloat16)
            d.out_bf16, d.out_bf16_lo, d.ld_bf16 = self.ob.data_ptr(), self.obl.data_ptr(), ld
            d.flags = 1 if kind == "fwd" else 0
        else:
            self.out = torch.zeros(M, N, device="cuda")
            d.out_f32, d.ld_f32 = self.out.data_ptr(), N
        if kind == "dgrad":
            y = rn(M, rup(N, 8)) * 0.7
            self.ref, self.refl = split(y)
            d.dact_ref, d.dact_ref_lo, d.ld_ref = self.ref.data_ptr(), self.refl.data_ptr(), rup(N, 8)
            self.colsum = torch.zeros(N, device="cuda")
            d.colsum = self.colsum.data_ptr()
            d.block_n = min(256, rup(N, 64))
        if kind == "wgrad":
            d.flags, d.k_splits, d.block_n = 2, 7, min(256, rup(N, 64))

    def want(self):
        w = self.Af.double() @ self.Bf.double().t()
        if self.bias is not None:
            w = w + self.bias.double()
        if self.kind == "fwd":
            w = torch.nn.functional.elu(w)
        if self.kind == "dgrad":
            y = (self.ref.double() + self.refl.double())[:, :self.N]
            w = w * torch.where(y > 0, torch.ones_like(y), y + 1)
        return w

    def result(self):
        if self.out is not None:
            return self.out.clone()
        return (self.ob.float() + self.obl.float())[:, :self.N].clone()

    def clear(self):
        for t in (self.out, self.ob, self.obl, self.colsum):
            if t is not None:
                t.zero_()


SHAPES = [("fwd", 1000, 256, 48), ("dgrad", 777, 128, 256), ("wgrad", 256, 48, 3000), ("head", 1000, 12, 128)]


def make():
    return [Problem(k, M, N, K, 10 + i) for i, (k, M, N, K) in enumerate(SHAPES)]


def check(p, got, tol=2e-5):
    w = p.want()
    scale = max(w.abs().max().item(), 1.0)
    assert (got.double() - w).abs().max().item() <= tol * scale, p.kind
    if p.colsum is not None:
        cs = w.sum(0)
        assert (p.colsum.double() - cs).abs().max().item() <= 1e-4 * max(cs.abs().max().item(), 1.0)


def test_single_launches_match_fp64():
    from unitree_rl_gym_b200 import _lib
    for p in make():
        _lib.check(_lib.lib().urlg_tc_gemm(C.byref(p.d), _lib.cur_stream()))
        torch.cuda.synchronize()
        check(p, p.result())


@pytest.mark.parametrize("order", [(0, 1, 2, 3), (2, 3, 1, 0), (1, 0), (3, 2, 0)])
def test_group_equals_single_launches(order):
    from unitree_rl_gym_b200 import _lib
    L = _lib.lib()
    ps = make()
    single = {}
    for i in order:
        _lib.check(L.urlg_tc_gemm(C.byref(ps[i].d), _lib.cur_stream()))
        torch.cuda.synchronize()
        single[i] = (ps[i].result(), None if ps[i].colsum is None else ps[i].colsum.clone())
        ps[i].clear()
    arr = (_lib.GemmDesc * len(order))()
    for k, i in enumerate(order):
        C.memmove(C.byref(arr[k]), C.byref(ps[i].d), C.sizeof(_lib.GemmDesc))
    _lib.check(L.urlg_tc_gemm_group(arr, len(order), _lib.cur_stream()))
    torch.cuda.synchronize()
    for i in order:
        p = ps[i]
        got = p.result()
        check(p, got)
        if p.kind == "wgrad":
            assert torch.allclose(got, single[i][0], rtol=1e-5, atol=1e-4)
        else:
            assert torch.equal(got, single[i][0]), p.kind


def test_group_rejects_bad_counts_and_mixed_modes():
    from unitree_rl_gym_b200 import _lib
    L = _lib.lib()
    ps = make()
    arr = (_lib.GemmDesc * 5)()
    assert L.urlg_tc_gemm_group(arr, 0, _lib.cur_stream()) != 0
    assert L.urlg_tc_gemm_group(arr, 5, _lib.cur_stream()) != 0
    two = (_lib.GemmDesc * 2)()
    C.memmove(C.byref(two[0]), C.byref(ps[0].d), C.sizeof(_lib.GemmDesc))
    C.memmove(C.byref(two[1]), C.byref(ps[3].d), C.sizeof(_lib.GemmDesc))
    two[1].A_lo, two[1].B_lo = None, None                           # plain bf16 next to a hi+lo problem
    assert L.urlg_tc_gemm_group(two, 2, _lib.cur_stream()) != 0
    assert b"operand mode" in L.urlg_last_error()


def test_split_k_partial_slabs_sum_to_the_product_and_are_reproducible():
    """out_part mode: every K split stores its fp32 partial tile into slab z of out_part[splits][M][ld] (no atomics, no
    flag needed); the slabs summed in split order equal the fp64 product, and two launches give bit-identical slabs."""
    from unitree_rl_gym_b200 import _lib
    L = _lib.lib()
    for (M, N, K, ks) in ((256, 48, 3000, 7), (12, 128, 1000, 96), (130, 200, 640, 3)):
        p = Problem("wgrad", M, N, K, 31)
        eff = L.urlg_tc_gemm_splits(K, ks)
        assert 1 <= eff <= ks
        part = torch.full((eff, M, N), float("nan"), device="cuda")
        p.d.flags, p.d.k_splits, p.d.out_f32, p.d.out_part = 0, ks, None, part.data_ptr()
        _lib.check(L.urlg_tc_gemm(C.byref(p.d), _lib.cur_stream()))
        torch.cuda.synchronize()
        first = part.clone()
        assert not torch.isnan(first).any(), "every slab element must be written"
        check(p, first.double().sum(0))
        part.fill_(float("nan"))
        _lib.check(L.urlg_tc_gemm(C.byref(p.d), _lib.cur_stream()))
        torch.cuda.synchronize()
        assert torch.equal(part, first)
