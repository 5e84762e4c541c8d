"""GPU: the tcgen05/TMA GEMM (indnet_tc_gemm) against a float64 matmul of the same bf16-rounded
operands, for all four operand-major combinations, ragged sizes, batching, split-K, epilogues."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _tc(a_mn, b_mn, M, N, K, A, B, out_f32=True, out_bf16=False, bias=None, act=0, accumulate=None,
        batch=1, splitk=1):
    from induction_b200._lib import LIB, current_stream, ptr
    lda = A.shape[-1]
    ldb = B.shape[-1]
    C = (accumulate.clone() if accumulate is not None else torch.full((batch, M, N), float("nan"), device="cuda")) if out_f32 else None
    Cb = torch.zeros(batch, M, N, dtype=torch.bfloat16, device="cuda") if out_bf16 else None
    sA = A.shape[-2] * A.shape[-1] if batch > 1 else 0
    sB = B.shape[-2] * B.shape[-1] if batch > 1 else 0
    LIB.call("indnet_tc_gemm", current_stream(), int(a_mn), int(b_mn), M, N, K, ptr(A), lda, ptr(B), ldb,
             ptr(C), ptr(Cb), N, ptr(bias), act, int(accumulate is not None), batch, sA, sB, M * N, splitk)
    torch.cuda.synchronize()
    return C, Cb


def _operands(a_mn, b_mn, M, N, K, batch, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    A = torch.randn(batch, *((K, M) if a_mn else (M, K)), generator=g, device="cuda").to(torch.bfloat16)
    B = torch.randn(batch, *((K, N) if b_mn else (N, K)), generator=g, device="cuda").to(torch.bfloat16)
    Am = A.double().transpose(1, 2) if a_mn else A.double()
    Bm = B.double() if b_mn else B.double().transpose(1, 2)
    return A.contiguous(), B.contiguous(), Am @ Bm


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 384, 256), (320, 200, 136), (25, 504, 256), (192, 512, 4096)])
def test_tc_gemm_matches_float64(a_mn, b_mn, M, N, K):
    if a_mn and M % 8:
        pytest.skip("MN-major A needs a leading dimension that is a multiple of 8 elements")
    A, B, want = _operands(a_mn, b_mn, M, N, K, 1, seed=M + N + K)
    C, _ = _tc(a_mn, b_mn, M, N, K, A, B)
    err = float((C.double() - want).abs().max() / want.abs().max())
    assert err < 2e-5, err


def test_tc_gemm_batched_bias_act_bf16_out_and_accumulate():
    A, B, want = _operands(0, 0, 25, 500, 256, 7, seed=1)       # NTN stage 2 shape: per-episode [n_q,H]x[H,N*S]
    C, Cb = _tc(0, 0, 25, 500, 256, A, B, out_bf16=True, batch=7)
    assert float((C.double() - want).abs().max() / want.abs().max()) < 2e-5
    assert float((Cb.double() - want).abs().max() / want.abs().max()) < 1e-2
    A, B, want = _operands(0, 1, 300, 256, 256, 1, seed=2)      # attention Dense: bias + tanh epilogue
    bias = torch.randn(256, device="cuda")
    C, _ = _tc(0, 1, 300, 256, 256, A, B, bias=bias, act=1)
    ref = torch.tanh(want / 1.0 + bias.double())
    assert float((C.double() - ref).abs().max()) < 2e-3      # epilogue tanh is MUFU tanh.approx
    base = torch.randn(1, 300, 256, device="cuda")
    C, _ = _tc(0, 1, 300, 256, 256, A, B, accumulate=base)
    assert float((C.double() - (want + base.double())).abs().max() / want.abs().max()) < 2e-5


@pytest.mark.parametrize("a_mn,b_mn", [(1, 1), (0, 0)])
def test_tc_gemm_split_k(a_mn, b_mn):
    M, N, K = 192, 512, 76800                                     # LSTM weight-gradient shape at cfg1
    A, B, want = _operands(a_mn, b_mn, M, N, K, 1, seed=3)
    C, _ = _tc(a_mn, b_mn, M, N, K, A, B, splitk=0)
    assert float((C.double() - want).abs().max() / want.abs().max()) < 5e-5


@pytest.mark.parametrize("M,N,rows,group,split", [(400, 256, 40, 5, 2), (203, 128, 7, 3, 3), (96, 48, 12, 4, 1)])
def test_tc_gemm_rank1_epilogue(M, N, rows, group, split):
    """C = A*B + rs[m] * v(m / rows)[n] with v taken from the (support, query) pair of matrices in the
    split layout of the attention kernels - the fused d(loss)/d(h) of the self-attention block.  Covers the
    16-byte row path (N % 32 == 0) and the transposing path (N = 48), fp32 and bf16 outputs."""
    from induction_b200._lib import LIB, current_stream, ptr
    K = 256
    A, B, want = _operands(0, 0, M, N, K, 1, seed=M + N)
    X = (M + rows - 1) // rows                       # vectors touched ("sentences")
    X = (X + group - 1) // group * group
    g = torch.Generator(device="cuda").manual_seed(9)
    rs = torch.randn(M, generator=g, device="cuda")
    va = torch.randn(X // group * split, N, generator=g, device="cuda")
    vb = torch.randn(max(X // group * (group - split), 1), N, generator=g, device="cuda")
    C = torch.full((M, N), float("nan"), device="cuda")
    Cb = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    LIB.call("indnet_tc_gemm_rank1", current_stream(), 0, 0, M, N, K, ptr(A), K, ptr(B), K, ptr(C), ptr(Cb), N,
             ptr(rs), ptr(va), ptr(vb) if split < group else None, rows, group, split)
    torch.cuda.synchronize()
    v = torch.empty(X, N, device="cuda")
    for x in range(X):
        e, r = divmod(x, group)
        v[x] = va[e * split + r] if r < split else vb[e * (group - split) + (r - split)]
    ref = want[0] + rs.double()[:, None] * v[torch.arange(M, device="cuda") // rows].double()
    scale = float(ref.abs().max())
    assert float((C.double() - ref).abs().max()) / scale < 2e-5
    assert float((Cb.double() - ref).abs().max()) / scale < 1e-2


@pytest.mark.parametrize("T,D", [(4096, 60), (1000, 18), (77, 64)])
def test_tc_gemm_lstm_dw_stacked_operand(T, D):
    """dkernel[D+u][4u] = [x | h_prev]^T dz in one pass over dz (two accumulators per CTA, split-K over tokens,
    16-byte vector reductions): rows < D from x, rows D.. from h_prev, the padding rows of x are dropped."""
    from induction_b200._lib import LIB, current_stream, ptr
    u = 128
    g = torch.Generator(device="cuda").manual_seed(T + D)
    x = torch.zeros(T, 64, device="cuda")
    x[:, :D] = torch.randn(T, D, generator=g, device="cuda")
    if D <= 62:
        x[:, D:D + 2] = 1.0          # the bias columns indnet_embed_fwd_bf16 writes
    x = x.to(torch.bfloat16)
    h = torch.randn(T, u, generator=g, device="cuda").to(torch.bfloat16)
    dz = torch.randn(T, 4 * u, generator=g, device="cuda").to(torch.bfloat16)
    out = torch.full((D + u, 4 * u), float("nan"), device="cuda")
    dbias = torch.full((4 * u,), float("nan"), device="cuda") if D <= 62 else None
    LIB.call("indnet_tc_gemm_lstm_dw", current_stream(), ptr(x), ptr(h), ptr(dz), ptr(out), ptr(dbias), T, D, u, 0, 0)
    torch.cuda.synchronize()
    ref = torch.cat([x[:, :D].double().t() @ dz.double(), h.double().t() @ dz.double()], 0)
    assert float((out.double() - ref).abs().max() / ref.abs().max()) < 2e-5
    if dbias is not None:            # row D of the stacked product = column sums of dz = the LSTM bias gradient
        want = dz.double().sum(0)
        assert float((dbias.double() - want).abs().max() / want.abs().max()) < 2e-5


def _split3(x, ld_out):
    """fp32 [rows][cols] -> three bf16 [rows][ld_out] matrices through indnet_cast_bf16_split3."""
    from induction_b200._lib import LIB, current_stream, ptr
    rows, cols = x.shape
    parts = [torch.full((rows, ld_out), 7.0, dtype=torch.bfloat16, device="cuda") for _ in range(3)]
    LIB.call("indnet_cast_bf16_split3", current_stream(), ptr(x), x.stride(0), rows, cols, ld_out,
             ptr(parts[0]), ptr(parts[1]), ptr(parts[2]))
    return parts


def test_split3_reconstructs_fp32_and_zero_pads():
    g = torch.Generator(device="cuda").manual_seed(4)
    x = torch.randn(37, 53, generator=g, device="cuda") * torch.logspace(-6, 6, 53, device="cuda")
    hi, mid, lo = _split3(x, 56)
    torch.cuda.synchronize()
    back = hi[:, :53].double() + mid[:, :53].double() + lo[:, :53].double()
    rel = ((back - x.double()).abs() / x.double().abs().clamp_min(1e-30)).max()
    assert float(rel) < 2 ** -22, float(rel)                  # 3 x 8 mantissa bits cover fp32's 24
    for p in (hi, mid, lo):
        assert float(p[:, 53:].abs().max()) == 0.0


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("M,N,K,act", [(128, 128, 64, 0), (320, 200, 136, 0), (25, 504, 256, 1), (192, 512, 4096, 0)])
def test_tc_gemm3_is_fp32_accurate(a_mn, b_mn, M, N, K, act):
    """The split GEMM against a float64 product of the SAME fp32 operands: the error budget is fp32
    accumulation (K * 2^-24 relative), not bf16 rounding of the inputs (2^-9)."""
    from induction_b200._lib import LIB, current_stream, ptr
    if a_mn and M % 8:
        pytest.skip("MN-major A needs a leading dimension that is a multiple of 8 elements")
    g = torch.Generator(device="cuda").manual_seed(M * 3 + N + K)
    A = torch.randn(*((K, M) if a_mn else (M, K)), generator=g, device="cuda")
    B = torch.randn(*((K, N) if b_mn else (N, K)), generator=g, device="cuda")
    if act:
        A = A / K ** 0.5
    bias = torch.randn(N, generator=g, device="cuda") if act else None
    want = (A.double().t() if a_mn else A.double()) @ (B.double() if b_mn else B.double().t())
    lda = (A.shape[1] + 7) // 8 * 8
    ldb = (B.shape[1] + 7) // 8 * 8
    Ah, Am, Al = _split3(A, lda)
    Bh, Bm, Bl = _split3(B, ldb)
    C = torch.full((M, N), float("nan"), device="cuda")
    LIB.call("indnet_tc_gemm3", current_stream(), a_mn, b_mn, M, N, K, ptr(Ah), ptr(Am), ptr(Al), lda,
             ptr(Bh), ptr(Bm), ptr(Bl), ldb, ptr(C), N, ptr(bias), act, 0, 1, 0, 0, M * N, 1)
    torch.cuda.synchronize()
    if act:
        want = torch.tanh(want + bias.double())
        # arguments up to ~5 in magnitude: the TMEM accumulator truncates, so ~1e-6 relative on the sum
        assert float((C.double() - want).abs().max()) < 1e-5
    else:
        err = float((C.double() - want).abs().max() / want.abs().max())
        assert err < 2e-6, err
        bf = (A.to(torch.bfloat16).double().t() if a_mn else A.to(torch.bfloat16).double()) @ \
             (B.to(torch.bfloat16).double() if b_mn else B.to(torch.bfloat16).double().t())
        assert float((bf - want).abs().max() / want.abs().max()) > 50 * err     # and far better than one bf16 pass


def test_tc_gemm3_split_k_and_accumulate():
    from induction_b200._lib import LIB, current_stream, ptr
    M, N, K = 192, 512, 20000
    g = torch.Generator(device="cuda").manual_seed(8)
    A = torch.randn(K, M, generator=g, device="cuda")
    B = torch.randn(K, N, generator=g, device="cuda")
    Ah, Am, Al = _split3(A, M)
    Bh, Bm, Bl = _split3(B, N)
    base = torch.randn(M, N, generator=g, device="cuda")
    C = base.clone()
    LIB.call("indnet_tc_gemm3", current_stream(), 1, 1, M, N, K, ptr(Ah), ptr(Am), ptr(Al), M,
             ptr(Bh), ptr(Bm), ptr(Bl), N, ptr(C), N, None, 0, 1, 1, 0, 0, M * N, 0)
    torch.cuda.synchronize()
    want = A.double().t() @ B.double() + base.double()
    assert float((C.double() - want).abs().max() / want.abs().max()) < 3e-6
