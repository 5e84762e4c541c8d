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
    LIB.call("indnet_tc_gemm_lstm_dw", current_stream(), ptr(x), ptr(h), ptr(dz), ptr(out), ptr(dbias), T, D, u, 0)
    torch.cuda.synchronize()
    ref = torch.cat([x[:, :D].double().t() @ dz.double(), h.double().t() @ dz.double()], 0)
    assert float((out.double() - ref).abs().max() / ref.abs().max()) < 2e-5
    if dbias is not None:            # row D of the stacked product = column sums of dz = the LSTM bias gradient
        want = dz.double().sum(0)
        assert float((dbias.double() - want).abs().max() / want.abs().max()) < 2e-5
