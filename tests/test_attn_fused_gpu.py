"""GPU: the fused attention kernels (indnet_attn_fused_fwd / _bwd, csrc/attn_fused.cu) on their own, against a
float64 torch restatement of model/layer_module.py:49-98 on the same bf16-rounded operands.  Covers both forward
variants (H <= 256: persistent, Wa1 resident; H = 512: two passes through the operand ring), every pass width, ragged
sentence counts, padded sentences, and the support/query split of the outputs."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(h, wa1, ba1, w2, mask):
    """h [X, L, H] (f64), Wa1 [H, H] ([in][out]), mask [X, L] -> proj, alphas, enc."""
    proj = torch.tanh(h @ wa1 + ba1)
    score = proj @ w2
    score = score + (1.0 - mask) * (-2147483648.0)
    alphas = torch.softmax(score, dim=1)
    enc = (alphas.unsqueeze(-1) * h).sum(dim=1)
    return proj, alphas, enc


def _case(X, L, H, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    h = (torch.randn(X, L, H, generator=g, device="cuda") * 0.5).to(torch.bfloat16)
    wa1 = (torch.randn(H, H, generator=g, device="cuda") / H ** 0.5).to(torch.bfloat16)
    ba1 = torch.randn(H, generator=g, device="cuda") * 0.1
    w2 = torch.randn(H, generator=g, device="cuda") / H ** 0.5
    lens = torch.randint(1, L + 1, (X,), generator=g, device="cuda")
    mask = (torch.arange(L, device="cuda")[None, :] < lens[:, None]).to(torch.int32)
    return h, wa1, ba1, w2, mask


def _split(enc, group, split):
    """[X, H] in sentence order -> (a, b) in the kernels' support / query layout."""
    X, H = enc.shape
    e = enc.view(X // group, group, H)
    return e[:, :split].reshape(-1, H), e[:, split:].reshape(-1, H)


@pytest.mark.parametrize("X,L,H,group,split", [
    (30, 40, 256, 10, 5),       # the training shape: 3 sentences per CTA, two 128-column passes
    (7, 40, 256, 7, 7),         # one output, a ragged last tile
    (12, 16, 128, 6, 2),        # 8 sentences per CTA, one pass
    (9, 25, 192, 3, 1),         # three 64-column passes
    (6, 128, 64, 2, 1),         # one sentence per CTA
    (8, 128, 512, 4, 2),        # config-5 shape: the ring variant, two passes over the same accumulator
])
def test_attn_fused_forward(X, L, H, group, split):
    from induction_b200._lib import LIB, current_stream, ptr
    h, wa1, ba1, w2, mask = _case(X, L, H, seed=X + L + H)
    proj = torch.zeros(X * L, H, dtype=torch.bfloat16, device="cuda")
    na, nb = X // group * split, X // group * (group - split)
    ea = torch.full((na, H), float("nan"), device="cuda")
    eb = torch.full((max(nb, 1), H), float("nan"), device="cuda")
    ea_bf = torch.zeros(na, H, dtype=torch.bfloat16, device="cuda")
    eb_bf = torch.zeros(max(nb, 1), H, dtype=torch.bfloat16, device="cuda")
    alphas = torch.full((X, L), float("nan"), device="cuda")
    LIB.call("indnet_attn_fused_fwd", current_stream(), ptr(h), ptr(wa1), ptr(ba1), ptr(w2), ptr(mask), ptr(proj),
             ptr(ea), ptr(eb) if nb else None, ptr(ea_bf), ptr(eb_bf) if nb else None, ptr(alphas), X, L, H, group, split)
    torch.cuda.synchronize()
    rp, ra, re = _ref(h.double(), wa1.double(), ba1.double(), w2.double(), mask.double())
    assert float((proj.view(X, L, H).double() - rp).abs().max()) < 1.2e-2          # bf16 store of tanh.approx
    assert float((alphas.double() - ra).abs().max()) < 2e-3
    assert float(alphas.sum(dim=1).sub(1).abs().max()) < 1e-5
    assert bool((alphas[mask == 0] == 0).all())                                     # padded positions get no weight
    wa, wb = _split(re, group, split)
    scale = float(re.abs().max())
    assert float((ea.double() - wa).abs().max()) < 5e-3 * scale
    assert float((ea_bf.double() - wa).abs().max()) < 1.5e-2 * scale
    if nb:
        assert float((eb.double() - wb).abs().max()) < 5e-3 * scale


@pytest.mark.parametrize("X,L,H,group,split", [(30, 40, 256, 10, 5), (7, 40, 256, 7, 7), (12, 16, 128, 6, 2), (9, 25, 192, 3, 1)])
def test_attn_fused_backward(X, L, H, group, split):
    from induction_b200._lib import LIB, current_stream, ptr
    h, wa1, ba1, w2, mask = _case(X, L, H, seed=2 * X + L + H)
    g = torch.Generator(device="cuda").manual_seed(5)
    denc = torch.randn(X, H, generator=g, device="cuda")
    hd = h.double().requires_grad_(True)
    w1d, bd, w2d = wa1.double().requires_grad_(True), ba1.double().requires_grad_(True), w2.double().requires_grad_(True)
    rp, ra, re = _ref(hd, w1d, bd, w2d, mask.double())
    (re * denc.double()).sum().backward()
    # the kernel's inputs: proj / alphas as the forward kernel leaves them
    proj = rp.detach().to(torch.bfloat16).reshape(X * L, H).contiguous()
    alphas = ra.detach().float().contiguous()
    da, db = _split(denc, group, split)
    da, db = da.contiguous(), db.contiguous()
    dpre = torch.zeros(X * L, H, dtype=torch.bfloat16, device="cuda")
    dh = torch.zeros(X * L, H, dtype=torch.bfloat16, device="cuda")
    dw2 = torch.full((H,), float("nan"), device="cuda")
    dba = torch.full((H,), float("nan"), device="cuda")
    LIB.call("indnet_attn_fused_bwd", current_stream(), ptr(da), ptr(db) if db.numel() else None, ptr(h), ptr(proj),
             ptr(alphas), ptr(w2), ptr(wa1), ptr(dpre), ptr(dh), ptr(dw2), ptr(dba), X, L, H, group, split, 0)
    torch.cuda.synchronize()
    want_dh = hd.grad.reshape(X * L, H)
    assert float((dh.double() - want_dh).abs().max()) < 2e-2 * float(want_dh.abs().max())
    e_w2 = float((dw2.double() - w2d.grad).abs().max()) / float(w2d.grad.abs().max())
    e_ba = float((dba.double() - bd.grad).abs().max()) / float(bd.grad.abs().max())
    print("attn bwd", (X, L, H), "dw2 %.1e dba %.1e" % (e_w2, e_ba))
    # sums over all tokens of terms proportional to dscore_t (which sums to zero inside a sentence): what survives the
    # cancellation is compared with the noise of the bf16-stored proj - see tests/_cases.py BF16_ATTN_GRAD_TOL
    assert e_w2 < 5e-2 and e_ba < 1e-1
    # dpre is the operand of the Dense-kernel gradient GEMM: h^T dpre must reproduce dWa1
    dw1 = h.double().reshape(X * L, H).t() @ dpre.double()
    assert float((dw1 - w1d.grad).abs().max()) < 5e-2 * float(w1d.grad.abs().max())
    # accumulate = 1 adds to what the caller put there
    dw2b = torch.ones(H, device="cuda")
    dbab = torch.full((H,), 2.0, device="cuda")
    LIB.call("indnet_attn_fused_bwd", current_stream(), ptr(da), ptr(db) if db.numel() else None, ptr(h), ptr(proj),
             ptr(alphas), ptr(w2), ptr(wa1), ptr(dpre), ptr(dh), ptr(dw2b), ptr(dbab), X, L, H, group, split, 1)
    torch.cuda.synchronize()
    assert float((dw2b - 1.0 - dw2).abs().max()) < 1e-4 * float(dw2.abs().max()) + 1e-6
    assert float((dbab - 2.0 - dba).abs().max()) < 1e-4 * float(dba.abs().max()) + 1e-6
