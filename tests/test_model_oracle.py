"""CPU: the model oracle against (a) an independent explicit-loop NumPy restatement, (b) closed-form
known answers derivable from the reference code, (c) the frozen golden fixtures, (d) finite differences."""
import os

import numpy as np
import pytest
import torch

from _cases import CASES, make_case, oracle_forward, oracle_grads
from oracle import induction_numpy as onp
from oracle import induction_oracle as orc


@pytest.mark.parametrize("name", ["smoke", "k1", "ragged"])
def test_torch_oracle_matches_numpy_restatement(name):
    cfg, tables, inputs, label, params = make_case(name, seed=3)
    out = oracle_forward(cfg, tables, inputs, label, params)
    pnp = {k: v.numpy() for k, v in params.items()}
    losses = []
    for b in range(cfg.B):
        ep = onp.forward_episode(pnp, tables["word_vec_mat"], inputs["word"][b], inputs["pos1"][b],
                                 inputs["pos2"][b], inputs["mask"][b], label[b], cfg.N, cfg.K, cfg.Q)
        np.testing.assert_allclose(out["probs"][b].numpy(), ep["probs"], rtol=1e-10, atol=1e-12)
        np.testing.assert_allclose(out["class_vector"][b].numpy(), ep["class_vector"], rtol=1e-10, atol=1e-12)
        np.testing.assert_allclose(out["encoder"][b].numpy(), ep["encoder"], rtol=1e-10, atol=1e-12)
        np.testing.assert_allclose(out["alphas"][b].numpy(), ep["alphas"], rtol=1e-10, atol=1e-14)
        assert np.array_equal(out["predict_test"][b].numpy(), ep["predict_test"])
        losses.append(ep["loss"])
    assert abs(float(out["loss"]) - np.mean(losses)) < 1e-12


@pytest.mark.parametrize("name", ["smoke", "k1", "ragged"])
def test_oracle_matches_frozen_golden(name, golden_dir):
    z = np.load(os.path.join(golden_dir, "model_%s.npz" % name))
    cfg, tables, inputs, label, params = make_case(name, seed=3)
    for k, v in inputs.items():
        assert np.array_equal(z["in_" + k], v)
    out, grads = oracle_grads(cfg, tables, inputs, label, params)
    for k in ("hout", "encoder", "alphas", "class_vector", "probs", "loss"):
        np.testing.assert_allclose(out[k].detach().numpy(), z["o_" + k], rtol=1e-9, atol=1e-12)
    for k, g in grads.items():
        np.testing.assert_allclose(g.numpy(), z["g_" + k.replace("/", "__")], rtol=1e-7, atol=1e-12)


def test_known_answers():
    cfg, tables, inputs, label, params = make_case("k1", seed=3)
    out = oracle_forward(cfg, tables, inputs, label, params)
    # K = 1: softmax over one element == 1, so c = squash(e W_s) for any number of iterations (:127)
    e_s = out["encoder"][:, : cfg.N * cfg.K].reshape(cfg.B, cfg.N, cfg.H)
    want = orc.squash(e_s @ params["W_s"].double())
    np.testing.assert_allclose(out["class_vector"].numpy(), want.numpy(), rtol=1e-12)
    # |c| < 1 (:160), probs in (0,1) (:46), masked positions carry exactly zero attention (:84-87)
    assert float(out["class_vector"].norm(dim=-1).max()) < 1.0
    assert 0.0 < float(out["probs"].min()) and float(out["probs"].max()) < 1.0
    m = torch.from_numpy(inputs["mask"]).reshape(cfg.B, cfg.n_s, cfg.L, 1)
    assert float(out["alphas"][m == 0].abs().max()) == 0.0
    np.testing.assert_allclose(out["alphas"].sum(dim=2).numpy(), 1.0, rtol=1e-12)
    # loss is a mean, not the paper's sum (graph.py:94)
    onehot = torch.nn.functional.one_hot(torch.from_numpy(label), cfg.N).double()
    assert abs(float(out["loss"]) - float(((out["probs"] - onehot) ** 2).sum() / onehot.numel())) < 1e-15


def test_routing_first_iteration_is_uniform_coupling():
    torch.manual_seed(0)
    e = torch.randn(4, 3, 6, dtype=torch.float64)
    W = torch.randn(6, 6, dtype=torch.float64)
    c1 = orc.dynamic_routing(W, e, torch.zeros(4, 3, dtype=torch.float64), iter_routing=1)
    np.testing.assert_allclose(c1.numpy(), orc.squash((e @ W).mean(dim=1)).numpy(), rtol=1e-12)


def test_gradients_match_finite_differences():
    cfg, tables, inputs, label, params = make_case("smoke", seed=5)
    _, grads = oracle_grads(cfg, tables, inputs, label, params)
    rs = np.random.RandomState(0)
    for name in ("W_s", "M", "self_attn/W_a2", "bidirectional_rnn/bw/basic_lstm_cell/kernel", "pos1_embedding", "w"):
        base = params[name].double()
        for _ in range(3):
            idx = tuple(int(rs.randint(0, s)) for s in base.shape)
            eps = 1e-6
            vals = []
            for sgn in (+1, -1):
                p2 = {k: v.double().clone() for k, v in params.items()}
                p2[name][idx] += sgn * eps
                vals.append(float(oracle_forward(cfg, tables, inputs, label, p2)["loss"]))
            fd = (vals[0] - vals[1]) / (2 * eps)
            assert abs(fd - float(grads[name][idx])) < 1e-7 + 1e-5 * abs(fd), (name, idx)


def test_tf_adam_form():
    p = {"a": torch.tensor([1.0, -2.0], dtype=torch.float64)}
    g = {"a": torch.tensor([0.5, -0.25], dtype=torch.float64)}
    m = {"a": torch.zeros(2, dtype=torch.float64)}
    v = {"a": torch.zeros(2, dtype=torch.float64)}
    lr = orc.tf_adam_step(p, g, m, v, global_step=0)
    # first step: m = .1 g, v = .001 g^2, lr_t = lr*sqrt(.001)/.1 -> update = lr * g/(|g| + eps*sqrt(1000)...)
    lr_t = 1e-3 * np.sqrt(1 - 0.999) / (1 - 0.9)
    want = np.array([1.0, -2.0]) - lr_t * (0.1 * np.array([0.5, -0.25])) / (np.sqrt(0.001 * np.array([0.25, 0.0625])) + 1e-8)
    np.testing.assert_allclose(p["a"].numpy(), want, rtol=1e-14)
    assert lr == 1e-3
    assert orc.decayed_lr(1e-3, 250) == pytest.approx(1e-3 * 0.9 ** 2)


# ---------------------------------------------------------------------------------------------------
# External pins: code nobody in this repository wrote (torch.nn / torch.nn.functional, CPU, float64)
# must reproduce the oracle's restatement of the TF ops.  What these cannot pin is listed in DESIGN.md
# section 2 (the routing loop and the slice-major FC ordering have no third-party implementation).
def _tf_to_torch_lstm(lstm, kernel, bias, D, u, suffix=""):
    """TF BasicLSTMCell (model/graph.py:34-47): one kernel [D+u, 4u], column blocks i, j, f, o, forget
    bias 1.0 added at run time.  torch.nn.LSTM: weight_ih [4u, D] / weight_hh [4u, u], row blocks
    i, f, g, o, two bias vectors."""
    blocks = torch.split(kernel, u, dim=1)                     # i, j, f, o
    re = torch.cat([blocks[0], blocks[2], blocks[1], blocks[3]], dim=1)      # i, f, g(=j), o
    b = torch.split(bias, u)
    rb = torch.cat([b[0], b[2], b[1], b[3]])
    fb = torch.zeros(4 * u, dtype=kernel.dtype)
    fb[u:2 * u] = 1.0                                          # forget_bias = 1.0 (rnn_cell_impl default)
    with torch.no_grad():
        getattr(lstm, "weight_ih_l0" + suffix).copy_(re[:D].t())
        getattr(lstm, "weight_hh_l0" + suffix).copy_(re[D:].t())
        getattr(lstm, "bias_ih_l0" + suffix).copy_(rb)
        getattr(lstm, "bias_hh_l0" + suffix).copy_(fb)


@pytest.mark.parametrize("name", ["smoke", "ragged", "u24"])
def test_bilstm_pinned_to_torch_nn_lstm(name):
    cfg, tables, inputs, label, params = make_case(name, seed=9)
    p = {k: v.double() for k, v in params.items()}
    wt = torch.from_numpy(tables["word_vec_mat"]).double()
    w = {k: torch.from_numpy(np.ascontiguousarray(v)).reshape(cfg.X, cfg.L) for k, v in inputs.items()}
    x = orc.get_embedding(p, wt, w["word"], w["pos1"], w["pos2"])
    want = orc.bilstm(p, x)
    lstm = torch.nn.LSTM(cfg.D, cfg.u, batch_first=True, bidirectional=True).double()
    _tf_to_torch_lstm(lstm, p["bidirectional_rnn/fw/basic_lstm_cell/kernel"], p["bidirectional_rnn/fw/basic_lstm_cell/bias"], cfg.D, cfg.u)
    _tf_to_torch_lstm(lstm, p["bidirectional_rnn/bw/basic_lstm_cell/kernel"], p["bidirectional_rnn/bw/basic_lstm_cell/bias"], cfg.D, cfg.u, "_reverse")
    with torch.no_grad():
        got, _ = lstm(x)
    # no sequence_length in the reference (graph.py:44-47): both run all L steps, the backward direction
    # starts on the padding - exactly torch's unpacked bidirectional LSTM
    np.testing.assert_allclose(got.numpy(), want.numpy(), rtol=0, atol=1e-12)


def test_ntn_pinned_to_torch_bilinear():
    torch.manual_seed(3)
    C, n_q, H, S = 4, 7, 6, 5
    cv, qe = torch.randn(C, H, dtype=torch.float64), torch.randn(n_q, H, dtype=torch.float64)
    p = {"M": torch.randn(H, H, S, dtype=torch.float64), "w": torch.randn(C * S, C, dtype=torch.float64),
         "b": torch.randn(C, dtype=torch.float64)}
    want = orc.neural_tensor_layer(p, cv, qe)
    # layer_module.py:20-32: slice k gives cv M[:,:,k] qe^T  ==  F.bilinear with weight[k] = M[:,:,k]
    x1 = cv.unsqueeze(0).expand(n_q, C, H).reshape(n_q * C, H)
    x2 = qe.unsqueeze(1).expand(n_q, C, H).reshape(n_q * C, H)
    bil = torch.nn.functional.bilinear(x1, x2, p["M"].permute(2, 0, 1).contiguous()).reshape(n_q, C, S)
    # concat(axis=0) of the S [C, n_q] slices then transpose (:35-37): column index k*C + c
    v = torch.relu(bil.permute(0, 2, 1).reshape(n_q, S * C))
    got = torch.sigmoid(torch.nn.functional.linear(v, p["w"].t(), p["b"]))
    np.testing.assert_allclose(got.numpy(), want.numpy(), rtol=1e-12, atol=1e-14)


def test_attention_pinned_to_torch_functional():
    torch.manual_seed(4)
    X, L, H = 5, 9, 8
    h = torch.randn(X, L, H, dtype=torch.float64)
    mask = torch.randint(0, 4, (X, L))
    mask[:, 0] = 1                                             # every sentence keeps at least one position
    p = {"self_attn/dense/kernel": torch.randn(H, H, dtype=torch.float64), "self_attn/dense/bias": torch.randn(H, dtype=torch.float64),
         "self_attn/W_a2": torch.randn(H, 1, dtype=torch.float64)}
    want_e, want_a = orc.self_attention(p, h, mask)
    F = torch.nn.functional
    score = F.linear(torch.tanh(F.linear(h, p["self_attn/dense/kernel"].t(), p["self_attn/dense/bias"])), p["self_attn/W_a2"].t())
    # adding (1-m)*(-2^31) before the softmax (layer_module.py:84-87) == masking with -inf whenever one
    # position survives: exp underflows to exactly 0 either way
    a = torch.softmax(score.masked_fill((mask == 0).unsqueeze(-1), float("-inf")), dim=1)
    e = torch.bmm(h.transpose(1, 2), a).squeeze(-1)
    np.testing.assert_allclose(a.numpy(), want_a.numpy(), rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(e.numpy(), want_e.numpy(), rtol=1e-12, atol=1e-14)


def test_loss_pinned_to_torch_mse():
    cfg, tables, inputs, label, params = make_case("k1", seed=3)
    out = oracle_forward(cfg, tables, inputs, label, params)
    onehot = torch.nn.functional.one_hot(torch.from_numpy(label), cfg.N).double()
    assert abs(float(torch.nn.functional.mse_loss(out["probs"], onehot)) - float(out["loss"])) < 1e-15


def test_tf_adam_pinned_to_torch_adam_up_to_epsilon_placement():
    """torch.optim.Adam puts eps inside the bias correction (sqrt(v_hat) + eps), TF outside
    (sqrt(v) + eps with lr_t carrying the correction): with eps -> 0 both are the same
    update, which pins m, v, the bias corrections and the staircase learning rate."""
    torch.manual_seed(0)
    p0 = torch.randn(7, dtype=torch.float64)
    pt = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([pt], lr=1.0, betas=(0.9, 0.999), eps=1e-300)
    p = {"a": p0.clone()}
    m = {"a": torch.zeros(7, dtype=torch.float64)}
    v = {"a": torch.zeros(7, dtype=torch.float64)}
    for step in range(205):
        g = torch.sin(torch.arange(7, dtype=torch.float64) + step) + 0.1
        for grp in opt.param_groups:
            grp["lr"] = orc.decayed_lr(1e-3, step)
        pt.grad = g.clone()
        opt.step()
        orc.tf_adam_step(p, {"a": g}, m, v, global_step=step, eps=1e-300)
    np.testing.assert_allclose(p["a"].numpy(), pt.detach().numpy(), rtol=1e-12)
    assert orc.decayed_lr(1e-3, 99) == 1e-3 and orc.decayed_lr(1e-3, 100) == pytest.approx(9e-4)
