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
