"""GPU: the reference-facing functional API (model/layer_module.py signatures) against the oracle,
forward and autograd, including the implicit variable store."""
import numpy as np
import pytest
import torch

from _cases import rel_err

pytestmark = pytest.mark.gpu


def _store(shapes_seed=0):
    from induction_b200.model import layer_module as lm
    return lm, lm.VariableStore(seed=shapes_seed)


def test_smoke_block_shapes_like_reference_main():
    """The shapes of the reference's __main__ smoke block (layer_module.py:165-202)."""
    from oracle import induction_oracle as orc
    lm, store = _store(1)
    rs = np.random.RandomState(0)
    inputs = torch.tensor(rs.random_sample((24, 5, 10)), dtype=torch.float32, device="cuda")
    mask = np.ones((24, 5), dtype=np.int32)
    mask[0][4:] = 0; mask[1][3:] = 0; mask[2][2:] = 0
    mask_t = torch.from_numpy(mask).cuda()
    with lm.variable_scope(store):
        encoder, att = lm.self_attention(inputs, mask_t)
        support = encoder[:9].reshape(3, 3, -1)
        query = encoder[9:24]
        cv = lm.dynamic_routing(support, torch.zeros(3, 3, device="cuda"))
        probs = lm.neural_tensor_layer(cv, query, out_size=10)
    assert encoder.shape == (24, 10) and att.shape == (24, 5, 1) and cv.shape == (3, 10) and probs.shape == (15, 3)
    p = {k: v.detach().cpu().double() for k, v in store.vars.items()}
    enc_ref, att_ref = orc.self_attention(p, inputs.cpu().double(), mask_t.cpu())
    cv_ref = orc.dynamic_routing(p["W_s"], enc_ref[:9].reshape(3, 3, -1), torch.zeros(3, 3, dtype=torch.float64))
    probs_ref = orc.neural_tensor_layer(p, cv_ref, enc_ref[9:])
    assert rel_err(encoder, enc_ref) < 1e-5 and rel_err(att, att_ref) < 1e-5
    assert rel_err(cv, cv_ref) < 1e-5 and rel_err(probs, probs_ref) < 1e-5
    assert float(att[0, 4:].abs().max()) == 0.0 and float(att[2, 2:].abs().max()) == 0.0


def test_autograd_through_the_functional_ops():
    from oracle import induction_oracle as orc
    lm, store = _store(2)
    rs = np.random.RandomState(1)
    x = torch.tensor(rs.standard_normal((14, 7, 12)), dtype=torch.float32, device="cuda", requires_grad=True)
    mask = torch.from_numpy((rs.random_sample((14, 7)) > 0.2).astype(np.int32)).cuda()
    mask[:, 0] = 1
    y = torch.tensor(rs.random_sample((8, 2)), dtype=torch.float32, device="cuda")
    with lm.variable_scope(store):
        enc, _ = lm.self_attention(x, mask)
        cv = lm.dynamic_routing(enc[:6].reshape(2, 3, 12), torch.zeros(2, 3, device="cuda"))
        probs = lm.neural_tensor_layer(cv, enc[6:], out_size=9)
        loss = ((probs - y) ** 2).mean() + lm.squash(cv).sum() * 0.1
    loss.backward()
    p = {k: v.detach().cpu().double().requires_grad_(True) for k, v in store.vars.items()}
    xr = x.detach().cpu().double().requires_grad_(True)
    enc_r, _ = orc.self_attention(p, xr, mask.cpu())
    cv_r = orc.dynamic_routing(p["W_s"], enc_r[:6].reshape(2, 3, 12), torch.zeros(2, 3, dtype=torch.float64))
    probs_r = orc.neural_tensor_layer(p, cv_r, enc_r[6:])
    loss_r = ((probs_r - y.cpu().double()) ** 2).mean() + orc.squash(cv_r).sum() * 0.1
    loss_r.backward()
    assert abs(float(loss) - float(loss_r)) < 1e-5 * abs(float(loss_r))
    assert rel_err(x.grad, xr.grad) < 2e-4
    for k, v in store.vars.items():
        assert rel_err(v.grad, p[k].grad) < 2e-4, k


def test_variable_store_reuse_and_shape_clash():
    lm, store = _store(3)
    with lm.variable_scope(store):
        a = lm.get_variable("W_s", (8, 8))
        assert lm.get_variable("W_s", (8, 8)) is a
        with pytest.raises(ValueError, match="Trying to share variable"):
            lm.get_variable("W_s", (4, 4))


def test_error_conventions():
    lm, _ = _store(4)
    with pytest.raises(Exception, match="CUDA"):
        lm.squash(torch.zeros(2, 3))
    with pytest.raises(ValueError):
        lm.dynamic_routing(torch.zeros(2, 3, 4, device="cuda"), torch.zeros(2, 4, device="cuda"))
    from induction_b200._lib import IndnetError
    with pytest.raises(IndnetError, match="limit of 32"):
        lm.dynamic_routing(torch.zeros(2, 33, 4, device="cuda"), torch.zeros(2, 33, device="cuda"))


def test_empty_batch_is_a_noop():
    from induction_b200._lib import LIB, current_stream
    LIB.call("indnet_squash_fwd", current_stream(), None, None, 0, 8)
    LIB.call("indnet_lstm_fwd", current_stream(), None, None, None, None, None, None, None, 0, 4, 8, 1.0, 0, None)
    torch.cuda.synchronize()
