"""GPU parity tests proper: the CUDA path (through the C-ABI) against the CPU oracle on the same seeded
inputs.  Tolerances (BASELINE.json north_star): FP32 mode - logits/probs and loss within 1e-5
relative of the reference graph restatement; gradients within 2e-4 of the largest gradient entry
(float32 accumulation over up to X*L-long reductions, float64 oracle)."""
import os

import numpy as np
import pytest
import torch

from _cases import CASES, bf16_grad_check, make_case, oracle_forward, oracle_grads, rel_err

pytestmark = pytest.mark.gpu

FWD_TOL = 1e-5
GRAD_TOL = 2e-4


def build_pipeline(cfg, tables, params, seed=0):
    from induction_b200.engine import EpisodePipeline, ParamStore, tf_param_shapes
    store = ParamStore(tf_param_shapes(cfg.N, cfg.L, cfg.E, cfg.u, cfg.S))
    store.load_state_dict(params)
    wt = torch.from_numpy(tables["word_vec_mat"])
    return EpisodePipeline(cfg, store, wt, seed=seed)


@pytest.mark.parametrize("name", list(CASES))
def test_forward_parity(name):
    cfg, tables, inputs, label, params = make_case(name, seed=3)
    pipe = build_pipeline(cfg, tables, params)
    pipe.set_inputs(inputs, label)
    pipe.forward(train=False)
    torch.cuda.synchronize()
    ref = oracle_forward(cfg, tables, inputs, label, params, dtype=torch.float64)
    enc = torch.cat([pipe.es, pipe.eq], dim=1)
    errs = {
        "hout": rel_err(pipe.hout.view(cfg.X, cfg.L, cfg.H), ref["hout"]),
        "alphas": rel_err(pipe.alphas.view(cfg.B, cfg.n_s, cfg.L, 1), ref["alphas"]),
        "encoder": rel_err(enc, ref["encoder"]),
        "class_vector": rel_err(pipe.cvec.view(cfg.B, cfg.N, cfg.H), ref["class_vector"]),
        "probs": rel_err(pipe.probs, ref["probs"]),
    }
    loss_err = abs(pipe.loss() - float(ref["loss"])) / float(ref["loss"])
    print(name, errs, "loss_err", loss_err)
    for k, e in errs.items():
        assert e < FWD_TOL, (name, k, e)
    assert loss_err < FWD_TOL
    assert abs(pipe.accuracy() - float(ref["accuracy"])) < 1e-6
    # masked positions: exactly zero attention weight
    m = torch.from_numpy(inputs["mask"]).reshape(cfg.X, cfg.L)
    assert float(pipe.alphas.cpu()[m == 0].abs().max() if (m == 0).any() else 0.0) == 0.0


@pytest.mark.parametrize("name", ["smoke", "k1", "ragged"])
def test_forward_matches_frozen_golden(name, golden_dir):
    z = np.load(os.path.join(golden_dir, "model_%s.npz" % name))
    cfg, tables, inputs, label, params = make_case(name, seed=3)
    pipe = build_pipeline(cfg, tables, params)
    pipe.set_inputs(inputs, label)
    pipe.forward(train=False)
    assert rel_err(pipe.probs, z["o_probs"]) < FWD_TOL
    assert abs(pipe.loss() - float(z["o_loss"])) / float(z["o_loss"]) < FWD_TOL


@pytest.mark.parametrize("name", list(CASES))
def test_backward_parity(name):
    cfg, tables, inputs, label, params = make_case(name, seed=3)
    pipe = build_pipeline(cfg, tables, params)
    pipe.set_inputs(inputs, label)
    pipe.forward(train=True)
    pipe.backward()
    torch.cuda.synchronize()
    _, gref = oracle_grads(cfg, tables, inputs, label, params, dtype=torch.float64)
    got = pipe.p.grad_dict()
    errs = {k: rel_err(got[k], gref[k]) for k in gref}
    print(name, {k: "%.2e" % v for k, v in errs.items()})
    for k, e in errs.items():
        assert e < GRAD_TOL, (name, k, e)


def test_dropout_mask_replays_in_oracle():
    from induction_b200.engine import dropout_keep_mask
    cfg, tables, inputs, label, params = make_case("ragged", seed=4, keep_prob=0.9)
    pipe = build_pipeline(cfg, tables, params, seed=1234)
    pipe.set_inputs(inputs, label)
    pipe.forward(train=True)
    pipe.backward()
    mask = dropout_keep_mask(1234, (cfg.X, cfg.L, cfg.H), 0.9, step=0)
    assert 0.85 < mask.mean() < 0.95
    dm = torch.from_numpy(mask)
    ref, gref = oracle_grads(cfg, tables, inputs, label, params, drop_mask=dm)
    assert rel_err(pipe.hout.view(cfg.X, cfg.L, cfg.H), ref["hout"]) < FWD_TOL
    assert rel_err(pipe.probs, ref["probs"]) < FWD_TOL
    got = pipe.p.grad_dict()
    for k in gref:
        assert rel_err(got[k], gref[k]) < GRAD_TOL, k
    # eval mode ignores keep_prob (model/base.py:199 feeds 1)
    pipe.forward(train=False)
    ref1 = oracle_forward(cfg, tables, inputs, label, params)
    assert rel_err(pipe.probs, ref1["probs"]) < FWD_TOL


def test_train_steps_follow_tf_adam_trajectory():
    from oracle import induction_oracle as orc
    cfg, tables, inputs, label, params = make_case("toy_train_py", seed=6)
    pipe = build_pipeline(cfg, tables, params)
    pipe.set_inputs(inputs, label)
    p = {k: v.double().clone() for k, v in params.items()}
    m = {k: torch.zeros_like(v) for k, v in p.items()}
    v = {k: torch.zeros_like(x) for k, x in p.items()}
    losses_ref, losses = [], []
    for step in range(3):
        out, g = oracle_grads(cfg, tables, inputs, label, p)
        losses_ref.append(float(out["loss"]))
        orc.tf_adam_step(p, g, m, v, global_step=step)
        pipe.train_step()
        losses.append(pipe.loss())
    assert int(pipe.p.global_step.item()) == 3
    np.testing.assert_allclose(losses, losses_ref, rtol=2e-5)
    got = pipe.p.state_dict()
    for k in p:
        # Adam's first steps move every weight by ~lr regardless of gradient size, so compare in
        # absolute terms against the step size (3 steps x 1e-3)
        assert float((got[k].cpu().double() - p[k]).abs().max()) < 3e-3 * 2e-2, k


def test_trainable_word_embedding_gradient():
    from induction_b200.engine import EpisodePipeline, ParamStore, init_tf_params, tf_param_shapes
    cfg, tables, inputs, label, params = make_case("smoke", seed=8, vocab=40)
    shapes = tf_param_shapes(cfg.N, cfg.L, cfg.E, cfg.u, cfg.S, vocab_rows=42)
    full = dict(params)
    full["word_embedding"] = torch.from_numpy(tables["word_vec_mat"]).clone()
    store = ParamStore(shapes)
    store.load_state_dict(full)
    pipe = EpisodePipeline(cfg, store, None)
    pipe.set_inputs(inputs, label)
    pipe.forward(train=True)
    pipe.backward()
    # oracle: make the word table a leaf
    from oracle import induction_oracle as orc
    from _cases import to_torch_inputs
    p = {k: v.double().requires_grad_(True) for k, v in full.items()}
    ti, tl = to_torch_inputs(inputs, label)
    out = orc.forward(p, p["word_embedding"], ti["word"], ti["pos1"], ti["pos2"], ti["mask"], tl, cfg.N, cfg.K, cfg.Q)
    out["loss"].backward()
    got = store.grad_dict()
    for k in p:
        assert rel_err(got[k], p[k].grad) < GRAD_TOL, k


def test_trainable_word_embedding_gradient_bf16_tensor_core_path():
    """No pred_embed (model/base.py:62-65): the word table is a trainable, so the BPTT kernel cannot keep only the
    position columns of dx - the two dz * W_x^T GEMMs and the full embedding scatter run instead (also the
    INDNET_NO_FUSED_DX path)."""
    from induction_b200.engine import EpisodePipeline, ParamStore, tf_param_shapes
    cfg, tables, inputs, label, params = make_case("cfg0_5w5s", seed=9, vocab=300, mode="bf16")
    rows = tables["word_vec_mat"].shape[0]
    shapes = tf_param_shapes(cfg.N, cfg.L, cfg.E, cfg.u, cfg.S, vocab_rows=rows)
    full = dict(params)
    full["word_embedding"] = torch.from_numpy(tables["word_vec_mat"]).clone()
    store = ParamStore(shapes)
    store.load_state_dict(full)
    pipe = EpisodePipeline(cfg, store, None)
    assert pipe.tc_lstm and pipe.trainable_words
    pipe.set_inputs(inputs, label)
    pipe.forward(train=True)
    pipe.backward()
    torch.cuda.synchronize()
    from oracle import induction_oracle as orc
    from _cases import to_torch_inputs
    p = {k: v.double().requires_grad_(True) for k, v in full.items()}
    ti, tl = to_torch_inputs(inputs, label)
    out = orc.forward(p, p["word_embedding"], ti["word"], ti["pos1"], ti["pos2"], ti["mask"], tl, cfg.N, cfg.K, cfg.Q)
    out["loss"].backward()
    assert rel_err(pipe.probs, out["probs"]) < BF16_TOL
    errs, worst, cos = bf16_grad_check(store.grad_dict(), {k: v.grad for k, v in p.items()}, "trainable words")
    print("trainable words bf16: worst", worst, errs[worst], "cos", cos)


def test_packed_inputs_equal_the_five_separate_copies():
    """pack_inputs(): one host tensor in the layout of the device input buffer = one H2D copy per step."""
    cfg, tables, inputs, label, params = make_case("smoke", seed=4)
    pipe = build_pipeline(cfg, tables, params)
    pipe.set_inputs(inputs, label)
    ref = [t.clone() for t in (pipe.word, pipe.pos1, pipe.pos2, pipe.mask, pipe.label)]
    pipe.inbuf.zero_()
    packed = pipe.pack_inputs(inputs, label)
    assert packed.is_pinned() and packed.dtype == torch.int32
    pipe.set_inputs(packed, None)
    torch.cuda.synchronize()
    for a, b in zip(ref, (pipe.word, pipe.pos1, pipe.pos2, pipe.mask, pipe.label)):
        assert torch.equal(a, b)
    with pytest.raises(ValueError):
        pipe.set_inputs(packed[:-1], None)


def test_large_batch_properties():
    """BASELINE configs[1] at full size (5-way 1-shot, 64 episodes, L=40, u=128): too slow for the
    float64 oracle in a unit test, so check size-independent properties: episode independence (a
    batch equals its episodes run alone), alphas sum to one, |c| < 1, K=1 closed form."""
    from induction_b200.engine import EpisodeConfig
    from induction_b200.util.synthetic import make_tables
    from induction_b200.engine import init_tf_params, tf_param_shapes
    cfg = EpisodeConfig(N=5, K=1, Q=5, B=64, L=40, E=50, u=128, S=100)
    tables = make_tables(16, 40, cfg.L, 3000, cfg.E, seed=0)
    rs = np.random.RandomState(0)
    rows = rs.randint(0, 16 * 40, size=(cfg.B, cfg.n_s))
    inputs = {k: tables[k][rows] for k in ("word", "pos1", "pos2", "mask")}
    label = rs.randint(0, cfg.N, size=(cfg.B, cfg.n_q))
    params = init_tf_params(tf_param_shapes(cfg.N, cfg.L, cfg.E, cfg.u, cfg.S), seed=0)
    pipe = build_pipeline(cfg, tables, params)
    pipe.set_inputs(inputs, label)
    pipe.forward()
    probs = pipe.probs.clone()
    assert torch.isfinite(probs).all()
    np.testing.assert_allclose(pipe.alphas.sum(dim=1).cpu().numpy(), 1.0, rtol=1e-5)
    assert float(pipe.cvec.norm(dim=1).max()) < 1.0
    # K = 1 closed form: c = squash(e_s W_s)
    from oracle import induction_oracle as orc
    want = orc.squash(pipe.es.cpu().double() @ params["W_s"].double()).reshape(cfg.B * cfg.N, cfg.H)
    assert rel_err(pipe.cvec, want) < FWD_TOL
    # episode independence: episodes 5 and 41 alone (B = 1 pipelines) reproduce their rows
    cfg1 = EpisodeConfig(N=5, K=1, Q=5, B=1, L=40, E=50, u=128, S=100)
    for b in (5, 41):
        p1 = build_pipeline(cfg1, tables, params)
        p1.set_inputs({k: v[b:b + 1] for k, v in inputs.items()}, label[b:b + 1])
        p1.forward()
        assert rel_err(p1.probs[0], probs[b]) < FWD_TOL
    # and the oracle agrees on those two episodes (float32 oracle is fast enough for 2 episodes)
    sub = {k: v[[5, 41]] for k, v in inputs.items()}
    cfg2 = EpisodeConfig(N=5, K=1, Q=5, B=2, L=40, E=50, u=128, S=100)
    ref = oracle_forward(cfg2, tables, sub, label[[5, 41]], params, dtype=torch.float64)
    assert rel_err(probs[[5, 41]], ref["probs"]) < FWD_TOL


# ---------------------------------------------------------------------------------------------------
# bf16 tensor-core mode (tcgen05, fp32 accumulation): north_star tolerance 2e-2 on probs / loss.
BF16_TOL = 2e-2
# gradients: per-tensor max-norm bounds + whole-gradient cosine, see tests/_cases.py bf16_grad_check
BF16_CASES = ["u24", "k1", "cfg0_5w5s"]


@pytest.mark.parametrize("name", BF16_CASES)
def test_bf16_mode_forward_and_backward_parity(name):
    cfg, tables, inputs, label, params = make_case(name, seed=3, mode="bf16")
    pipe = build_pipeline(cfg, tables, params)
    pipe.set_inputs(inputs, label)
    pipe.forward(train=True)
    pipe.backward()
    torch.cuda.synchronize()
    ref, gref = oracle_grads(cfg, tables, inputs, label, params, dtype=torch.float64)
    e_probs = rel_err(pipe.probs, ref["probs"])
    e_loss = abs(pipe.loss() - float(ref["loss"])) / float(ref["loss"])
    assert e_probs < BF16_TOL and e_loss < BF16_TOL
    errs, worst, cos = bf16_grad_check(pipe.p.grad_dict(), gref, name)
    print(name, "probs", e_probs, "loss", e_loss, "cos", cos, {k: "%.1e" % v for k, v in errs.items()})


@pytest.mark.parametrize("name", ["cfg1_5w1s_b64", "cfg2_10w5s", "cfg3_10w10s"])
def test_bf16_mode_parity_at_benchmarked_shapes(name):
    """The shapes bench.py times (BASELINE configs[1..3]; configs[1] at its full 64-episode batch = 60
    sentence tiles per direction and the K = 1 routing path) through the tcgen05 recurrence / BPTT /
    stacked weight-gradient kernels, against the float64 oracle on the whole batch."""
    cfg, tables, inputs, label, params = make_case(name, seed=5, mode="bf16", relations=16, per_relation=40, vocab=3000)
    pipe = build_pipeline(cfg, tables, params)
    assert pipe.tc_lstm
    pipe.set_inputs(inputs, label)
    pipe.forward(train=True)
    pipe.backward()
    torch.cuda.synchronize()
    ref, gref = oracle_grads(cfg, tables, inputs, label, params, dtype=torch.float64)
    e_probs = rel_err(pipe.probs, ref["probs"])
    e_hout = rel_err(pipe.hout_bf.float().view(cfg.X, cfg.L, cfg.H), ref["hout"])
    e_loss = abs(pipe.loss() - float(ref["loss"])) / float(ref["loss"])
    assert e_probs < BF16_TOL and e_loss < BF16_TOL and e_hout < BF16_TOL
    assert abs(pipe.accuracy() - float(ref["accuracy"])) < 2.0 / (cfg.B * cfg.n_q)   # at most one near-tie flips
    errs, worst, cos = bf16_grad_check(pipe.p.grad_dict(), gref, name)
    print(name, "probs", e_probs, "hout", e_hout, "loss", e_loss, "cos", cos, {k: "%.1e" % v for k, v in errs.items()})


def test_bf16_mode_dropout_replays_in_oracle():
    """keep_prob = 0.9 (the reference's training value, model/base.py:26,149) through the DROP
    instantiations of the tcgen05 recurrence and BPTT kernels; the device mask is replayed in the oracle."""
    from induction_b200.engine import dropout_keep_mask
    cfg, tables, inputs, label, params = make_case("cfg0_5w5s", seed=4, keep_prob=0.9, mode="bf16")
    pipe = build_pipeline(cfg, tables, params, seed=77)
    assert pipe.tc_lstm
    pipe.set_inputs(inputs, label)
    pipe.forward(train=True)
    pipe.backward()
    torch.cuda.synchronize()
    mask = dropout_keep_mask(77, (cfg.X, cfg.L, cfg.H), 0.9, step=0)
    assert 0.88 < mask.mean() < 0.92
    hb = pipe.hout_bf.float().view(cfg.X, cfg.L, cfg.H).cpu()
    assert bool(((hb == 0) | torch.from_numpy(mask)).all())      # every dropped position is exactly zero
    ref, gref = oracle_grads(cfg, tables, inputs, label, params, drop_mask=torch.from_numpy(mask))
    assert rel_err(hb, ref["hout"]) < BF16_TOL
    assert rel_err(pipe.probs, ref["probs"]) < BF16_TOL
    assert abs(pipe.loss() - float(ref["loss"])) / float(ref["loss"]) < BF16_TOL
    errs, worst, cos = bf16_grad_check(pipe.p.grad_dict(), gref, "dropout")
    print("dropout bf16 cos", cos, {k: "%.1e" % v for k, v in errs.items()})
    # the mask of the next step differs (the seed is offset by global_step) and eval ignores keep_prob
    pipe.adam()
    pipe.forward(train=True)
    mask1 = dropout_keep_mask(77, (cfg.X, cfg.L, cfg.H), 0.9, step=1)
    hb1 = pipe.hout_bf.float().view(cfg.X, cfg.L, cfg.H).cpu()
    assert bool(((hb1 == 0) | torch.from_numpy(mask1)).all()) and not np.array_equal(mask, mask1)


def test_adam_staircase_is_crossed():
    """101+ steps so global_step passes decay_steps = 100 and the learning rate drops to 0.9e-3
    (model/base.py:96-97, staircase=True); trajectory against the float64 oracle."""
    from oracle import induction_oracle as orc
    cfg, tables, inputs, label, params = make_case("smoke", seed=6)
    pipe = build_pipeline(cfg, tables, params)
    pipe.set_inputs(inputs, label)
    p = {k: v.double().clone() for k, v in params.items()}
    m = {k: torch.zeros_like(v) for k, v in p.items()}
    v = {k: torch.zeros_like(x) for k, x in p.items()}
    lrs = []
    for step in range(103):
        _, g = oracle_grads(cfg, tables, inputs, label, p)
        lrs.append(orc.tf_adam_step(p, g, m, v, global_step=step))
        pipe.train_step()
        if step in (98, 99, 100, 102):
            assert float(pipe.p.hyper[1].item()) == pytest.approx(lrs[-1], rel=1e-6), step
    assert lrs[99] == 1e-3 and lrs[100] == pytest.approx(9e-4)
    assert int(pipe.p.global_step.item()) == 103
    got = pipe.p.state_dict()
    for k in p:
        # Adam moves a weight whose gradient is at rounding-noise level by +-lr per step whatever the noise
        # says, so the worst entry is not meaningful over 100 steps: bound the bulk of the entries
        d = (got[k].cpu().double() - p[k]).abs().flatten()
        assert float(d.quantile(0.9)) < 103e-3 * 2e-2, k
    out = oracle_forward(cfg, tables, inputs, label, {k: v.float() for k, v in p.items()})
    pipe.forward()
    assert abs(pipe.loss() - float(out["loss"])) / float(out["loss"]) < 1e-3


def test_bf16_mode_train_steps_track_fp32_mode():
    cfg, tables, inputs, label, params = make_case("u24", seed=6, mode="bf16")
    cfg32, *_ = make_case("u24", seed=6)
    pb, pf = build_pipeline(cfg, tables, params), build_pipeline(cfg32, tables, params)
    for p in (pb, pf):
        p.set_inputs(inputs, label)
    lb, lf = [], []
    for _ in range(5):
        pb.train_step(); pf.train_step()
        lb.append(pb.loss()); lf.append(pf.loss())
    np.testing.assert_allclose(lb, lf, rtol=BF16_TOL)
    assert lb[-1] < lb[0]      # the bf16 shadow weights are refreshed by Adam: the loss goes down


def test_bf16_mode_rejects_unsupported_shapes():
    cfg, tables, inputs, label, params = make_case("ragged", seed=3, mode="bf16")   # u = 13
    with pytest.raises(ValueError, match="bf16 tensor-core mode needs"):
        build_pipeline(cfg, tables, params)


@pytest.mark.parametrize("shape", [dict(N=3, K=2, Q=2, B=2, L=9), dict(N=5, K=1, Q=5, B=40, L=6), dict(N=4, K=3, Q=5, B=3, L=33)])
def test_hidden_size_256_cluster_recurrence_inference_parity(shape):
    """BASELINE configs[4] width (u = 256): the four-CTA cluster recurrence (tc_lstm256.cu) - both tile sizes,
    partial tiles, several clusters - through encoder + induction against the float64 oracle."""
    from induction_b200.engine import EpisodeConfig, EpisodePipeline, ParamStore, init_tf_params, tf_param_shapes
    from induction_b200.util.synthetic import make_tables
    from oracle import induction_oracle as orc
    cfg = EpisodeConfig(E=50, u=256, S=8, mode="bf16", inference=True, **shape)
    tables = make_tables(12, 30, cfg.L, 500, cfg.E, seed=2, min_len=min(3, cfg.L))
    rs = np.random.RandomState(4)
    rows = rs.randint(0, 12 * 30, size=(cfg.B, cfg.n_s))
    inputs = {k: tables[k][rows] for k in ("word", "pos1", "pos2", "mask")}
    label = rs.randint(0, cfg.N, size=(cfg.B, cfg.n_q))
    params = init_tf_params(tf_param_shapes(cfg.N, cfg.L, cfg.E, cfg.u, cfg.S), seed=1)
    g = torch.Generator().manual_seed(9)
    for k in params:
        if k.endswith("bias"):
            params[k] = (torch.randn(params[k].shape, generator=g) * 0.05).float()
    store = ParamStore(tf_param_shapes(cfg.N, cfg.L, cfg.E, cfg.u, cfg.S))
    store.load_state_dict(params)
    pipe = EpisodePipeline(cfg, store, torch.from_numpy(tables["word_vec_mat"]))
    assert pipe.tc_lstm256
    pipe.set_inputs(inputs, label)
    pipe.encode(train=False)
    pipe.induction()
    torch.cuda.synchronize()
    ref = oracle_forward(cfg, tables, inputs, label, params, dtype=torch.float64)
    e_h = rel_err(pipe.hout_bf.float().view(cfg.X, cfg.L, cfg.H), ref["hout"])
    e_enc = rel_err(torch.cat([pipe.es, pipe.eq], dim=1), ref["encoder"])
    e_cv = rel_err(pipe.cvec.view(cfg.B, cfg.N, cfg.H), ref["class_vector"])
    print(shape, "hout", e_h, "encoder", e_enc, "class_vector", e_cv)
    assert e_h < BF16_TOL and e_enc < BF16_TOL and e_cv < BF16_TOL
    with pytest.raises(Exception, match="inference"):
        pipe.relation()


def test_branched_backward_and_early_adam_match_the_serial_step():
    """The backward pass issued as a DAG over side streams, with M's Adam update hoisted under the encoder's
    BPTT (engine._backward_bf16, adam_m), against the same step issued serially with one whole-buffer Adam."""
    cfg, tables, inputs, label, params = make_case("cfg0_5w5s", seed=8, keep_prob=0.9, mode="bf16")
    out = []
    for branched in (True, False):
        pipe = build_pipeline(cfg, tables, params, seed=5)
        pipe.branches = branched
        pipe.set_inputs(inputs, label)
        pipe.forward(train=True)
        pipe.backward(early_adam=branched)
        g = pipe.p.grads.clone()
        pipe.adam()
        torch.cuda.synchronize()
        first = (g, pipe.p.params.clone(), pipe.p.m.clone(), pipe.p.v.clone(), pipe.p.params_bf.clone(), pipe.loss())
        pipe.forward(train=True)
        pipe.backward(early_adam=branched)
        pipe.adam()
        torch.cuda.synchronize()
        out.append(first + (int(pipe.p.global_step.item()), pipe.p.hyper.clone(), pipe.loss()))
    a, b = out
    assert a[6] == b[6] == 2 and torch.equal(a[7], b[7])
    assert a[5] == b[5]                                          # the forward is deterministic
    # split-K reductions use float atomics, so two runs agree to rounding, not bitwise
    for i in (0, 2, 3):
        assert float((a[i] - b[i]).abs().max()) <= 1e-4 * float(b[i].abs().max()), i
    # the first Adam step moves every weight by ~lr * sign(g): only entries whose gradient is at noise level may differ
    moved = (a[1] - b[1]).abs() > 1e-6
    assert float(moved.float().mean()) < 0.02
    assert torch.equal(a[4].float(), a[1].to(torch.bfloat16).float())      # the bf16 shadow tracks both halves
    assert abs(a[8] - b[8]) < 2e-2 * abs(b[8])                   # second step: same trajectory
