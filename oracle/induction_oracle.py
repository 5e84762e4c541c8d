"""CPU oracle for the Induction-Network episode pipeline (torch-CPU, float32 or float64).

TEST INFRASTRUCTURE ONLY.  Nothing in the product package may import this file; only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` do, and only as the checker / timed CPU baseline.

PARITY UNPINNED (model path): the arithmetic of the reference lives in TensorFlow 1.x, which is not
vendored under /root/reference and cannot be installed here (no wheel for Python 3.12, no network),
and the reference ships no tests, golden vectors or checkpoints.  This
file therefore restates the reference call sites line by line together with the TF-1.x op semantics
listed in SURVEY.md Appendix A, and is cross-checked against an independent explicit-loop NumPy
restatement (``oracle/induction_numpy.py``) and against closed-form known answers in the tests.
The *sampler* path is pinned: see ``oracle/sampler_oracle.py``.

Every function cites the reference file:line it follows (paths relative to /root/reference).
Tensors use the reference's layouts and TF variable names.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch

POS_DIM = 5  # model/base.py:25


# ----------------------------------------------------------------------------------------------
# parameters
# ----------------------------------------------------------------------------------------------
def param_shapes(N: int, L: int, E: int, u: int, S: int = 100) -> Dict[str, tuple]:
    """Trainable variable names and shapes of the reference graph.

    model/base.py:68-74 (pos tables), model/graph.py:34-35 (LSTM cells),
    model/layer_module.py:65-72 (attention), :116-119 (W_s), :16-17 (M), :39-42 (w, b).
    """
    D, H = E + 2 * POS_DIM, 2 * u
    return {
        "pos1_embedding": (2 * L, POS_DIM),
        "pos2_embedding": (2 * L, POS_DIM),
        "bidirectional_rnn/fw/basic_lstm_cell/kernel": (D + u, 4 * u),
        "bidirectional_rnn/fw/basic_lstm_cell/bias": (4 * u,),
        "bidirectional_rnn/bw/basic_lstm_cell/kernel": (D + u, 4 * u),
        "bidirectional_rnn/bw/basic_lstm_cell/bias": (4 * u,),
        "self_attn/dense/kernel": (H, H),
        "self_attn/dense/bias": (H,),
        "self_attn/W_a2": (H, 1),
        "W_s": (H, H),
        "M": (H, H, S),
        "w": (N * S, N),
        "b": (N,),
    }


# ----------------------------------------------------------------------------------------------
# encoder
# ----------------------------------------------------------------------------------------------
def get_embedding(params, word_table, word, pos1, pos2):
    """model/base.py:76-86 - three lookups concatenated on the feature axis."""
    ew = word_table[word.long()]
    e1 = params["pos1_embedding"][pos1.long()]
    e2 = params["pos2_embedding"][pos2.long()]
    return torch.cat([ew.to(e1.dtype), e1, e2], dim=2)


def basic_lstm_direction(x, kernel, bias, reverse: bool):
    """rnn.BasicLSTMCell unrolled over all L steps (model/graph.py:34-47).

    One kernel [D+u, 4u] applied to concat([x_t, h]); gate order i, j, f, o; forget_bias = 1.0
    added at run time; zero initial state; NO sequence_length => padding steps are processed and
    the backward direction starts on the padding.  Returns un-dropped outputs [X, L, u].
    """
    X, L, _ = x.shape
    u = kernel.shape[1] // 4
    h = x.new_zeros(X, u)
    c = x.new_zeros(X, u)
    outs = [None] * L
    order = range(L - 1, -1, -1) if reverse else range(L)
    for t in order:
        z = torch.cat([x[:, t, :], h], dim=1) @ kernel + bias
        i, j, f, o = torch.split(z, u, dim=1)
        c = c * torch.sigmoid(f + 1.0) + torch.sigmoid(i) * torch.tanh(j)
        h = torch.tanh(c) * torch.sigmoid(o)
        outs[t] = h
    return torch.stack(outs, dim=1)


def bilstm(params, x, drop_mask=None, keep_prob: float = 1.0):
    """model/graph.py:34-49.  DropoutWrapper(output_keep_prob) scales each step's *output* by
    mask/keep_prob and leaves the recurrent state alone.  ``drop_mask`` is an explicit
    {0,1} tensor [X, L, 2u] so a device-generated mask can be replayed here."""
    fw = basic_lstm_direction(x, params["bidirectional_rnn/fw/basic_lstm_cell/kernel"],
                              params["bidirectional_rnn/fw/basic_lstm_cell/bias"], False)
    bw = basic_lstm_direction(x, params["bidirectional_rnn/bw/basic_lstm_cell/kernel"],
                              params["bidirectional_rnn/bw/basic_lstm_cell/bias"], True)
    out = torch.cat([fw, bw], dim=2)
    if drop_mask is not None:
        out = out * drop_mask.to(out.dtype) / keep_prob
    return out


def self_attention(params, inputs, mask):
    """model/layer_module.py:49-98.  ``mask`` is the raw {0,1,2,3} mask; model/graph.py:52 turns it
    into {0,1} first.  Dense has a bias (:65); adder is (1-m)*(-2**31) (:84); softmax over time."""
    m = (mask > 0).to(inputs.dtype)
    proj = torch.tanh(inputs @ params["self_attn/dense/kernel"] + params["self_attn/dense/bias"])
    score = proj @ params["self_attn/W_a2"]                      # [X, L, 1]
    adder = (1.0 - m.unsqueeze(-1)) * float(-2 ** 31)
    alphas = torch.softmax(score + adder, dim=1)                  # [X, L, 1]
    output = torch.matmul(inputs.transpose(1, 2), alphas).squeeze(-1)
    return output, alphas


# ----------------------------------------------------------------------------------------------
# induction
# ----------------------------------------------------------------------------------------------
def squash(vector):
    """model/layer_module.py:153-162 (axis=1 of a [C, H] matrix; here the last axis)."""
    sq = (vector * vector).sum(dim=-1, keepdim=True)
    return sq / (1 + sq) / torch.sqrt(sq + 1e-9) * vector


def dynamic_routing(W_s, inp, b_IJ, iter_routing: int = 3):
    """model/layer_module.py:101-150.  inp [..., C, K, H], b_IJ [..., C, K] -> c_I [..., C, H].
    e_hat = inp @ W_s with no bias and no squash (:132); 3 iterations of softmax over K (:127),
    weighted sum (:136), squash (:139), agreement update (:143-147)."""
    e_hat = inp @ W_s
    b = b_IJ
    c = None
    for _ in range(iter_routing):
        d = torch.softmax(b, dim=-1).unsqueeze(-1)
        c = squash((d * e_hat).sum(dim=-2))
        b = b + (e_hat * c.unsqueeze(-2)).sum(dim=-1)
    return c


# ----------------------------------------------------------------------------------------------
# relation
# ----------------------------------------------------------------------------------------------
def neural_tensor_layer(params, class_vector, query_encoder):
    """model/layer_module.py:8-47.  class_vector [..., C, H], query_encoder [..., n_q, H].
    v[q, k*C + c] = relu(sum_ij cv[c,i] M[i,j,k] qe[q,j])  (slice-major rows from concat(axis=0)
    at :35, transposed at :37), probs = sigmoid(v @ w + b) (:46)."""
    M, w, b = params["M"], params["w"], params["b"]
    S = M.shape[2]
    C = class_vector.shape[-2]
    v = torch.einsum("...ci,ijk,...qj->...qkc", class_vector, M, query_encoder)
    v = torch.relu(v.reshape(*v.shape[:-2], S * C))
    return torch.sigmoid(v @ w + b)


# ----------------------------------------------------------------------------------------------
# whole graph
# ----------------------------------------------------------------------------------------------
def forward(params, word_table, word, pos1, pos2, mask, query_label, N: int, K: int, Q: int,
            drop_mask=None, keep_prob: float = 1.0, iter_routing: int = 3):
    """InductionGraph.forward + build_predict/accuracy/loss for a batch of B episodes.

    model/graph.py:29-109, model/base.py:101-107.  Inputs are [B, n_s, L] int tensors (support rows
    class-major first, then the permuted queries - util/data_loader.py:303-307) and query_label
    [B, n_q].  The reference runs one episode per sess.run (model/base.py:137); a batch is B
    independent episodes sharing the weights, loss/accuracy averaged over the batch.
    """
    B, n_s, L = word.shape
    n_q = N * Q
    x = get_embedding(params, word_table, word.reshape(B * n_s, L), pos1.reshape(B * n_s, L),
                      pos2.reshape(B * n_s, L))
    hout = bilstm(params, x, drop_mask, keep_prob)
    enc, alphas = self_attention(params, hout, mask.reshape(B * n_s, L))
    H = enc.shape[1]
    enc = enc.reshape(B, n_s, H)
    support = enc[:, : N * K, :].reshape(B, N, K, H)          # graph.py:57-58, :74
    query = enc[:, N * K:, :]                                   # graph.py:60-62
    b_IJ = enc.new_zeros(B, N, K)                               # graph.py:71
    class_vector = dynamic_routing(params["W_s"], support, b_IJ, iter_routing)
    probs = neural_tensor_layer(params, class_vector, query)    # [B, n_q, N]
    onehot = torch.nn.functional.one_hot(query_label.long(), N).to(probs.dtype)
    # tf.losses.mean_squared_error == mean over all elements; L2 term is dead code
    loss = ((probs - onehot) ** 2).mean()
    predict_test = probs.argmax(dim=-1)                         # first max wins (base.py:102)
    accuracy = (predict_test == query_label.long()).to(probs.dtype).mean()
    return {
        "x": x, "hout": hout, "encoder": enc, "alphas": alphas.reshape(B, n_s, L, 1),
        "class_vector": class_vector, "probs": probs, "loss": loss, "accuracy": accuracy,
        "predict_test": predict_test, "predict": torch.round(probs),
    }


def loss_and_grads(params, word_table, word, pos1, pos2, mask, query_label, N, K, Q,
                   drop_mask=None, keep_prob: float = 1.0):
    """Backward oracle: torch autograd through ``forward`` (TF autodiff has no stop_gradient on this
    path, model/layer_module.py:123-147, so gradients flow through routing and BPTT)."""
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
    out = forward(leaves, word_table, word, pos1, pos2, mask, query_label, N, K, Q, drop_mask, keep_prob)
    out["loss"].backward()
    grads = {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in leaves.items()}
    return out, grads


# ----------------------------------------------------------------------------------------------
# optimizer
# ----------------------------------------------------------------------------------------------
def decayed_lr(lr0: float, global_step: int, decay_steps: int = 100, decay_rate: float = 0.9) -> float:
    """tf.train.exponential_decay(..., staircase=True) - model/base.py:96-97."""
    return lr0 * decay_rate ** math.floor(global_step / decay_steps)


def tf_adam_step(params, grads, m, v, global_step: int, lr0: float = 1e-3, decay_steps: int = 100,
                 decay_rate: float = 0.9, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8):
    """tf.train.AdamOptimizer as driven by optimize_loss (model/base.py:96-99):
    lr_t = lr*sqrt(1-b2^t)/(1-b1^t); theta -= lr_t*m/(sqrt(v)+eps)  (eps outside the correction).
    ``global_step`` is the value *before* this step; t = global_step + 1.  In place."""
    lr = decayed_lr(lr0, global_step, decay_steps, decay_rate)
    t = global_step + 1
    lr_t = lr * math.sqrt(1.0 - beta2 ** t) / (1.0 - beta1 ** t)
    with torch.no_grad():
        for k in params:
            g = grads[k]
            m[k].mul_(beta1).add_(g, alpha=1 - beta1)
            v[k].mul_(beta2).addcmul_(g, g, value=1 - beta2)
            params[k].sub_(lr_t * m[k] / (v[k].sqrt() + eps))
    return lr
