"""Independent explicit-index NumPy (float64) restatement of the reference forward graph.

TEST INFRASTRUCTURE ONLY (see oracle/induction_oracle.py for the rules).  PARITY UNPINNED for the
model path: no reference golden vectors exist; this second restatement exists so the two oracles can
be checked against each other.  It deliberately mirrors the reference's *loop structure* (the
100-slice Python loop of model/layer_module.py:20-35, the per-iteration recomputation of e_hat at
:132, one episode at a time as in model/base.py:137) rather than the vectorised torch form.
Small cases only - it is slow on purpose.
"""
from __future__ import annotations

import numpy as np


def _sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def lstm_cell_step(x_t, h, c, kernel, bias):
    """rnn.BasicLSTMCell.call: gates i, j, f, o; forget_bias 1.0."""
    u = h.shape[1]
    z = np.concatenate([x_t, h], axis=1).dot(kernel) + bias
    i, j, f, o = z[:, 0:u], z[:, u:2 * u], z[:, 2 * u:3 * u], z[:, 3 * u:4 * u]
    new_c = c * _sigmoid(f + 1.0) + _sigmoid(i) * np.tanh(j)
    new_h = np.tanh(new_c) * _sigmoid(o)
    return new_h, new_c


def bidirectional_dynamic_rnn(x, kf, bf, kb, bb):
    """model/graph.py:44-49 without sequence_length: bw runs over the fully reversed input and its
    outputs are reversed back."""
    X, L, _ = x.shape
    u = kf.shape[1] // 4
    fw = np.zeros((X, L, u))
    h = np.zeros((X, u)); c = np.zeros((X, u))
    for t in range(L):
        h, c = lstm_cell_step(x[:, t], h, c, kf, bf)
        fw[:, t] = h
    xr = x[:, ::-1]
    bwr = np.zeros((X, L, u))
    h = np.zeros((X, u)); c = np.zeros((X, u))
    for t in range(L):
        h, c = lstm_cell_step(xr[:, t], h, c, kb, bb)
        bwr[:, t] = h
    return np.concatenate([fw, bwr[:, ::-1]], axis=2)


def self_attention(inputs, mask01, Wa1, ba1, Wa2):
    """model/layer_module.py:49-98 with an already-binarised mask (graph.py:52)."""
    X, L, H = inputs.shape
    out = np.zeros((X, H)); alphas = np.zeros((X, L, 1))
    for s in range(X):
        x_proj = np.tanh(inputs[s].dot(Wa1) + ba1)               # :65-67
        sc = x_proj.dot(Wa2)                                      # :76  [L,1]
        adder = (1.0 - mask01[s].astype(np.float64))[:, None] * (-2.0 ** 31)   # :84
        z = sc + adder
        z = z - z.max(axis=0, keepdims=True)
        a = np.exp(z); a = a / a.sum(axis=0, keepdims=True)       # :87
        alphas[s] = a
        out[s] = inputs[s].T.dot(a)[:, 0]                         # :90-96
    return out, alphas


def squash(v):
    """model/layer_module.py:153-162."""
    n2 = np.sum(np.square(v), axis=1, keepdims=True)
    return n2 / (1 + n2) / np.sqrt(n2 + 1e-9) * v


def dynamic_routing(inp, b_IJ, W_s, iter_routing=3):
    """model/layer_module.py:101-150, one episode: inp [C,K,H], b_IJ [C,K]."""
    C, K, H = inp.shape
    b = b_IJ.copy()
    c_I = None
    for _ in range(iter_routing):
        eb = np.exp(b - b.max(axis=1, keepdims=True))
        d_I = (eb / eb.sum(axis=1, keepdims=True)).reshape(C, K, 1)          # :127
        e_IJ = inp.reshape(-1, H).dot(W_s).reshape(C, K, -1)                 # :132
        c_I = squash(np.sum(d_I * e_IJ, axis=1))                             # :136-139
        c_produce_e = np.matmul(e_IJ, c_I.reshape(C, H, 1))                  # :143
        b = b + c_produce_e.reshape(C, K)                                    # :147
    return c_I


def neural_tensor_layer(class_vector, query_encoder, M, w, b):
    """model/layer_module.py:8-47 with the literal per-slice loop."""
    C, H = class_vector.shape
    out_size = M.shape[2]
    mid_pro = []
    for sl in range(out_size):
        class_m = class_vector.dot(M[:, :, sl])                   # :24
        mid_pro.append(class_m.dot(query_encoder.T))              # :30
    tensor_bi_product = np.concatenate(mid_pro, axis=0)           # :35  [C*out_size, n_q]
    V = np.maximum(tensor_bi_product.T, 0.0)                      # :37
    return _sigmoid(V.dot(w) + b)                                 # :46


def forward_episode(p, word_table, word, pos1, pos2, mask, query_label, N, K, Q):
    """model/graph.py:29-109 + model/base.py:76-86,101-107 for ONE episode ([n_s, L] inputs)."""
    p = {k: np.asarray(v, dtype=np.float64) for k, v in p.items()}
    wt = np.asarray(word_table, dtype=np.float64)
    x = np.concatenate([wt[word], p["pos1_embedding"][pos1], p["pos2_embedding"][pos2]], axis=2)
    hout = bidirectional_dynamic_rnn(
        x, p["bidirectional_rnn/fw/basic_lstm_cell/kernel"], p["bidirectional_rnn/fw/basic_lstm_cell/bias"],
        p["bidirectional_rnn/bw/basic_lstm_cell/kernel"], p["bidirectional_rnn/bw/basic_lstm_cell/bias"])
    mask01 = (mask > 0).astype(np.int32)                          # graph.py:52
    enc, alphas = self_attention(hout, mask01, p["self_attn/dense/kernel"], p["self_attn/dense/bias"],
                                 p["self_attn/W_a2"])
    H = enc.shape[1]
    support = enc[: N * K].reshape(N, K, H)
    query = enc[N * K: N * K + N * Q]
    cv = dynamic_routing(support, np.zeros((N, K)), p["W_s"])
    probs = neural_tensor_layer(cv, query, p["M"], p["w"], p["b"])
    onehot = np.eye(N)[np.asarray(query_label)]
    loss = np.mean((probs - onehot) ** 2)                         # graph.py:92-94 (mean, A.4)
    pred = np.argmax(probs, axis=1)
    acc = np.mean((pred == np.asarray(query_label)).astype(np.float64))
    return {"hout": hout, "encoder": enc, "alphas": alphas, "class_vector": cv, "probs": probs,
            "loss": loss, "accuracy": acc, "predict_test": pred}
