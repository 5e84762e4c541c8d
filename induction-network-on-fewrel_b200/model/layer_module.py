"""Layer ops of the Induction Network - host mirror of the reference's ``model/layer_module.py``.

Same names, argument order and return arity as the reference (``self_attention(inputs, mask)``,
``dynamic_routing(input, b_IJ, iter_routing=3)``, ``squash(vector)``,
``neural_tensor_layer(class_vector, query_encoder, out_size=100)``), on CUDA ``torch.Tensor``s instead
of ``tf.Tensor``s.  Every op is a ``torch.autograd.Function`` whose forward and backward are the
hand-written kernels of ``csrc/`` reached through the C-ABI; there is no torch-op or CPU fallback.

The reference creates its weights implicitly with ``tf.get_variable`` / ``tf.layers.Dense`` inside the
op (layer_module.py:16-17, 39-42, 61-72, 116-119).  torch has no implicit variable store, so this
module keeps one: ``get_variable(name, shape)`` creates a leaf tensor on first use under the TF name
 in the current ``VariableStore`` and returns it afterwards, which is what
``tf.get_variable`` with reuse does.  ``variable_scope(store)`` swaps the store.
"""
from __future__ import annotations

import contextlib
from typing import Dict

import torch

from .._lib import LIB, IndnetError, current_stream, ptr, require_cuda
from ..engine import ATT_B, ATT_K, ATT_W2, init_tf_params


class VariableStore:
    """Named trainable leaves (the analogue of TF's global variable collection)."""

    def __init__(self, seed: int = 0, device="cuda"):
        self.vars: Dict[str, torch.Tensor] = {}
        self.seed = seed
        self.device = device

    def get_variable(self, name, shape):
        v = self.vars.get(name)
        if v is None:
            require_cuda()
            init = init_tf_params({name: tuple(int(s) for s in shape)}, seed=self.seed + len(self.vars))[name]
            v = init.to(self.device).requires_grad_(True)
            self.vars[name] = v
        elif tuple(v.shape) != tuple(int(s) for s in shape):
            raise ValueError("Trying to share variable %s, but specified shape %s and found shape %s."
                             % (name, tuple(shape), tuple(v.shape)))
        return v

    def trainable_variables(self):
        return list(self.vars.values())


_STORE = VariableStore()


def default_store() -> VariableStore:
    return _STORE


@contextlib.contextmanager
def variable_scope(store: VariableStore):
    global _STORE
    prev, _STORE = _STORE, store
    try:
        yield store
    finally:
        _STORE = prev


def get_variable(name, shape):
    return _STORE.get_variable(name, shape)


def _check(t, name, ndim=None):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise IndnetError("%s must be a CUDA torch.Tensor (no CPU fallback)" % name)
    if ndim is not None and t.dim() != ndim:
        raise ValueError("%s must have %d dimensions, got shape %s" % (name, ndim, tuple(t.shape)))


def _f32(t):
    return t.detach().to(torch.float32).contiguous()


def _call(name, *args):
    LIB.call(name, current_stream(), *args)


def _gemm(tA, tB, M, N, K, A, lda, B, ldb, C, ldc, beta=0.0, bias=None, act=0, batch=1, sA=0, sB=0, sC=0):
    _call("indnet_sgemm", int(tA), int(tB), M, N, K, 1.0, ptr(A), lda, ptr(B), ldb, beta, ptr(C), ldc,
          ptr(bias), act, batch, 1, sA, 0, sB, 0, sC, 0, 0)


# --------------------------------------------------------------------------------------------------
class _SelfAttention(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inputs, mask, Wa1, ba1, Wa2):
        X, L, H = inputs.shape
        h, m = _f32(inputs), mask.detach().to(torch.int32).contiguous()
        k1, b1, w2 = _f32(Wa1), _f32(ba1), _f32(Wa2)
        proj = torch.empty_like(h)
        _gemm(0, 0, X * L, H, H, h, H, k1, H, proj, H, bias=b1, act=1)
        enc = torch.empty(X, H, device=h.device)
        alphas = torch.empty(X, L, device=h.device)
        _call("indnet_attn_pool_fwd", ptr(h), ptr(proj), ptr(m), ptr(w2), ptr(enc), None, ptr(alphas), X, L, H, X, X)
        ctx.save_for_backward(h, proj, alphas, k1, w2)
        ctx.mark_non_differentiable(alphas)
        return enc, alphas.view(X, L, 1)

    @staticmethod
    def backward(ctx, denc, _dalphas):
        h, proj, alphas, k1, w2 = ctx.saved_tensors
        X, L, H = h.shape
        denc = _f32(denc)
        dh = torch.empty_like(h)
        dpre = torch.empty_like(h)
        dw2_part = torch.empty(X, H, device=h.device)
        _call("indnet_attn_pool_bwd", ptr(denc), None, ptr(h), ptr(proj), ptr(alphas), ptr(w2), ptr(dh), ptr(dpre),
              ptr(dw2_part), X, L, H, X, X)
        dw2 = torch.empty(H, 1, device=h.device)
        _call("indnet_colsum", ptr(dw2_part), X, H, H, ptr(dw2), 0)
        dk1 = torch.empty(H, H, device=h.device)
        _gemm(1, 0, H, H, X * L, h, H, dpre, H, dk1, H)
        db1 = torch.empty(H, device=h.device)
        _call("indnet_colsum", ptr(dpre), X * L, H, H, ptr(db1), 0)
        _gemm(0, 1, X * L, H, H, dpre, H, k1, H, dh, H, beta=1.0)
        return dh, None, dk1, db1, dw2


def self_attention(inputs, mask):
    """a = softmax(W_a2 tanh(W_a1 H^T)), e = sum_t a_t h_t  (reference layer_module.py:49-98).

    inputs [X, L, H] float32 CUDA; mask [X, L] - the reference call site (model/graph.py:52) passes the
    binarised mask, any value > 0 counts as "attend".  Returns ``(output [X, H], alphas [X, L, 1])``.
    Weights ``self_attn/dense/kernel|bias`` and ``self_attn/W_a2`` come from the variable store.
    ``alphas`` is returned for inspection (model/base.py:144) and carries no gradient."""
    _check(inputs, "inputs", 3)
    _check(mask, "mask", 2)
    H = inputs.shape[2]
    Wa1 = get_variable(ATT_K, (H, H))
    ba1 = get_variable(ATT_B, (H,))
    Wa2 = get_variable(ATT_W2, (H, 1))
    return _SelfAttention.apply(inputs, mask, Wa1, ba1, Wa2)


# --------------------------------------------------------------------------------------------------
class _Squash(torch.autograd.Function):
    @staticmethod
    def forward(ctx, vector):
        v = _f32(vector)
        out = torch.empty_like(v)
        _call("indnet_squash_fwd", ptr(v), ptr(out), v.shape[0], v.shape[1])
        ctx.save_for_backward(v)
        return out

    @staticmethod
    def backward(ctx, dout):
        (v,) = ctx.saved_tensors
        dv = torch.empty_like(v)
        _call("indnet_squash_bwd", ptr(v), ptr(_f32(dout)), ptr(dv), v.shape[0], v.shape[1])
        return dv


def squash(vector):
    """f(x) = |x|^2/(1+|x|^2) * x/sqrt(|x|^2+1e-9) along axis 1 of [C, H] (layer_module.py:153-162)."""
    _check(vector, "vector", 2)
    return _Squash.apply(vector)


# --------------------------------------------------------------------------------------------------
class _DynamicRouting(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inp, b_IJ, W, iters):
        *lead, K, H = inp.shape
        BC = 1
        for d in lead:
            BC *= int(d)
        e, w, b = _f32(inp).view(BC * K, H), _f32(W), _f32(b_IJ).view(BC, K)
        ehat = torch.empty_like(e)
        _gemm(0, 0, BC * K, H, H, e, H, w, H, ehat, H)
        cvec = torch.empty(BC, H, device=e.device)
        _call("indnet_routing_fwd", ptr(ehat), ptr(b), ptr(cvec), None, BC, K, H, iters)
        ctx.save_for_backward(e, w, b, ehat)
        ctx.meta = (tuple(inp.shape), BC, K, H, iters)
        return cvec.view(*lead, H)

    @staticmethod
    def backward(ctx, dc):
        e, w, b, ehat = ctx.saved_tensors
        shape, BC, K, H, iters = ctx.meta
        dehat = torch.empty_like(ehat)
        _call("indnet_routing_bwd", ptr(ehat), ptr(b), ptr(_f32(dc).view(BC, H)), ptr(dehat), None, BC, K, H, iters)
        de = torch.empty_like(e)
        _gemm(0, 1, BC * K, H, H, dehat, H, w, H, de, H)
        dW = torch.empty_like(w)
        _gemm(1, 0, H, H, BC * K, e, H, dehat, H, dW, H)
        return de.view(shape), None, dW, None


def dynamic_routing(input, b_IJ, iter_routing=3):
    """The routing algorithm (reference layer_module.py:101-150).

    input [C, K, H] (a leading batch of episodes [B, C, K, H] is also accepted), b_IJ [C, K] coupling
    logits (zeros at the call site, model/graph.py:71; treated as a constant, it is not returned).
    e_hat = input W_s without bias or squash (:132); returns the last c_I [C, H].  K <= 32."""
    _check(input, "input")
    _check(b_IJ, "b_IJ")
    if input.dim() not in (3, 4) or tuple(b_IJ.shape) != tuple(input.shape[:-1]):
        raise ValueError("input must be [C, K, H] and b_IJ [C, K]; got %s and %s" % (tuple(input.shape), tuple(b_IJ.shape)))
    H = input.shape[-1]
    W = get_variable("W_s", (H, H))
    return _DynamicRouting.apply(input, b_IJ, W, int(iter_routing))


# --------------------------------------------------------------------------------------------------
class _NeuralTensorLayer(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cv, qe, M, w, b):
        batched = cv.dim() == 3
        cvb = _f32(cv) if batched else _f32(cv).unsqueeze(0)
        qeb = _f32(qe) if batched else _f32(qe).unsqueeze(0)
        B, C, H = cvb.shape
        nq = qeb.shape[1]
        S = M.shape[2]
        dev = cvb.device
        Mp = torch.empty(H, S, H, device=dev)
        _call("indnet_permute021", ptr(_f32(M)), ptr(Mp), H, H, S)
        wc, bc = _f32(w), _f32(b)
        Tp = torch.empty(B * C, S * H, device=dev)
        _gemm(0, 0, B * C, S * H, H, cvb, H, Mp, S * H, Tp, S * H)
        V = torch.empty(B * nq, C * S, device=dev)
        _gemm(0, 1, nq, C * S, H, qeb, H, Tp, H, V, C * S, batch=B, sA=nq * H, sB=C * S * H, sC=nq * C * S)
        # head without labels: probs only (labels = 0 feed the fused loss/accuracy, ignored here)
        probs = torch.empty(B, nq, C, device=dev)
        labels = torch.zeros(B * nq, dtype=torch.int32, device=dev)
        lp = torch.empty(B * nq, device=dev)
        cr = torch.empty(B * nq, dtype=torch.int32, device=dev)
        _call("indnet_ntn_head_fwd", ptr(V), ptr(wc), ptr(bc), ptr(labels), ptr(probs), ptr(lp), ptr(cr), B * nq, C, S,
              None, None, None, None, None, None, 0, 0.0)
        ctx.save_for_backward(cvb, qeb, Mp, wc, Tp, V, probs)
        ctx.batched = batched
        return probs if batched else probs[0]

    @staticmethod
    def backward(ctx, dprobs):
        cvb, qeb, Mp, wc, Tp, V, probs = ctx.saved_tensors
        B, C, H = cvb.shape
        nq = qeb.shape[1]
        S = Mp.shape[1]
        dev = cvb.device
        NS = C * S
        dp = _f32(dprobs).reshape(B * nq, C)
        dlogit = torch.empty(B * nq, C, device=dev)
        dV = torch.empty(B * nq, NS, device=dev)
        dw = torch.empty_like(wc)
        db = torch.empty(C, device=dev)
        _call("indnet_ntn_head_bwd", ptr(V), ptr(wc), ptr(probs), None, ptr(dp), ptr(dlogit), ptr(dV), None, 0,
              ptr(dw), ptr(db), B * nq, C, S, 1.0)
        dqe = torch.empty_like(qeb)
        _gemm(0, 0, nq, H, NS, dV, NS, Tp, H, dqe, H, batch=B, sA=nq * NS, sB=NS * H, sC=nq * H)
        dTp = torch.empty_like(Tp)
        _gemm(1, 0, NS, H, nq, dV, NS, qeb, H, dTp, H, batch=B, sA=nq * NS, sB=nq * H, sC=NS * H)
        dMp = torch.empty_like(Mp)
        _gemm(1, 0, H, S * H, B * C, cvb, H, dTp, S * H, dMp, S * H)
        dM = torch.empty(H, H, S, device=dev)
        _call("indnet_permute021", ptr(dMp), ptr(dM), H, S, H)
        dcv = torch.empty_like(cvb)
        _gemm(0, 1, B * C, H, S * H, dTp, S * H, Mp, S * H, dcv, H)
        if not ctx.batched:
            dcv, dqe = dcv[0], dqe[0]
        return dcv, dqe, dM, dw, db


def neural_tensor_layer(class_vector, query_encoder, out_size=100):
    """Neural tensor layer (reference layer_module.py:8-47): class_vector [C, H], query_encoder
    [n_q, H] -> probs [n_q, C] = sigmoid(relu(v) w + b), v[q, k*C + c] = c_c^T M[:, :, k] e_q.
    Weights ``M [H, H, out_size]``, ``w [C*out_size, C]``, ``b [C]`` come from the variable store.
    (A leading batch of episodes is accepted on both inputs.)"""
    _check(class_vector, "class_vector")
    _check(query_encoder, "query_encoder")
    C, H = class_vector.shape[-2:]
    M = get_variable("M", (H, H, out_size))
    w = get_variable("w", (C * out_size, C))
    b = get_variable("b", (C,))
    return _NeuralTensorLayer.apply(class_vector, query_encoder, M, w, b)
