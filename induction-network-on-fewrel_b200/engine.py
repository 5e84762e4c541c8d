"""Episode pipeline engine: parameter store + the fused forward / backward / update sequence.

This is the host-side orchestration of the reference's ``sess.run([loss, accuracy, optimize, ...])``
(model/base.py:143-152): one call enqueues every kernel of InductionGraph.forward
(model/graph.py:29-86), build_loss (:88-109), build_predict/accuracy (model/base.py:101-107), the
hand-written backward of each, and the TF-form Adam update (model/base.py:94-99) on the current CUDA
stream through the C-ABI.  Shapes are static per pipeline, buffers are preallocated, and nothing in
a step synchronises with the host, so a whole step can be captured into one CUDA graph.

torch is used for device memory and streams only; every arithmetic kernel is in csrc/.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
import contextlib
from typing import Dict, Optional

import numpy as np
import torch

from ._lib import LIB, IndnetError, current_stream, ptr, require_cuda

POS_DIM = 5  # model/base.py:25


@dataclass
class EpisodeConfig:
    N: int            # classes per episode            (train.py:11)
    K: int            # support instances per class    (train.py:12)
    Q: int = 5        # query instances per class      (train.py:33)
    B: int = 1        # episodes per step (reference: 1, model/base.py:137)
    L: int = 40       # sequence_length                (model/base.py:13)
    E: int = 50       # embed_size                     (model/base.py:15)
    u: int = 100      # hidden_size                    (model/base.py:16)
    S: int = 100      # NTN slices (out_size)          (model/layer_module.py:8)
    iters: int = 3    # routing iterations             (model/layer_module.py:101)
    keep_prob: float = 1.0
    mode: str = "fp32"   # "fp32" parity mode | "bf16" tensor-core mode
    inference: bool = False   # encoder + induction forward only (BASELINE configs[4]): no BPTT / relation buffers

    @property
    def D(self): return self.E + 2 * POS_DIM
    @property
    def H(self): return 2 * self.u
    @property
    def n_s(self): return self.N * (self.K + self.Q)
    @property
    def n_q(self): return self.N * self.Q
    @property
    def NK(self): return self.N * self.K
    @property
    def X(self): return self.B * self.n_s
    @property
    def T(self): return self.X * self.L


# TF variable names
FW_K = "bidirectional_rnn/fw/basic_lstm_cell/kernel"
FW_B = "bidirectional_rnn/fw/basic_lstm_cell/bias"
BW_K = "bidirectional_rnn/bw/basic_lstm_cell/kernel"
BW_B = "bidirectional_rnn/bw/basic_lstm_cell/bias"
ATT_K = "self_attn/dense/kernel"
ATT_B = "self_attn/dense/bias"
ATT_W2 = "self_attn/W_a2"


def tf_param_shapes(N, L, E, u, S, vocab_rows=None):
    """Names and TF-layout shapes of the trainables (model/base.py:62-74, model/graph.py:34-35,
    model/layer_module.py:16-17,39-42,65-72,116-119).  ``vocab_rows`` is set only when the word table
    is trainable (no ``pred_embed``, base.py:62-65)."""
    D, H = E + 2 * POS_DIM, 2 * u
    shapes = {
        "M": (H, H, S), "w": (N * S, N), "b": (N,), "W_s": (H, H),
        ATT_K: (H, H), ATT_B: (H,), ATT_W2: (H, 1),
        FW_K: (D + u, 4 * u), FW_B: (4 * u,), BW_K: (D + u, 4 * u), BW_B: (4 * u,),
        "pos1_embedding": (2 * L, POS_DIM), "pos2_embedding": (2 * L, POS_DIM),
    }
    if vocab_rows is not None:
        shapes["word_embedding"] = (vocab_rows, E)
    return shapes


def _fans(shape):
    if len(shape) == 1:
        return shape[0], shape[0]
    if len(shape) == 2:
        return shape[0], shape[1]
    rf = int(np.prod(shape[:-2]))
    return shape[-2] * rf, shape[-1] * rf


def _trunc_normal(gen, shape, std):
    out = torch.empty(shape, dtype=torch.float64)
    flat = out.view(-1)
    todo = torch.ones(flat.numel(), dtype=torch.bool)
    while todo.any():
        n = int(todo.sum())
        draw = torch.randn(n, generator=gen, dtype=torch.float64)
        flat[todo] = draw
        todo[todo.clone()] = draw.abs() > 2.0
    return (out * std).float()


def init_tf_params(shapes: Dict[str, tuple], seed: int = 0, pos_std: float = 0.1) -> Dict[str, torch.Tensor]:
    """CPU float32 initial values in TF layouts with the reference's initialisers (A.8):
    glorot-uniform for the LSTM and Dense kernels (scope default), glorot-normal (truncated) for
    W_a2 / W_s / M / w / b, N(0, 0.1^2) for the embedding tables (base.py:19), zero biases."""
    gen = torch.Generator().manual_seed(seed)
    out = {}
    for name, shape in shapes.items():
        fi, fo = _fans(shape)
        if name in (FW_B, BW_B, ATT_B):
            t = torch.zeros(shape)
        elif name in (FW_K, BW_K, ATT_K):
            lim = math.sqrt(6.0 / (fi + fo))
            t = (torch.rand(shape, generator=gen, dtype=torch.float64) * 2 - 1).mul(lim).float()
        elif name in ("pos1_embedding", "pos2_embedding", "word_embedding"):
            t = (torch.randn(shape, generator=gen, dtype=torch.float64) * pos_std).float()
        else:
            t = _trunc_normal(gen, shape, math.sqrt(2.0 / (fi + fo)) / 0.87962566103423978)
        out[name] = t.contiguous()
    return out


class ParamStore:
    """All trainables in ONE flat fp32 device buffer (+ grads, Adam m/v), addressed by TF name.
    ``M`` is stored as M' = M.permute(0, 2, 1) ([H_i][S][H_j]) so both NTN contractions are plain
    row-major GEMMs (see csrc/ntn.cu); ``state_dict`` / ``load_state_dict`` speak TF layouts."""

    ALIGN = 64  # floats (256 B)

    def __init__(self, shapes: Dict[str, tuple], device="cuda"):
        require_cuda()
        self.tf_shapes = dict(shapes)
        self.offsets = {}
        off = 0
        for name, shape in shapes.items():
            self.offsets[name] = off
            off += (int(np.prod(shape)) + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        self.numel = off
        self.device = torch.device(device)
        self.params = torch.zeros(off, dtype=torch.float32, device=self.device)
        self.grads = torch.zeros_like(self.params)
        self.m = torch.zeros_like(self.params)
        self.v = torch.zeros_like(self.params)
        self.global_step = torch.zeros(1, dtype=torch.int64, device=self.device)
        self.hyper = torch.zeros(2, dtype=torch.float32, device=self.device)   # {lr_t, lr}
        self.params_bf = None      # bf16 shadow of ``params`` (tensor-core mode), same indexing

    def enable_bf16_shadow(self):
        if self.params_bf is None:
            self.params_bf = torch.zeros(self.numel, dtype=torch.bfloat16, device=self.device)
        self.refresh_bf16()

    def refresh_bf16(self):
        if self.params_bf is not None:
            LIB.call("indnet_cast_bf16", current_stream(), ptr(self.params), ptr(self.params_bf), self.numel)

    def view_bf(self, name):
        return self.view(name, self.params_bf)

    def internal_shape(self, name):
        s = self.tf_shapes[name]
        return (s[0], s[2], s[1]) if name == "M" else s

    def view(self, name, buf=None):
        buf = self.params if buf is None else buf
        n = int(np.prod(self.tf_shapes[name]))
        o = self.offsets[name]
        return buf[o:o + n].view(self.internal_shape(name))

    def grad(self, name):
        return self.view(name, self.grads)

    def load_state_dict(self, tf_params: Dict[str, torch.Tensor]):
        for name in self.tf_shapes:
            t = tf_params[name].to(torch.float32)
            if tuple(t.shape) != tuple(self.tf_shapes[name]):
                raise ValueError("parameter %s has shape %s, expected %s" % (name, tuple(t.shape), self.tf_shapes[name]))
            if name == "M":
                t = t.permute(0, 2, 1)
            self.view(name).copy_(t.contiguous().to(self.device))
        self.refresh_bf16()

    def _export(self, buf):
        out = {}
        for name in self.tf_shapes:
            t = self.view(name, buf).detach().clone()
            out[name] = t.permute(0, 2, 1).contiguous() if name == "M" else t
        return out

    def state_dict(self):
        return self._export(self.params)

    def grad_dict(self):
        return self._export(self.grads)


def dropout_keep_mask(seed: int, shape, keep_prob: float, step: int = 0) -> np.ndarray:
    """Host restatement of the device dropout stream (csrc/common.cuh ``dropout_keep``) so a mask the
    LSTM kernels drew can be replayed through the oracle.  ``shape`` = (X, L, 2u).  One 32-bit hash serves the
    sentence pair (2x, 2x+1) at the same (time, feature): it is taken at the flat index of the even sentence's
    element; the even sentence keeps iff its low 16 bits < uint32(keep_prob * 2^16), the odd one uses the high
    16 bits.  key = dropout_key(seed + step * 1000003)."""
    X, L, F = (int(v) for v in shape)
    n = X * L * F
    M64 = np.uint64(0xFFFFFFFFFFFFFFFF)
    with np.errstate(over="ignore"):
        z = (np.uint64((seed + step * 1000003) & 0xFFFFFFFFFFFFFFFF) + np.uint64(0x9E3779B97F4A7C15)) & M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & M64
        z = z ^ (z >> np.uint64(31))
        key = np.uint32(int(z) >> 32)
        idx = np.arange(n, dtype=np.uint64)
        odd = ((idx // np.uint64(L * F)) & np.uint64(1)).astype(bool)
        idx = idx - odd.astype(np.uint64) * np.uint64(L * F)
        h = ((idx & np.uint64(0xFFFFFFFF)).astype(np.uint32) * np.uint32(0x9E3779B1) + key
             + (idx >> np.uint64(32)).astype(np.uint32) * np.uint32(0x7FEB352D))
        h ^= h >> np.uint32(16); h *= np.uint32(0x85EBCA6B)
        h ^= h >> np.uint32(13); h *= np.uint32(0xC2B2AE35)
        h ^= h >> np.uint32(16)
    thr = np.uint32(int(np.float32(keep_prob) * np.float32(65536.0)))
    field = np.where(odd, h >> np.uint32(16), h & np.uint32(0xFFFF))
    return (field < thr).reshape(shape)


class EpisodePipeline:
    """Forward / backward / Adam over a batch of B episodes with preallocated buffers.

    ``word_table``: float32 [V+2, E] frozen embedding (``pred_embed``, model/base.py:56-59) or None
    when ``word_embedding`` is trainable and lives in the parameter store.
    """

    def __init__(self, cfg: EpisodeConfig, store: ParamStore, word_table: Optional[torch.Tensor],
                 lr: float = 1e-3, decay_steps: int = 100, decay_rate: float = 0.9, seed: int = 0):
        require_cuda()
        LIB.load()
        if cfg.mode not in ("fp32", "bf16"):
            raise ValueError("mode must be 'fp32' or 'bf16'")
        if cfg.K > 32:
            raise ValueError("K=%d exceeds the warp-per-class routing limit of 32" % cfg.K)
        self.cfg, self.p = cfg, store
        self.dev = store.device
        self.word_table = None if word_table is None else word_table.to(self.dev, torch.float32).contiguous()
        self.trainable_words = word_table is None
        if self.trainable_words and "word_embedding" not in store.tf_shapes:
            raise ValueError("no pred_embed given and the store has no word_embedding")
        self.lr0, self.decay_steps, self.decay_rate = float(lr), int(decay_steps), float(decay_rate)
        self.seed = int(seed)
        self.launches = 0
        self.profile = None      # set to [] to collect per-call CUDA-event timings
        self.grad_ready_hook = None   # callable(name): fired when a gradient slot is final ("M" is 95 % of the bytes
                                      # and is produced first, so its all-reduce can hide under the encoder BPTT)
        self.step_hook = None         # callable(where): "begin" / "before_M" of a training forward pass (trainer.EpisodeTrainer)
        self._tag = None
        # The backward pass is a DAG, not a chain: the weight-gradient GEMMs and M's Adam update hang off the chain
        # of activation gradients.  They are issued on two side streams (forked from / joined back to the caller's
        # stream with events, which a CUDA-graph capture turns into parallel branches).  INDNET_SERIAL_BWD=1 keeps
        # everything on one stream.
        import os as _os
        self.branches = not _os.environ.get("INDNET_SERIAL_BWD")
        self._side = None             # [stream, stream], created on first use (same priority as the caller's)
        self._adam_m_done = False     # M's slice has had its update in this step already
        self._dirty = set()           # side streams holding work the caller's stream has not waited for yet
        c = cfg
        f = lambda *s: torch.empty(*s, dtype=torch.float32, device=self.dev)  # noqa: E731
        i32 = lambda *s: torch.zeros(*s, dtype=torch.int32, device=self.dev)  # noqa: E731
        # inputs ("placeholders", model/base.py:44-52)
        # the five placeholders are slices of ONE buffer, so a host that keeps its episodes in the same layout
        # (pack_inputs) feeds a step with a single H2D copy
        n_in, n_lab = c.B * c.n_s * c.L, c.B * c.n_q
        seg = (n_in + 63) // 64 * 64
        self._in_seg, self._in_n, self._in_lab = seg, n_in, n_lab
        self.inbuf = i32(4 * seg + (n_lab + 63) // 64 * 64)
        self.word, self.pos1, self.pos2, self.mask = (self.inbuf[k * seg: k * seg + n_in].view(c.B, c.n_s, c.L) for k in range(4))
        self.label = self.inbuf[4 * seg: 4 * seg + n_lab].view(c.B, c.n_q)
        # forward activations
        inf = bool(cfg.inference)
        if inf and not (cfg.mode == "bf16" and c.u in (128, 256) and c.D <= 62):
            raise ValueError("inference=True pipelines run the tcgen05 encoder only: mode='bf16', hidden_size 128 or 256, "
                             "E + 10 <= 62")
        if not inf:      # fp32 activations: the FFMA parity mode, and the generic-width fallback of the bf16 mode
            self.x = f(c.T, c.D)
            self.xproj = f(2, c.T, 4 * c.u)
            self.hout = f(c.T, c.H)
            self.gates = f(2, c.T, 4 * c.u)
            self.cstate = f(2, c.T, c.u)
            self.hprev = f(2, c.T, c.u)
            self.proj = f(c.T, c.H)
        self.alphas = f(c.X, c.L)
        self.es = f(c.B, c.NK, c.H)          # support encodings (model/graph.py:57-58)
        self.eq = f(c.B, c.n_q, c.H)         # query encodings   (model/graph.py:60-62)
        self.ehat = f(c.B * c.NK, c.H)
        self.cvec = f(c.B * c.N, c.H)
        if not inf:
            self.Tp = f(c.B * c.N, c.S * c.H)
            self.V = f(c.B * c.n_q, c.N * c.S)
            self.probs = f(c.B, c.n_q, c.N)
            self.loss_part = f(c.B * c.n_q)
            self.correct = i32(c.B * c.n_q)
        self.out2 = torch.zeros(2, dtype=torch.float32, device=self.dev)   # {loss, accuracy}
        self.ticket = torch.zeros(1, dtype=torch.int32, device=self.dev)   # last-block counter of the head kernel
        self.running = torch.zeros(3, dtype=torch.float32, device=self.dev)   # training-step sums {loss, accuracy, steps}
        self._bwd_ready = False
        self.bf16 = cfg.mode == "bf16"
        import os as _os0
        self.tc3 = cfg.mode == "fp32" and not _os0.environ.get("INDNET_FP32_FFMA")     # FP32-accurate tensor-core GEMMs
        self._tc3_buf = {}
        if self.bf16:
            if c.D > 64 or c.u % 8:
                raise ValueError("bf16 tensor-core mode needs E + 10 <= 64 and hidden_size % 8 == 0 "
                                 "(16-byte aligned bf16 rows for TMA)")
            store.enable_bf16_shadow()
            h = lambda *s: torch.zeros(*s, dtype=torch.bfloat16, device=self.dev)  # noqa: E731
            self.x_bf = h(c.T, 64)
            self.hout_bf = h(c.T, c.H)
            import os as _os
            self.fused_attn = (c.L <= 128 and c.H % 64 == 0 and (c.H <= 256 or c.H % 256 == 0) and c.H <= 1024
                               and not _os.environ.get("INDNET_NO_FUSED_ATTN"))
            if not (inf and self.fused_attn):
                self.proj_bf = h(c.T, c.H)
            self.es_bf = h(c.B * c.NK, c.H)
            self.cvec_bf = h(c.B * c.N, c.H)
            self.eq_bf = h(c.B * c.n_q, c.H)
            if not inf:
                self.hprev_bf = h(2, c.T, c.u)
                self.Tp_bf = h(c.B * c.N, c.S * c.H)
            self.NSp = (c.N * c.S + 7) // 8 * 8          # dV rows padded to 16 bytes for TMA
            # the persistent tcgen05 recurrence is specialised for u = 128; its bias rides in two spare x columns
            self.tc_lstm = (c.u == 128 and c.D <= 62)
            # hidden_size 256: forward-only cluster kernel (eval / inference); training at 256 takes the generic path
            self.tc_lstm256 = (c.u == 256 and c.D <= 62)
            if self.tc_lstm or self.tc_lstm256:
                self.wt_bf = h(2, 4 * c.u, 64 + c.u)
            if self.tc_lstm and not inf:
                self.gates_bf = h(2, c.T, 4 * c.u)
                self.cstate_bf = h(2, c.T, c.u)

    # ------------------------------------------------------------------ helpers
    def _call(self, name, *args):
        if self.profile is None:
            LIB.call(name, current_stream(), *args)
        else:   # per-call CUDA-event timing on the launching stream (bench.py roofline leg)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            LIB.call(name, current_stream(), *args)
            e1.record()
            self.profile.append((self._tag or name, e0, e1))
        self.launches += 1

    class _Branch:
        """``with pipe._branch(i):`` issues the enclosed launches on side stream i, ordered after everything the
        caller's stream holds at entry (and after the branch's own earlier work).  ``_join`` makes the caller's
        stream wait for the branches."""

        def __init__(self, pipe, i):
            self.pipe, self.i = pipe, i

        def __enter__(self):
            pipe = self.pipe
            if not pipe.branches:
                return self
            if pipe._side is None:
                cur = torch.cuda.current_stream()
                pipe._side = [torch.cuda.Stream(device=pipe.dev, priority=cur.priority) for _ in range(2)]
            side = pipe._side[self.i]
            side.wait_stream(torch.cuda.current_stream())
            self._ctx = torch.cuda.stream(side)
            self._ctx.__enter__()
            pipe._dirty.add(self.i)
            return self

        def __exit__(self, *exc):
            if self.pipe.branches:
                self._ctx.__exit__(*exc)
            return False

    def _branch(self, i):
        return EpisodePipeline._Branch(self, i)

    def _join(self, *which):
        if not self.branches or self._side is None:
            return
        cur = torch.cuda.current_stream()
        for i in (which or tuple(self._dirty)):
            if i in self._dirty:
                cur.wait_stream(self._side[i])
                self._dirty.discard(i)

    def profile_summary(self):
        """{tag: (calls, total_ms)} from the events collected while ``self.profile`` was a list."""
        torch.cuda.synchronize()
        out = {}
        for tag, e0, e1 in self.profile or []:
            n, ms = out.get(tag, (0, 0.0))
            out[tag] = (n + 1, ms + e0.elapsed_time(e1))
        return out

    def _gemm(self, tA, tB, M, N, K, A, lda, Bm, ldb, C, ldc, alpha=1.0, beta=0.0, bias=None, act=0,
              batch=1, inner=1, sA=(0, 0), sB=(0, 0), sC=(0, 0), splitk=0, tag=None):
        self._tag = ("sgemm:" + tag) if tag else None
        try:
            self._gemm_call(tA, tB, M, N, K, A, lda, Bm, ldb, C, ldc, alpha, beta, bias, act, batch, inner,
                            sA, sB, sC, splitk)
        finally:
            self._tag = None

    def _split3(self, key, base, ld_in, rows, cols):
        """fp32 matrix -> its three bf16 parts (head, middle, tail) in persistent scratch with a padded row stride."""
        ld = (cols + 7) // 8 * 8
        buf = self._tc3_buf.get(key)
        if buf is None or buf[0].numel() < rows * ld:
            buf = tuple(torch.zeros(rows * ld, dtype=torch.bfloat16, device=self.dev) for _ in range(3))
            self._tc3_buf[key] = buf
        self._call("indnet_cast_bf16_split3", base, ld_in, rows, cols, ld, ptr(buf[0]), ptr(buf[1]), ptr(buf[2]))
        return buf, ld

    def _gemm_call(self, tA, tB, M, N, K, A, lda, Bm, ldb, C, ldc, alpha, beta, bias, act, batch, inner,
                   sA, sB, sC, splitk):
        # FP32-accurate mode on the tensor cores: operands split into three bf16 parts, six MMAs per K step
        # (csrc/tc_gemm.cu NS = 3).  Anything the split path does not cover (alpha != 1, non-contiguous batches)
        # stays on the FFMA kernel.
        ra, ca = (K, M) if tA else (M, K)          # stored shape of A
        rb, cb = (N, K) if tB else (K, N)          # stored shape of B
        ok = (self.tc3 and alpha == 1.0 and beta in (0.0, 1.0) and inner == 1 and sA[1] == 0 and sB[1] == 0 and sC[1] == 0
              and (batch == 1 or (sA[0] == ra * lda and sB[0] == rb * ldb)))
        if ok:
            tag = self._tag or "gemm"
            (ah, am, al), la = self._split3((tag, "A"), A, lda, batch * ra, ca)
            (bh, bm, bl), lb = self._split3((tag, "B"), Bm, ldb, batch * rb, cb)
            self._call("indnet_tc_gemm3", int(tA), int(not tB), M, N, K, ptr(ah), ptr(am), ptr(al), la, ptr(bh), ptr(bm),
                       ptr(bl), lb, C, ldc, ptr(bias) if isinstance(bias, torch.Tensor) else bias, act,
                       1 if beta == 1.0 else 0, batch, ra * la, rb * lb, sC[0], splitk)
            return
        self._call("indnet_sgemm", int(tA), int(tB), M, N, K, alpha, A, lda, Bm, ldb, beta, C, ldc,
                   ptr(bias) if isinstance(bias, torch.Tensor) else bias, act, batch, inner,
                   sA[0], sA[1], sB[0], sB[1], sC[0], sC[1], splitk)

    def _tc(self, a_mn, b_mn, M, N, K, A, lda, Bm, ldb, C=None, Cb=None, ldc=None, bias=None, act=0,
            accumulate=0, batch=1, sA=0, sB=0, sC=0, splitk=0, tag=None):
        """tcgen05 GEMM (csrc/tc_gemm.cu).  A/Bm/C/Cb: tensors or raw device addresses."""
        self._tag = ("tc_gemm:" + tag) if tag else None
        ad = lambda t: t.data_ptr() if isinstance(t, torch.Tensor) else t  # noqa: E731
        try:
            self._call("indnet_tc_gemm", int(a_mn), int(b_mn), M, N, K, ad(A), lda, ad(Bm), ldb, ad(C), ad(Cb),
                       ldc if ldc is not None else N, ptr(bias) if isinstance(bias, torch.Tensor) else bias, act,
                       accumulate, batch, sA, sB, sC, splitk)
        finally:
            self._tag = None

    def _cast(self, src, dst, tag=None):
        self._tag = ("cast_bf16:" + tag) if tag else None
        try:
            self._call("indnet_cast_bf16", ptr(src), ptr(dst), src.numel())
        finally:
            self._tag = None

    def set_inputs(self, inputs, query_label):
        """Copies one batch into the static input buffers.  Accepts torch tensors (any device) or
        numpy arrays shaped [B, n_s, L] (or [n_s, L] when B == 1)."""
        c = self.cfg
        if isinstance(inputs, torch.Tensor) and query_label is None:       # a batch in the layout of pack_inputs
            if inputs.dtype != torch.int32 or inputs.numel() != self.inbuf.numel():
                raise ValueError("packed inputs must be the int32 tensor pack_inputs() returns")
            self.inbuf.copy_(inputs.reshape(-1), non_blocking=True)
            return
        for k, buf in (("word", self.word), ("pos1", self.pos1), ("pos2", self.pos2), ("mask", self.mask)):
            t = torch.as_tensor(inputs[k])
            buf.copy_(t.reshape(c.B, c.n_s, c.L).to(torch.int32), non_blocking=True)
        self.label.copy_(torch.as_tensor(query_label).reshape(c.B, c.n_q).to(torch.int32), non_blocking=True)

    def pack_inputs(self, inputs, query_label, pin: bool = True) -> torch.Tensor:
        """One batch ({word, pos1, pos2, mask} [B, n_s, L] + labels [B, n_q]) as ONE host int32 tensor in the layout
        of the device input buffer (pinned by default): ``set_inputs(packed)`` / ``EpisodeTrainer.step_on(packed)`` then
        move it with a single copy."""
        c = self.cfg
        out = torch.zeros(self.inbuf.numel(), dtype=torch.int32)
        for i, k in enumerate(("word", "pos1", "pos2", "mask")):
            out[i * self._in_seg: i * self._in_seg + self._in_n] = torch.as_tensor(inputs[k]).reshape(-1).to(torch.int32)
        out[4 * self._in_seg: 4 * self._in_seg + self._in_lab] = torch.as_tensor(query_label).reshape(-1).to(torch.int32)
        return out.pin_memory() if pin else out

    # ------------------------------------------------------------------ forward (model/graph.py:29-86)
    def forward(self, train: bool = False):
        """InductionGraph.forward + loss / accuracy (model/graph.py:29-109, model/base.py:101-107)."""
        if train and self.step_hook is not None:
            self.step_hook("begin")             # e.g. a deferred part of the previous step's update of M starts here ...
        self.encode(train)
        self.induction()
        if train and self.step_hook is not None:
            self.step_hook("before_M")          # ... and must be complete here: the relation module reads M
        self.relation(train)

    def encode(self, train: bool = False):
        """Encoder module: embedding -> BiLSTM -> self-attention pooling, split into support | query
        (model/graph.py:29-62).  The `encoder sentences/sec` leg of bench.py times exactly this."""
        if self.bf16:
            return self._encode_bf16(train)
        c, p = self.cfg, self.p
        keep = c.keep_prob if train else 1.0
        wt = p.view("word_embedding") if self.trainable_words else self.word_table
        u4 = 4 * c.u
        # a2: embedding lookups + concat (model/base.py:76-86)
        self._call("indnet_embed_fwd", ptr(self.word), ptr(self.pos1), ptr(self.pos2), ptr(wt),
                   ptr(p.view("pos1_embedding")), ptr(p.view("pos2_embedding")), ptr(self.x), c.T, c.E, POS_DIM)
        # a3: input projection for all timesteps (first D rows of the TF LSTM kernel) + recurrence
        for d, (kn, bn) in enumerate(((FW_K, FW_B), (BW_K, BW_B))):
            self._gemm(0, 0, c.T, u4, c.D, ptr(self.x), c.D, ptr(p.view(kn)), u4, ptr(self.xproj[d]), u4,
                       bias=p.view(bn), tag="inproj")
        wh_fw = p.view(FW_K)[c.D:]
        wh_bw = p.view(BW_K)[c.D:]
        self._call("indnet_lstm_fwd", ptr(self.xproj), ptr(wh_fw), ptr(wh_bw), ptr(self.hout), ptr(self.gates),
                   ptr(self.cstate), ptr(self.hprev), c.X, c.L, c.u, keep, self.seed, ptr(p.global_step))
        # a4: Dense+tanh, then score/softmax/pool, split into support | query (a5)
        self._gemm(0, 0, c.T, c.H, c.H, ptr(self.hout), c.H, ptr(p.view(ATT_K)), c.H, ptr(self.proj), c.H,
                   bias=p.view(ATT_B), act=1, tag="attn_dense")
        self._call("indnet_attn_pool_fwd", ptr(self.hout), ptr(self.proj), ptr(self.mask), ptr(p.view(ATT_W2)),
                   ptr(self.es), ptr(self.eq), ptr(self.alphas), c.X, c.L, c.H, c.n_s, c.NK)
        self._last_keep = keep

    def induction(self):
        """Induction module: e_hat = e_s W_s once, then the routing iterations (model/graph.py:64-80)."""
        if self.bf16:
            return self._induction_bf16()
        c, p = self.cfg, self.p
        self._gemm(0, 0, c.B * c.NK, c.H, c.H, ptr(self.es), c.H, ptr(p.view("W_s")), c.H, ptr(self.ehat), c.H, tag="routing_transform")
        self._call("indnet_routing_fwd", ptr(self.ehat), None, ptr(self.cvec), None, c.B * c.N, c.K, c.H, c.iters)

    def relation(self, train: bool = False):
        """Relation module + loss + accuracy (model/graph.py:82-109, model/base.py:101-107)."""
        if self.cfg.inference:
            raise IndnetError("this pipeline was built with inference=True (encoder + induction only)")
        if self.bf16:
            return self._relation_bf16(train)
        c, p = self.cfg, self.p
        # a8: T'[b,c][k][j] = cv M' ; V[b][q][c*S+k] = qe_b T'_b^T
        self._gemm(0, 0, c.B * c.N, c.S * c.H, c.H, ptr(self.cvec), c.H, ptr(p.view("M")), c.S * c.H,
                   ptr(self.Tp), c.S * c.H, tag="ntn_stage1")
        NS = c.N * c.S
        self._gemm(0, 1, c.n_q, NS, c.H, ptr(self.eq), c.H, ptr(self.Tp), c.H, ptr(self.V), NS,
                   batch=c.B, sA=(c.n_q * c.H, 0), sB=(NS * c.H, 0), sC=(c.n_q * NS, 0), tag="ntn_stage2")
        # a8-a10: relu, FC, sigmoid, squared error, argmax hit
        self._call("indnet_ntn_head_fwd", ptr(self.V), ptr(p.view("w")), ptr(p.view("b")), ptr(self.label),
                   ptr(self.probs), ptr(self.loss_part), ptr(self.correct), c.B * c.n_q, c.N, c.S,
                   ptr(self.out2), ptr(self.running) if train else None, ptr(self.ticket),   # + loss / accuracy means
                   *self._head_bwd_args(train, False))

    def _head_bwd_args(self, train: bool, bf: bool):
        """Training: the head kernel also emits its own backward (dlogit, dV[, bf16 copy]) - it is local to a row - and
        ``backward`` skips the launch that would have produced them."""
        c = self.cfg
        self._head_bwd_done = bool(train and c.N <= 128)
        if not self._head_bwd_done:
            return None, None, None, 0, 0.0
        if not self._bwd_ready:
            self._alloc_bwd()
        rows = c.B * c.n_q
        return (ptr(self.dlogit), ptr(self.dV), ptr(self.dV_bf) if bf else None, self.NSp if bf else 0,
                1.0 / (rows * c.N))

    def _encode_bf16(self, train: bool):
        """bf16 tensor-core mode: every large contraction runs on tcgen05 (fp32 accumulation); the
        reductions, softmaxes, routing and the loss stay in fp32."""
        c, p = self.cfg, self.p
        keep = c.keep_prob if train else 1.0
        wt = p.view("word_embedding") if self.trainable_words else self.word_table
        u4 = 4 * c.u
        if self.tc_lstm:
            with self._branch(0):       # the weight packing depends on the weights only: beside the embedding gather
                self._call("indnet_lstm_pack_tc", ptr(p.view(FW_K)), ptr(p.view(FW_B)), ptr(p.view(BW_K)), ptr(p.view(BW_B)),
                           ptr(self.wt_bf), c.D, c.u)
        self._call("indnet_embed_fwd_bf16", ptr(self.word), ptr(self.pos1), ptr(self.pos2), ptr(wt),
                   ptr(p.view("pos1_embedding")), ptr(p.view("pos2_embedding")), ptr(self.x_bf), c.T, c.E, POS_DIM)
        if self.tc_lstm:
            # input projection + recurrence + gates in one persistent tcgen05 kernel; eval / inference runs skip the
            # BPTT streams (gates, c, h_prev)
            save = train and not c.inference
            self._join(0)
            self._call("indnet_lstm_fwd_tc", ptr(self.x_bf), None, None, None, None, ptr(self.wt_bf), None, ptr(self.hout_bf),
                       ptr(self.gates_bf) if save else None, ptr(self.cstate_bf) if save else None,
                       ptr(self.hprev_bf) if save else None, c.X, c.L, c.D, c.u, keep, self.seed, ptr(p.global_step))
        elif self.tc_lstm256 and not train:
            self._call("indnet_lstm256_fwd_tc", ptr(self.x_bf), ptr(p.view(FW_K)), ptr(p.view(FW_B)),
                       ptr(p.view(BW_K)), ptr(p.view(BW_B)), ptr(self.wt_bf), ptr(self.hout_bf), c.X, c.L, c.D, c.u)
        else:
            for d, (kn, bn) in enumerate(((FW_K, FW_B), (BW_K, BW_B))):
                # K = D exactly: TMA zero-fills the padded x columns and never reads kernel rows past W_x
                self._tc(0, 1, c.T, u4, c.D, self.x_bf, 64, p.view_bf(kn), u4, C=self.xproj[d], bias=p.view(bn),
                         tag="inproj")
            wh_fw = p.view(FW_K)[c.D:]
            wh_bw = p.view(BW_K)[c.D:]
            self._call("indnet_lstm_fwd", ptr(self.xproj), ptr(wh_fw), ptr(wh_bw), ptr(self.hout), ptr(self.gates),
                       ptr(self.cstate), ptr(self.hprev), c.X, c.L, c.u, keep, self.seed, ptr(p.global_step))
            self._cast(self.hout, self.hout_bf, "hout")
            self._cast(self.hprev, self.hprev_bf, "hprev")
        if self.fused_attn:
            # Dense + tanh + score + masked softmax + weighted sum in one kernel; proj is kept only for the backward
            self._call("indnet_attn_fused_fwd", ptr(self.hout_bf), ptr(p.view_bf(ATT_K)), ptr(p.view(ATT_B)),
                       ptr(p.view(ATT_W2)), ptr(self.mask), ptr(self.proj_bf) if (train and not c.inference) else None,
                       ptr(self.es), ptr(self.eq), ptr(self.es_bf), ptr(self.eq_bf), ptr(self.alphas),
                       c.X, c.L, c.H, c.n_s, c.NK)
        else:
            self._tc(0, 1, c.T, c.H, c.H, self.hout_bf, c.H, p.view_bf(ATT_K), c.H, Cb=self.proj_bf, bias=p.view(ATT_B),
                     act=1, tag="attn_dense")
            self._call("indnet_attn_pool_fwd_bf16", ptr(self.hout_bf), ptr(self.proj_bf), ptr(self.mask),
                       ptr(p.view(ATT_W2)), ptr(self.es), ptr(self.eq), ptr(self.es_bf), ptr(self.eq_bf), ptr(self.alphas),
                       c.X, c.L, c.H, c.n_s, c.NK)
        self._last_keep = keep

    def _induction_bf16(self):
        c, p = self.cfg, self.p
        self._tc(0, 1, c.B * c.NK, c.H, c.H, self.es_bf, c.H, p.view_bf("W_s"), c.H, C=self.ehat,
                 tag="routing_transform")
        self._call("indnet_routing_fwd", ptr(self.ehat), None, ptr(self.cvec), ptr(self.cvec_bf), c.B * c.N, c.K, c.H,
                   c.iters)

    def _relation_bf16(self, train: bool):
        c, p = self.cfg, self.p
        self._tc(0, 1, c.B * c.N, c.S * c.H, c.H, self.cvec_bf, c.H, p.view_bf("M"), c.S * c.H, Cb=self.Tp_bf,
                 tag="ntn_stage1")
        NS = c.N * c.S
        # V_b[n_q, N*S] = qe_b [n_q, H] * T'_b^T (T'_b stored [N*S][H]: K-major) - one batch entry per episode
        self._tc(0, 0, c.n_q, NS, c.H, self.eq_bf, c.H, self.Tp_bf, c.H, C=self.V, ldc=NS, batch=c.B,
                 sA=c.n_q * c.H, sB=NS * c.H, sC=c.n_q * NS, tag="ntn_stage2")
        self._call("indnet_ntn_head_fwd", ptr(self.V), ptr(p.view("w")), ptr(p.view("b")), ptr(self.label),
                   ptr(self.probs), ptr(self.loss_part), ptr(self.correct), c.B * c.n_q, c.N, c.S,
                   ptr(self.out2), ptr(self.running) if train else None, ptr(self.ticket),   # + loss / accuracy means
                   *self._head_bwd_args(train, True))

    # ------------------------------------------------------------------ backward
    def _alloc_bwd(self):
        c = self.cfg
        f = lambda *s: torch.empty(*s, dtype=torch.float32, device=self.dev)  # noqa: E731
        self.dlogit = f(c.B * c.n_q, c.N)
        self.dV = f(c.B * c.n_q, c.N * c.S)
        self.dTp = f(c.B * c.N, c.S * c.H)
        self.dcvec = f(c.B * c.N, c.H)
        self.dehat = f(c.B * c.NK, c.H)
        self.des = f(c.B, c.NK, c.H)
        self.deq = f(c.B, c.n_q, c.H)
        self.dhout = f(c.T, c.H)
        self.dpre = self.proj                  # in place: attn_pool_bwd reads proj[i] then writes dpre[i]
        self.dw2_part = f(c.X, c.H)
        self.dz = self.xproj                   # xproj is dead after the forward recurrence
        self.dx = f(c.T, c.D)
        self.whT = f(2, 4 * c.u, c.u)
        if self.bf16:
            h = lambda *s: torch.zeros(*s, dtype=torch.bfloat16, device=self.dev)  # noqa: E731
            self.dTp_bf = h(c.B * c.N, c.S * c.H)
            self.dV_bf = h(c.B * c.n_q, self.NSp)
            self.dba_part = f(c.X, c.H)
            self.dehat_bf = h(c.B * c.NK, c.H)
            self.dpre_bf = h(c.T, c.H)
            self.dz_bf = h(2, c.T, 4 * c.u)
            self.dhout_bf = h(c.T, c.H) if self.tc_lstm else None
            self.dx2 = f(2, c.T, 16) if self.tc_lstm else None
        self._bwd_ready = True

    def backward(self, early_adam: bool = False):
        """Hand-written backward of every forward op, in reverse order; fills ``self.p.grads``.
        ``early_adam`` (tensor-core mode, single process): M's Adam update is issued as soon as its gradient
        and the last reader of M are done, underneath the rest of the pass; ``adam()`` then finishes the other
        slots.  Data-parallel callers do the same from their ``grad_ready_hook`` with ``adam_m``."""
        if not self._bwd_ready:
            self._alloc_bwd()
        if self.bf16:
            return self._backward_bf16(early_adam)
        c, p = self.cfg, self.p
        u4, NS = 4 * c.u, c.N * c.S
        rows = c.B * c.n_q
        # loss + head (dlogit / dV came out of the forward's head kernel when that ran in training mode)
        if getattr(self, "_head_bwd_done", False):
            self._head_bwd_done = False
            self._call("indnet_ntn_head_dw", ptr(self.V), ptr(self.dlogit), ptr(p.grad("w")), ptr(p.grad("b")), rows, c.N, c.S)
        else:
            self._call("indnet_ntn_head_bwd", ptr(self.V), ptr(p.view("w")), ptr(self.probs), ptr(self.label),
                       None, ptr(self.dlogit), ptr(self.dV), None, 0, ptr(p.grad("w")), ptr(p.grad("b")), rows, c.N, c.S,
                       1.0 / (rows * c.N))
        # dqe_b = dV_b T'_b ; dT'_b = dV_b^T qe_b
        self._gemm(0, 0, c.n_q, c.H, NS, ptr(self.dV), NS, ptr(self.Tp), c.H, ptr(self.deq), c.H,
                   batch=c.B, sA=(c.n_q * NS, 0), sB=(NS * c.H, 0), sC=(c.n_q * c.H, 0), tag="ntn_dqe")
        self._gemm(1, 0, NS, c.H, c.n_q, ptr(self.dV), NS, ptr(self.eq), c.H, ptr(self.dTp), c.H,
                   batch=c.B, sA=(c.n_q * NS, 0), sB=(c.n_q * c.H, 0), sC=(NS * c.H, 0), tag="ntn_dT")
        # dM' = cv^T dT' ; dcv = dT' M'^T
        self._gemm(1, 0, c.H, c.S * c.H, c.B * c.N, ptr(self.cvec), c.H, ptr(self.dTp), c.S * c.H,
                   ptr(p.grad("M")), c.S * c.H, tag="ntn_dM")
        self._gemm(0, 1, c.B * c.N, c.H, c.S * c.H, ptr(self.dTp), c.S * c.H, ptr(p.view("M")), c.S * c.H,
                   ptr(self.dcvec), c.H, tag="ntn_dcv")
        if self.grad_ready_hook is not None:     # after the last reader of M: the hook may update M in place
            self.grad_ready_hook("M")
        # routing
        self._call("indnet_routing_bwd", ptr(self.ehat), None, ptr(self.dcvec), ptr(self.dehat), None,
                   c.B * c.N, c.K, c.H, c.iters)
        self._gemm(0, 1, c.B * c.NK, c.H, c.H, ptr(self.dehat), c.H, ptr(p.view("W_s")), c.H, ptr(self.des), c.H, tag="routing_des")
        self._gemm(1, 0, c.H, c.H, c.B * c.NK, ptr(self.es), c.H, ptr(self.dehat), c.H, ptr(p.grad("W_s")), c.H, tag="routing_dWs")
        # attention
        self._call("indnet_attn_pool_bwd", ptr(self.des), ptr(self.deq), ptr(self.hout), ptr(self.proj),
                   ptr(self.alphas), ptr(p.view(ATT_W2)), ptr(self.dhout), ptr(self.dpre), ptr(self.dw2_part),
                   c.X, c.L, c.H, c.n_s, c.NK)
        self._call("indnet_colsum", ptr(self.dw2_part), c.X, c.H, c.H, ptr(p.grad(ATT_W2)), 0)
        self._gemm(1, 0, c.H, c.H, c.T, ptr(self.hout), c.H, ptr(self.dpre), c.H, ptr(p.grad(ATT_K)), c.H, tag="attn_dW")
        self._call("indnet_colsum", ptr(self.dpre), c.T, c.H, c.H, ptr(p.grad(ATT_B)), 0)
        self._gemm(0, 1, c.T, c.H, c.H, ptr(self.dpre), c.H, ptr(p.view(ATT_K)), c.H, ptr(self.dhout), c.H, beta=1.0, tag="attn_dh")
        # BPTT
        wh_fw = p.view(FW_K)[c.D:]
        wh_bw = p.view(BW_K)[c.D:]
        self._call("indnet_lstm_bwd", ptr(self.dhout), ptr(self.gates), ptr(self.cstate), ptr(wh_fw), ptr(wh_bw),
                   ptr(self.whT), ptr(self.dz), c.X, c.L, c.u, self._last_keep, self.seed, ptr(p.global_step))
        n0 = 0 if self.trainable_words else c.E     # dx columns actually needed
        for d, (kn, bn) in enumerate(((FW_K, FW_B), (BW_K, BW_B))):
            gk = p.grad(kn)
            self._gemm(1, 0, c.D, u4, c.T, ptr(self.x), c.D, ptr(self.dz[d]), u4, ptr(gk), u4, tag="lstm_dWx")
            self._gemm(1, 0, c.u, u4, c.T, ptr(self.hprev[d]), c.u, ptr(self.dz[d]), u4, ptr(gk[c.D:]), u4, tag="lstm_dWh")
            self._call("indnet_colsum", ptr(self.dz[d]), c.T, u4, u4, ptr(p.grad(bn)), 0)
            self._gemm(0, 1, c.T, c.D - n0, u4, ptr(self.dz[d]), u4, ptr(p.view(kn)[n0:]), u4,
                       self.dx.data_ptr() + 4 * n0, c.D, beta=(0.0 if d == 0 else 1.0), tag="lstm_dx")
        # embedding tables
        g1, g2 = p.grad("pos1_embedding"), p.grad("pos2_embedding")
        self._call("indnet_memset", ptr(g1), 0, g1.numel() * 4)
        self._call("indnet_memset", ptr(g2), 0, g2.numel() * 4)
        gw = None
        if self.trainable_words:
            gw = p.grad("word_embedding")
            self._call("indnet_memset", ptr(gw), 0, gw.numel() * 4)
        self._call("indnet_embed_bwd", ptr(self.word), ptr(self.pos1), ptr(self.pos2), ptr(self.dx), ptr(gw),
                   ptr(g1), ptr(g2), c.T, c.E, POS_DIM, 2 * c.L)

    def _backward_bf16(self, early_adam: bool = False):
        c, p = self.cfg, self.p
        u4, NS = 4 * c.u, c.N * c.S
        rows = c.B * c.n_q
        # dV rows are N*S wide (1000 B in bf16): the bf16 copy has rows padded to a 16-byte multiple for TMA.  In a
        # training step the forward's head kernel has written dlogit / dV already.
        if getattr(self, "_head_bwd_done", False):
            self._head_bwd_done = False
        else:
            self._call("indnet_ntn_head_bwd", ptr(self.V), ptr(p.view("w")), ptr(self.probs), ptr(self.label),
                       None, ptr(self.dlogit), ptr(self.dV), ptr(self.dV_bf), self.NSp, None, None,
                       rows, c.N, c.S, 1.0 / (rows * c.N))
        # every gradient slot that is reduced with atomics later in the pass (attention bias / W_a2, the LSTM kernels
        # and biases, the position tables: adjacent slots of the flat buffer) is zeroed by ONE memset here, on a
        # side branch, instead of seven memset nodes in front of the kernels on the critical path
        self._pre0 = bool(self.tc_lstm and self.fused_attn and c.H <= 256 and self._fused_dx())
        with self._branch(0):
            if self._pre0:
                z0 = p.offsets[ATT_B]
                z1 = p.offsets["pos2_embedding"] + p.grad("pos2_embedding").numel()
                self._call("indnet_memset", ptr(p.grads[z0:]), 0, (z1 - z0) * 4)
            # dqe_b = dV_b [n_q, N*S] * T'_b [N*S, H]; FC weight / bias gradients (read V and dlogit only)
            self._tc(0, 1, c.n_q, c.H, NS, self.dV_bf, self.NSp, self.Tp_bf, c.H, C=self.deq, ldc=c.H, batch=c.B,
                     sA=c.n_q * self.NSp, sB=NS * c.H, sC=c.n_q * c.H, splitk=1, tag="ntn_dqe")
            self._call("indnet_ntn_head_dw", ptr(self.V), ptr(self.dlogit), ptr(p.grad("w")), ptr(p.grad("b")),
                       rows, c.N, c.S)
        # dT'_b = dV_b^T [N*S, n_q] * qe_b [n_q, H]
        self._tc(1, 1, NS, c.H, c.n_q, self.dV_bf, self.NSp, self.eq_bf, c.H, Cb=self.dTp_bf, ldc=c.H, batch=c.B,
                 sA=c.n_q * self.NSp, sB=c.n_q * c.H, sC=NS * c.H, splitk=1, tag="ntn_dT")
        with self._branch(1):
            # dM' = cv^T dT' (both operands stored [K][.]) - 96 % of the gradient bytes, off the critical path
            self._tc(1, 1, c.H, c.S * c.H, c.B * c.N, self.cvec_bf, c.H, self.dTp_bf, c.S * c.H, C=p.grad("M"),
                     splitk=1, tag="ntn_dM")
        # dcv = dT' M'^T (both K-major)
        self._tc(0, 0, c.B * c.N, c.H, c.S * c.H, self.dTp_bf, c.S * c.H, p.view_bf("M"), c.S * c.H, C=self.dcvec,
                 tag="ntn_dcv")
        if self.grad_ready_hook is not None:
            with self._branch(1):
                self.grad_ready_hook("M")          # after ntn_dM: the data-parallel all-reduce of M's gradient starts here
        self._call("indnet_routing_bwd", ptr(self.ehat), None, ptr(self.dcvec), ptr(self.dehat), ptr(self.dehat_bf),
                   c.B * c.N, c.K, c.H, c.iters)
        with self._branch(1):
            self._tc(1, 1, c.H, c.H, c.B * c.NK, self.es_bf, c.H, self.dehat_bf, c.H, C=p.grad("W_s"), tag="routing_dWs")
        self._tc(0, 0, c.B * c.NK, c.H, c.H, self.dehat_bf, c.H, p.view_bf("W_s"), c.H, C=self.des, tag="routing_des")
        self._join(0)                 # deq
        # d(loss)/d(h): pooling path alpha_t * denc + Dense path dpre * W_a1^T.  With the tensor-core LSTM the
        # sum is formed once in the GEMM epilogue and written in bf16; otherwise it is accumulated in fp32.
        fused = self.tc_lstm
        if fused and self.fused_attn and c.H <= 256:
            # pooling / softmax / score / tanh backward + dpre * Wa1^T + alpha_t * denc in ONE kernel
            self._call("indnet_attn_fused_bwd", ptr(self.des), ptr(self.deq), ptr(self.hout_bf), ptr(self.proj_bf),
                       ptr(self.alphas), ptr(p.view(ATT_W2)), ptr(p.view_bf(ATT_K)), ptr(self.dpre_bf), ptr(self.dhout_bf),
                       ptr(p.grad(ATT_W2)), ptr(p.grad(ATT_B)), c.X, c.L, c.H, c.n_s, c.NK, int(self._pre0))
            with self._branch(0):
                self._tc(1, 1, c.H, c.H, c.T, self.hout_bf, c.H, self.dpre_bf, c.H, C=p.grad(ATT_K), tag="attn_dW")
            if early_adam:
                # M's Adam update (207 MB of HBM traffic) goes under the BPTT kernel, whose input tiles are prefetched
                # two steps ahead - beside the latency-bound routing / attention kernels it cost them 2.5x
                with self._branch(1):
                    if self.grad_ready_hook is not None:
                        self.grad_ready_hook("M:update")
                    else:
                        self.adam_m()
            return self._backward_bf16_lstm()
        # with the 16-byte kernel variant (H % 8 == 0) the W_a2 / Dense-bias gradients are summed over sentences
        # inside the pooling kernel; otherwise per-sentence partials + two column sums
        insum = c.H % 8 == 0 and c.H <= 512
        self._call("indnet_attn_pool_bwd_bf16", ptr(self.des), ptr(self.deq), ptr(self.hout_bf), ptr(self.proj_bf),
                   ptr(self.alphas), ptr(p.view(ATT_W2)), None if fused else ptr(self.dhout), ptr(self.dpre_bf),
                   None if insum else ptr(self.dw2_part), None if insum else ptr(self.dba_part),
                   ptr(p.grad(ATT_W2)) if insum else None, ptr(p.grad(ATT_B)) if insum else None,
                   c.X, c.L, c.H, c.n_s, c.NK)
        if not insum:
            self._call("indnet_colsum", ptr(self.dw2_part), c.X, c.H, c.H, ptr(p.grad(ATT_W2)), 0)
        self._tc(1, 1, c.H, c.H, c.T, self.hout_bf, c.H, self.dpre_bf, c.H, C=p.grad(ATT_K), tag="attn_dW")
        if not insum:
            self._call("indnet_colsum", ptr(self.dba_part), c.X, c.H, c.H, ptr(p.grad(ATT_B)), 0)
        if fused:
            self._tag = "tc_gemm:attn_dh"
            try:
                self._call("indnet_tc_gemm_rank1", 0, 0, c.T, c.H, c.H, ptr(self.dpre_bf), c.H, ptr(p.view_bf(ATT_K)),
                           c.H, None, ptr(self.dhout_bf), c.H, ptr(self.alphas), ptr(self.des), ptr(self.deq), c.L,
                           c.n_s, c.NK)
            finally:
                self._tag = None
        else:
            self._tc(0, 0, c.T, c.H, c.H, self.dpre_bf, c.H, p.view_bf(ATT_K), c.H, C=self.dhout, accumulate=1,
                     splitk=1, tag="attn_dh")
        return self._backward_bf16_lstm()

    def _fused_dx(self):
        import os as _os
        c = self.cfg
        return bool(self.tc_lstm and not self.trainable_words and c.D - c.E <= 16
                    and not _os.environ.get("INDNET_NO_FUSED_DX"))

    def _backward_bf16_lstm(self):
        c, p = self.cfg, self.p
        u4 = 4 * c.u
        # with a frozen word table only the 2 x 5 position columns of dx are needed: the BPTT kernel forms
        # them itself (per direction) instead of two GEMM launches that would re-read all of dz
        fused_dx = self._fused_dx()
        pre0 = getattr(self, "_pre0", False)
        if self.tc_lstm:
            self._call("indnet_lstm_bwd_tc", ptr(self.dhout_bf), ptr(self.gates_bf), ptr(self.cstate_bf),
                       ptr(p.view_bf(FW_K)), ptr(p.view_bf(BW_K)), ptr(self.dz_bf),
                       ptr(self.dx2) if fused_dx else None, c.E if fused_dx else 0,
                       c.X, c.L, c.D, c.u, self._last_keep, self.seed, ptr(p.global_step))
        else:
            wh_fw = p.view(FW_K)[c.D:]
            wh_bw = p.view(BW_K)[c.D:]
            self._call("indnet_lstm_bwd", ptr(self.dhout), ptr(self.gates), ptr(self.cstate), ptr(wh_fw), ptr(wh_bw),
                       ptr(self.whT), ptr(self.dz), c.X, c.L, c.u, self._last_keep, self.seed, ptr(p.global_step))
            self._cast(self.dz, self.dz_bf, "dz")
        n0 = 0 if self.trainable_words else c.E
        g1, g2 = p.grad("pos1_embedding"), p.grad("pos2_embedding")
        if fused_dx:
            # side branches first (a branch forks from what the caller's stream holds at that moment)
            with self._branch(1):
                if not pre0:
                    self._call("indnet_memset", ptr(g1), 0, g1.numel() * 4)
                    self._call("indnet_memset", ptr(g2), 0, g2.numel() * 4)
                self._call("indnet_embed_pos_bwd2", ptr(self.pos1), ptr(self.pos2), ptr(self.dx2[0]), ptr(self.dx2[1]), 16, 0,
                           ptr(g1), ptr(g2), c.T, POS_DIM, 2 * c.L)
        for nth, d in enumerate((1, 0) if self.tc_lstm else (0, 1)):
            kn, bn = ((FW_K, FW_B), (BW_K, BW_B))[d]
            gk = p.grad(kn)
            if self.tc_lstm:
                # both weight gradients of the direction in one pass over dz: one CTA multiplies every dz tile
                # with the stacked [x | h_prev] operand (two accumulators); the two directions run side by side
                self._tag = "tc_gemm:lstm_dW"
                try:
                    with self._branch(0) if d == 1 else contextlib.nullcontext():
                        # the two directions run side by side, one CTA per SM each (160 KB of operand ring): half of the
                        # SMs per direction so that both grids are resident at once instead of queueing as two waves
                        sk = max(1, 148 // (4 * c.u // 128) // 2) if self.branches else 0
                        self._call("indnet_tc_gemm_lstm_dw", ptr(self.x_bf), ptr(self.hprev_bf[d]), ptr(self.dz_bf[d]),
                                   ptr(gk), ptr(p.grad(bn)), c.T, c.D, c.u, sk, int(pre0))
                finally:
                    self._tag = None
            else:
                self._tc(1, 1, c.D, u4, c.T, self.x_bf, 64, self.dz_bf[d], u4, C=gk, tag="lstm_dWx")
                self._tc(1, 1, c.u, u4, c.T, self.hprev_bf[d], c.u, self.dz_bf[d], u4, C=gk[c.D:], tag="lstm_dWh")
            if not self.tc_lstm:
                self._call("indnet_colsum_bf16", ptr(self.dz_bf[d]), c.T, u4, ptr(p.grad(bn)))
            if not fused_dx:
                self._tc(0, 0, c.T, c.D - n0, u4, self.dz_bf[d], u4, p.view_bf(kn)[n0:], u4,
                         C=self.dx.data_ptr() + 4 * n0, ldc=c.D, accumulate=(0 if nth == 0 else 1), splitk=1, tag="lstm_dx")
        if not fused_dx:
            self._call("indnet_memset", ptr(g1), 0, g1.numel() * 4)
            self._call("indnet_memset", ptr(g2), 0, g2.numel() * 4)
            gw = None
            if self.trainable_words:
                gw = p.grad("word_embedding")
                self._call("indnet_memset", ptr(gw), 0, gw.numel() * 4)
            self._call("indnet_embed_bwd", ptr(self.word), ptr(self.pos1), ptr(self.pos2), ptr(self.dx), ptr(gw),
                       ptr(g1), ptr(g2), c.T, c.E, POS_DIM, 2 * c.L)
        self._join()

    # ------------------------------------------------------------------ update (model/base.py:94-99)
    def adam_m(self, grad_scale: float = 1.0, lo: int = 0, hi: Optional[int] = None, final: Optional[bool] = None,
               enable=None):
        """Adam update of M's slice alone (or of elements [lo, hi) of it), on the current stream, as soon as that
        stream holds M's final gradient (after ntn_dM, or after its all-reduce) AND is ordered after the last reader
        of M's bf16 shadow (ntn_dcv); leaves global_step alone.  ``final`` (default: hi reaches the end of M): the
        next ``adam()`` updates only the other slots and advances the step.  A call with lo == 0 computes this step's
        lr_t first; one with lo > 0 uses the lr_t in ``hyper`` as it stands (a deferred slice is applied with the
        lr_t of the step its gradient belongs to).  ``enable``: device int32, the update is skipped when it is 0."""
        p = self.p
        n = p.offsets["w"]                     # M is the first slot of the flat buffer
        hi = n if hi is None else int(hi)
        if lo == 0:
            self._call("indnet_adam_prep", ptr(p.global_step), ptr(p.hyper), self.lr0, self.decay_steps, self.decay_rate,
                       0.9, 0.999, 0)
        bf = p.params_bf[lo:] if p.params_bf is not None else None
        self._call("indnet_adam_apply", ptr(p.params[lo:]), ptr(p.grads[lo:]), ptr(p.m[lo:]), ptr(p.v[lo:]), hi - lo,
                   ptr(p.hyper), 0.9, 0.999, 1e-8, grad_scale, ptr(bf), None, ptr(enable))
        if (hi == n) if final is None else final:
            self._adam_m_done = True

    def adam(self, grad_scale: float = 1.0):
        p = self.p
        self._join()
        if self._adam_m_done:
            n0 = p.offsets["w"]
            self._adam_m_done = False
            bf = p.params_bf[n0:] if p.params_bf is not None else None
            self._call("indnet_adam_apply", ptr(p.params[n0:]), ptr(p.grads[n0:]), ptr(p.m[n0:]), ptr(p.v[n0:]),
                       p.numel - n0, ptr(p.hyper), 0.9, 0.999, 1e-8, grad_scale, ptr(bf), ptr(p.global_step), None)
            return
        self._call("indnet_adam_step", ptr(p.params), ptr(p.grads), ptr(p.m), ptr(p.v), p.numel,
                   ptr(p.global_step), ptr(p.hyper), self.lr0, self.decay_steps, self.decay_rate,
                   0.9, 0.999, 1e-8, grad_scale, ptr(p.params_bf))

    def train_step(self):
        """fwd + bwd + Adam on the inputs currently in the static buffers (no host sync)."""
        self.forward(train=True)
        self.backward(early_adam=self.grad_ready_hook is None)
        self.adam()

    # ------------------------------------------------------------------ results
    def loss(self) -> float:
        return float(self.out2[0].item())

    def accuracy(self) -> float:
        return float(self.out2[1].item())
