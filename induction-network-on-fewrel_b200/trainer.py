"""EpisodeTrainer: the train step the reference runs as ``next_one_tf`` + ``sess.run([loss, accuracy,
optimize, ...])`` (model/base.py:136-152), as ONE replayable CUDA graph per step.

  gather(step t picks) -> [fork: stream walk for step t+1 on a low-priority side stream] -> forward ->
  loss -> backward -> gradient all-reduce (world > 1) -> TF Adam -> [join]

* The device sampler (csrc/sampler.cu) draws the same episodes the reference would for the same seeds; the
  serial walk of the NEXT step's draws hides under this step's kernels.
* Data parallel, one process per GPU: every rank walks the SAME two streams for the global batch
  of ``world * B`` episodes and gathers only its own episodes ``[rank*B, (rank+1)*B)``, so the global batch is
  bit-identical for any world size.  Gradients: the slot of ``M`` (95 % of the bytes, final right after the
  relation-module backward) is all-reduced on a communication stream under the encoder backward; the other
  slots go in one trailing all-reduce; Adam applies the sum scaled by 1/world (loss = mean over the global
  batch).  No other collective exists on this path.
* Loss / accuracy stay on the device (``pipe.out2``; running sums in ``pipe.running``): a loop reads them
  every n steps instead of synchronising each step.

Used by ``Base.train`` (the product loop), ``train.py`` and ``bench.py`` - the timed code is the code users call.
"""
from __future__ import annotations

from typing import Optional

import torch

from .engine import EpisodePipeline


class EpisodeTrainer:
    M_DEFER = 0.0     # fraction of M whose all-reduce + update moves into the next step's forward window (see __init__)
    # Pieces of M's gradient all-reduced / updated as a pipeline (data parallel).  Measured on 8 B200 at config 1:
    # 1 piece 0.442 ms / step, 2 pieces 0.510, 4 pieces 0.597 - every extra NCCL call costs more (ring, LL protocol,
    # ~40 us each at 8 ranks) than the overlap of the update with the next piece's reduction returns.
    M_CHUNKS = 1

    def __init__(self, pipe: EpisodePipeline, sampler=None, rank: int = 0, world: int = 1, process_group=None,
                 cuda_graph: bool = True, overlap_allreduce: bool = True, stream_mode: str = "global"):
        self.pipe, self.store, self.cfg = pipe, pipe.p, pipe.cfg
        self.smp = sampler
        self.rank, self.world = int(rank), int(world)
        self.group = process_group
        self.use_graph = bool(cuda_graph)
        self.overlap = bool(overlap_allreduce) and self.world > 1
        # "global": one episode stream for the whole job - every rank walks the draws of all world * B episodes and
        # gathers its own B (the global batch is bit-identical for any world size, the serial walk grows
        # with the global batch).  "per-rank": every rank owns a stream (seed it differently per rank) and walks only
        # its own B episodes - the sampler's cost stays constant under weak scaling.
        if stream_mode not in ("global", "per-rank"):
            raise ValueError("stream_mode must be 'global' or 'per-rank'")
        self.stream_mode = stream_mode
        self.global_B = self.cfg.B * (self.world if stream_mode == "global" else 1)
        dev = pipe.dev
        # the step's own stream (high priority), the sampler's side stream (low: its table kernel fills the
        # GPU and would otherwise take SMs from the training kernels) and the communication stream
        self.stream = torch.cuda.Stream(device=dev, priority=-1)
        self.side = torch.cuda.Stream(device=dev, priority=0)
        self.comm = torch.cuda.Stream(device=dev, priority=-1)
        self.upd = torch.cuda.Stream(device=dev, priority=-1)      # M's Adam update behind its all-reduce (data parallel)
        self._upd_used = False
        self.m_end = self.store.offsets["w"]        # slots are laid out M, w, b, ... (engine.ParamStore)
        self.inputs = dict(word=pipe.word, pos1=pipe.pos1, pos2=pipe.pos2, mask=pipe.mask, label=pipe.label)
        self.graph_step: Optional[torch.cuda.CUDAGraph] = None     # sampler in the loop
        self.graph_train: Optional[torch.cuda.CUDAGraph] = None    # host-fed / resident inputs
        self._primed = False
        self._need_walk = False
        self.steps_done = 0
        if sampler is not None:      # stream states as they were BEFORE the look-ahead walk (see suspend())
            self._py_prev = torch.zeros_like(sampler.py_state)
            self._np_prev = torch.zeros_like(sampler.np_state)
        # Data parallel, optional: the tail [m_split, m_end) of M's gradient is all-reduced and applied at the START of
        # the next step's graph, underneath the encoder's forward pass (which does not read M), the head [0, m_split)
        # underneath the backward pass as before - two windows instead of one for the 26 MB that dominate the step's
        # communication.  The update is exact (M is complete before the relation module reads it); between steps the
        # tail is one step behind until flush() (suspend / close call it).  INDNET_M_DEFER = fraction deferred.
        import os
        frac = float(os.environ.get("INDNET_M_DEFER", self.M_DEFER))
        ok = self.world > 1 and self.overlap and pipe.bf16 and pipe.tc_lstm and pipe.fused_attn and pipe.cfg.H <= 256
        self.m_split = self.m_end - (int(frac * self.m_end) // 256 * 256 if ok and frac > 0 else 0)
        self.defer = self.m_split < self.m_end
        self.pending = torch.zeros(1, dtype=torch.int32, device=dev)     # 1: the deferred tail holds an un-applied gradient
        self.comm2 = torch.cuda.Stream(device=dev, priority=-1)
        self._comm2_used = False
        if self.world > 1:
            pipe.grad_ready_hook = self._on_grad_ready
            if self.defer:
                pipe.step_hook = self._on_step
            import os
            n, k = self.m_split, int(os.environ.get("INDNET_M_CHUNKS", self.M_CHUNKS))
            step = (n // k + 255) // 256 * 256
            self._m_chunks = [(lo, min(n, lo + step)) for lo in range(0, n, step)]
            self._m_events = [torch.cuda.Event() for _ in self._m_chunks]

    # ------------------------------------------------------------------ collectives
    def _dist(self):
        import torch.distributed as dist
        return dist

    def _on_grad_ready(self, name):
        if not self.overlap:
            return
        if name == "M":
            # M's gradient is final: its all-reduce runs under the rest of the backward pass, in M_CHUNKS pieces so that
            # the update of piece i (on the pipeline's side stream) can run while piece i + 1 is still being reduced
            self.comm.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.comm):
                for (lo, hi), ev in zip(self._m_chunks, self._m_events):
                    self._dist().all_reduce(self.store.grads[lo:hi], group=self.group)
                    ev.record(self.comm)
        elif name == "M:update":
            # ... and so does its Adam update, placed under the encoder's BPTT kernel (the caller's current stream is a
            # side branch forked there)
            # on a stream of its own: the backward pass must not wait for it, nor the trailing all-reduce
            self._upd_used = True
            self.upd.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.upd):
                for (lo, hi), ev in zip(self._m_chunks, self._m_events):
                    self.upd.wait_event(ev)
                    self.pipe.adam_m(1.0 / self.world, lo, hi, final=(hi == self.m_split))

    def _on_step(self, where):
        if where == "begin":                    # the previous step's tail of M: reduce + apply under the encoder forward
            self._comm2_used = True
            self.comm2.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.comm2):
                self._deferred_tail()
        elif where == "before_M" and self._comm2_used:
            torch.cuda.current_stream().wait_stream(self.comm2)
            self._comm2_used = False

    def _deferred_tail(self):
        self._dist().all_reduce(self.store.grads[self.m_split: self.m_end], group=self.group)
        self.pipe.adam_m(1.0 / self.world, self.m_split, self.m_end, final=False, enable=self.pending)

    def flush(self):
        """Applies what the deferred part of M's update still owes (collective: every rank calls it).  After it the
        weights are exactly those of the synchronous step; the next step's graph finds nothing pending."""
        if not self.defer or not self._primed:
            return
        from ._lib import LIB, current_stream, ptr
        with torch.cuda.stream(self.stream):
            self._deferred_tail()
            LIB.call("indnet_memset", current_stream(), ptr(self.pending), 0, 4)
        self.stream.synchronize()

    def _grad_sync(self):
        if self.world <= 1:
            return
        if self.overlap:
            self._dist().all_reduce(self.store.grads[self.m_end:], group=self.group)
            torch.cuda.current_stream().wait_stream(self.comm)
            if self._upd_used:                 # (a stream that took no part in this step must not be waited on under capture)
                torch.cuda.current_stream().wait_stream(self.upd)
                self._upd_used = False
        else:
            self._dist().all_reduce(self.store.grads, group=self.group)

    # ------------------------------------------------------------------ step bodies (enqueue only, no host sync)
    def train_body(self):
        """fwd + bwd + all-reduce + Adam on the episodes in the pipeline's input buffers."""
        self.pipe.forward(train=True)
        self.pipe.backward(early_adam=self.world == 1 or self.overlap)
        self._grad_sync()
        self.pipe.adam(grad_scale=1.0 / self.world)
        if self.defer:
            from ._lib import LIB, current_stream, ptr
            LIB.call("indnet_memset", current_stream(), ptr(self.pending), 1, 4)     # this step's tail of M is now owed

    def _walk(self):
        c = self.cfg
        self._py_prev.copy_(self.smp.py_state, non_blocking=True)
        self._np_prev.copy_(self.smp.np_state, non_blocking=True)
        self.smp.enqueue_walk(self.global_B, c.N, c.K, c.Q)

    def _gather(self):
        c = self.cfg
        b0 = self.rank * c.B if self.stream_mode == "global" else 0
        self.smp.enqueue_gather(self.global_B, c.N, c.K, c.Q, out=self.inputs, b0=b0, nb=c.B)

    def step_body(self):
        """This step's episodes = the picks drawn during the previous step; the walk for the NEXT step runs
        on the side stream underneath this step's forward / backward / update."""
        self._gather()
        cur = torch.cuda.current_stream()
        self.side.wait_stream(cur)
        with torch.cuda.stream(self.side):
            self._walk()
        self.train_body()
        cur.wait_stream(self.side)

    # ------------------------------------------------------------------ set-up
    def prime(self, warm: int = 2):
        """Runs ``warm`` eager steps on the trainer's stream (allocates the backward buffers, creates the
        sampler's internal stream, warms NCCL - all of which must happen outside stream capture), then
        captures the CUDA graphs.  With a sampler the first walk happens here, so step 1 trains on the first
        episodes the seeded streams produce."""
        if self._primed:
            return
        self.stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.stream):
            # warm-up steps must not move the weights or the streams' episode sequence out from under the
            # caller: they run on throw-away state that is restored afterwards
            snap = self._snapshot()
            if self.smp is not None:
                self._walk()
            for _ in range(max(1, warm)):
                if self.smp is not None:
                    self.step_body()
                else:
                    self.train_body()
            self.stream.synchronize()
            self._restore(snap)
            if self.smp is not None:
                self._walk()
            self.stream.synchronize()
            if self.use_graph:
                self._capture()
        self._primed = True

    def _snapshot(self):
        s = self.store
        snap = dict(params=s.params.clone(), m=s.m.clone(), v=s.v.clone(), step=s.global_step.clone(),
                    running=self.pipe.running.clone())
        if self.smp is not None:
            snap.update(py=self.smp.py_state.clone(), np=self.smp.np_state.clone())
        return snap

    def _restore(self, snap):
        s = self.store
        s.params.copy_(snap["params"]); s.m.copy_(snap["m"]); s.v.copy_(snap["v"])
        s.global_step.copy_(snap["step"])
        self.pipe.running.copy_(snap["running"])
        self.pending.zero_()
        s.refresh_bf16()
        if self.smp is not None:
            self.smp.py_state.copy_(snap["py"]); self.smp.np_state.copy_(snap["np"])

    def _capture(self):
        try:
            if self.smp is not None:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=self.stream):
                    self.step_body()
                self.graph_step = g
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=self.stream):
                self.train_body()
            self.graph_train = g
        except Exception as e:      # e.g. a collective that refuses capture: eager launches still work
            import sys
            print("EpisodeTrainer: CUDA graph capture failed, using eager launches: %r" % (e,), file=sys.stderr)
            self.graph_step = self.graph_train = None
            torch.cuda.synchronize()

    # ------------------------------------------------------------------ stepping
    def step(self):
        """One train step on freshly sampled episodes (device sampler in the loop).  Asynchronous."""
        if not self._primed:
            self.prime()
        if self.smp is None:
            raise RuntimeError("EpisodeTrainer.step() needs a device sampler; use step_on(inputs, label)")
        if self._need_walk:
            self.resume()
        with torch.cuda.stream(self.stream):
            if self.graph_step is not None:
                self.graph_step.replay()
            else:
                self.step_body()
        self.steps_done += 1

    def step_on(self, inputs=None, query_label=None):
        """One train step on caller-supplied episodes (host or device tensors / arrays, [B, n_s, L]); with no
        arguments, on whatever is in the pipeline's input buffers.  Asynchronous."""
        if not self._primed:
            self.prime()
        with torch.cuda.stream(self.stream):
            if inputs is not None:
                self.pipe.set_inputs(inputs, query_label)
            if self.graph_train is not None:
                self.graph_train.replay()
            else:
                self.train_body()
        self.steps_done += 1

    def read(self):
        """(loss, accuracy) of the last step - synchronises the trainer's stream (an 8-byte D2H copy)."""
        with torch.cuda.stream(self.stream):
            v = self.pipe.out2.cpu()
        return float(v[0]), float(v[1])

    class _Pending:
        """Result of ``read_async``: ``get()`` waits for that step's copy only (not for later steps)."""

        def __init__(self, buf, ev):
            self.buf, self.ev = buf, ev

        def get(self):
            self.ev.synchronize()
            return float(self.buf[0]), float(self.buf[1])

    def read_async(self):
        """Enqueues the 8-byte D2H copy of the last step's {loss, accuracy} into a pinned slot and returns a handle;
        a loop that calls ``get()`` on the PREVIOUS step's handle logs every step without ever draining the GPU
        (the reference's ``sess.run`` returns the same two numbers synchronously, model/base.py:148-152).  At most
        ``len(self._rd)`` handles may be outstanding."""
        if not hasattr(self, "_rd"):
            self._rd = [(torch.zeros(2, dtype=torch.float32).pin_memory(), torch.cuda.Event()) for _ in range(8)]
            self._rd_i = 0
        buf, ev = self._rd[self._rd_i % len(self._rd)]
        self._rd_i += 1
        with torch.cuda.stream(self.stream):
            buf.copy_(self.pipe.out2, non_blocking=True)
            ev.record(self.stream)
        return EpisodeTrainer._Pending(buf, ev)

    def read_running(self, reset: bool = False):
        """(mean loss, mean accuracy, steps) since the last reset - model/base.py:156-163's running means,
        kept on the device.  Synchronises."""
        with torch.cuda.stream(self.stream):
            v = self.pipe.running.cpu()
            if reset:
                self.pipe.running.zero_()
        n = max(float(v[2]), 1.0)
        return float(v[0]) / n, float(v[1]) / n, int(v[2])

    def suspend(self):
        """Hands the two host streams back exactly where the reference's loop would have them after the steps
        run so far: the look-ahead walk (the picks of the step that has not run yet) is discarded and the
        streams rewound to just before it.  Call before anything else draws from ``random`` / ``numpy.random``
        (the evaluation loop, model/base.py:168-170); ``step`` re-walks on its own afterwards."""
        self.flush()
        if self.smp is None or self._need_walk or not self._primed:
            return
        self.stream.synchronize()
        self.smp.check()
        self.smp.py_state.copy_(self._py_prev)
        self.smp.np_state.copy_(self._np_prev)
        self.smp.store_host_streams()
        self._need_walk = True

    def resume(self):
        self.smp.load_host_streams()
        with torch.cuda.stream(self.stream):
            self._walk()
        self._need_walk = False

    def check_sampler(self):
        if self.smp is not None:
            self.stream.synchronize()
            self.smp.check()

    def close(self):
        """Ordered teardown: drain, drop the graphs (they hold references to the NCCL communicator's work),
        drain again.  The caller destroys the process group afterwards."""
        self.flush()
        torch.cuda.synchronize()
        self.graph_step = self.graph_train = None
        self.pipe.grad_ready_hook = None
        self.pipe.step_hook = None
        torch.cuda.synchronize()
