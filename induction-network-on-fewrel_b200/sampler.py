"""Device episode sampler: host wrapper of ``indnet_sample_episodes`` (csrc/sampler.cu).

Keeps the two MT19937 streams of the reference's sampler (``random`` and NumPy's legacy global
stream, util/data_loader.py:260,267,296) on the GPU as (key[624], pos) arrays.  ``load_host_streams``
/ ``store_host_streams`` move them from / to the ``random`` and ``numpy.random`` modules, so a program
may mix host draws and device draws and get exactly the sequence the reference would.
"""
from __future__ import annotations

import random

import numpy as np
import torch

from ._lib import LIB, IndnetError, current_stream, ptr, require_cuda

KEYS = ("word", "pos1", "pos2", "mask")


class DeviceEpisodeSampler:
    def __init__(self, loader, device="cuda"):
        require_cuda()
        LIB.load()
        self.device = torch.device(device)
        self.L = int(loader.max_length)
        sc = loader.scopes()
        if len(sc) > 1024:
            raise ValueError("at most 1024 relations are supported by the device sampler")
        self.R = len(sc)
        self.n_max = max(b - a for a, b in sc)
        self.scopes = torch.tensor(sc, dtype=torch.int32, device=self.device).contiguous()
        host = {k: np.ascontiguousarray(getattr(loader, "data_" + k), dtype=np.int32) for k in KEYS}
        self._id_ranges = {k: ((int(host[k].min()), int(host[k].max())) if host[k].size else (0, 0)) for k in ("word", "pos1", "pos2")}
        self.tables = {k: torch.from_numpy(host[k]).to(self.device) for k in KEYS}
        self.py_state = torch.zeros(625, dtype=torch.int32, device=self.device)
        self.np_state = torch.zeros(625, dtype=torch.int32, device=self.device)
        self.status = torch.zeros(16, dtype=torch.int32, device=self.device)  # {status, cycles/16, jumps, rows, timeline build: 6 more}
        self._bufs = {}
        self.load_host_streams()

    def validate_ids(self, vocab_rows: int, sequence_length: int):
        """The embedding kernels index their tables with the sampled ids unchecked (TF raises on an out-of-range id):
        a model checks the instance tables once against ITS table sizes before training on them."""
        lo, hi = self._id_ranges["word"]
        if lo < 0 or hi >= vocab_rows:
            raise ValueError("word ids [%d, %d] outside the embedding table of %d rows" % (lo, hi, vocab_rows))
        for k in ("pos1", "pos2"):
            lo, hi = self._id_ranges[k]
            if lo < 0 or hi >= 2 * sequence_length:
                raise ValueError("%s ids [%d, %d] outside [0, %d)" % (k, lo, hi, 2 * sequence_length))

    # ------------------------------------------------------------------ stream hand-over
    def load_host_streams(self):
        st = random.getstate()[1]
        self.py_state.copy_(torch.from_numpy(np.array(st, dtype=np.uint32).view(np.int32)))
        ns = np.random.get_state()
        arr = np.concatenate([np.asarray(ns[1], dtype=np.uint32), np.array([ns[2]], dtype=np.uint32)])
        self.np_state.copy_(torch.from_numpy(arr.view(np.int32)))

    def store_host_streams(self):
        py = self.py_state.cpu().numpy().view(np.uint32)
        random.setstate((3, tuple(int(x) for x in py), None))
        ns = self.np_state.cpu().numpy().view(np.uint32)
        np.random.set_state(("MT19937", ns[:624].copy(), int(ns[624]), 0, 0.0))

    def seed(self, seed: int):
        """random.seed(seed); np.random.seed(seed) - then move both streams to the device."""
        random.seed(seed)
        np.random.seed(seed)
        self.load_host_streams()

    # ------------------------------------------------------------------ sampling
    def buffers(self, B, N, K, Q):
        key = (B, N, K, Q)
        if key not in self._bufs:
            n_s, n_q = N * (K + Q), N * Q
            nbytes = LIB.load().indnet_sampler_scratch_bytes(B, N, K, Q, self.n_max)
            z = lambda *s: torch.zeros(*s, dtype=torch.int32, device=self.device)  # noqa: E731
            self._bufs[key] = dict(scratch=torch.zeros(nbytes, dtype=torch.uint8, device=self.device),
                                   rows=z(B, n_s), label=z(B, n_q), **{k: z(B, n_s, self.L) for k in KEYS})
        return self._bufs[key]

    def enqueue(self, B, N, K, Q, out=None):
        """Enqueues sampling + gather of B episodes on the current stream (no host sync).  ``out``:
        optional dict of preallocated int32 CUDA tensors (word/pos1/pos2/mask [B,n_s,L], label [B,n_q],
        rows [B,n_s]); the sampler's own static buffers otherwise."""
        if N > self.R:
            raise ValueError("Sample larger than population or is negative")
        bufs = self.buffers(B, N, K, Q)
        o = dict(bufs)
        if out is not None:
            o.update(out)
        LIB.call("indnet_sample_episodes", current_stream(), ptr(self.py_state), ptr(self.np_state),
                 ptr(self.scopes), self.R, ptr(self.tables["word"]), ptr(self.tables["pos1"]),
                 ptr(self.tables["pos2"]), ptr(self.tables["mask"]), self.L, B, N, K, Q, self.n_max,
                 ptr(bufs["scratch"]), ptr(o["rows"]), ptr(o["label"]), ptr(o["word"]), ptr(o["pos1"]),
                 ptr(o["pos2"]), ptr(o["mask"]), ptr(self.status))
        return o

    def enqueue_walk(self, B, N, K, Q):
        """First half: walk both streams for B episodes and leave the picks in the scratch buffer."""
        if N > self.R:
            raise ValueError("Sample larger than population or is negative")
        bufs = self.buffers(B, N, K, Q)
        LIB.call("indnet_sample_walk", current_stream(), ptr(self.py_state), ptr(self.np_state), ptr(self.scopes),
                 self.R, B, N, K, Q, self.n_max, ptr(bufs["scratch"]), ptr(self.status))

    def enqueue_gather(self, B, N, K, Q, out=None, b0=0, nb=None):
        """Second half: picks -> row ids, labels and the gathered [nb, n_s, L] tables of episodes
        [b0, b0 + nb) of the walked batch (default: all B).  Data-parallel ranks walk the same global batch
        and gather only their own episodes, so episode b is the same whatever the world size."""
        bufs = self.buffers(B, N, K, Q)
        o = dict(bufs)
        if out is not None:
            o.update(out)
        nb = B if nb is None else nb
        LIB.call("indnet_sample_gather_range", current_stream(), ptr(self.scopes), ptr(self.tables["word"]),
                 ptr(self.tables["pos1"]), ptr(self.tables["pos2"]), ptr(self.tables["mask"]), self.L, B, N, K, Q,
                 self.n_max, ptr(bufs["scratch"]), b0, nb, ptr(o["rows"]), ptr(o["label"]), ptr(o["word"]),
                 ptr(o["pos1"]), ptr(o["pos2"]), ptr(o["mask"]))
        return o

    def check(self):
        """Synchronising status check (mirrors the ValueError np.random.choice raises on the host)."""
        s = int(self.status[0].item())
        if s == 1:
            raise ValueError("Cannot take a larger sample than population when 'replace=False'")
        if s == 2:
            raise IndnetError("device sampler: a relation has more instances than n_max=%d" % self.n_max)
        if s == 3:
            raise IndnetError("device sampler: the batch consumed more NumPy draws than the pre-generated stream "
                              "budget (expected draws + 3 % + 8192); both streams are unchanged - sample() retries "
                              "in smaller batches, an asynchronous caller must re-issue the step")
        if s != 0:
            raise IndnetError("device sampler failed with status %d" % s)

    def sample(self, B, N, K, Q):
        """Synchronous sampling.  A batch that outruns the pre-generated stream budget (status 3; the call
        then leaves both streams untouched) is redone as two half batches - consecutive draws from the same
        streams, i.e. the same episodes - each with the full fixed slack of its own."""
        o = self.enqueue(B, N, K, Q)
        if int(self.status[0].item()) == 3 and B > 1:
            self.status.zero_()
            h = B // 2
            first = {k: v.clone() for k, v in self.sample(h, N, K, Q).items() if k != "scratch"}
            second = self.sample(B - h, N, K, Q)
            return {k: torch.cat([first[k], second[k]], dim=0) for k in first}
        self.check()
        return o
