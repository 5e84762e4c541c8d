"""CPU oracle for the episode sampler (util/data_loader.py:259-309, ``next_one_tf``).

TEST INFRASTRUCTURE ONLY.  PARITY PINNED: ``tests/golden/sampler_*.npz`` were produced by importing
the reference's own ``util/data_loader.py`` in the build container (``tests/golden/make_sampler_golden.py``)
and both restatements below are checked against them bit for bit.

Two restatements:

* ``next_one_tf_stdlib`` - the reference algorithm driven by the *real* ``random`` / ``numpy.random``
  global streams (exactly the calls the reference makes), usable wherever CPython + NumPy exist.
* ``MT19937`` + ``EpisodeSampler`` - a from-scratch restatement of the two RNG streams and of the
  algorithms CPython's ``random.sample`` and NumPy's legacy ``RandomState.permutation`` run on them
  (SURVEY.md Appendix A.7).  This is the algorithm the CUDA sampler implements, state-compatible with
  ``random.getstate()`` / ``np.random.get_state()``.
"""
from __future__ import annotations

import random
from math import ceil, log
from typing import List, Sequence, Tuple

import numpy as np

_N, _M = 624, 397
_MASK32 = 0xFFFFFFFF


class MT19937:
    """Mersenne Twister with the (key[624], pos) state layout CPython and NumPy both expose."""

    def __init__(self, key: Sequence[int], pos: int):
        self.key = [int(k) & _MASK32 for k in key]
        assert len(self.key) == _N
        self.pos = int(pos)

    @classmethod
    def from_python_random(cls) -> "MT19937":
        st = random.getstate()[1]
        return cls(st[:_N], st[_N])

    @classmethod
    def from_numpy_global(cls) -> "MT19937":
        st = np.random.get_state()
        return cls(st[1].tolist(), int(st[2]))

    def to_python_random(self) -> None:
        random.setstate((3, tuple(self.key) + (self.pos,), None))

    def to_numpy_global(self) -> None:
        np.random.set_state(("MT19937", np.array(self.key, dtype=np.uint32), self.pos, 0, 0.0))

    def _twist(self) -> None:
        k = self.key
        for i in range(_N):
            y = (k[i] & 0x80000000) | (k[(i + 1) % _N] & 0x7FFFFFFF)
            k[i] = k[(i + _M) % _N] ^ (y >> 1) ^ (0x9908B0DF if (y & 1) else 0)
        self.pos = 0

    def next_uint32(self) -> int:
        if self.pos >= _N:
            self._twist()
        y = self.key[self.pos]
        self.pos += 1
        y ^= y >> 11
        y ^= (y << 7) & 0x9D2C5680
        y ^= (y << 15) & 0xEFC60000
        y ^= y >> 18
        return y & _MASK32


def py_randbelow(mt: MT19937, n: int) -> int:
    """CPython Random._randbelow_with_getrandbits: k = n.bit_length(); draw top-k bits until < n."""
    k = n.bit_length()
    r = mt.next_uint32() >> (32 - k)
    while r >= n:
        r = mt.next_uint32() >> (32 - k)
    return r


def py_sample_indices(mt: MT19937, n: int, k: int) -> List[int]:
    """CPython random.sample(population, k) index selection for a population of length n
    (reference call site: util/data_loader.py:260, ``random.sample(self.rel2scope.keys(), N)``)."""
    if not 0 <= k <= n:
        raise ValueError("Sample larger than population or is negative")
    result = [0] * k
    setsize = 21
    if k > 5:
        setsize += 4 ** ceil(log(k * 3, 4))
    if n <= setsize:
        pool = list(range(n))
        for i in range(k):
            j = py_randbelow(mt, n - i)
            result[i] = pool[j]
            pool[j] = pool[n - i - 1]
    else:
        selected = set()
        for i in range(k):
            j = py_randbelow(mt, n)
            while j in selected:
                j = py_randbelow(mt, n)
            selected.add(j)
            result[i] = j
    return result


def np_random_interval(mt: MT19937, mx: int) -> int:
    """NumPy legacy ``random_interval`` (32-bit branch): smallest all-ones mask >= mx, rejection."""
    if mx == 0:
        return 0
    mask = mx
    for s in (1, 2, 4, 8, 16):
        mask |= mask >> s
    v = mt.next_uint32() & mask
    while v > mx:
        v = mt.next_uint32() & mask
    return v


def np_permutation(mt: MT19937, n: int) -> List[int]:
    """NumPy legacy RandomState.permutation(n): Fisher-Yates on arange(n), i from n-1 down to 1."""
    a = list(range(n))
    for i in range(n - 1, 0, -1):
        j = np_random_interval(mt, i)
        a[i], a[j] = a[j], a[i]
    return a


class EpisodeSampler:
    """From-scratch restatement of ``JSONFileDataLoader.next_one_tf`` (util/data_loader.py:259-309)
    returning *row ids* plus labels; ``gather`` applies them to the four tables."""

    def __init__(self, scopes: Sequence[Tuple[int, int]], py_state: MT19937, np_state: MT19937):
        self.scopes = [(int(a), int(b)) for a, b in scopes]   # rel2scope values in dict order
        self.py = py_state
        self.np = np_state

    def next_rows(self, N: int, K: int, Q: int):
        classes = py_sample_indices(self.py, len(self.scopes), N)          # :260
        support_rows, query_rows, query_label = [], [], []
        for i, ci in enumerate(classes):                                    # :265
            lo, hi = self.scopes[ci]
            perm = np_permutation(self.np, hi - lo)[: K + Q]                # :267 (choice w/o replacement)
            idx = [lo + p for p in perm]
            support_rows += idx[:K]                                         # :272-275
            query_rows += idx[K:K + Q]
            query_label += [i] * Q                                          # :284
        perm = np_permutation(self.np, N * Q)                               # :296
        query_rows = [query_rows[p] for p in perm]                          # :297-300
        query_label = [query_label[p] for p in perm]                        # :301
        rows = np.array(support_rows + query_rows, dtype=np.int64)          # :303-307
        return rows, np.array(query_label, dtype=np.int64), classes


def gather(tables, rows):
    """util/data_loader.py:268-271: fancy-index the four [instances, L] int32 tables."""
    return {k: np.asarray(tables[k])[rows] for k in ("word", "pos1", "pos2", "mask")}


def next_one_tf_stdlib(tables, scopes: Sequence[Tuple[int, int]], N: int, K: int, Q: int):
    """The reference algorithm on the real global ``random`` / ``np.random`` streams, statement for
    statement (util/data_loader.py:259-309).  ``scopes`` = list(rel2scope.values()) in dict order."""
    target = random.sample(range(len(scopes)), N)           # same draws as sample(keys, N): len + order only
    sup = {k: [] for k in ("word", "pos1", "pos2", "mask")}
    qry = {k: [] for k in ("word", "pos1", "pos2", "mask")}
    query_label = []
    rows_s, rows_q = [], []
    for i, ci in enumerate(target):
        lo, hi = scopes[ci]
        indices = np.random.choice(list(range(lo, hi)), K + Q, False)
        rows_s.append(indices[:K]); rows_q.append(indices[K:K + Q])
        for k in sup:
            t = np.asarray(tables[k])[indices]
            sup[k].append(t[:K]); qry[k].append(t[K:K + Q])
        query_label += [i] * Q
    perm = np.random.permutation(N * Q)
    out = {}
    for k in sup:
        out[k] = np.concatenate([np.concatenate(sup[k], 0), np.concatenate(qry[k], 0)[perm]])
    rows = np.concatenate([np.concatenate(rows_s), np.concatenate(rows_q)[perm]])
    return out, np.array(query_label)[perm], rows, target
