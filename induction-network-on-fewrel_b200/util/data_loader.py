"""Episode data loader - host mirror of the reference's ``util/data_loader.py``.

Keeps the reference's public surface (``JSONFileDataLoader(file_name, word_vec_file_name,
max_length=40, case_sensitive=False, reprocess=False, cuda=True)``, ``next_one_tf(N, K, Q)``,
``next_one(N, K, Q)``, the ``next_batch(B, N, K, Q)`` hook of ``FileDataLoader``, attributes
``data_word/pos1/pos2/mask/length``, ``rel2scope``, ``word_vec_mat``, ``word2id``) and its
``_processed_data/`` cache format (data_loader.py:21-58, 197-211), so arrays pre-processed by the
reference load here unchanged and vice versa.

Episode *sampling and row gathering* run on the GPU (``csrc/sampler.cu`` through the C-ABI entry
point ``indnet_sample_episodes``): the two Mersenne-Twister streams the reference consumes
(``random`` for the class draw, data_loader.py:260, and NumPy's global legacy stream for the
instance permutations, :267/:296) are advanced on the device bit-for-bit, and handed back to the
host modules after each call so host and device draws stay interchangeable.  There is no CPU
sampling path in this package: without the CUDA library the sampling calls raise.
"""
from __future__ import annotations

import json
import os
import random

import numpy as np


class FileDataLoader:
    def next_batch(self, B, N, K, Q):
        """B consecutive ``next_one_tf`` episodes stacked on a new leading axis (data_loader.py:9-17
        documents the signature; the reference leaves it abstract)."""
        raise NotImplementedError


_CACHE_DIR = "_processed_data"
_TABLE_KEYS = ("word", "pos1", "pos2", "mask")


def _prefix(path):
    return ".".join(os.path.basename(path).split(".")[:-1])


class JSONFileDataLoader(FileDataLoader):
    def __init__(self, file_name, word_vec_file_name, max_length=40, case_sensitive=False,
                 reprocess=False, cuda=True):
        self.file_name = file_name
        self.word_vec_file_name = word_vec_file_name
        self.case_sensitive = case_sensitive
        self.max_length = max_length
        self.cuda = cuda
        self._dev = None
        if reprocess or not self._load_preprocessed_file():
            if file_name is None or not os.path.isfile(file_name):
                raise Exception("[ERROR] Data file doesn't exist")
            if word_vec_file_name is None or not os.path.isfile(word_vec_file_name):
                raise Exception("[ERROR] Word vector file doesn't exist")
            self._preprocess()
            self._store_preprocessed_file()

    # ------------------------------------------------------------------ construction from arrays
    @classmethod
    def from_tables(cls, tables, max_length=None):
        """Builds a loader directly from pre-processed arrays (``util.synthetic.make_tables``)."""
        self = cls.__new__(cls)
        self.file_name = self.word_vec_file_name = None
        self.case_sensitive = False
        self.cuda = True
        self._dev = None
        self.data_word = np.ascontiguousarray(tables["word"], dtype=np.int32)
        self.data_pos1 = np.ascontiguousarray(tables["pos1"], dtype=np.int32)
        self.data_pos2 = np.ascontiguousarray(tables["pos2"], dtype=np.int32)
        self.data_mask = np.ascontiguousarray(tables["mask"], dtype=np.int32)
        self.data_length = np.ascontiguousarray(tables["length"], dtype=np.int32)
        self.rel2scope = dict(tables["rel2scope"])
        self.word_vec_mat = np.ascontiguousarray(tables["word_vec_mat"], dtype=np.float32)
        self.word2id = tables.get("word2id", {})
        self.max_length = int(max_length or self.data_word.shape[1])
        return self

    # ------------------------------------------------------------------ cache (data_loader.py:21-58)
    def _cache_names(self):
        p, w = _prefix(self.file_name), _prefix(self.word_vec_file_name)
        j = lambda n: os.path.join(_CACHE_DIR, n)  # noqa: E731
        return {"word": j(p + "_word.npy"), "pos1": j(p + "_pos1.npy"), "pos2": j(p + "_pos2.npy"),
                "mask": j(p + "_mask.npy"), "length": j(p + "_length.npy"),
                "rel2scope": j(p + "_rel2scope.json"), "mat": j(w + "_mat.npy"),
                "word2id": j(w + "_word2id.json")}

    def _load_preprocessed_file(self):
        if self.file_name is None or self.word_vec_file_name is None or not os.path.isdir(_CACHE_DIR):
            return False
        names = self._cache_names()
        if not all(os.path.exists(n) for n in names.values()):
            return False
        print("Pre-processed files exist. Loading them...")
        self.data_word = np.load(names["word"])
        self.data_pos1 = np.load(names["pos1"])
        self.data_pos2 = np.load(names["pos2"])
        self.data_mask = np.load(names["mask"])
        self.data_length = np.load(names["length"])
        with open(names["rel2scope"]) as f:
            self.rel2scope = json.load(f)
        self.word_vec_mat = np.load(names["mat"])
        with open(names["word2id"]) as f:
            self.word2id = json.load(f)
        if self.data_word.shape[1] != self.max_length:
            print("Pre-processed files don't match current settings. Reprocessing...")
            return False
        print("Finish loading")
        return True

    def _store_preprocessed_file(self):
        names = self._cache_names()
        os.makedirs(_CACHE_DIR, exist_ok=True)
        np.save(names["word"], self.data_word)
        np.save(names["pos1"], self.data_pos1)
        np.save(names["pos2"], self.data_pos2)
        np.save(names["mask"], self.data_mask)
        np.save(names["length"], self.data_length)
        with open(names["rel2scope"], "w") as f:
            json.dump(self.rel2scope, f)
        np.save(names["mat"], self.word_vec_mat)
        with open(names["word2id"], "w") as f:
            json.dump(self.word2id, f)

    # ------------------------------------------------------------------ JSON -> arrays (data_loader.py:104-193)
    def _preprocess(self):
        L = self.max_length
        with open(self.file_name, "r", encoding="utf-8") as f:
            ori_data = json.load(f)
        with open(self.word_vec_file_name, "r", encoding="utf-8") as f:
            ori_vec = json.load(f)
        fold = (lambda s: s) if self.case_sensitive else (lambda s: s.lower())

        tot = len(ori_vec)
        unk_id, blank_id = tot, tot + 1
        dim = len(ori_vec[0]["vec"])
        mat = np.zeros((tot + 2, dim), dtype=np.float32)
        word2id = {}
        for cur, entry in enumerate(ori_vec):
            word2id[fold(entry["word"])] = cur            # later duplicates win, as in the reference
            row = np.asarray(entry["vec"], dtype=np.float32)
            mat[cur] = row / np.sqrt(np.sum(row ** 2))    # float32 arithmetic, row by row (:138)
        word2id["UNK"] = unk_id
        word2id["BLANK"] = blank_id

        total = sum(len(v) for v in ori_data.values())
        word = np.full((total, L), 0, dtype=np.int32)
        pos1 = np.zeros((total, L), dtype=np.int32)
        pos2 = np.zeros((total, L), dtype=np.int32)
        mask = np.zeros((total, L), dtype=np.int32)
        length = np.zeros((total,), dtype=np.int32)
        rel2scope = {}
        t = np.arange(L, dtype=np.int32)
        row = 0
        for relation, instances in ori_data.items():
            start = row
            for ins in instances:
                toks = [fold(w) for w in ins["tokens"]]
                n = min(len(toks), L)
                ids = [word2id.get(w, unk_id) for w in toks[:n]]
                word[row, :n] = ids
                # the reference pads with BLANK from (last token index + 1), i.e. only when the
                # sentence is shorter than L (:170-171)
                if len(toks) < L:
                    word[row, len(toks):] = blank_id
                length[row] = n
                p1 = min(ins["h"][2][0][0], L - 1)
                p2 = min(ins["t"][2][0][0], L - 1)
                pos1[row] = t - p1 + L
                pos2[row] = t - p2 + L
                lo, hi = min(p1, p2), max(p1, p2)
                m = np.where(t <= lo, 1, np.where(t <= hi, 2, 3))
                m[t >= n] = 0
                mask[row] = m
                row += 1
            rel2scope[relation] = [start, row]
        self.data_word, self.data_pos1, self.data_pos2 = word, pos1, pos2
        self.data_mask, self.data_length = mask, length
        self.rel2scope, self.word_vec_mat, self.word2id = rel2scope, mat, word2id
        self.word_vec_tot, self.word_vec_dim, self.instance_tot = tot, dim, total

    # ------------------------------------------------------------------ device sampler
    def scopes(self):
        """[(lo, hi)] in ``rel2scope`` dict order - the order ``random.sample(keys, N)`` indexes."""
        return [(int(v[0]), int(v[1])) for v in self.rel2scope.values()]

    def _device(self):
        if self._dev is None:
            from ..sampler import DeviceEpisodeSampler
            self._dev = DeviceEpisodeSampler(self)
        return self._dev

    def next_batch(self, B, N, K, Q):
        """B consecutive ``next_one_tf`` draws, sampled and gathered on the GPU.  Returns
        ``(inputs, query_label)`` with CUDA tensors ``inputs[k]`` int32 [B, N*(K+Q), L] and
        ``query_label`` int32 [B, N*Q].  Consumes (and hands back) the global ``random`` and
        ``numpy.random`` streams exactly as B reference calls would."""
        dev = self._device()
        dev.load_host_streams()
        out = dev.sample(B, N, K, Q)
        dev.store_host_streams()
        return {k: out[k] for k in _TABLE_KEYS}, out["label"]

    def next_one_tf(self, N, K, Q):
        """Reference ``next_one_tf`` (data_loader.py:259-309): numpy int32 arrays [n_s, L] with the
        class-major support rows first and the permuted query rows after, plus ``query_label``."""
        inputs, label = self.next_batch(1, N, K, Q)
        return ({k: v[0].cpu().numpy() for k, v in inputs.items()}, label[0].cpu().numpy().astype(np.int64))

    def next_one(self, N, K, Q):
        """Reference ``next_one`` (data_loader.py:213-257): same draws as ``next_one_tf`` but the
        support set stays stacked [N, K, L] and the query set is returned separately."""
        inputs, label = self.next_one_tf(N, K, Q)
        L = self.max_length
        sup = {k: v[: N * K].reshape(N, K, L) for k, v in inputs.items()}
        qry = {k: v[N * K:] for k, v in inputs.items()}
        return sup, qry, label
