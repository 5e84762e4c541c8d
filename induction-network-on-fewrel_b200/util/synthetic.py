"""Synthetic FewRel-shaped inputs: there is no network for FewRel / GloVe, so
benchmarks and tests use tables with the layout ``util/data_loader.py:148-193`` of the reference
produces, generated from ``np.random.RandomState(seed)`` (a private stream - the global streams the
sampler consumes are untouched)."""
from __future__ import annotations

import json
import os

import numpy as np


def make_tables(num_relations: int = 64, per_relation: int = 700, L: int = 40, vocab: int = 400000,
                E: int = 50, seed: int = 0, min_len: int = 8):
    """Returns dict(word, pos1, pos2, mask [I, L] int32, length [I] int32, rel2scope {name: [lo, hi]},
    word_vec_mat [vocab+2, E] float32) shaped like the reference's pre-processed arrays.

    word ids ~ U[0, vocab) before ``length`` and BLANK_ID = vocab+1 after (data_loader.py:170-171);
    pos = t - head/tail + L (:182-183); mask in {1,2,3} before length, 0 after (:184-191);
    embedding rows N(0,1) L2-normalised (:138) with the two extra rows zero (:130)."""
    rs = np.random.RandomState(seed)
    I = num_relations * per_relation
    lo_len = min(min_len, L)
    length = rs.randint(lo_len, L + 1, size=I).astype(np.int32)
    t = np.arange(L, dtype=np.int32)[None, :]
    valid = t < length[:, None]
    word = rs.randint(0, vocab, size=(I, L)).astype(np.int32)
    word[~valid] = vocab + 1
    head = (rs.random_sample(I) * length).astype(np.int32)
    tail = (rs.random_sample(I) * length).astype(np.int32)
    pos1 = (t - head[:, None] + L).astype(np.int32)
    pos2 = (t - tail[:, None] + L).astype(np.int32)
    pmin = np.minimum(head, tail)[:, None]
    pmax = np.maximum(head, tail)[:, None]
    mask = np.where(t <= pmin, 1, np.where(t <= pmax, 2, 3)).astype(np.int32)
    mask[~valid] = 0
    rel2scope = {"P%d" % r: [r * per_relation, (r + 1) * per_relation] for r in range(num_relations)}
    mat = rs.standard_normal((vocab + 2, E)).astype(np.float32)
    mat[:vocab] /= np.sqrt((mat[:vocab] ** 2).sum(axis=1, keepdims=True))
    mat[vocab:] = 0.0
    return {"word": word, "pos1": pos1, "pos2": pos2, "mask": mask, "length": length,
            "rel2scope": rel2scope, "word_vec_mat": mat}


def write_fewrel_json(dirname: str, num_relations: int = 6, per_relation: int = 12, vocab: int = 30,
                      E: int = 5, max_tokens: int = 14, seed: int = 0, name: str = "train"):
    """Writes a tiny FewRel-schema data file and a GloVe-schema vector file (the schemas documented at
    data_loader.py:62-84; the data key the code actually reads is 'tokens', :117).  Returns paths."""
    rs = np.random.RandomState(seed)
    words = ["w%d" % i for i in range(vocab)]
    vecs = [{"word": w, "vec": [float(x) for x in rs.standard_normal(E)]} for w in words]
    data = {}
    for r in range(num_relations):
        inst = []
        for _ in range(per_relation):
            n = int(rs.randint(3, max_tokens + 1))
            toks = [words[int(i)] if rs.random_sample() > 0.1 else "OOV%d" % int(i)
                    for i in rs.randint(0, vocab, size=n)]
            toks = [w.upper() if rs.random_sample() < 0.2 else w for w in toks]
            h, tl = int(rs.randint(0, n)), int(rs.randint(0, n))
            inst.append({"h": ["head", "Q1", [[h]]], "t": ["tail", "Q2", [[tl]]], "tokens": toks})
        data["P%d" % (100 + r)] = inst
    os.makedirs(dirname, exist_ok=True)
    data_path = os.path.join(dirname, name + ".json")
    vec_path = os.path.join(dirname, "glove.tiny.%dd.json" % E)
    with open(data_path, "w", encoding="utf-8") as f:
        json.dump(data, f)
    with open(vec_path, "w", encoding="utf-8") as f:
        json.dump(vecs, f)
    return data_path, vec_path
