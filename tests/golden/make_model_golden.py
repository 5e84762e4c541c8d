"""Generates tests/golden/model_<case>.npz from the float64 torch oracle (oracle/induction_oracle.py).

PARITY UNPINNED: the reference ships no golden vectors and TensorFlow 1.x cannot run here, so these
fixtures freeze the *oracle's* outputs (inputs, parameters in TF layouts, forward tensors, loss,
gradients).  They guard against silent oracle drift and give the GPU tests a fixed target; they are
not an external pin.  Run:  python tests/golden/make_model_golden.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
from _cases import make_case, oracle_grads  # noqa: E402

for name in ("smoke", "k1", "ragged"):
    cfg, tables, inputs, label, params = make_case(name, seed=3)
    out, grads = oracle_grads(cfg, tables, inputs, label, params)
    used = np.unique(inputs["word"])
    blob = {"label": label, "word_vec_mat": tables["word_vec_mat"]}
    blob.update({"in_" + k: v for k, v in inputs.items()})
    blob.update({"p_" + k.replace("/", "__"): v.numpy() for k, v in params.items()})
    blob.update({"g_" + k.replace("/", "__"): v.numpy() for k, v in grads.items()})
    for k in ("hout", "encoder", "alphas", "class_vector", "probs", "loss", "accuracy"):
        blob["o_" + k] = out[k].detach().numpy()
    np.savez_compressed(os.path.join(HERE, "model_%s.npz" % name), **blob)
    print(name, "loss", float(out["loss"]), "acc", float(out["accuracy"]))
