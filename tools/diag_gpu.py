"""Prints per-tensor forward/backward errors of the CUDA path vs the float64 oracle for every test
case (no asserts) - the first thing to run on a GPU box after touching a kernel."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _cases import CASES, make_case, oracle_forward, oracle_grads, rel_err  # noqa: E402
from induction_b200.engine import EpisodePipeline, ParamStore, tf_param_shapes  # noqa: E402

mode = "fp32"
args = sys.argv[1:]
if args and args[0] in ("fp32", "bf16"):
    mode, args = args[0], args[1:]
names = args or list(CASES)
for name in names:
    cfg, tables, inputs, label, params = make_case(name, seed=3, mode=mode)
    store = ParamStore(tf_param_shapes(cfg.N, cfg.L, cfg.E, cfg.u, cfg.S))
    store.load_state_dict(params)
    pipe = EpisodePipeline(cfg, store, torch.from_numpy(tables["word_vec_mat"]))
    pipe.set_inputs(inputs, label)
    try:
        pipe.forward(train=True)
        torch.cuda.synchronize()
        t0 = time.time()
        ref, gref = oracle_grads(cfg, tables, inputs, label, params)
        enc = torch.cat([pipe.es, pipe.eq], dim=1)
        print("== %s  (oracle %.1fs)" % (name, time.time() - t0))
        print("  fwd: x %.2e hout %.2e alphas %.2e enc %.2e cvec %.2e probs %.2e loss %.8f/%.8f acc %.4f/%.4f" % (
            rel_err(pipe.x.view(cfg.X, cfg.L, cfg.D), ref["x"]),
            rel_err((pipe.hout_bf.float() if pipe.bf16 else pipe.hout).view(cfg.X, cfg.L, cfg.H), ref["hout"]),
            rel_err(pipe.alphas.view(cfg.B, cfg.n_s, cfg.L, 1), ref["alphas"]),
            rel_err(enc, ref["encoder"]), rel_err(pipe.cvec.view(cfg.B, cfg.N, cfg.H), ref["class_vector"]),
            rel_err(pipe.probs, ref["probs"]), pipe.loss(), float(ref["loss"]), pipe.accuracy(), float(ref["accuracy"])))
        pipe.backward()
        torch.cuda.synchronize()
        got = store.grad_dict()
        print("  bwd: " + " ".join("%s=%.1e" % (k.split("/")[-2][:2] + "/" + k.split("/")[-1] if "/" in k else k,
                                                  rel_err(got[k], gref[k])) for k in gref))
    except Exception as e:  # keep going: one broken case should not hide the others
        print("== %s FAILED: %r" % (name, e))
