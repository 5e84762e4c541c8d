"""Per-parameter gradient error of the bf16 tensor-core mode against the float64 oracle over several seeds
(which tensor sets the BF16_GRAD_TOL bound).  Test infrastructure: imports the oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from _cases import make_case, oracle_grads, rel_err, bf16_grad_report
from test_pipeline_gpu import build_pipeline
from induction_b200.engine import dropout_keep_mask

name = sys.argv[1] if len(sys.argv) > 1 else "cfg0_5w5s"
for keep in (1.0, 0.9):
    for seed in (3, 4, 5, 11, 12, 13):
        cfg, tables, inputs, label, params = make_case(name, seed=seed, keep_prob=keep, mode="bf16")
        pipe = build_pipeline(cfg, tables, params, seed=3)
        pipe.set_inputs(inputs, label)
        pipe.forward(train=True)
        pipe.backward()
        torch.cuda.synchronize()
        mask = torch.from_numpy(dropout_keep_mask(3, (cfg.X, cfg.L, cfg.H), keep, step=0)) if keep < 1 else None
        ref, gref = oracle_grads(cfg, tables, inputs, label, params, drop_mask=mask, dtype=torch.float64)
        errs, worst, cos = bf16_grad_report(pipe.p.grad_dict(), gref)
        print("keep %.1f seed %2d probs %.1e cos %.6f worst %s %.1e | %s" % (
            keep, seed, rel_err(pipe.probs, ref["probs"]), cos, worst, errs[worst],
            " ".join("%s=%.0e" % (k.split("/")[-1][:10], v) for k, v in errs.items())), flush=True)
