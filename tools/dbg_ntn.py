import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from _cases import make_case
from induction_b200.engine import EpisodePipeline, ParamStore, tf_param_shapes
import gc
for name in (sys.argv[1:] or ["k1"]):
  gc.collect()
  print("=====", name)
  cfg, tables, inputs, label, params = make_case(name, seed=3, mode="bf16")
  store = ParamStore(tf_param_shapes(cfg.N, cfg.L, cfg.E, cfg.u, cfg.S)); store.load_state_dict(params)
  pipe = EpisodePipeline(cfg, store, torch.from_numpy(tables["word_vec_mat"]))
  pipe.set_inputs(inputs, label); pipe.forward(train=True); pipe.backward(); torch.cuda.synchronize()
  c = cfg; NS = c.N * c.S
  for nm in ("x_bf", "xproj", "hout", "hout_bf", "proj", "es", "eq", "es_bf", "ehat", "cvec", "cvec_bf", "Tp_bf", "eq_bf", "V"):
      t = getattr(pipe, nm)
      print("   ", nm, "finite" if torch.isfinite(t.float()).all() else "NAN count %d of %d" % (int((~torch.isfinite(t.float())).sum()), t.numel()))
  print("    params_bf", bool(torch.isfinite(store.params_bf.float()).all()))
  eq = pipe.eq.double(); Tp = pipe.Tp_bf.double().view(c.B, NS, c.H)
  V_ref = torch.einsum("bqh,bnh->bqn", eq, Tp)
  print("V err", float((pipe.V.view(c.B, c.n_q, NS).double() - V_ref).abs().max()), "V max", float(V_ref.abs().max()), "finite", bool(torch.isfinite(pipe.V).all()))
  dV = pipe.dV.double().view(c.B, c.n_q, NS)
  print("dV finite", bool(torch.isfinite(pipe.dV).all()), "dV_bf finite", bool(torch.isfinite(pipe.dV_bf.float()).all()))
  deq_ref = torch.einsum("bqn,bnh->bqh", dV, Tp)
  print("deq err", float((pipe.deq.double() - deq_ref).abs().max()), "max", float(deq_ref.abs().max()), "finite", bool(torch.isfinite(pipe.deq).all()))
  dTp_ref = torch.einsum("bqn,bqh->bnh", dV, eq)
  print("dTp err", float((pipe.dTp_bf.double().view(c.B, NS, c.H) - dTp_ref).abs().max()), "max", float(dTp_ref.abs().max()), "finite", bool(torch.isfinite(pipe.dTp_bf.float()).all()))
  g = store.grad_dict()
  for k, v in g.items():
      print(k, "finite" if torch.isfinite(v).all() else "NAN", float(v.abs().max()) if torch.isfinite(v).all() else "")
  print("dlogit finite", bool(torch.isfinite(pipe.dlogit).all()), "probs finite", bool(torch.isfinite(pipe.probs).all()))
