"""torchrun worker for tests/test_trainer_gpu.py: one data-parallel train step of the global batch, one rank
per GPU.  Writes this rank's view (episodes, averaged gradient slots, parameters after Adam, loss) to
``<out>.<rank>.npz``."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main(out, global_B, mode, steps):
    import induction_b200  # noqa: F401
    from induction_b200 import dist as idist
    from induction_b200.engine import EpisodeConfig, EpisodePipeline, ParamStore, init_tf_params, tf_param_shapes
    from induction_b200.trainer import EpisodeTrainer
    from induction_b200.util.data_loader import JSONFileDataLoader
    from induction_b200.util.synthetic import make_tables
    rank, local, world = idist.init_from_env()
    torch.cuda.set_device(local)
    B = global_B // world
    u = 128 if mode == "bf16" else 16
    cfg = EpisodeConfig(N=5, K=2, Q=5, B=B, L=12, E=50 if mode == "bf16" else 8, u=u, S=12, keep_prob=1.0, mode=mode)
    tables = make_tables(16, 40, cfg.L, 500, cfg.E, seed=0, min_len=3)
    store = ParamStore(tf_param_shapes(cfg.N, cfg.L, cfg.E, cfg.u, cfg.S))
    store.load_state_dict(init_tf_params(store.tf_shapes, seed=0))
    pipe = EpisodePipeline(cfg, store, torch.from_numpy(tables["word_vec_mat"]), seed=5)
    smp = JSONFileDataLoader.from_tables(tables)._device()
    smp.seed(3)
    tr = EpisodeTrainer(pipe, smp, rank=rank, world=world, cuda_graph=True)
    losses = []
    for _ in range(steps):
        tr.step()
        losses.append(tr.read()[0])
    tr.check_sampler()
    tr.flush()                       # a deferred tail of M's update (INDNET_M_DEFER) is applied before the weights are read
    torch.cuda.synchronize()
    np.savez(out + ".%d.npz" % rank, word=pipe.word.cpu().numpy(), label=pipe.label.cpu().numpy(),
             grads=store.grads.cpu().numpy(), params=store.params.cpu().numpy(), losses=np.array(losses),
             adam_m=store.m.cpu().numpy(), adam_v=store.v.cpu().numpy(), step=store.global_step.cpu().numpy(),
             deferred=np.array(bool(tr.defer)),
             graph=np.array(tr.graph_step is not None))
    tr.close()
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]), sys.argv[3], int(sys.argv[4]))
