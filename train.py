"""N-way K-shot few-shot relation classification on B200 - same CLI as the reference's train.py:

    python train.py [model_name] [N] [K] [--flags]

Positional arguments and banner lines follow reference train.py:10-21.  The reference hard-codes toy
constants (max_length 37, 5-d GloVe, hidden_size 20, train.py:24-36); they are the defaults here and
every one is a flag.  Without FewRel/GloVe files on disk (there is no network) ``--synthetic`` trains
on FewRel-shaped synthetic tables.
"""
import argparse
import sys

import induction_b200  # noqa: F401
from induction_b200.model.graph import InductionGraph
from induction_b200.util.data_loader import JSONFileDataLoader
from induction_b200.util.synthetic import make_tables


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("model_name", nargs="?", default="induction")
    ap.add_argument("N", nargs="?", type=int, default=5)
    ap.add_argument("K", nargs="?", type=int, default=2)
    ap.add_argument("--Q", type=int, default=5)
    ap.add_argument("--max-length", type=int, default=37)
    ap.add_argument("--hidden-size", type=int, default=20)
    ap.add_argument("--batch-size", type=int, default=1)
    ap.add_argument("--embedding-file", default="./data/glove.6B.5d.json")
    ap.add_argument("--train-file", default="./data/train.json")
    ap.add_argument("--val-file", default="./data/val.json")
    ap.add_argument("--synthetic", action="store_true")
    ap.add_argument("--train-iter", type=int, default=30000)
    ap.add_argument("--val-iter", type=int, default=1000)
    ap.add_argument("--val-step", type=int, default=2000)
    ap.add_argument("--test-iter", type=int, default=3000)
    ap.add_argument("--mode", default="fp32", choices=["fp32", "bf16"])
    ap.add_argument("--keep-prob", type=float, default=0.9)
    ap.add_argument("--seed", type=int, default=None, help="random.seed / np.random.seed before training (the reference "
                    "never seeds); REQUIRED to be the same on every rank under torchrun")
    ap.add_argument("--cuda-graph", action="store_true", help="device sampler + whole step as one CUDA graph "
                    "(EpisodeTrainer); implied under torchrun with WORLD_SIZE > 1, where --batch-size is per GPU")
    ap.add_argument("--scalar-log", default=None, help="JSONL file for the loss / accuracy / learning_rate scalars")
    ap.add_argument("--model-dir", default="checkpoints/inductionNetwork_test")
    a = ap.parse_args(argv)

    import os
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    if world > 1:
        import torch
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        if a.seed is None:
            a.seed = 0          # one global episode stream: every rank must walk the same draws
    if a.seed is not None:
        import random
        import numpy as np
        random.seed(a.seed); np.random.seed(a.seed)
    if rank == 0:
        print("{}-way-{}-shot Few-Shot Relation Classification".format(a.N, a.K))
        print("Model: {}".format(a.model_name))

    if a.synthetic:
        train_data_loader = JSONFileDataLoader.from_tables(make_tables(64, 700, a.max_length, 4000, 5, seed=0))
        val_data_loader = JSONFileDataLoader.from_tables(make_tables(16, 700, a.max_length, 4000, 5, seed=1))
        val_data_loader.word_vec_mat = train_data_loader.word_vec_mat
    else:
        train_data_loader = JSONFileDataLoader(a.train_file, a.embedding_file, max_length=a.max_length)
        val_data_loader = JSONFileDataLoader(a.val_file, a.embedding_file, max_length=a.max_length)

    if a.model_name == "induction":
        model = InductionGraph(N=a.N, K=a.K, Q=a.Q, pred_embed=train_data_loader.word_vec_mat,
                               sequence_length=a.max_length, hidden_size=a.hidden_size,
                               batch_size=a.batch_size, mode=a.mode, keep_prob=a.keep_prob, cuda_graph=a.cuda_graph,
                               scalar_log=a.scalar_log)
        model.train(dataloader=(train_data_loader, val_data_loader), model_dir_path=a.model_dir,
                    train_iter=a.train_iter, val_iter=a.val_iter, val_step=a.val_step, test_iter=a.test_iter)
    else:
        print("unknown model: {}".format(a.model_name))
        return 1
    if world > 1:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
