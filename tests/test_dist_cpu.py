"""CPU, gloo, world_size 2: the data-parallel host logic - episode sharding, one flat all-reduce, and
shard-weighted gradient averaging reproduce the single-process gradient of the global batch (checked
with the CPU oracle as the per-rank 'pipeline')."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    import induction_b200  # noqa: F401
    from induction_b200 import dist as idist
    from _cases import make_case, oracle_grads
    r, _, w = idist.init_from_env("gloo")
    cfg, tables, inputs, label, params = make_case("ragged", seed=3)      # B = 5 episodes: uneven shards
    lo, hi = idist.episode_shard(cfg.B, r, w)
    sub = {k: v[lo:hi] for k, v in inputs.items()}
    cfg_r = type(cfg)(**{**cfg.__dict__, "B": hi - lo})
    _, g = oracle_grads(cfg_r, tables, sub, label[lo:hi], params)
    names = sorted(g)
    flat = torch.cat([g[k].reshape(-1) for k in names]) * idist.grad_scale(hi - lo, cfg.B)
    idist.allreduce_sum_(flat)
    p0 = torch.full((3,), float(r))
    idist.broadcast_(p0, 0)
    if r == 0:
        _, g_full = oracle_grads(cfg, tables, inputs, label, params)
        want = torch.cat([g_full[k].reshape(-1) for k in names])
        q.put((float((flat - want).abs().max() / want.abs().max()), (lo, hi)))
    else:
        q.put((float(p0.abs().max()), (lo, hi)))
    torch.distributed.destroy_process_group()


def test_two_rank_gloo_gradient_equals_global_batch_gradient():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    shards = sorted(r[1] for r in res)
    assert shards == [(0, 3), (3, 5)]
    err = [r[0] for r in res if r[1] == (0, 3)][0]
    assert err < 1e-12
    assert [r[0] for r in res if r[1] == (3, 5)][0] == 0.0       # broadcast from rank 0 arrived


def test_episode_shard_partitions_exactly():
    from induction_b200.dist import episode_shard, grad_scale
    for B in (1, 5, 64, 1000, 1024):
        for world in (1, 2, 3, 4, 8):
            parts = [episode_shard(B, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == B
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1
            assert abs(sum(grad_scale(s, B) for s in sizes) - 1.0) < 1e-12
    with pytest.raises(ValueError):
        episode_shard(8, 2, 2)
