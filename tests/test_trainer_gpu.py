"""GPU: EpisodeTrainer - the step Base.train / train.py / bench.py all run (device sampler + fwd + bwd +
all-reduce + Adam as one CUDA graph) - draws the reference's episodes, equals the eager path, hands the host
streams back in the reference's order, and (>= 2 GPUs) a G-rank step equals the 1-rank step on the same
global batch."""
import os
import random
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import sampler_oracle as so

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ("word", "pos1", "pos2", "mask")


def _make(B=3, mode="fp32", keep_prob=0.9, graph=True, seed=11):
    from induction_b200.engine import EpisodeConfig, EpisodePipeline, ParamStore, init_tf_params, tf_param_shapes
    from induction_b200.trainer import EpisodeTrainer
    from induction_b200.util.data_loader import JSONFileDataLoader
    from induction_b200.util.synthetic import make_tables
    cfg = EpisodeConfig(N=4, K=2, Q=3, B=B, L=10, E=8, u=16, S=9, keep_prob=keep_prob, mode=mode)
    tables = make_tables(10, 30, cfg.L, 400, cfg.E, seed=0, min_len=3)
    store = ParamStore(tf_param_shapes(cfg.N, cfg.L, cfg.E, cfg.u, cfg.S))
    store.load_state_dict(init_tf_params(store.tf_shapes, seed=0))
    pipe = EpisodePipeline(cfg, store, torch.from_numpy(tables["word_vec_mat"]), seed=2)
    ld = JSONFileDataLoader.from_tables(tables)
    smp = ld._device()
    random.seed(seed); np.random.seed(seed)
    smp.load_host_streams()
    return cfg, tables, ld, smp, pipe, EpisodeTrainer(pipe, smp, cuda_graph=graph)


def test_trainer_steps_on_the_reference_episode_sequence_and_hands_streams_back():
    cfg, tables, ld, smp, pipe, tr = _make()
    random.seed(11); np.random.seed(11)
    want = [[so.next_one_tf_stdlib(tables, ld.scopes(), cfg.N, cfg.K, cfg.Q) for _ in range(cfg.B)] for _ in range(4)]
    tail = (random.random(), float(np.random.rand()))
    for step in range(4):
        tr.step()
        tr.stream.synchronize()
        for b in range(cfg.B):
            for k in KEYS:
                assert np.array_equal(getattr(pipe, k)[b].cpu().numpy(), want[step][b][0][k]), (step, b, k)
            assert np.array_equal(pipe.label[b].cpu().numpy(), want[step][b][1])
    assert tr.graph_step is not None and int(pipe.p.global_step.item()) == 4
    mean_loss, mean_acc, n = tr.read_running()
    assert n == 4 and mean_loss > 0
    # suspend() discards the look-ahead walk: the host streams are where 4 reference iterations leave them
    tr.suspend()
    assert (random.random(), float(np.random.rand())) == tail
    # ... and stepping again resumes from whatever the host streams now are (here: one draw further on each)
    st_py, st_np = random.getstate(), np.random.get_state()
    nxt = [so.next_one_tf_stdlib(tables, ld.scopes(), cfg.N, cfg.K, cfg.Q) for _ in range(cfg.B)]
    random.setstate(st_py); np.random.set_state(st_np)
    tr.step()
    tr.stream.synchronize()
    for b in range(cfg.B):
        assert np.array_equal(pipe.word[b].cpu().numpy(), nxt[b][0]["word"])
    tr.close()


def test_read_async_returns_every_step_one_step_behind():
    """read_async(): the host reads step t's {loss, accuracy} while step t + 1 runs - same numbers as read()."""
    cfg, tables, ld, smp, pipe, tr = _make(mode="bf16")
    cfg2, tables2, ld2, smp2, pipe2, tr2 = _make(mode="bf16")
    sync, lagged, pend = [], [], None
    for _ in range(6):
        tr.step()
        sync.append(tr.read())
        tr2.step()
        nxt = tr2.read_async()
        if pend is not None:
            lagged.append(pend.get())
        pend = nxt
    lagged.append(pend.get())
    assert len(lagged) == 6
    np.testing.assert_allclose(np.array(lagged), np.array(sync), rtol=2e-2)      # float atomics: not bitwise
    assert all(l > 0 for l, _ in lagged)
    tr.close(); tr2.close()


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_graph_replay_equals_eager_launches(mode):
    res = []
    for graph in (True, False):
        cfg, tables, ld, smp, pipe, tr = _make(mode=mode, graph=graph)
        losses = []
        for _ in range(5):
            tr.step()
            losses.append(tr.read()[0])
        res.append((losses, pipe.p.params.clone()))
        assert (tr.graph_step is not None) == graph
        tr.close()
    # same kernels in the same order; split-K / scatter reductions use float atomics, so equality is up to
    # their summation order
    np.testing.assert_allclose(res[0][0], res[1][0], rtol=2e-4 if mode == "fp32" else 2e-2)
    assert float((res[0][1] - res[1][1]).abs().max()) < (2e-5 if mode == "fp32" else 5e-3)


def test_base_train_fast_path_prints_the_reference_lines_and_logs_scalars(tmp_path, capsys):
    import json
    from induction_b200.model.graph import InductionGraph
    from induction_b200.util.data_loader import JSONFileDataLoader
    from induction_b200.util.synthetic import make_tables
    trl = JSONFileDataLoader.from_tables(make_tables(12, 30, 12, 300, 8, seed=0, min_len=3))
    val = JSONFileDataLoader.from_tables(make_tables(8, 30, 12, 300, 8, seed=1, min_len=3))
    val.word_vec_mat = trl.word_vec_mat
    outs = []
    for fast in (False, True):
        random.seed(0); np.random.seed(0)
        log = str(tmp_path / ("scalars%d.jsonl" % fast))
        model = InductionGraph(N=3, K=2, Q=5, pred_embed=trl.word_vec_mat, sequence_length=12, hidden_size=8,
                               out_size=10, keep_prob=1.0, cuda_graph=fast, scalar_log=log)
        acc = model.train(dataloader=(trl, val), model_dir_path=str(tmp_path / ("ck%d" % fast)), model_name="m",
                          train_iter=12, val_iter=3, val_step=6, test_iter=4)
        out = capsys.readouterr().out
        assert "[train] step:   1 | loss:" in out and "evaluate start .." in out and "Finish training :m" in out
        rows = [json.loads(l) for l in open(log)]
        assert rows and set(rows[0]) == {"step", "loss", "accuracy", "learning_rate"}
        outs.append((acc, [l for l in out.splitlines() if l.startswith("[train]")], model.store.state_dict()))
    # same seeds, same draw order (train step, eval, train step ...): the graph path IS the reference loop
    def nums(line):
        return [float(x.strip("%")) for x in line.replace("|", " ").replace(",", " ").split() if x.strip("%").replace(".", "").isdigit()]
    assert len(outs[0][1]) == len(outs[1][1])
    for a, b in zip(outs[0][1], outs[1][1]):
        np.testing.assert_allclose(nums(a), nums(b), rtol=1e-3, atol=1e-4)
    assert abs(outs[0][0] - outs[1][0]) < 1e-6
    for k in outs[0][2]:
        assert torch.allclose(outs[0][2][k], outs[1][2][k], rtol=0, atol=1e-4), k


def _run_dp(tmp_path, world, global_B, mode, steps=2, env=None, tag=""):
    out = str(tmp_path / ("dp_w%d%s" % (world, tag)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(29500 + world + os.getpid() % 500),
           os.path.join(ROOT, "tests", "_dp_worker.py"), out, str(global_B), mode, str(steps)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, **(env or {})))
    assert r.returncode == 0, r.stderr[-3000:]
    return [np.load(out + ".%d.npz" % k) for k in range(world)]


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_two_gpu_step_equals_one_gpu_step_on_the_same_global_batch(tmp_path, mode):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    one = _run_dp(tmp_path, 1, 4, mode)[0]
    two = _run_dp(tmp_path, 2, 4, mode)
    # the global batch is bit-identical: rank r holds episodes [2r, 2r+2) of the 1-GPU batch
    for r in range(2):
        assert np.array_equal(two[r]["word"], one["word"][2 * r: 2 * r + 2])
        assert np.array_equal(two[r]["label"], one["label"][2 * r: 2 * r + 2])
        assert bool(two[r]["graph"])
    # all-reduced gradient sum / world == the 1-GPU gradient of the global batch; both ranks agree bitwise
    assert np.array_equal(two[0]["grads"], two[1]["grads"]) and np.array_equal(two[0]["params"], two[1]["params"])
    g1, g2 = one["grads"], two[0]["grads"] / 2.0
    tol = 1e-4 if mode == "fp32" else 3e-2
    assert np.abs(g1 - g2).max() <= tol * np.abs(g1).max()
    # each rank's loss is the mean over ITS episodes; their mean is the global-batch loss
    l1, l2 = float(one["losses"][0]), float(np.mean([t["losses"][0] for t in two]))
    assert abs(l1 - l2) < (1e-5 if mode == "fp32" else 2e-2) * l1
    assert np.abs(one["params"] - two[0]["params"]).max() < 2 * 1e-3 * (2e-2 if mode == "fp32" else 1.0)


def test_deferred_tail_of_M_is_the_synchronous_update(tmp_path):
    """INDNET_M_DEFER: part of M's all-reduce + Adam update moves into the next step's graph (under the encoder's
    forward pass).  After flush() the weights and Adam moments are those of the run that does it all inside the step."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    sync = _run_dp(tmp_path, 2, 4, "bf16", steps=4, tag="s")
    late = _run_dp(tmp_path, 2, 4, "bf16", steps=4, env={"INDNET_M_DEFER": "0.4"}, tag="d")
    assert bool(late[0]["deferred"]) and not bool(sync[0]["deferred"]) and bool(late[0]["graph"])
    assert int(late[0]["step"][0]) == int(sync[0]["step"][0]) == 4
    assert np.array_equal(late[0]["params"], late[1]["params"])              # the ranks stay bitwise in step
    for k in ("adam_m", "adam_v"):                                              # one application per step, no more, no less
        a, b = late[0][k], sync[0][k]
        assert np.abs(a - b).max() <= 5e-2 * np.abs(b).max(), k
    moved = np.abs(late[0]["params"] - sync[0]["params"]) > 1e-6              # Adam: +-lr flips where the gradient is noise
    assert moved.mean() < 0.05
    np.testing.assert_allclose(late[0]["losses"], sync[0]["losses"], rtol=3e-2)
