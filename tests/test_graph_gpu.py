"""GPU: the reference-facing graph object (model/graph.py + model/base.py mirrors): InductionGraph
construction with the reference kwargs, run_step == one sess.run, train / eval loops with the reference
print format and best-checkpoint policy, checkpoint save / restore keyed by the TF variable names, and
the train.py CLI."""
import os
import random
import subprocess
import sys

import numpy as np
import pytest
import torch

from _cases import make_case, oracle_forward, rel_err

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _loaders(L=12, E=8):
    from induction_b200.util.data_loader import JSONFileDataLoader
    from induction_b200.util.synthetic import make_tables
    tr = JSONFileDataLoader.from_tables(make_tables(12, 30, L, 300, E, seed=0, min_len=3))
    va = JSONFileDataLoader.from_tables(make_tables(8, 30, L, 300, E, seed=1, min_len=3))
    va.word_vec_mat = tr.word_vec_mat
    return tr, va


def test_run_step_matches_oracle_and_exposes_reference_attributes():
    from induction_b200.model.graph import InductionGraph
    cfg, tables, inputs, label, params = make_case("k1", seed=3)
    model = InductionGraph(N=cfg.N, K=cfg.K, Q=cfg.Q, pred_embed=tables["word_vec_mat"], sequence_length=cfg.L,
                           hidden_size=cfg.u, out_size=cfg.S, batch_size=cfg.B, keep_prob=1.0)
    model.store.load_state_dict(params)
    assert model.embed_size == cfg.E and model.num_classes == cfg.N
    loss, acc = model.run_step(inputs, label, train=False)
    ref = oracle_forward(cfg, tables, inputs, label, params)
    assert abs(loss - float(ref["loss"])) < 1e-5 * float(ref["loss"]) and abs(acc - float(ref["accuracy"])) < 1e-6
    assert rel_err(model.probs, ref["probs"]) < 1e-5
    assert model.alphas.shape == (cfg.X, cfg.L, 1)
    assert torch.equal(model.predict_test.cpu(), ref["predict_test"])
    assert torch.equal(model.predict.cpu(), torch.round(model.probs).cpu())
    assert model.get_embedding().shape == (cfg.X, cfg.L, cfg.D)
    # a train step moves global_step, the decayed learning rate and the loss
    l0, _ = model.run_step(inputs, label, train=True)
    for _ in range(20):
        l1, _ = model.run_step(inputs, label, train=True)
    assert model.global_step == 21 and abs(model.learning_rate - 1e-3) < 1e-9 and l1 < l0
    assert set(model.summary) == {"loss", "accuracy", "learning_rate"}


def test_train_eval_loops_checkpoint_and_restore(tmp_path, capsys):
    from induction_b200.model.graph import InductionGraph
    tr, va = _loaders()
    random.seed(0); np.random.seed(0)
    model = InductionGraph(N=3, K=2, Q=5, pred_embed=tr.word_vec_mat, sequence_length=12, hidden_size=8,
                           out_size=10, keep_prob=0.9)
    acc = model.train(dataloader=(tr, va), model_dir_path=str(tmp_path / "ckpt"), model_name="m",
                      train_iter=12, val_iter=3, val_step=6, test_iter=4)
    out = capsys.readouterr().out
    assert "training start .." in out and "[train] step:   1 | loss:" in out and "evaluate start .." in out
    assert "Finish training :m" in out and "Test accuracy:" in out
    assert 0.0 <= acc <= 1.0
    files = sorted(os.listdir(tmp_path / "ckpt" / "checkpoint")) if os.path.isdir(tmp_path / "ckpt" / "checkpoint") else []
    # best-only policy: a checkpoint exists iff some eval beat 0.0 accuracy
    model.save(str(tmp_path / "manual.npz"))
    z = np.load(tmp_path / "manual.npz")
    assert "bidirectional_rnn/fw/basic_lstm_cell/kernel" in z and z["M"].shape == (16, 16, 10)
    assert int(z["global_step"]) == 12
    model2 = InductionGraph(N=3, K=2, Q=5, pred_embed=tr.word_vec_mat, sequence_length=12, hidden_size=8,
                            out_size=10, keep_prob=0.9, seed=99)
    model2.restore(str(tmp_path / "manual.npz"))
    assert model2.global_step == 12
    a, b = model.store.state_dict(), model2.store.state_dict()
    for k in a:
        assert torch.equal(a[k], b[k]), k
    random.seed(5); np.random.seed(5)
    e1 = model.eval(va, None, 3)
    random.seed(5); np.random.seed(5)
    e2 = model2.eval(va, None, 3)
    assert e1 == e2
    assert files == [] or files[0].startswith("m-")
    # the frozen word table is not in the checkpoint: its fingerprint is, and a different table is refused
    assert tuple(z["pred_embed/shape"]) == tuple(tr.word_vec_mat.shape)
    other = tr.word_vec_mat.copy()
    other[0, 0] += 1.0
    model3 = InductionGraph(N=3, K=2, Q=5, pred_embed=other, sequence_length=12, hidden_size=8, out_size=10)
    with pytest.raises(ValueError, match="different word embedding"):
        model3.restore(str(tmp_path / "manual.npz"))


def test_missing_vocab_size_is_an_error():
    from induction_b200.model.graph import InductionGraph
    with pytest.raises(ValueError, match="vocab_size"):
        InductionGraph(N=3, K=2, Q=5, sequence_length=8, hidden_size=8)


def test_train_py_cli_banner_and_run(tmp_path):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "train.py"), "induction", "4", "2", "--synthetic",
                        "--train-iter", "4", "--val-iter", "2", "--val-step", "2", "--test-iter", "2",
                        "--max-length", "16", "--hidden-size", "8", "--model-dir", str(tmp_path / "out")],
                       capture_output=True, text=True, timeout=300, cwd=str(tmp_path))
    assert r.returncode == 0, r.stderr[-2000:]
    lines = r.stdout.splitlines()
    assert lines[0] == "4-way-2-shot Few-Shot Relation Classification" and lines[1] == "Model: induction"
    assert any(l.startswith("Test accuracy:") for l in lines)
