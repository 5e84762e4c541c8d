"""GPU: the device sampler + gather (indnet_sample_episodes through JSONFileDataLoader) is bit-exact
with fixtures produced by the reference's own next_one_tf, hands the streams back exactly, and at
BASELINE sizes equals the stdlib-driven oracle on the same seeds."""
import json
import os
import random

import numpy as np
import pytest
import torch

from oracle import sampler_oracle as so

pytestmark = pytest.mark.gpu
KEYS = ("word", "pos1", "pos2", "mask")


def _loader_from_golden(z):
    from induction_b200.util.data_loader import JSONFileDataLoader
    tables = {"word": z["data_word"], "pos1": z["data_pos1"], "pos2": z["data_pos2"], "mask": z["data_mask"],
              "length": z["data_length"], "rel2scope": json.loads(str(z["rel2scope"])),
              "word_vec_mat": z["word_vec_mat"]}
    return JSONFileDataLoader.from_tables(tables)


@pytest.mark.parametrize("name", ["tiny", "wide"])
def test_next_one_tf_bit_exact_vs_reference_fixtures(golden_dir, name):
    z = np.load(os.path.join(golden_dir, "sampler_%s.npz" % name))
    ld = _loader_from_golden(z)
    seed = int(z["seed"])
    random.seed(seed); np.random.seed(seed)
    e = 0
    for (N, K, Q, count) in z["plan"]:
        for _ in range(int(count)):
            inputs, label = ld.next_one_tf(int(N), int(K), int(Q))
            for k in KEYS:
                assert inputs[k].dtype == np.int32
                assert np.array_equal(inputs[k], z["ep%d_%s" % (e, k)]), (name, e, k)
            assert np.array_equal(label, z["ep%d_label" % e]), (name, e)
            e += 1
    # streams handed back to the host modules are where the reference left them
    assert np.array_equal(np.array(random.getstate()[1], dtype=np.uint64), z["py_state_after"])
    a = np.random.randint(0, 2 ** 31 - 1, size=8)
    np.random.set_state(("MT19937", z["np_key_after"], int(z["np_pos_after"]), 0, 0.0))
    assert np.array_equal(a, np.random.randint(0, 2 ** 31 - 1, size=8))
    np.random.set_state(("MT19937", z["np_key_after"], int(z["np_pos_after"]), 0, 0.0))
    # next_one (reference data_loader.py:213-257) continues on the same streams
    N1, K1, Q1 = (int(v) for v in z["plan"][0][:3])
    for r in range(2):
        sup, qry, lab = ld.next_one(N1, K1, Q1)
        for k in KEYS:
            assert sup[k].shape == (N1, K1, ld.max_length) and np.array_equal(sup[k], z["one%d_sup_%s" % (r, k)]), (name, r, k)
            assert np.array_equal(qry[k], z["one%d_qry_%s" % (r, k)]), (name, r, k)
        assert np.array_equal(lab, z["one%d_label" % r])
    assert np.array_equal(np.array(random.getstate()[1], dtype=np.uint64), z["py_state_after_one"])
    st = np.random.get_state()
    assert np.array_equal(st[1], z["np_key_after_one"]) and int(st[2]) == int(z["np_pos_after_one"])


@pytest.mark.parametrize("B,N,K,Q", [(64, 5, 1, 5), (16, 10, 5, 5), (3, 20, 10, 5)])
def test_next_batch_equals_consecutive_reference_draws_at_baseline_sizes(B, N, K, Q):
    from induction_b200.util.data_loader import JSONFileDataLoader
    from induction_b200.util.synthetic import make_tables
    tables = make_tables(64, 700, 40, 5000, 8, seed=0)      # FewRel train layout: 64 relations x 700
    ld = JSONFileDataLoader.from_tables(tables)
    scopes = ld.scopes()
    for seed in (0, 123):
        random.seed(seed); np.random.seed(seed)
        want = [so.next_one_tf_stdlib(tables, scopes, N, K, Q) for _ in range(B)]
        host_next = (random.random(), float(np.random.rand()))
        random.seed(seed); np.random.seed(seed)
        inputs, label = ld.next_batch(B, N, K, Q)
        assert inputs["word"].is_cuda and inputs["word"].shape == (B, N * (K + Q), 40)
        for b in range(B):
            for k in KEYS:
                assert np.array_equal(inputs[k][b].cpu().numpy(), want[b][0][k]), (seed, b, k)
            assert np.array_equal(label[b].cpu().numpy(), want[b][1]), (seed, b)
        rows = ld._device().buffers(B, N, K, Q)["rows"].cpu().numpy()
        assert np.array_equal(rows, np.stack([w[2] for w in want]))
        # host draws after the device batch continue the same streams
        assert (random.random(), float(np.random.rand())) == host_next


def test_ragged_scopes_and_two_batches_in_a_row():
    from induction_b200.util.data_loader import JSONFileDataLoader
    rs = np.random.RandomState(5)
    sizes = [7, 33, 64, 9, 128, 700, 12, 1023, 1024, 1025, 8, 31, 32, 65, 17, 300, 6, 21, 22, 40, 41, 513]
    I, L = sum(sizes), 6
    lo = np.cumsum([0] + sizes)
    tables = {k: rs.randint(0, 50, size=(I, L)).astype(np.int32) for k in KEYS}
    tables.update(length=np.full(I, L, np.int32), word_vec_mat=np.zeros((52, 4), np.float32),
                  rel2scope={"R%d" % i: [int(lo[i]), int(lo[i + 1])] for i in range(len(sizes))})
    ld = JSONFileDataLoader.from_tables(tables)
    random.seed(9); np.random.seed(9)
    want = [so.next_one_tf_stdlib(tables, ld.scopes(), 6, 2, 4) for _ in range(40)]
    random.seed(9); np.random.seed(9)
    got = []
    for _ in range(2):
        inputs, label = ld.next_batch(20, 6, 2, 4)
        got += [({k: inputs[k][b].cpu().numpy() for k in KEYS}, label[b].cpu().numpy()) for b in range(20)]
    for b in range(40):
        for k in KEYS:
            assert np.array_equal(got[b][0][k], want[b][0][k]), (b, k)
        assert np.array_equal(got[b][1], want[b][1])


def test_too_few_instances_raises_like_numpy_choice():
    from induction_b200.util.data_loader import JSONFileDataLoader
    tables = {k: np.zeros((12, 4), np.int32) for k in KEYS}
    tables.update(length=np.full(12, 4, np.int32), word_vec_mat=np.zeros((3, 2), np.float32),
                  rel2scope={"a": [0, 4], "b": [4, 8], "c": [8, 12]})
    ld = JSONFileDataLoader.from_tables(tables)
    with pytest.raises(ValueError, match="larger sample than population"):
        ld.next_one_tf(2, 3, 2)
    with pytest.raises(ValueError, match="Sample larger than population"):
        ld.next_one_tf(4, 1, 1)


def test_stream_rings_survive_many_batches_shape_switches_and_reseeds():
    """The stream / table rings in the sampler scratch persist across calls (the next window is built while
    the current one is walked).  Eight batches in a row wrap the rings several times; switching between two
    batch shapes and re-seeding in between must be noticed (the scratch no longer continues the stream handed
    in) and rebuilt - every episode still equals the host reference drawing from the same streams."""
    from induction_b200.util.data_loader import JSONFileDataLoader
    from induction_b200.util.synthetic import make_tables
    tables = make_tables(24, 300, 16, 2000, 8, seed=3)
    ld = JSONFileDataLoader.from_tables(tables)
    scopes = ld.scopes()
    plan = [(12, 5, 2, 3)] * 8 + [(5, 4, 1, 2), (12, 5, 2, 3), (5, 4, 1, 2), (5, 4, 1, 2)]

    def check(seed):
        random.seed(seed); np.random.seed(seed)
        want = [[so.next_one_tf_stdlib(tables, scopes, N, K, Q) for _ in range(B)] for (B, N, K, Q) in plan]
        tail = (random.random(), float(np.random.rand()))
        random.seed(seed); np.random.seed(seed)
        for step, (B, N, K, Q) in enumerate(plan):
            inputs, label = ld.next_batch(B, N, K, Q)
            for b in range(B):
                for k in KEYS:
                    assert np.array_equal(inputs[k][b].cpu().numpy(), want[step][b][0][k]), (seed, step, b, k)
                assert np.array_equal(label[b].cpu().numpy(), want[step][b][1]), (seed, step, b)
        assert (random.random(), float(np.random.rand())) == tail

    check(5)
    check(5)        # same seed again: the scratch holds a later part of this very stream
    check(77)


def test_stream_budget_overrun_leaves_streams_unchanged_and_retries():
    """Status 3 (the batch needs more NumPy draws than were pre-generated): the failing call must not consume
    anything, and ``sample`` redoes the batch in halves - same episodes as one uninterrupted run."""
    from induction_b200._lib import LIB
    from induction_b200.util.data_loader import JSONFileDataLoader
    from induction_b200.util.synthetic import make_tables
    tables = make_tables(24, 300, 16, 2000, 8, seed=3)
    B, N, K, Q = 32, 5, 2, 3
    random.seed(9); np.random.seed(9)
    ld0 = JSONFileDataLoader.from_tables(tables)
    want = [so.next_one_tf_stdlib(tables, ld0.scopes(), N, K, Q) for _ in range(B)]
    tail = (random.random(), float(np.random.rand()))
    LIB.call("indnet_sampler_set_budget", -0.5, 4096)        # half the expected draws + a little: B = 32 cannot fit
    try:
        ld = JSONFileDataLoader.from_tables(tables)
        dev = ld._device()
        random.seed(9); np.random.seed(9)
        dev.load_host_streams()
        py0, np0 = dev.py_state.clone(), dev.np_state.clone()
        dev.enqueue(B, N, K, Q)
        assert int(dev.status[0].item()) == 3
        assert torch.equal(dev.py_state, py0) and torch.equal(dev.np_state, np0)     # nothing consumed
        dev.status.zero_()
        inputs, label = ld.next_batch(B, N, K, Q)            # retries in halves until every piece fits
        for b in range(B):
            for k in KEYS:
                assert np.array_equal(inputs[k][b].cpu().numpy(), want[b][0][k]), (b, k)
            assert np.array_equal(label[b].cpu().numpy(), want[b][1])
        assert (random.random(), float(np.random.rand())) == tail
    finally:
        LIB.call("indnet_sampler_set_budget", 0.03, 8192)


def test_parallel_and_serial_sampler_paths_draw_the_same_episodes():
    """The jump-table / chain / replay path and the serial walker are two evaluations of the same draws: same
    episodes, same stream positions afterwards - with one relation size (parallel path), with three sizes (parallel
    path, three tables + the query permutation) and with five (serial walker by necessity)."""
    from induction_b200._lib import LIB
    from induction_b200.util.data_loader import JSONFileDataLoader
    from induction_b200.util.synthetic import make_tables
    for sizes in ([300] * 12, [300, 120, 77] * 4, [300, 120, 77, 64, 51] * 3):
        base = make_tables(len(sizes), max(sizes), 16, 2000, 8, seed=3)
        keep = np.concatenate([np.arange(r * max(sizes), r * max(sizes) + n) for r, n in enumerate(sizes)])
        tables = {k: base[k][keep] for k in ("word", "pos1", "pos2", "mask", "length")}
        lo = np.concatenate([[0], np.cumsum(sizes)])
        tables["rel2scope"] = {"P%d" % r: [int(lo[r]), int(lo[r + 1])] for r in range(len(sizes))}
        tables["word_vec_mat"] = base["word_vec_mat"]
        B, N, K, Q = 9, 5, 3, 4
        got = []
        for parallel in (1, 2, 0):
            LIB.call("indnet_sampler_set_path", parallel)
            try:
                ld = JSONFileDataLoader.from_tables(tables)
                random.seed(21); np.random.seed(21)
                ins, labs = [], []
                for _ in range(2):           # the sampler hands out its static buffers: copy before the next call
                    x = ld.next_batch(B, N, K, Q)
                    ins.append({k: v.cpu().numpy().copy() for k, v in x[0].items()})
                    labs.append(x[1].cpu().numpy().copy())
                got.append((ins, labs, random.random(), float(np.random.rand())))
            finally:
                LIB.call("indnet_sampler_set_path", 1)
        random.seed(21); np.random.seed(21)
        ld = JSONFileDataLoader.from_tables(tables)
        want = [so.next_one_tf_stdlib(tables, ld.scopes(), N, K, Q) for _ in range(2 * B)]
        for path in got:
            for half in range(2):
                for b in range(B):
                    for k in KEYS:
                        assert np.array_equal(path[0][half][k][b], want[half * B + b][0][k]), (sizes[:5], half, b, k)
                    assert np.array_equal(path[1][half][b], want[half * B + b][1])
        assert got[0][2:] == got[1][2:] == got[2][2:]
