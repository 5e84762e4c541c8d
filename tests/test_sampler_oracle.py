"""CPU: the sampler oracle (both restatements) against fixtures produced by the reference's own
util/data_loader.py, and the host-side JSON preprocessing / cache format against the same fixtures."""
import json
import os
import random

import numpy as np
import pytest

import induction_b200  # noqa: F401
from induction_b200.util.data_loader import JSONFileDataLoader
from induction_b200.util.synthetic import write_fewrel_json
from oracle import sampler_oracle as so

CASES = ["tiny", "wide"]


def _load(golden_dir, name):
    z = np.load(os.path.join(golden_dir, "sampler_%s.npz" % name), allow_pickle=False)
    tables = {"word": z["data_word"], "pos1": z["data_pos1"], "pos2": z["data_pos2"], "mask": z["data_mask"]}
    scopes = [tuple(v) for v in json.loads(str(z["rel2scope"])).values()]
    return z, tables, scopes


@pytest.mark.parametrize("name", CASES)
def test_stdlib_restatement_matches_reference(golden_dir, name):
    z, tables, scopes = _load(golden_dir, name)
    seed = int(z["seed"])
    random.seed(seed); np.random.seed(seed)
    e = 0
    for (N, K, Q, count) in z["plan"]:
        for _ in range(int(count)):
            out, label, _, _ = so.next_one_tf_stdlib(tables, scopes, int(N), int(K), int(Q))
            for k in ("word", "pos1", "pos2", "mask"):
                assert np.array_equal(out[k], z["ep%d_%s" % (e, k)]), (name, e, k)
            assert np.array_equal(label, z["ep%d_label" % e])
            e += 1
    assert e == int(z["num_episodes"])


@pytest.mark.parametrize("name", CASES)
def test_mt_restatement_matches_reference_and_hands_back_state(golden_dir, name):
    z, tables, scopes = _load(golden_dir, name)
    seed = int(z["seed"])
    random.seed(seed); np.random.seed(seed)
    smp = so.EpisodeSampler(scopes, so.MT19937.from_python_random(), so.MT19937.from_numpy_global())
    e = 0
    for (N, K, Q, count) in z["plan"]:
        for _ in range(int(count)):
            rows, label, _ = smp.next_rows(int(N), int(K), int(Q))
            got = so.gather(tables, rows)
            for k in ("word", "pos1", "pos2", "mask"):
                assert np.array_equal(got[k], z["ep%d_%s" % (e, k)]), (name, e, k)
            assert np.array_equal(label, z["ep%d_label" % e])
            e += 1
    # the restated streams end exactly where the reference's real streams ended
    smp.py.to_python_random(); smp.np.to_numpy_global()
    py_after = np.array(random.getstate()[1], dtype=np.uint64)
    assert np.array_equal(py_after, z["py_state_after"])
    # numpy: compare the *next draws* (key layout may be mid-twist equivalent)
    a = np.random.randint(0, 2 ** 31 - 1, size=8)
    np.random.set_state(("MT19937", z["np_key_after"], int(z["np_pos_after"]), 0, 0.0))
    b = np.random.randint(0, 2 ** 31 - 1, size=8)
    assert np.array_equal(a, b)


def test_mt19937_matches_cpython_and_numpy_streams():
    random.seed(12345); np.random.seed(54321)
    mp, mn = so.MT19937.from_python_random(), so.MT19937.from_numpy_global()
    assert [mp.next_uint32() for _ in range(1500)] == [random.getrandbits(32) for _ in range(1500)]
    want = np.random.randint(0, 2 ** 32, size=1500, dtype=np.uint64)  # legacy: one uint32 per draw
    np.random.seed(54321)
    ref_perm = np.random.permutation(701)
    assert so.np_permutation(mn, 701) == ref_perm.tolist()
    assert want is not None


@pytest.mark.parametrize("n,k", [(64, 5), (64, 10), (64, 20), (16, 5), (16, 10), (6, 3), (100, 6)])
def test_py_sample_indices_matches_random_sample(n, k):
    random.seed(99)
    mt = so.MT19937.from_python_random()
    for _ in range(20):
        assert so.py_sample_indices(mt, n, k) == random.sample(range(n), k)


@pytest.mark.parametrize("name", CASES)
def test_preprocessing_and_cache_match_reference(golden_dir, name, tmp_path, monkeypatch):
    z, _, _ = _load(golden_dir, name)
    jkw = json.loads(str(z["json_kwargs"]))
    data_path, vec_path = write_fewrel_json(str(tmp_path / "in"), **jkw)
    monkeypatch.chdir(tmp_path)
    L = int(z["max_length"])
    ld = JSONFileDataLoader(data_path, vec_path, max_length=L)
    for attr, key in [("data_word", "data_word"), ("data_pos1", "data_pos1"), ("data_pos2", "data_pos2"),
                      ("data_mask", "data_mask"), ("data_length", "data_length")]:
        got = getattr(ld, attr)
        assert got.dtype == z[key].dtype and np.array_equal(got, z[key]), attr
    assert ld.word_vec_mat.dtype == np.float32
    assert np.array_equal(ld.word_vec_mat.view(np.uint32), z["word_vec_mat"].view(np.uint32))
    assert ld.rel2scope == json.loads(str(z["rel2scope"]))
    assert ld.word2id == json.loads(str(z["word2id"]))
    # cache files exist under the reference's names and reload without the JSON inputs
    assert sorted(os.listdir("_processed_data")) == sorted(
        ["train_%s" % s for s in ("word.npy", "pos1.npy", "pos2.npy", "mask.npy", "length.npy", "rel2scope.json")]
        + ["glove.tiny.5d_mat.npy", "glove.tiny.5d_word2id.json"])
    ld2 = JSONFileDataLoader(data_path, vec_path, max_length=L)
    assert np.array_equal(ld2.data_word, ld.data_word) and ld2.rel2scope == ld.rel2scope


def test_missing_files_raise_like_reference(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    with pytest.raises(Exception, match=r"\[ERROR\] Data file doesn't exist"):
        JSONFileDataLoader("nope.json", "nope_vec.json")
