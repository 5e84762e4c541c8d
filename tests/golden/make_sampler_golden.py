"""Generates tests/golden/sampler_*.npz by running the REFERENCE's own util/data_loader.py.

Run in the build container only (needs /root/reference):  python tests/golden/make_sampler_golden.py
The reference file is imported unmodified.  Two run-time accommodations, neither edits it:
  * it reads/writes ./_processed_data relative to the CWD (data_loader.py:24,200) -> run in a temp dir;
  * on Python >= 3.11 ``random.sample(dict.keys(), N)`` (:260) raises TypeError -> after construction
    ``loader.rel2scope`` is replaced by a dict subclass whose ``keys()`` returns a list; ``sample``
    depends only on len() and order, so the draws are the ones Python <= 3.10 would have made.
Fixtures hold the pre-processed tables (pins the JSON -> npy preprocessing too), the seeds, and the
outputs of consecutive ``next_one_tf`` calls.
"""
import importlib.util
import json
import os
import random
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import induction_b200  # noqa: E402
from induction_b200.util.synthetic import write_fewrel_json  # noqa: E402

REF = "/root/reference/util/data_loader.py"


class _ListKeys(dict):
    def keys(self):
        return list(super().keys())


def load_reference():
    spec = importlib.util.spec_from_file_location("ref_data_loader", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


CASES = {
    # name: (json kwargs, max_length, seed, [(N, K, Q, count), ...])
    "tiny": (dict(num_relations=6, per_relation=12, vocab=30, E=5, max_tokens=14, seed=1), 10, 7,
             [(3, 2, 3, 5), (5, 1, 5, 3)]),
    "wide": (dict(num_relations=64, per_relation=20, vocab=50, E=5, max_tokens=16, seed=2), 12, 0,
             [(5, 5, 5, 4), (10, 5, 5, 3), (20, 10, 5, 2), (5, 1, 5, 3)]),
}


def main():
    ref = load_reference()
    for name, (jkw, L, seed, plan) in CASES.items():
        with tempfile.TemporaryDirectory() as tmp:
            data_path, vec_path = write_fewrel_json(tmp, **jkw)
            cwd = os.getcwd()
            os.chdir(tmp)
            try:
                loader = ref.JSONFileDataLoader(data_path, vec_path, max_length=L)
                out = {
                    "json_kwargs": json.dumps(jkw), "max_length": L, "seed": seed,
                    "plan": np.array(plan, dtype=np.int64),
                    "data_word": loader.data_word, "data_pos1": loader.data_pos1,
                    "data_pos2": loader.data_pos2, "data_mask": loader.data_mask,
                    "data_length": loader.data_length, "word_vec_mat": loader.word_vec_mat,
                    "rel2scope": json.dumps(loader.rel2scope), "word2id": json.dumps(loader.word2id),
                }
                loader.rel2scope = _ListKeys(loader.rel2scope)
                random.seed(seed)
                np.random.seed(seed)
                e = 0
                for (N, K, Q, count) in plan:
                    for _ in range(count):
                        inputs, label = loader.next_one_tf(N, K, Q)
                        for k in ("word", "pos1", "pos2", "mask"):
                            out["ep%d_%s" % (e, k)] = inputs[k]
                        out["ep%d_label" % e] = label
                        e += 1
                out["num_episodes"] = e
                # stream positions after the run, so state hand-back can be checked too
                out["py_state_after"] = np.array(random.getstate()[1], dtype=np.uint64)
                st = np.random.get_state()
                out["np_key_after"] = st[1]
                out["np_pos_after"] = st[2]
                # the reference's other sampler, next_one (data_loader.py:213-257): stacked support [N, K, L],
                # separate permuted query set; drawn right after the next_one_tf episodes on the same streams
                N1, K1, Q1 = plan[0][:3]
                for r in range(2):
                    sup, qry, lab = loader.next_one(N1, K1, Q1)
                    for k in ("word", "pos1", "pos2", "mask"):
                        out["one%d_sup_%s" % (r, k)] = sup[k]
                        out["one%d_qry_%s" % (r, k)] = qry[k]
                    out["one%d_label" % r] = lab
                out["py_state_after_one"] = np.array(random.getstate()[1], dtype=np.uint64)
                st = np.random.get_state()
                out["np_key_after_one"] = st[1]
                out["np_pos_after_one"] = st[2]
            finally:
                os.chdir(cwd)
        np.savez_compressed(os.path.join(HERE, "sampler_%s.npz" % name), **out)
        print("wrote sampler_%s.npz (%d episodes)" % (name, e))


if __name__ == "__main__":
    main()
