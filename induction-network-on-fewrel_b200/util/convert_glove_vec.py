"""GloVe truncation tool - mirror of the reference's ``util/convert_glove_vec.py:1-23``: keep the first K
dimensions of every vector of a GloVe JSON file (``[{"word": w, "vec": [...]}, ...]``), lower-case the words,
and write ``glove.6B.<K>d.json`` next to it.  The reference hard-codes K = 5 and the paths; here they are
arguments with the reference's values as defaults.

    python -m induction_b200.util.convert_glove_vec [--k 5] [--input ../data/glove.6B.50d.json] [--output ...]
"""
from __future__ import annotations

import argparse
import json
import os


def convert(input_file: str, out_file: str, k: int = 5) -> int:
    with open(input_file, "r", encoding="utf-8") as f:
        ori_word_vec = json.load(f)
    all_vec_list = [{"word": item["word"].lower(), "vec": item["vec"][0:k]} for item in ori_word_vec]
    with open(out_file, "w", encoding="utf-8") as f:
        json.dump(all_vec_list, f)
    return len(all_vec_list)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--k", type=int, default=5)
    ap.add_argument("--input", default="../data/glove.6B.50d.json")
    ap.add_argument("--output", default=None)
    a = ap.parse_args(argv)
    out = a.output or os.path.join(os.path.dirname(a.input), "glove.6B.{}d.json".format(a.k))
    convert(a.input, out, a.k)
    print("outfile:", out)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
