#!/usr/bin/env python3
"""Synthetic word2vec corpus: one sentence of space-separated integer word ids per line.

ref: /root/reference/src/tools/gen-word2vec-data.py:4-19 (10 000 lines x 6-15 ids in [0, 300],
Python 2).  Same default shape; optional Zipf-distributed ids for the 1M-vocabulary config.
usage: gen_word2vec_data.py [--lines N] [--min-len A] [--max-len B] [--vocab V] [--zipf S] [--seed K] > data.txt
"""
import argparse
import sys

import numpy as np


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lines", type=int, default=10000)
    ap.add_argument("--min-len", type=int, default=6)
    ap.add_argument("--max-len", type=int, default=15)
    ap.add_argument("--vocab", type=int, default=301)
    ap.add_argument("--zipf", type=float, default=0.0, help="Zipf exponent (0 = uniform like the reference)")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--out", default="-")
    a = ap.parse_args()
    rng = np.random.default_rng(a.seed)
    lens = rng.integers(a.min_len, a.max_len + 1, size=a.lines)
    total = int(lens.sum())
    if a.zipf > 0:
        w = 1.0 / np.power(np.arange(1, a.vocab + 1, dtype=np.float64), a.zipf)
        cdf = np.cumsum(w) / w.sum()
        ranks = np.minimum(np.searchsorted(cdf, rng.random(total), side="right"), a.vocab - 1)
        ids = np.random.default_rng(12345).permutation(a.vocab)[ranks]
    else:
        ids = rng.integers(0, a.vocab, size=total)
    out = sys.stdout if a.out == "-" else open(a.out, "w")
    pos = 0
    for n in lens:
        out.write(" ".join(map(str, ids[pos:pos + n])) + "\n")
        pos += n
    if out is not sys.stdout:
        out.close()


if __name__ == "__main__":
    main()
