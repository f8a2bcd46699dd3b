"""word2vec SGNS device ops (pair generation + tile kernels).

The reference's word2vec app is absent from its snapshot; these implement the
standard SGNS objective on the worker's pulled cache — the "first consumer compute" between
pull and push in /root/reference/src/core/framework/SwiftWorker.h:88-113's loop."""
from __future__ import annotations

import torch

from .. import _lib


def make_pairs(inv_in, inv_out, sent_len, window, seed, offset, pairs_in, pairs_out, n_pairs):
    n_tok = inv_in.numel()
    _lib.call("ssb_sgns_make_pairs", inv_in.data_ptr(), inv_out.data_ptr(), n_tok, sent_len, window, seed,
              offset, pairs_in.data_ptr(), pairs_out.data_ptr(), n_pairs.data_ptr(), _lib.stream_ptr())


def sgns_tiles(params, grads, pairs_in, pairs_out, n_pairs, max_pairs, neg_idx, n_neg, neg_weight, loss_out,
               impl="cuda_core"):
    dim = params.shape[-1]
    name = "ssb_sgns_tiles" if impl == "cuda_core" else "ssb_sgns_tiles_tc"
    _lib.call(name, params.data_ptr(), grads.data_ptr(), dim, pairs_in.data_ptr(), pairs_out.data_ptr(),
              n_pairs.data_ptr(), max_pairs, neg_idx.data_ptr(), n_neg, neg_weight, loss_out.data_ptr(),
              _lib.stream_ptr())
