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


def sgns_tiles(params, grads, pairs_in, pairs_out, n_pairs, max_pairs, neg_idx, n_neg, neg_weight, loss_out):
    dim = params.shape[-1]
    _lib.call("ssb_sgns_tiles", params.data_ptr(), grads.data_ptr(), dim, pairs_in.data_ptr(), pairs_out.data_ptr(),
              n_pairs.data_ptr(), max_pairs, neg_idx.data_ptr(), n_neg, neg_weight, loss_out.data_ptr(),
              _lib.stream_ptr())


def sgns_block_tc(params, grads, inv_in, inv_out, sent_len, window, seed, offset, neg_idx, neg_k, loss_out,
                  mode=1, box_rows=1, debug=None):
    """Block-formulation SGNS step on tcgen05 tensor cores (csrc/kernels/sgns_tc.cu).
    params: bf16 cache [n_rows, 128]; grads: fp32.  mode 0 = vector-load gather, 1 = TMA tile::gather4."""
    import torch

    assert params.dtype == torch.bfloat16 and grads.dtype == torch.float32
    n_rows, dim = params.shape
    n_tok = inv_in.numel()
    _lib.call("ssb_sgns_block_tc", params.data_ptr(), grads.data_ptr(), n_rows, dim, inv_in.data_ptr(),
              inv_out.data_ptr(), n_tok, sent_len, window, seed, offset, neg_idx.data_ptr(), float(neg_k),
              loss_out.data_ptr(), mode, box_rows, None if debug is None else debug.data_ptr(),
              _lib.stream_ptr())


def sgns_block_tc_direct(table, gslot, grads, inv_in, inv_out, sent_len, window, seed, offset, neg_idx, neg_k, loss_out,
                         mode=1):
    """Fused-pull variant: the TMA gathers operand rows straight from the table's bf16 row mirror
    (own HBM or peers over NVLink) instead of from a pulled cache.  ``table`` must have a mirror laid
    out in one virtual range (``make_table(..., with_shadow=True)`` or a single local shard)."""
    mm = getattr(table, "mirror_map", None)
    if mm is None:
        raise ValueError("table has no single-range bf16 row mirror (create it with with_shadow=True)")
    base, total_rows, shard_rows, mirror_row_off = mm
    _lib.call("ssb_sgns_block_tc_direct", base, total_rows, table.capacity, shard_rows, mirror_row_off, gslot.data_ptr(),
              grads.data_ptr(), table.dim, inv_in.data_ptr(), inv_out.data_ptr(), inv_in.numel(), sent_len, window, seed,
              offset, neg_idx.data_ptr(), float(neg_k), loss_out.data_ptr(), mode, _lib.stream_ptr())
