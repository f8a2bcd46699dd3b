"""K6 — sparse logistic regression forward/backward on the pulled cache."""
from __future__ import annotations

from .. import _lib


def lr_fwd_bwd(params, grads, inv, x, labels, n_feat, loss_out):
    n_ex = labels.numel()
    _lib.call("ssb_lr_fwd_bwd", params.data_ptr(), grads.data_ptr(), inv.data_ptr(),
              None if x is None else x.data_ptr(), labels.data_ptr(), n_ex, n_feat, loss_out.data_ptr(),
              _lib.stream_ptr())
