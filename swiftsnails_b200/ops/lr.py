"""K6 — sparse logistic regression forward/backward on the pulled cache."""
from __future__ import annotations

from .. import _lib


def lr_fwd_bwd(params, grads, inv, x, labels, n_feat, loss_out):
    n_ex = labels.numel()
    _lib.call("ssb_lr_fwd_bwd", params.data_ptr(), grads.data_ptr(), inv.data_ptr(),
              None if x is None else x.data_ptr(), labels.data_ptr(), n_ex, n_feat, loss_out.data_ptr(),
              _lib.stream_ptr())


RULE_IDS = {"sgd": 0, "adagrad": 1, "ftrl": 2}


class FusedLRState:
    """Worker-side scratch of the fused sparse-LR step (csrc/kernels/lr.cu, K6 fused), allocated once: one 16-byte
    entry {tag, idx, slot} per feature id (``feature_bits`` > 0: 4 GB at 2^28) or per global table slot
    (``feature_bits`` = 0: any 32-bit key, every instance probes the table), the per-instance index stash and the
    compact per-step arrays (slot, weight, gradient sum of every touched key)."""

    def __init__(self, table, batch_size, n_feat, feature_bits, device, max_touched=None):
        import torch

        self.table, self.device = table, device
        self.batch_size, self.n_feat = int(batch_size), int(n_feat)
        i32 = dict(dtype=torch.int32, device=device)
        self.dense_keys = (1 << feature_bits) if feature_bits else 0
        n_ent = self.dense_keys or table.num_shards * table.capacity
        self.entries = torch.zeros((n_ent, 4), **i32)
        self.entries[:, 2] = -1                                   # slot not known yet
        self.stash = torch.empty(self.batch_size * self.n_feat, **i32)
        self.max_touched = int(max_touched or self.batch_size * self.n_feat)
        self.touched = torch.empty(self.max_touched, **i32)
        self.wc = torch.empty(self.max_touched, dtype=torch.float32, device=device)
        self.gc = torch.zeros(self.max_touched, dtype=torch.float32, device=device)
        self.counters = torch.zeros(16, **i32)
        self.step = 0

    def phase_us(self):
        """Device time of the last step's phases (resolve, gather, compute, push) in microseconds, from the
        kernel's own globaltimer stamps (block 0)."""
        t = [x & 0xFFFFFFFF for x in self.counters[8:13].tolist()]
        d = [((t[i + 1] - t[i]) & 0xFFFFFFFF) / 1e3 for i in range(4)]
        return dict(zip(("resolve", "gather", "compute", "push"), d))

    def nbytes(self):
        return sum(t.numel() * 4 for t in (self.entries, self.stash, self.touched, self.wc, self.gc))


def lr_fused_step(state: FusedLRState, keys, x, labels, loss_out, rule="adagrad", blocks_per_sm=0, feature_major=False, **hyper):
    """One training step. keys int32 [B, n_feat] (uint32 bits, -1 = empty position), x float32 [B, n_feat] or None,
    labels float32 [B]; hyper = lr, eps, a, b (ops.rules.Hyper). ``feature_major``: keys / x are [n_feat, B] instead
    (the minibatch stored feature by feature: a warp's 32 examples are contiguous, its loads coalesced)."""
    from .rules import Hyper
    import ctypes as C

    n_ex = labels.numel()
    if n_ex > state.batch_size or keys.numel() != n_ex * state.n_feat:
        raise ValueError("batch of %d x %d does not fit the state (%d x %d)" % (n_ex, keys.numel() // max(n_ex, 1),
                                                                                 state.batch_size, state.n_feat))
    state.step += 1
    hy = Hyper(float(hyper.get("lr", 0.05)), float(hyper.get("eps", 1e-8)), float(hyper.get("a", 0.0)),
               float(hyper.get("b", 0.0)), float(hyper.get("c", 0.0)), state.step)
    t = state.table
    _lib.call("ssb_lr_fused_step", t._view_ref(), keys.data_ptr(), None if x is None else x.data_ptr(), labels.data_ptr(),
              n_ex, state.n_feat, state.entries.data_ptr(), state.dense_keys, state.stash.data_ptr(), state.touched.data_ptr(),
              state.wc.data_ptr(), state.gc.data_ptr(), state.max_touched, state.counters.data_ptr(), state.step,
              loss_out.data_ptr(), RULE_IDS[rule], C.addressof(hy), int(blocks_per_sm) | ((1 << 30) if feature_major else 0),
              _lib.stream_ptr())
