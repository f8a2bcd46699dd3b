"""The user-facing framework layer on the GPU path: role objects, the algorithm base class, the
worker cache and the pull / push calls — the API a user of the reference programs against.

ref: /root/reference/src/core/framework/{SwiftMaster,SwiftServer,SwiftWorker}.h,
core/parameter/{global_param_cache,global_pull_access,global_push_access,sparse_access_method}.h
and SURVEY.md §2.12 (the user-extension contract).

Mapping (one process per GPU hosts server shard g *and* worker g):

=================================  ==================================================================
reference                          here
=================================  ==================================================================
``GlobalParamCache<K,V,G>``        :class:`DeviceParamCache` — dense ``params[U, D]`` / ``grads[U, D]`` plus the
                                   ``unique`` / ``inverse`` / ``gslot`` index arrays, all in HBM
``pull_with_barrier(keys, cache)`` :func:`pull_with_barrier` — dedup + one pull kernel (route, probe/insert, gather)
``push_with_barrier(keys, cache)`` :func:`push_with_barrier` — one push kernel (scatter-add fused with the update)
``PushAccessMethod`` (SGD/AdaGrad) :class:`SgdPushMethod`, :class:`AdaGradPushMethod` (update rule run on the owner shard)
``PullAccessMethod::init_param``   deterministic lazy init inside the pull kernel (``seed``, ``init_scale`` of the table)
``BaseAlgorithm``                  :class:`BaseAlgorithm` (``train()``, ``parse_record()``, ``set_data_path()``)
``SwiftWorker<Alg>``               :class:`SwiftWorker`
``SwiftServer<…>``                 :class:`SwiftServer` — owns the shard; periodic backup and final text dump
``SwiftMaster``                    :class:`SwiftMaster` — rendezvous (``torch.distributed``) + fragment map
=================================  ==================================================================

The CPU / ZeroMQ roles are the native binaries (``swiftsnails_b200/bin``) started by
``swiftsnails_b200.parallel.launcher``.
"""
from __future__ import annotations

import os

import torch

from .ops.dedup import Deduper
from .ops.table import ShardedTable
from .utils import dump as dump_utils
from .utils.config import ConfigParser


class DeviceParamCache:
    """Worker-side parameter + gradient cache for one minibatch (ref: global_param_cache.h:28-118)."""

    def __init__(self, max_keys: int, dim: int, device, bf16: bool = False):
        i32 = dict(dtype=torch.int32, device=device)
        self.max_keys, self.dim, self.device = int(max_keys), int(dim), device
        self.unique = torch.empty(self.max_keys, **i32)
        self.inverse = torch.empty(self.max_keys, **i32)
        self.n_unique = torch.zeros(1, **i32)
        self.gslot = torch.empty(self.max_keys, **i32)
        self.params = torch.empty((self.max_keys, dim), dtype=torch.float32, device=device)
        self.params_bf16 = torch.empty((self.max_keys, dim), dtype=torch.bfloat16, device=device) if bf16 else None
        self.grads = torch.zeros((self.max_keys, dim), dtype=torch.float32, device=device)
        self._dedup = Deduper(self.max_keys, device)
        self.n_keys = 0

    def size(self) -> int:
        return int(self.n_unique.item())

    def rows_of(self, positions=None):
        """Parameter rows of the original key positions (through the inverse index)."""
        inv = self.inverse[:self.n_keys] if positions is None else self.inverse[positions]
        return self.params[inv.long()]


class SgdPushMethod:
    """w -= lr * g on the owner shard (ref: the SGD PushAccessMethod of the reference's apps)."""

    def __init__(self, lr: float):
        self.lr = float(lr)

    def apply(self, table: ShardedTable, cache: DeviceParamCache):
        table.push_sgd(cache.gslot, cache.grads, self.lr, n_dev=cache.n_unique, zero_grads=True, n=cache.n_keys)


class AdaGradPushMethod:
    """acc += g^2; w -= lr * g / sqrt(acc + eps) on the owner shard; the accumulator never leaves it."""

    def __init__(self, lr: float, eps: float = 1e-8):
        self.lr, self.eps = float(lr), float(eps)

    def apply(self, table: ShardedTable, cache: DeviceParamCache):
        inbox = getattr(table, "inbox", None)
        if inbox is not None:
            inbox.push(cache.gslot, cache.grads, self.lr, self.eps, adagrad=True, n_dev=cache.n_unique, n=cache.n_keys)
            cache.grads[:cache.n_keys].zero_()
        else:
            table.push_adagrad(cache.gslot, cache.grads, self.lr, self.eps, n_dev=cache.n_unique, zero_grads=True,
                               n=cache.n_keys)


class PushMethod:
    """A push method chosen BY NAME: any device-side update rule of ``include/ssb200_rule.cuh`` — the built-in
    "sgd", "adagrad", "ftrl", or a user's rule loaded with ``ops.rules.Rule.from_source`` — with its
    hyper-parameters (lr, eps, a, b, c; FTRL: lr = alpha, eps = beta, a = l1, b = l2).
    ref: PushAccessMethod::apply_push_value, parameter/sparse_access_method.h:36-46 — the user's class there, the
    user's ``__device__`` struct here. ``PushMethod("adagrad", lr=0.05).apply(table, cache)``."""

    def __init__(self, name: str, **hyper):
        from .ops import rules

        self.rule = name if isinstance(name, rules.Rule) else rules.Rule.get(name)
        self.name = self.rule.name
        self.hyper = dict(hyper)
        if "lr" not in self.hyper:
            raise ValueError("PushMethod needs lr=")
        self.step = 0

    @property
    def state_words(self) -> int:
        """Optimizer-state words per weight the table must be created with (``with_state=``)."""
        return self.rule.state_words

    @property
    def init_kind(self) -> int:
        """How the rule wants missing rows created (``init_kind=`` of the table; ops.rules.INIT_*)."""
        return self.rule.init_kind

    def apply(self, table: ShardedTable, cache: DeviceParamCache):
        self.step += 1
        inbox = getattr(table, "inbox", None)
        if inbox is not None:
            inbox.push(cache.gslot, cache.grads, n_dev=cache.n_unique, n=cache.n_keys, rule=self.rule, step=self.step,
                       **self.hyper)
            cache.grads[:cache.n_keys].zero_()
        else:
            table.push_rule(self.rule, cache.gslot, cache.grads, n_dev=cache.n_unique, zero_grads=True, n=cache.n_keys,
                            step=self.step, **self.hyper)


def pull_with_barrier(table: ShardedTable, keys: torch.Tensor, cache: DeviceParamCache):
    """Fetch the rows of ``keys`` (duplicates allowed) into ``cache`` and reset its gradients.
    ref: global_pull_access.h:40-55.  Returns when the kernels are enqueued; stream order is the barrier."""
    n = keys.numel()
    if n > cache.max_keys:
        raise ValueError("%d keys exceed the cache capacity %d" % (n, cache.max_keys))
    cache.n_keys = n
    cache._dedup(keys, unique=cache.unique[:n], inverse=cache.inverse[:n], n_unique=cache.n_unique)
    table.pull(cache.unique, n_dev=cache.n_unique, rows_out=cache.params, rows_bf16=cache.params_bf16,
               gslot_out=cache.gslot, grads_zero=cache.grads, n=n)
    return cache


def push_with_barrier(table: ShardedTable, cache: DeviceParamCache, method):
    """Send the cache's accumulated gradients to the owner shards and apply ``method`` there; the
    local gradients are reset (ref: global_push_access.h:36-53, Grad::reset :78)."""
    method.apply(table, cache)


class BaseAlgorithm:
    """ref: SwiftWorker.h:19-57.  Subclass, implement ``train()`` (and ``parse_record()`` if the
    algorithm reads text records); ``self.table`` / ``self.config`` are set by :class:`SwiftWorker`."""

    def __init__(self):
        self._data_path = None
        self.table = None
        self.config = None
        self.local_train = False

    def set_data_path(self, path):
        self._data_path = path

    def data_path(self):
        if not self._data_path:
            raise RuntimeError("set_data_path first")
        return self._data_path

    def parse_record(self, line: str):
        raise NotImplementedError

    def train(self):
        raise NotImplementedError


class SwiftMaster:
    """Rendezvous + fragment map.  On the GPU path the registration / route broadcast of the
    reference's master (system/master/init.h) is the ``torch.distributed`` rendezvous followed by
    the VMM file-descriptor exchange in ``parallel.peer``; this object just carries the result."""

    def __init__(self, config: ConfigParser | None = None):
        self.config = config or ConfigParser()

    def __call__(self, rows_per_shard, dim, device, with_state=False, **kw):
        from .parallel.cluster import make_table

        frag_num = self.config.to_int32("frag_num", 1024)
        return make_table(rows_per_shard, dim, device, with_state=with_state, frag_num=frag_num, **kw)


class SwiftServer:
    """Owner of this process's shard: periodic parameter backup and the final text dump.
    ref: SwiftServer.h:17-53, system/server/init.h:138-149 (backup), server/terminate.h:33-41 (dump)."""

    def __init__(self, table: ShardedTable, config: ConfigParser | None = None):
        self.table = table
        cfg = config or ConfigParser()
        self.backup_period = cfg.to_int32("param_backup_period", 0)
        self.backup_root = cfg.to_string("param_backup_root", ".")
        self.push_counter = 0

    def on_push(self):
        """Call once per push request; writes <root>/param-<rank>-<n>.txt every backup_period pushes."""
        self.push_counter += 1
        if self.backup_period > 0 and self.push_counter % self.backup_period == 0:
            self.backup()

    def backup(self):
        path = os.path.join(self.backup_root, "param-%d-%d.txt" % (self.table.rank, self.push_counter))
        torch.cuda.synchronize(self.table.device)
        dump_utils.save_table(self.table, path)
        return path

    def terminate(self, path=None):
        """Final dump (the reference prints it to stdout; a path keeps multi-GB dumps off the terminal)."""
        torch.cuda.synchronize(self.table.device)
        if path is None:
            import sys
            import tempfile

            with tempfile.NamedTemporaryFile(suffix=".txt", delete=False) as f:
                tmp = f.name
            dump_utils.save_table(self.table, tmp)
            with open(tmp) as f:
                for line in f:
                    sys.stdout.write(line)
            os.unlink(tmp)
            return None
        return dump_utils.save_table(self.table, path)


class SwiftWorker:
    """ref: SwiftWorker.h:61-153 — reads ``num_iters`` / ``learning_rate`` / ``local_train`` and runs
    the algorithm's ``train()``."""

    def __init__(self, alg: BaseAlgorithm, table: ShardedTable, config: ConfigParser | None = None):
        self.alg, self.table = alg, table
        self.config = config or ConfigParser()
        self.num_iters = self.config.to_int32("num_iters", 1)
        self.learning_rate = self.config.to_float("learning_rate", 0.01)
        alg.table, alg.config = table, self.config
        alg.local_train = self.config.to_int32("local_train", 0) > 0

    def __call__(self):
        self.alg.train()
        torch.cuda.synchronize(self.table.device)
