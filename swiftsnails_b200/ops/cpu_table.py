"""CPU sparse table (the C++ ``SparseTable`` with the stock SGD / AdaGrad access methods) —
the CPU fallback data plane and the oracle the HBM table is checked against.
ref: /root/reference/src/core/parameter/sparsetable.h:69-198."""
from __future__ import annotations

import ctypes as C

import numpy as np

from .. import _lib


class CpuSparseTable:
    def __init__(self, dim, shard_num=8, optimizer="sgd", lr=0.05, eps=1e-8, seed=0x5EED, init_scale=None):
        lib = _lib.core()
        lib.ssb_cpu_table_new.restype = C.c_void_p
        lib.ssb_cpu_table_new.argtypes = [C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_uint64, C.c_float]
        lib.ssb_cpu_table_free.argtypes = [C.c_void_p]
        lib.ssb_cpu_table_pull.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        lib.ssb_cpu_table_push.restype = C.c_int64
        lib.ssb_cpu_table_push.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
        lib.ssb_cpu_table_size.restype = C.c_int64
        lib.ssb_cpu_table_size.argtypes = [C.c_void_p]
        lib.ssb_cpu_table_dump.argtypes = [C.c_void_p, C.c_char_p]
        self._lib, self.dim = lib, int(dim)
        scale = float(init_scale if init_scale is not None else 1.0 / dim)
        self._h = lib.ssb_cpu_table_new(shard_num, dim, 1 if optimizer == "adagrad" else 0, lr, eps, seed, scale)

    def pull(self, keys):
        keys = np.ascontiguousarray(keys, dtype=np.uint32)
        out = np.empty((len(keys), self.dim), dtype=np.float32)
        self._lib.ssb_cpu_table_pull(self._h, keys.ctypes.data, len(keys), out.ctypes.data)
        return out

    def push(self, keys, grads):
        keys = np.ascontiguousarray(keys, dtype=np.uint32)
        grads = np.ascontiguousarray(grads, dtype=np.float32)
        return int(self._lib.ssb_cpu_table_push(self._h, keys.ctypes.data, grads.ctypes.data, len(keys)))

    def size(self):
        return int(self._lib.ssb_cpu_table_size(self._h))

    def dump(self, path):
        if self._lib.ssb_cpu_table_dump(self._h, path.encode()) != 0:
            raise OSError("cannot write %s" % path)

    def __del__(self):
        try:
            self._lib.ssb_cpu_table_free(self._h)
        except Exception:
            pass
