"""The text parameter-dump checkpoint format and its loader.

ref: /root/reference/src/core/parameter/sparsetable.h:49-56,100-104 and
system/server/init.h:138-149: one line per entry, ``<key> \\t <value> \\n``; a vector row
prints like the reference's ``Vec``: ``Vec:\\t v0 v1 … `` (utils/vec1.h:106-112).  The writer
and reader are native (``csrc/core/core_capi.cc``); reading a dump back is a new capability (the
reference never does).  ``save_table`` / ``load_table`` move a whole HBM table through the format;
``save_table_binary`` is the compact alternative for tables where text would be >100 GB.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .. import _lib


def _core():
    lib = _lib.core()
    lib.ssb_dump_write_text.restype = C.c_int64
    lib.ssb_dump_write_text.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int]
    lib.ssb_dump_read_text.restype = C.c_int64
    lib.ssb_dump_read_text.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int]
    return lib


def write_text(path, keys, rows, append=False):
    keys = np.ascontiguousarray(keys, dtype=np.uint32)
    rows = np.ascontiguousarray(rows, dtype=np.float32).reshape(len(keys), -1)
    n = _core().ssb_dump_write_text(path.encode(), keys.ctypes.data, rows.ctypes.data, len(keys), rows.shape[1],
                                    1 if append else 0)
    if n < 0:
        raise OSError("cannot write %s" % path)
    return n


def read_text(path, dim, max_rows=None):
    if max_rows is None:
        with open(path, "rb") as f:
            max_rows = sum(1 for _ in f)
    keys = np.empty(max_rows, dtype=np.uint32)
    rows = np.zeros((max_rows, dim), dtype=np.float32)
    n = _core().ssb_dump_read_text(path.encode(), keys.ctypes.data, rows.ctypes.data, max_rows, dim)
    if n < 0:
        raise OSError("cannot read %s" % path)
    return keys[:n], rows[:n]


def save_table(table, path):
    """Dump every shard this process owns to ``path`` (text, reference format)."""
    total = 0
    first = True
    for shard in table.owned:
        k, r = table.dump_shard(shard)
        total += write_text(path, k, r, append=not first)
        first = False
    return total


def load_table(table, path, chunk=1 << 20):
    """Insert a text dump into an HBM table (checkpoint resume)."""
    import torch

    from ..ops._util import u32_tensor

    keys, rows = read_text(path, table.dim)
    for s in range(0, len(keys), chunk):
        kd = u32_tensor(keys[s:s + chunk], table.device)
        rd = torch.from_numpy(rows[s:s + chunk]).to(table.device)
        table.load(kd, rd)
    return len(keys)


def save_table_binary(table, path):
    """Raw shard dump: <uint32 n><uint32 dim> keys[n] rows[n*dim] per owned shard (little endian)."""
    with open(path, "wb") as f:
        for shard in table.owned:
            k, r = table.dump_shard(shard)
            np.array([len(k), table.dim], dtype=np.uint32).tofile(f)
            k.tofile(f)
            r.astype(np.float32).tofile(f)
    return os.path.getsize(path)


def load_table_binary(table, path):
    import torch

    from ..ops._util import u32_tensor

    total = 0
    with open(path, "rb") as f:
        while True:
            hdr = np.fromfile(f, dtype=np.uint32, count=2)
            if hdr.size < 2:
                break
            n, dim = int(hdr[0]), int(hdr[1])
            k = np.fromfile(f, dtype=np.uint32, count=n)
            r = np.fromfile(f, dtype=np.float32, count=n * dim).reshape(n, dim)
            table.load(u32_tensor(k, table.device), torch.from_numpy(r).to(table.device))
            total += n
    return total
