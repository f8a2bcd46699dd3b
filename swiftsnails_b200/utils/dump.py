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


def write_text_param(path, keys, rows, state=None, append=False, digits=6):
    """Param-style lines: the row followed by ``\\tAcc:\\t`` and the optimizer state when ``state`` is given."""
    lib = _core()
    lib.ssb_dump_write_text3.restype = C.c_int64
    lib.ssb_dump_write_text3.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int]
    keys = np.ascontiguousarray(keys, dtype=np.uint32)
    rows = np.ascontiguousarray(rows, dtype=np.float32).reshape(len(keys), -1)
    st = None if state is None else np.ascontiguousarray(state, dtype=np.float32).reshape(len(keys), -1)
    n = lib.ssb_dump_write_text3(path.encode(), keys.ctypes.data, rows.ctypes.data, None if st is None else st.ctypes.data,
                                 len(keys), rows.shape[1], rows.shape[1] if st is None else st.shape[1],
                                 1 if append else 0, int(digits))
    if n < 0:
        raise OSError("cannot write %s" % path)
    return n


def read_text_param(path, dim, max_rows=None, state_dim=None):
    """Reads plain or Param-style lines. Returns (keys, rows, state or None). ``state_dim``: floats of optimizer
    state per entry (state words x dim), default ``dim``."""
    lib = _core()
    state_dim = int(state_dim or dim)
    lib.ssb_dump_read_text3.restype = C.c_int64
    lib.ssb_dump_read_text3.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p]
    if max_rows is None:
        with open(path, "rb") as f:
            max_rows = sum(1 for _ in f)
    keys = np.empty(max_rows, dtype=np.uint32)
    rows = np.zeros((max_rows, dim), dtype=np.float32)
    state = np.zeros((max_rows, state_dim), dtype=np.float32)
    has = C.c_int(0)
    n = lib.ssb_dump_read_text3(path.encode(), keys.ctypes.data, rows.ctypes.data, state.ctypes.data, max_rows, dim, state_dim,
                                C.byref(has))
    if n < 0:
        raise OSError("cannot read %s" % path)
    return keys[:n], rows[:n], (state[:n] if has.value else None)


def save_table(table, path, with_state=False, digits=6, max_bytes=256 << 20):
    """Dump every shard this process owns to ``path`` (text, reference format), streamed in bounded pieces.
    ``with_state``: Param-style lines that also carry the optimizer state (AdaGrad accumulator), so that a run
    resumed from the dump continues with the same step sizes; ``digits=9`` makes the text exact for float32."""
    total = 0
    first = True
    with_state = bool(with_state) and table.has_state
    for shard in table.owned:
        for k, r, st in table.iter_shard_chunks(shard, with_state=with_state, max_bytes=max_bytes):
            if with_state or digits != 6:
                total += write_text_param(path, k, r, st, append=not first, digits=digits)
            else:
                total += write_text(path, k, r, append=not first)
            first = False
    if first:
        open(path, "w").close()
    return total


def load_table(table, path, chunk=1 << 20):
    """Insert a text dump (plain or Param-style) into an HBM table (checkpoint resume)."""
    import torch

    from ..ops._util import u32_tensor

    keys, rows, state = read_text_param(path, table.dim, state_dim=table.dim * max(table.state_words, 1))
    for s in range(0, len(keys), chunk):
        kd = u32_tensor(keys[s:s + chunk], table.device)
        rd = torch.from_numpy(rows[s:s + chunk]).to(table.device)
        sd = torch.from_numpy(state[s:s + chunk]).to(table.device) if (state is not None and table.has_state) else None
        table.load(kd, rd, sd)
    return len(keys)


_BIN_MAGIC = 0x32425353          # "SSB2": records <magic><n><dim><flags> keys[n] rows[n*dim] (state[n*words*dim] if flags & 1)


def save_table_binary(table, path, with_state=True, max_bytes=256 << 20):
    """Raw dump, streamed in bounded pieces: a sequence of records
    ``<uint32 magic "SSB2"><uint32 n><uint32 dim><uint32 flags> keys[n] rows[n*dim] [state[n*words*dim] if flags & 1]``
    (little endian; ``words`` = optimizer-state words per weight = bits 8..15 of flags, 0 reads as 1). The optimizer
    state is included by default when the table has one."""
    with_state = bool(with_state) and table.has_state
    with open(path, "wb") as f:
        for shard in table.owned:
            for k, r, st in table.iter_shard_chunks(shard, with_state=with_state, max_bytes=max_bytes):
                words = (st.shape[1] // table.dim) if st is not None else 0
                np.array([_BIN_MAGIC, len(k), table.dim, (1 | (words << 8)) if st is not None else 0], dtype=np.uint32).tofile(f)
                k.tofile(f)
                r.astype(np.float32).tofile(f)
                if st is not None:
                    st.astype(np.float32).tofile(f)
    return os.path.getsize(path)


def load_table_binary(table, path):
    """Reads the record format above, and the round-1 format (``<n><dim> keys rows`` per shard, no state)."""
    import torch

    from ..ops._util import u32_tensor

    total = 0
    with open(path, "rb") as f:
        while True:
            hdr = np.fromfile(f, dtype=np.uint32, count=2)
            if hdr.size < 2:
                break
            flags = 0
            if int(hdr[0]) == _BIN_MAGIC:
                rest = np.fromfile(f, dtype=np.uint32, count=2)
                n, dim, flags = int(hdr[1]), int(rest[0]), int(rest[1])
            else:
                n, dim = int(hdr[0]), int(hdr[1])
            k = np.fromfile(f, dtype=np.uint32, count=n)
            r = np.fromfile(f, dtype=np.float32, count=n * dim).reshape(n, dim)
            words = max((flags >> 8) & 0xFF, 1)
            st = np.fromfile(f, dtype=np.float32, count=n * dim * words).reshape(n, dim * words) if flags & 1 else None
            sd = torch.from_numpy(st).to(table.device) if (st is not None and table.has_state
                                                           and words == max(table.state_words, 1)) else None
            table.load(u32_tensor(k, table.device), torch.from_numpy(r).to(table.device), sd)
            total += n
    return total
