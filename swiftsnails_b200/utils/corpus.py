"""Host-side data pipeline for the GPU workers: a native background reader
(``csrc/core/corpus_loader.h``) parses a text corpus (one sentence of integer word ids per line —
the reference's data format) into fixed-size token batches that are copied straight into pinned
memory, so parsing overlaps the GPU step."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _lib

PAD = 0xFFFFFFFF  # SSB_EMPTY_KEY: skipped by dedup / pull / the SGNS kernels


class CorpusLoader:
    def __init__(self, path, batch_tokens, sent_len, epochs=1, pin=True, split=None):
        """``split=(n, i)``: keep only the non-empty lines whose index % n == i — worker i's share of a
        corpus that n workers read together (the reference gets this split from Hadoop streaming)."""
        lib = _lib.core()
        lib.ssb_corpus_open_split.restype = C.c_void_p
        lib.ssb_corpus_open_split.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_uint32, C.c_int, C.c_int, C.c_int]
        lib.ssb_corpus_next.restype = C.c_long
        lib.ssb_corpus_next.argtypes = [C.c_void_p, C.c_void_p]
        lib.ssb_corpus_close.argtypes = [C.c_void_p]
        self._lib = lib
        n, i = split if split else (1, 0)
        self._h = lib.ssb_corpus_open_split(str(path).encode(), batch_tokens, sent_len, PAD, epochs, int(n), int(i))
        if not self._h:
            raise OSError("cannot open corpus %s" % path)
        self.batch_tokens = batch_tokens
        pinned = pin and torch.cuda.is_available()
        self._bufs = [torch.empty(batch_tokens, dtype=torch.int32, pin_memory=pinned) for _ in range(2)]
        self._i = 0

    def __iter__(self):
        return self

    def __next__(self):
        """Returns (tokens int32 [batch_tokens] in pinned host memory, number of real tokens)."""
        buf = self._bufs[self._i & 1]
        self._i += 1
        n = self._lib.ssb_corpus_next(self._h, buf.data_ptr())
        if n < 0:
            raise StopIteration
        return buf, int(n)

    def close(self):
        if self._h:
            self._lib.ssb_corpus_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
