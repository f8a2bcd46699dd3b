"""Host-side data pipeline for the GPU workers: a native background reader
(``csrc/core/corpus_loader.h``) parses a text corpus (one sentence of integer word ids per line —
the reference's data format) into fixed-size token batches that are copied straight into pinned
memory, so parsing overlaps the GPU step."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _lib

PAD = 0xFFFFFFFF  # SSB_EMPTY_KEY: skipped by dedup / pull / the SGNS kernels


def count_words(path, vocab, threads=0):
    """Word frequencies of a text corpus (numpy uint64 [vocab]) and the number of ids >= vocab in it — the
    counts the unigram^0.75 negative-sampling table is built from (native, parallel over byte ranges of the file)."""
    import numpy as np

    lib = _lib.core()
    lib.ssb_corpus_count_words.restype = C.c_longlong
    lib.ssb_corpus_count_words.argtypes = [C.c_char_p, C.c_uint32, C.c_void_p, C.c_int]
    counts = np.zeros(int(vocab), dtype=np.uint64)
    bad = lib.ssb_corpus_count_words(str(path).encode(), int(vocab), counts.ctypes.data, int(threads))
    if bad < 0:
        raise OSError("cannot open corpus %s" % path)
    return counts, int(bad)


class CorpusLoader:
    def __init__(self, path, batch_tokens, sent_len, epochs=1, pin=True, split=None, max_id=0, threads=0):
        """``split=(n, i)``: keep only the sentence lines whose index % n == i — worker i's share of a
        corpus that n workers read together (the reference gets this split from Hadoop streaming).
        ``max_id`` > 0: token ids >= max_id are turned into padding and counted (``bad_tokens``) — an id >= vocab
        would otherwise alias another word's output vector. ``threads``: parser threads (0 = chosen from the host)."""
        lib = _lib.core()
        lib.ssb_corpus_open_ex.restype = C.c_void_p
        lib.ssb_corpus_open_ex.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint32]
        lib.ssb_corpus_next.restype = C.c_long
        lib.ssb_corpus_next.argtypes = [C.c_void_p, C.c_void_p]
        lib.ssb_corpus_stats.argtypes = [C.c_void_p, C.c_void_p]
        lib.ssb_corpus_close.argtypes = [C.c_void_p]
        self._lib = lib
        n, i = split if split else (1, 0)
        if batch_tokens <= 0 or sent_len <= 0 or batch_tokens % sent_len != 0:
            raise ValueError("batch_tokens (%d) must be a positive multiple of sent_len (%d)" % (batch_tokens, sent_len))
        self._h = lib.ssb_corpus_open_ex(str(path).encode(), batch_tokens, sent_len, PAD, epochs, int(n), int(i), int(threads),
                                         int(max_id))
        if not self._h:
            raise OSError("cannot open corpus %s" % path)
        self.batch_tokens = batch_tokens
        pinned = pin and torch.cuda.is_available()
        self._bufs = [torch.empty(batch_tokens, dtype=torch.int32, pin_memory=pinned) for _ in range(2)]
        self._i = 0

    def _stats(self):
        out = (C.c_longlong * 3)()
        self._lib.ssb_corpus_stats(self._h, out)
        return int(out[0]), int(out[1]), int(out[2])

    @property
    def sentences(self):
        """Sentence lines of this reader's share of the corpus, per pass."""
        return self._stats()[0]

    @property
    def batches(self):
        """Minibatches this reader will deliver (all passes; the last one may be partly padding)."""
        return self._stats()[1]

    @property
    def bad_tokens(self):
        """Token ids >= max_id replaced by padding so far."""
        return self._stats()[2]

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
