"""Native parallel corpus loader (csrc/core/corpus_loader.h) against a plain Python restatement: batch order,
padding, per-worker split, passes that span minibatches, id bound, word counts, and the line-numbering rule it
shares with train.count_sentences. ref: the reference reads its shard with scan_file_by_line
(/root/reference/src/utils/file.h:14-33); Hadoop streaming gives each worker its split (tools/hadoop-worker.sh:8-21)."""
import numpy as np
import pytest

PAD = 0xFFFFFFFF


def _write(path, rng, n_lines, vocab, junk=True):
    lines, sents = [], []
    for i in range(n_lines):
        if junk and i % 7 == 3:
            lines.append("" if i % 2 else "# comment %d" % i)          # not sentences: skipped and not numbered
        toks = rng.integers(0, vocab, rng.integers(1, 12)).tolist()
        sents.append(toks)
        lines.append(("  " if i % 5 == 0 else "") + " ".join(map(str, toks)) + ("\r" if i % 11 == 0 else ""))
    with open(path, "w") as f:
        f.write("\n".join(lines))          # no trailing newline on purpose
    return sents


def _expected(sents, batch, sent_len, epochs=1, split=(1, 0), max_id=0):
    mine = [s for i, s in enumerate(sents) if i % split[0] == split[1]]
    rows, real = [], []
    for _ in range(epochs):
        for s in mine:
            t = [x if (max_id == 0 or x < max_id) else PAD for x in s[:sent_len]]
            real.append(sum(1 for x in t if x != PAD))
            rows.append(t + [PAD] * (sent_len - len(t)))
    per = batch // sent_len
    out = []
    for b0 in range(0, len(rows), per):
        chunk = rows[b0:b0 + per]
        n = sum(real[b0:b0 + per])
        chunk = chunk + [[PAD] * sent_len] * (per - len(chunk))
        out.append((np.array(chunk, dtype=np.uint32).reshape(-1), n))
    return out


@pytest.mark.parametrize("threads", [1, 3, 8])
@pytest.mark.parametrize("split", [(1, 0), (3, 1)])
def test_loader_matches_python_restatement(tmp_path, native_built, threads, split):
    from swiftsnails_b200.utils.corpus import CorpusLoader

    rng = np.random.default_rng(5)
    path = str(tmp_path / "c.txt")
    sents = _write(path, rng, 157, 50)
    exp = _expected(sents, batch=40, sent_len=8, epochs=2, split=split, max_id=45)
    loader = CorpusLoader(path, 40, 8, epochs=2, pin=False, split=split, max_id=45, threads=threads)
    assert loader.batches == len(exp)
    got = [(b.numpy().view(np.uint32).copy(), n) for b, n in loader]
    assert len(got) == len(exp)
    for (gb, gn), (eb, en) in zip(got, exp):
        assert gn == en
        assert np.array_equal(gb, eb)
    bad = sum(1 for _ in range(2) for i, s in enumerate(sents) if i % split[0] == split[1] for x in s[:8] if x >= 45)
    assert loader.bad_tokens == bad
    loader.close()


def test_sentence_count_rule_is_shared_with_the_trainer(tmp_path, native_built):
    from swiftsnails_b200.train import count_sentences
    from swiftsnails_b200.utils.corpus import CorpusLoader

    path = str(tmp_path / "c.txt")
    sents = _write(path, np.random.default_rng(1), 90, 30)
    loader = CorpusLoader(path, 16, 4, pin=False)
    assert loader.sentences == len(sents) == count_sentences(path)
    loader.close()


def test_word_counts(tmp_path, native_built):
    from swiftsnails_b200.utils.corpus import count_words

    path = str(tmp_path / "c.txt")
    sents = _write(path, np.random.default_rng(2), 400, 64)
    ref = np.zeros(60, dtype=np.uint64)
    bad = 0
    for s in sents:
        for x in s:
            if x < 60:
                ref[x] += 1
            else:
                bad += 1
    for threads in (1, 4):
        counts, nbad = count_words(path, 60, threads=threads)
        assert nbad == bad and np.array_equal(counts, ref)


def test_bad_arguments_are_refused_not_fatal(tmp_path, native_built):
    from swiftsnails_b200.utils.corpus import CorpusLoader

    path = str(tmp_path / "c.txt")
    _write(path, np.random.default_rng(3), 10, 5)
    with pytest.raises(ValueError):
        CorpusLoader(path, 10, 4, pin=False)            # batch not a multiple of sent_len
    with pytest.raises(OSError):
        CorpusLoader(str(tmp_path / "missing.txt"), 8, 4, pin=False)
    empty = str(tmp_path / "empty.txt")
    open(empty, "w").close()
    assert list(CorpusLoader(empty, 8, 4, pin=False)) == []
