"""Native parser of sparse-LR example files (`label key:value key:value …`, csrc/core/core_capi.cc: ssb_lr_parse_file),
the input of `python -m swiftsnails_b200.train -app lr`, against a plain Python restatement: padding, default value 1,
label sign, key modulo (hashed feature space), per-worker line split, dropped features, junk lines.
ref: the CPU worker's LogisticRegressionAlg::parse_record (csrc/apps/algorithms.h) reads the same format."""
import ctypes as C

import numpy as np


def _parse(lib, path, max_feat, key_mod, split=(1, 0)):
    stats = (C.c_longlong * 2)()
    n = lib.ssb_lr_parse_file(path.encode(), max_feat, None, None, None, 0, key_mod, split[0], split[1], stats)
    assert n >= 0
    keys = np.empty((n, max_feat), dtype=np.uint32)
    vals = np.empty((n, max_feat), dtype=np.float32)
    labels = np.empty(n, dtype=np.float32)
    got = lib.ssb_lr_parse_file(path.encode(), max_feat, keys.ctypes.data, vals.ctypes.data, labels.ctypes.data, n, key_mod,
                                split[0], split[1], stats)
    assert got == n
    return keys, vals, labels, int(stats[1])


def test_parse_matches_python(tmp_path, native_built):
    from swiftsnails_b200 import _lib

    lib = _lib.core()
    lib.ssb_lr_parse_file.restype = C.c_longlong
    lib.ssb_lr_parse_file.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_ulonglong,
                                      C.c_int, C.c_int, C.c_void_p]
    rng = np.random.default_rng(0)
    lines, recs = [], []
    for i in range(300):
        if i % 17 == 5:
            lines.append("# a comment" if i % 2 else "")
        nf = int(rng.integers(0, 9))
        ks = rng.integers(0, 5000, nf)
        vs = [1.0 if rng.random() < 0.5 else round(float(rng.random() * 3), 3) for _ in range(nf)]
        label = int(rng.integers(-1, 2))
        recs.append((label, list(zip(ks.tolist(), vs))))
        toks = [("%d" % k) if v == 1.0 else ("%d:%g" % (k, v)) for k, v in zip(ks, vs)]
        lines.append(("  " if i % 7 == 0 else "") + " ".join([str(label)] + toks) + ("\r" if i % 13 == 0 else ""))
    path = str(tmp_path / "ex.txt")
    with open(path, "w") as f:
        f.write("\n".join(lines))
    for max_feat, key_mod, split in [(8, 1 << 20, (1, 0)), (4, 1000, (1, 0)), (8, 1 << 20, (3, 1))]:
        keys, vals, labels, dropped = _parse(lib, path, max_feat, key_mod, split)
        mine = [r for i, r in enumerate(recs) if i % split[0] == split[1]]
        assert len(mine) == len(labels)
        exp_dropped = 0
        for row, (label, feats) in enumerate(mine):
            assert labels[row] == (1.0 if label > 0 else 0.0)
            exp_dropped += max(0, len(feats) - max_feat)
            for j in range(max_feat):
                if j < len(feats):
                    assert keys[row, j] == feats[j][0] % key_mod
                    assert abs(vals[row, j] - feats[j][1]) < 1e-6
                else:
                    assert keys[row, j] == 0xFFFFFFFF and vals[row, j] == 0.0
        assert dropped == exp_dropped
    assert lib.ssb_lr_parse_file(str(tmp_path / "missing.txt").encode(), 4, None, None, None, 0, 0, 1, 0, None) == -1
