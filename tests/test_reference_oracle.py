"""The reference's OWN code as the oracle (SURVEY.md §4 / §8 "Oracle" column): baseline/ref_apps/ref_oracle.cpp is
compiled against the unmodified headers of /root/reference (hashfrag.h, sparsetable.h, Buffer.h, ConfigParser.h) with the
shim headers, and its output is compared with this repository's native core and Python router:

* key -> node routing of BasicHashFrag (parameter/hashfrag.h:36-53) on 200 000 keys, several (nodes, fragments) shapes,
  including the remainder-goes-to-the-last-server rule, and the wire form of the fragment map (hashfrag.h:55-62);
* a SparseTable driven through PullAccessAgent / PushAccessAgent with an SGD rule (sparsetable.h:142-149,181-192): the
  update applied per key and the `key \\t Vec:\\t v…` dump (sparsetable.h:49-56), read back by this repository's loader."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


@pytest.fixture(scope="module")
def oracle(native_built):
    if not os.path.isdir("/root/reference/src") and not os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "src")):
        pytest.skip("reference sources not available")
    from baseline import run_reference as rr

    err = rr.ensure_oracle_built()
    if err:
        pytest.skip("reference oracle cannot be built here: " + err)
    return rr.ORACLE


def _lcg_stream(seed, n):
    out = np.empty(n, dtype=np.uint32)
    s = np.uint64(seed)
    for i in range(n):
        s = (s * np.uint64(1664525) + np.uint64(1013904223)) & np.uint64(0xFFFFFFFF)
        out[i] = s
    return out


@pytest.mark.parametrize("nodes,frags", [(2, 100), (7, 100), (8, 1024), (3, 10), (5, 5)])
def test_routing_matches_the_reference_hashfrag(oracle, nodes, frags):
    from swiftsnails_b200 import _lib
    from swiftsnails_b200.parallel.hashfrag import HashFrag

    n = 200_000 if frags == 1024 else 20_000
    p = subprocess.run([oracle, "route", str(nodes), str(frags), str(n), "12345"], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr[-500:]
    lines = p.stdout.splitlines()
    ref = np.array([l.split() for l in lines[:n]], dtype=np.int64)
    ref_keys, ref_nodes = ref[:, 0].astype(np.uint32), ref[:, 1].astype(np.int32)
    assert np.array_equal(ref_keys, _lcg_stream(12345, n))
    m = lines[n].split()
    assert m[0] == "MAP" and int(m[1]) == nodes and int(m[2]) == frags
    ref_map = np.array(m[3:], dtype=np.uint32)
    # Python router (the one the GPU kernels' fragment map comes from)
    hf = HashFrag(nodes, frags)
    assert np.array_equal(hf.map, ref_map)
    assert np.array_equal(hf.to_node_id(ref_keys), ref_nodes)
    assert hf.serialize() == np.array([nodes, frags], dtype="<i4").tobytes() + ref_map.astype("<u4").tobytes()
    # native core
    lib = _lib.core()
    lib.ssb_hashfrag_build.argtypes = [C.c_int, C.c_int, C.c_void_p]
    lib.ssb_hashfrag_route.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_void_p]
    cmap = np.zeros(frags, dtype=np.uint32)
    assert lib.ssb_hashfrag_build(nodes, frags, cmap.ctypes.data) == 0
    assert np.array_equal(cmap, ref_map)
    out = np.zeros(n, dtype=np.int32)
    lib.ssb_hashfrag_route(ref_keys.ctypes.data, n, cmap.ctypes.data, frags, out.ctypes.data)
    assert np.array_equal(out, ref_nodes)


def test_sgd_table_and_dump_match_the_reference_sparsetable(oracle, tmp_path):
    from swiftsnails_b200.ops.cpu_table import CpuSparseTable as CpuTable
    from swiftsnails_b200.utils import dump

    dim, lr, n, steps, seed = 6, 0.1, 3000, 3, 99
    p = subprocess.run([oracle, "sgd", "4", str(dim), str(lr), str(n), str(steps), str(seed)], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr[-500:]
    ref_dump = str(tmp_path / "ref_dump.txt")
    with open(ref_dump, "w") as f:
        f.write(p.stdout)
    # this repository's loader reads the reference's dump as it is
    rk, rr_, st = dump.read_text_param(ref_dump, dim)
    assert st is None and len(rk) == len(np.unique(rk))
    ref = {int(k): r for k, r in zip(rk, rr_)}
    # the same key / gradient streams, restated
    stream = _lcg_stream(seed, n + steps * n * dim)
    keys = (stream[:n] % np.uint32(100000)).astype(np.uint32)
    g = ((stream[n:] % np.uint32(2001)).astype(np.int64) - 1000).astype(np.float32) / np.float32(1000.0)
    g = g.reshape(steps, n, dim)
    assert set(ref) == set(int(k) for k in keys)
    # the oracle's init rule (ref_oracle.cpp OPull::init_param) -> what its SGD steps added to every key
    def ref_init(k):
        return np.array([((int(k) * 31 + j) % (1 << 32) % 17 - 8) / 16.0 for j in range(dim)], dtype=np.float32)
    ours = CpuTable(dim, shard_num=4, optimizer="sgd", lr=lr)
    init = {int(k): r.copy() for k, r in zip(np.unique(keys), ours.pull(np.unique(keys)))}
    for s in range(steps):
        for i in range(n):                      # one key at a time, in the oracle's order (duplicates included)
            ours.push(keys[i:i + 1], g[s, i:i + 1])
    fin = {int(k): r for k, r in zip(np.unique(keys), ours.pull(np.unique(keys)))}
    for k in list(ref)[:1500]:
        assert np.allclose(fin[k] - init[k], ref[k] - ref_init(k), rtol=0, atol=2e-4), k     # the reference prints 6 digits
    # and our dump of the same table has the reference's line format
    ours_dump = str(tmp_path / "ours.txt")
    ours.dump(ours_dump)
    first = open(ours_dump).readline()
    assert "\tVec:\t" in first and first.split("\t")[0].isdigit()
