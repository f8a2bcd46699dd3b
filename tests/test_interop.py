"""Wire interoperability with the UNMODIFIED reference: nodes built from /root/reference (baseline/_ref/bin/ref_ps,
its master / server / worker roles through the reference's own framework classes) share a cluster with nodes of this
repository's C++ core (swift_master / swift_server / swift_worker). Every mix of the three roles must register, route,
fetch the fragment map, pull, push, finish and terminate — meta message layout, route table, fragment map, value
wire form and the six message classes are therefore the reference's. `wire_compat: reference` switches this core's
pull requests from bare key arrays to the reference's (key, empty value) pairs (global_pull_access.h:84-88).
ref: /root/reference/src/core/Message.h:150-183, system/ServerWorkerRoute.h:40-71, parameter/hashfrag.h:55-62."""
import os
import subprocess
import sys
import time

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


@pytest.fixture(scope="module")
def ref_bin(native_built):
    if not os.path.isdir("/root/reference/src") and not os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "src")):
        pytest.skip("reference sources not available")
    from baseline import run_reference as rr

    err = rr.ensure_built()
    if err:
        pytest.skip("reference cannot be built here: " + err)
    return rr


def _run_cluster(rr, wd, roles):
    """roles: dict role -> 'ref' | 'ours'. Returns {role: (returncode, stdout, stderr)}."""
    rr._write_conf(wd, 2, {"dim": 16, "bench_keys": 1500, "bench_key_space": 1000000, "bench_iters": 3, "optimizer": "sgd",
                           "update": "sgd", "wire_compat": "reference"})
    ours_bin = os.path.join(ROOT, "swiftsnails_b200", "bin")
    cmds = {
        ("master", "ref"): [rr.BIN, "-role", "master", "-config", "master.conf"],
        ("server", "ref"): [rr.BIN, "-role", "server", "-config", "node.conf"],
        ("worker", "ref"): [rr.BIN, "-role", "worker", "-config", "node.conf", "-app", "bench"],
        ("master", "ours"): [os.path.join(ours_bin, "swift_master"), "-config", "master.conf"],
        ("server", "ours"): [os.path.join(ours_bin, "swift_server"), "-config", "node.conf"],
        ("worker", "ours"): [os.path.join(ours_bin, "swift_worker"), "-config", "node.conf", "-app", "bench"],
    }
    procs = {}
    for role in ("master", "server", "worker"):
        so, se = open(os.path.join(wd, role + ".out"), "w"), open(os.path.join(wd, role + ".err"), "w")
        procs[role] = subprocess.Popen(cmds[(role, roles[role])], cwd=wd, stdout=so, stderr=se)
        if role == "master":
            time.sleep(0.3)
    out = {}
    for role in ("worker", "server", "master"):
        try:
            rc = procs[role].wait(timeout=90)
        except subprocess.TimeoutExpired:
            procs[role].kill()
            rc = "timeout"
        out[role] = (rc, open(os.path.join(wd, role + ".out")).read(), open(os.path.join(wd, role + ".err")).read())
    return out


@pytest.mark.parametrize("mix", ["ref-ref-ours", "ours-ref-ours", "ref-ours-ours", "ours-ours-ref", "ref-ours-ref", "ours-ref-ref",
                                 "ours-ours-ours"])          # the last one: this core alone, speaking the reference's pull form
def test_mixed_cluster_runs_to_completion(ref_bin, tmp_path, mix):
    m, s, w = mix.split("-")
    res = _run_cluster(ref_bin, str(tmp_path), {"master": m, "server": s, "worker": w})
    for role, (rc, so, se) in res.items():
        assert rc == 0, (role, rc, so[-500:], se[-1500:])
    # the worker pulled and pushed its keys ...
    assert "BENCH pull_keys_per_s" in res["worker"][1]
    # ... and the server, told to terminate by the master after the worker finished, dumped one row per key
    rows = [l for l in res["server"][1].splitlines() if "\t" in l]
    assert len(rows) == 1500, len(rows)
    key, vec = rows[0].split("\t", 1)
    assert key.isdigit() and len(vec.replace("Vec:", "").split()) == 16
