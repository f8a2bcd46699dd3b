"""Launch a localhost parameter-server cluster on the CPU / ZeroMQ path: 1 master + S servers +
W workers as separate processes (the native ``swift_master`` / ``swift_server`` /
``swift_worker`` binaries built from ``csrc/apps``).

ref: /root/reference/src/tools/run_master.sh, run_server.sh, run_worker.sh, cluster_test.sh
(the reference launches the roles as Hadoop-streaming reducers; here a local process group) and
SURVEY.md Appendix A (the config shape: master.conf / node.conf importing common.conf).
"""
from __future__ import annotations

import os
import re
import socket
import subprocess
import tempfile
import time

_HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN_DIR = os.path.join(_HERE, "bin")


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


DEFAULTS = {
    "async_exec_num": 4, "listen_thread_num": 1, "init_timeout": 60, "master_time_out": 60,
    "master_longest_alive_duration": 0, "frag_num": 100, "shard_num": 8, "param_backup_period": 0,
    "param_backup_root": ".", "num_iters": 1, "learning_rate": 0.05, "async_channel_thread_num": 2,
    "local_train": 0, "rpc_timeout": 120,
}


def write_configs(workdir, n_servers, n_workers, overrides=None, master_port=None):
    cfg = dict(DEFAULTS)
    cfg.update(overrides or {})
    port = master_port or free_port()
    master_addr = "tcp://127.0.0.1:%d" % port
    with open(os.path.join(workdir, "common.conf"), "w") as f:
        f.write("# shared by every role\n")
        for k, v in cfg.items():
            f.write("%s: %s\n" % (k, v))
        f.write("master_addr: %s\n" % master_addr)
    with open(os.path.join(workdir, "master.conf"), "w") as f:
        f.write("import common.conf\nlisten_addr: %s\nexpected_node_num: %d\n" % (master_addr, n_servers + n_workers))
    with open(os.path.join(workdir, "node.conf"), "w") as f:
        f.write("import common.conf\nlisten_addr:\n")
    return master_addr


def launch_local_cluster(app, n_servers=2, n_workers=2, overrides=None, data_files=None, workdir=None,
                         timeout=300, env=None):
    """Run the cluster to completion. Returns dict(master=..., servers=[...], workers=[...]) of
    (returncode, stdout, stderr) triples plus ``workdir`` and ``seconds``."""
    for b in ("swift_master", "swift_server", "swift_worker"):
        if not os.path.exists(os.path.join(BIN_DIR, b)):
            raise RuntimeError("native binary %s missing: run `python -m swiftsnails_b200.build`" % b)
    workdir = workdir or tempfile.mkdtemp(prefix="ssb200_cluster_")
    write_configs(workdir, n_servers, n_workers, overrides)
    e = dict(os.environ)
    e.setdefault("SSB200_LOCAL_IP", "127.0.0.1")
    e.update(env or {})

    counter = [0]

    def spawn(binary, conf, extra=()):
        # stdout/stderr go to files: a server dumps its whole table to stdout on termination and
        # would block on a full pipe while the master waits for its ack
        tag = "%s.%d" % (binary, counter[0])
        counter[0] += 1
        so = open(os.path.join(workdir, tag + ".out"), "w")
        se = open(os.path.join(workdir, tag + ".err"), "w")
        p = subprocess.Popen([os.path.join(BIN_DIR, binary), "-config", conf] + list(extra), cwd=workdir,
                             stdout=so, stderr=se, env=e)
        p._ssb_files = (so, se)
        return p

    t0 = time.time()
    procs = {"master": spawn("swift_master", "master.conf"), "servers": [], "workers": []}
    time.sleep(0.2)
    for _ in range(n_servers):
        procs["servers"].append(spawn("swift_server", "node.conf"))
    for w in range(n_workers):
        extra = ["-app", app]
        if data_files:
            extra += ["-data", data_files[w % len(data_files)]]
        procs["workers"].append(spawn("swift_worker", "node.conf", extra))
    out = {"workdir": workdir}
    deadline = t0 + timeout

    def finish(p):
        note = ""
        try:
            p.wait(timeout=max(1.0, deadline - time.time()))
        except subprocess.TimeoutExpired:
            p.kill()
            p.wait()
            note = "\n[launcher] timed out"
        for f in p._ssb_files:
            f.close()
        with open(p._ssb_files[0].name) as f:
            so = f.read()
        with open(p._ssb_files[1].name) as f:
            se = f.read()
        return (p.returncode, so, se + note)

    out["workers"] = [finish(p) for p in procs["workers"]]
    out["master"] = finish(procs["master"])
    out["servers"] = [finish(p) for p in procs["servers"]]
    out["seconds"] = time.time() - t0
    return out


def parse_result_line(stdout, tag="RESULT"):
    """`RESULT k v k v …` / `BENCH k v …` line -> dict of floats."""
    for line in stdout.splitlines():
        if line.startswith(tag + " "):
            toks = line.split()[1:]
            return {toks[i]: float(toks[i + 1]) for i in range(0, len(toks) - 1, 2)}
    return None


def parse_dump(text):
    """Reference-format dump text -> {key: [floats]}."""
    out = {}
    for line in text.splitlines():
        if "\t" not in line:
            continue
        k, rest = line.split("\t", 1)
        if not re.fullmatch(r"\d+", k):
            continue
        rest = rest.replace("Vec:", " ")
        out[int(k)] = [float(x) for x in rest.split()]
    return out
