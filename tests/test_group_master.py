"""The master as the control plane of a GPU job (parallel/group.py, csrc/core/system/group_exchange.h), on CPU:
three processes register with a swift_master over ZeroMQ, get rank / world / fragment map from it, all-gather and
barrier through it, and shut down with WORKER_FINISH_WORK -> SERVER_TOLD_TO_TERMINATE -> dump -> ack.
ref: /root/reference/src/core/system/master/init.h:21-171, node_init.h:16-152, master/terminate.h:15-109,
server/terminate.h:16-54."""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = textwrap.dedent('''
    import os, sys
    sys.path.insert(0, %r)
    import numpy as np
    from swiftsnails_b200.parallel import group
    from swiftsnails_b200.parallel.hashfrag import HashFrag
    g = group.init("master", frag_num=96)
    assert g.kind == "master" and g.world == 3 and 0 <= g.rank < 3
    out = g.all_gather_object({"rank": g.rank, "env": int(os.environ["RANK"])})
    assert [o["rank"] for o in out] == [0, 1, 2] and sorted(o["env"] for o in out) == [0, 1, 2]
    assert g.broadcast_object("tag-%%d" %% g.rank if g.rank == 1 else None, src=1) == "tag-1"
    assert g.all_reduce(g.rank + 0.5, "max") == 2.5 and g.all_reduce(1, "sum") == 3
    big = bytes([g.rank]) * 70000                       # larger than one ZeroMQ small-message frame
    parts = g.all_gather_bytes(big)
    assert [len(p) for p in parts] == [70000] * 3 and all(p[0] == r for r, p in enumerate(parts))
    for _ in range(20):
        g.barrier()
    hf = g.hashfrag()
    ref = HashFrag(3, 96)                               # same rule on both sides (parameter/hashfrag.h:36-53)
    assert hf.num_nodes == 3 and hf.frag_num == 96 and np.array_equal(hf.map, ref.map)
    marker = os.path.join(sys.argv[1], "dump.%%d" %% g.rank)
    g.finish(dump=lambda: open(marker, "w").write("dumped"))
    g.close()
    print("CHILD_OK", g.rank)
''')


def _free_port():
    import socket

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_three_ranks_register_exchange_and_terminate_through_the_master(tmp_path, native_built):
    script = tmp_path / "child.py"
    script.write_text(CHILD % ROOT)
    port = _free_port() - 17                # group.init puts the master on MASTER_PORT + 17
    env = dict(os.environ, WORLD_SIZE="3", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
               SSB200_MASTER_LOG=str(tmp_path / "master.log"))
    procs = [subprocess.Popen([sys.executable, str(script), str(tmp_path)], env=dict(env, RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(3)]
    outs = [p.communicate(timeout=180)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0 and "CHILD_OK" in o, o[-2000:]
    assert sorted(os.listdir(tmp_path))[:5].count("child.py") == 1
    for r in range(3):
        assert (tmp_path / ("dump.%d" % r)).read_text() == "dumped"
    log = (tmp_path / "master.log").read_text()
    assert "cluster is up (3 servers" in log and "Master terminated normally" in log


def test_own_launcher_starts_one_process_per_rank(tmp_path, native_built):
    """`python -m swiftsnails_b200.parallel.launch_gpu N script`: RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* for every child,
    no torch.distributed rendezvous; a failing child takes the job down with a non-zero exit."""
    ok = tmp_path / "ok.py"
    ok.write_text("import os, sys\nsys.path.insert(0, %r)\nfrom swiftsnails_b200.parallel import group\n"
                  "g = group.init('master')\nv = g.all_gather_object(int(os.environ['LOCAL_RANK']))\n"
                  "assert sorted(v) == list(range(g.world)) and g.world == 3\ng.finish(); g.close(); print('LAUNCHED_OK')\n" % ROOT)
    env = dict(os.environ, MASTER_PORT=str(_free_port() - 17))
    p = subprocess.run([sys.executable, "-m", "swiftsnails_b200.parallel.launch_gpu", "3", str(ok)], capture_output=True, text=True,
                       timeout=180, cwd=ROOT, env=env)
    assert p.returncode == 0 and p.stdout.count("LAUNCHED_OK") == 3, p.stdout[-1500:] + p.stderr[-1500:]
    bad = tmp_path / "bad.py"
    bad.write_text("import os, sys, time\nif os.environ['RANK'] == '1': sys.exit(7)\ntime.sleep(60)\n")
    p = subprocess.run([sys.executable, "-m", "swiftsnails_b200.parallel.launch_gpu", "3", str(bad)], capture_output=True, text=True,
                       timeout=60, cwd=ROOT, env=env)
    assert p.returncode == 7
