"""Start one process per GPU for a job that the master runs (parallel/group.py) — the B200 analogue of the reference's
hadoop-server.sh / hadoop-worker.sh, which start the roles as Hadoop-streaming reducers (src/tools/hadoop-*.sh).

    python -m swiftsnails_b200.parallel.launch_gpu 8 -m swiftsnails_b200.train -config worker.conf -data corpus.txt -out params.txt
    python -m swiftsnails_b200.parallel.launch_gpu 4 bench.py --gpus 4

Every child gets RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT (what `group.init` reads; `torchrun` sets the
same variables, so either launcher works). No torch.distributed rendezvous is involved: the ranks find one another
through swift_master. The first child that fails takes the others down with it."""
from __future__ import annotations

import os
import signal
import subprocess
import sys
import time


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if len(argv) < 2 or not argv[0].isdigit():
        print(__doc__)
        return 2
    n, cmd = int(argv[0]), argv[1:]
    port = int(os.environ.get("MASTER_PORT", "29511"))
    procs = []
    for r in range(n):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(n), LOCAL_WORLD_SIZE=str(n),
                   MASTER_ADDR=os.environ.get("MASTER_ADDR", "127.0.0.1"), MASTER_PORT=str(port),
                   OMP_NUM_THREADS=os.environ.get("OMP_NUM_THREADS", "1"))
        procs.append(subprocess.Popen([sys.executable] + cmd, env=env))
    rc = 0
    try:
        alive = set(range(n))
        while alive:
            for r in list(alive):
                code = procs[r].poll()
                if code is None:
                    continue
                alive.discard(r)
                if code != 0 and rc == 0:
                    rc = code
                    for q in alive:                      # exact PIDs we started, nothing else
                        procs[q].send_signal(signal.SIGTERM)
            time.sleep(0.05)
    except KeyboardInterrupt:
        for p in procs:
            if p.poll() is None:
                p.send_signal(signal.SIGTERM)
        rc = 130
    return rc


if __name__ == "__main__":
    sys.exit(main())
