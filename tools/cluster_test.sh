#!/bin/bash
# Localhost integration run: 1 master + 2 servers + 2 workers over ZeroMQ (BASELINE config 1).
# ref: /root/reference/src/tools/cluster_test.sh (a manual run on a Hadoop cluster there).
cd "$(dirname "$0")/.." && exec python -m swiftsnails_b200.parallel.launch_cluster --app "${1:-bench}" --servers 2 --workers 2
