#!/bin/bash
# One process per GPU (server shard g + worker g): the ranks register with swift_master over ZeroMQ, data moves over
# NVLink peer memory. The B200 analogue of the reference's hadoop-server.sh / hadoop-worker.sh (which start the roles
# as Hadoop-streaming reducers): usage: launch_gpu_job.sh <n_gpus> <script.py | -m module> [args…]
cd "$(dirname "$0")/.." && exec python -m swiftsnails_b200.parallel.launch_gpu "$@"
