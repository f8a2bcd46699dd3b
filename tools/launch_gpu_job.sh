#!/bin/bash
# One process per GPU (server shard g + worker g), NCCL for plumbing, NVLink peer memory for data.
# The B200 analogue of the reference's hadoop-server.sh / hadoop-worker.sh (which start the roles
# as Hadoop-streaming reducers): usage: launch_gpu_job.sh <n_gpus> <script.py> [args…]
N="${1:?n_gpus}"; shift
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port "${MASTER_PORT:-29511}" "$@"
