#!/bin/bash
# Pack the source tree and copy it to another host. ref: /root/reference/src/tools/distribute.sh:3-12.
HOST="${1:?usage: distribute.sh user@host [remote_dir]}"; REMOTE="${2:-swiftsnails_b200}"
ROOT="$(cd "$(dirname "$0")/.." && pwd)"; TAR=/tmp/swiftsnails_b200_src.tar.gz
tar --exclude=.git --exclude=build --exclude=gpurun_out --exclude='*.so' -czf "$TAR" -C "$ROOT" . && scp "$TAR" "$HOST:/tmp/" && \
  ssh "$HOST" "mkdir -p $REMOTE && tar -xzf /tmp/$(basename $TAR) -C $REMOTE && cd $REMOTE && python -m swiftsnails_b200.build"
