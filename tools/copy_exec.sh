#!/bin/bash
# Copy the role binaries (and the native libraries they load) into a run directory.
# ref: /root/reference/src/tools/copy_exec.sh:4-9 (copies src/apps/$APP/bin/{worker,server,master}).
DST="${1:?usage: copy_exec.sh <run_dir>}"; ROOT="$(cd "$(dirname "$0")/.." && pwd)"
mkdir -p "$DST/bin" "$DST/lib"
cp "$ROOT"/swiftsnails_b200/bin/swift_{master,server,worker} "$DST/bin/" && cp "$ROOT"/swiftsnails_b200/lib/libssb200_core.so "$DST/lib/"
echo "copied to $DST (binaries resolve ../lib through their rpath)"
