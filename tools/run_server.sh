#!/bin/bash
# ref: /root/reference/src/tools/run_server.sh — `./swift_server -config server.conf`;
# the final parameter dump is the process's stdout.
BIN="$(dirname "$0")/../swiftsnails_b200/bin"
exec "$BIN/swift_server" -config "${1:-./server.conf}"
