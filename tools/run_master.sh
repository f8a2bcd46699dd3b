#!/bin/bash
# ref: /root/reference/src/tools/run_master.sh — `./swift_master -config ./master.conf`
BIN="$(dirname "$0")/../swiftsnails_b200/bin"
exec "$BIN/swift_master" -config "${1:-./master.conf}"
