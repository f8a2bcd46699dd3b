#!/bin/bash
# ref: /root/reference/src/tools/run_worker.sh — the worker's data shard arrives on stdin and is
# spooled to ./data.txt first, then `./swift_worker -config worker.conf -data ./data.txt`.
BIN="$(dirname "$0")/../swiftsnails_b200/bin"
CONF="${1:-./worker.conf}"; APP="${2:-word2vec}"
cat > ./data.txt
exec "$BIN/swift_worker" -config "$CONF" -data ./data.txt -app "$APP"
