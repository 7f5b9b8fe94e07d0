#!/bin/bash
# Start one worker in the background.  Environment contract of the reference script:
# WORKER_ID, COORDINATOR_ADDR (required), ITERATIONS, WORKER_ADDR, WORKER_PORT,
# CHECKPOINT_PATH (6th argument only when set), BINARY_PATH, LOG_FILE.  Additions: PID_FILE
# and WORKER_FLAGS for the trailing --mode= / --device= extensions.
set -euo pipefail

WORKER_ID="${WORKER_ID:-0}"
COORDINATOR_ADDR="${COORDINATOR_ADDR:?COORDINATOR_ADDR (host:port of the coordinator) is required}"
ITERATIONS="${ITERATIONS:-100}"
WORKER_ADDR="${WORKER_ADDR:-}"
WORKER_PORT="${WORKER_PORT:-0}"
CHECKPOINT_PATH="${CHECKPOINT_PATH:-}"
BINARY_PATH="${BINARY_PATH:-/opt/parameter-server/worker_main}"
LOG_FILE="${LOG_FILE:-/var/log/worker_${WORKER_ID}.log}"
PID_FILE="${PID_FILE:-/var/run/worker_${WORKER_ID}.pid}"
WORKER_FLAGS="${WORKER_FLAGS:-}"

ARGS=("$COORDINATOR_ADDR" "$WORKER_ID" "$ITERATIONS" "$WORKER_ADDR" "$WORKER_PORT")
if [ -n "$CHECKPOINT_PATH" ]; then ARGS+=("$CHECKPOINT_PATH"); fi
# shellcheck disable=SC2086
nohup "$BINARY_PATH" "${ARGS[@]}" $WORKER_FLAGS >"$LOG_FILE" 2>&1 &
echo $! >"$PID_FILE"
echo "worker ${WORKER_ID} started (pid $(cat "$PID_FILE"), ${ITERATIONS} iterations, log ${LOG_FILE})"
