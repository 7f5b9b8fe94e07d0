#!/bin/bash
# Start the coordinator in the background (log + pidfile).  Same environment contract as the
# reference script: COORDINATOR_PORT, BINARY_PATH, LOG_FILE.  Additions: PS_ADDRESS (the
# reference never passes one, so its workers are told "localhost:50051" — SURVEY D1) and
# PID_FILE.
set -euo pipefail

COORDINATOR_PORT="${COORDINATOR_PORT:-50052}"
BINARY_PATH="${BINARY_PATH:-/opt/parameter-server/coordinator}"
LOG_FILE="${LOG_FILE:-/var/log/coordinator.log}"
PID_FILE="${PID_FILE:-/var/run/coordinator.pid}"
PS_ADDRESS="${PS_ADDRESS:-}"

if [ -n "$PS_ADDRESS" ]; then
  nohup "$BINARY_PATH" "0.0.0.0:${COORDINATOR_PORT}" "$PS_ADDRESS" >"$LOG_FILE" 2>&1 &
else
  nohup "$BINARY_PATH" "0.0.0.0:${COORDINATOR_PORT}" >"$LOG_FILE" 2>&1 &
fi
echo $! >"$PID_FILE"
echo "coordinator started (pid $(cat "$PID_FILE"), port ${COORDINATOR_PORT}, log ${LOG_FILE})"
