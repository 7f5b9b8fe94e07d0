#!/bin/bash
# Start a parameter-server shard in the background.  Environment contract of the reference
# script: PS_PORT, TOTAL_WORKERS, CHECKPOINT_INTERVAL, BINARY_PATH, LOG_FILE.  Additions:
# PID_FILE, PS_WORKDIR (checkpoints are written to the server's cwd), PS_FLAGS for the
# trailing --mode= / --optimizer= / --lr= / --device= extensions, and COORDINATOR_ADDR: when
# set the server follows the coordinator's worker registry, so an evicted worker stops
# blocking the sync barrier (the reference's two services never talk); SHARD_ID
# (with COORDINATOR_ADDR): this server is one of several, it registers in the coordinator's shard
# map and workers started with WORKER_FLAGS=--shards=N split the model's tensors over them.
set -euo pipefail

PS_PORT="${PS_PORT:-50051}"
TOTAL_WORKERS="${TOTAL_WORKERS:-3}"
CHECKPOINT_INTERVAL="${CHECKPOINT_INTERVAL:-10}"
BINARY_PATH="${BINARY_PATH:-/opt/parameter-server/parameter_server}"
LOG_FILE="${LOG_FILE:-/var/log/parameter_server.log}"
PID_FILE="${PID_FILE:-/var/run/parameter_server.pid}"
PS_FLAGS="${PS_FLAGS:-}"
if [ -n "${COORDINATOR_ADDR:-}" ]; then PS_FLAGS="$PS_FLAGS --coordinator=${COORDINATOR_ADDR}"; fi
if [ -n "${SHARD_ID:-}" ]; then PS_FLAGS="$PS_FLAGS --shard-id=${SHARD_ID}"; fi
if [ -n "${ADVERTISE_ADDR:-}" ]; then PS_FLAGS="$PS_FLAGS --advertise=${ADVERTISE_ADDR}"; fi

if [ -n "${PS_WORKDIR:-}" ]; then cd "$PS_WORKDIR"; fi
# shellcheck disable=SC2086
nohup "$BINARY_PATH" "0.0.0.0:${PS_PORT}" "$TOTAL_WORKERS" "$CHECKPOINT_INTERVAL" $PS_FLAGS \
  >"$LOG_FILE" 2>&1 &
echo $! >"$PID_FILE"
echo "parameter server started (pid $(cat "$PID_FILE"), port ${PS_PORT}, workers ${TOTAL_WORKERS})"
