#!/bin/bash
# Elastic scaling on one box — the local analog of the reference's scripts/scale_workers.sh
# (`up N` / `down N` = Terraform resize + kill and restart the parameter server with a new
# TOTAL_WORKERS, losing the in-memory model).  Here the model survives: before the restart the
# server is asked for a checkpoint (SaveCheckpoint RPC) and the first new worker passes it back
# as its checkpoint argument, so training continues from the saved parameters.
#
#   scripts/scale_workers.sh up 5 | down 2 | status | stop
set -euo pipefail

ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
RUN_DIR="${RUN_DIR:-/tmp/psb_run}"
BIN="$ROOT/build/bin"
COORDINATOR_PORT="${COORDINATOR_PORT:-50052}"
PS_PORT="${PS_PORT:-50051}"
ITERATIONS="${ITERATIONS:-100}"
CHECKPOINT_INTERVAL="${CHECKPOINT_INTERVAL:-10}"

alive() { [ -f "$1" ] && kill -0 "$(cat "$1")" 2>/dev/null; }
stop_pid() {  # TERM, then wait until the process is gone: its port must be free for a successor
  if alive "$1"; then
    local pid; pid="$(cat "$1")"
    kill "$pid" 2>/dev/null || true
    # (a zombie nobody reaps — containers without an init — still answers kill -0: read its state)
    running() { kill -0 "$1" 2>/dev/null && ! grep -q '^State:[[:space:]]*Z' "/proc/$1/status" 2>/dev/null; }
    for _ in $(seq 1 100); do running "$pid" || break; sleep 0.05; done
    running "$pid" && kill -9 "$pid" 2>/dev/null || true
  fi
  rm -f "$1"
}
workers() { ls "$RUN_DIR"/worker_*.pid 2>/dev/null | sed 's/.*worker_\([0-9]*\)\.pid/\1/' | sort -n; }

save_checkpoint() {
  python3 - "$ROOT" "$PS_PORT" <<'PY'
import sys
sys.path.insert(0, sys.argv[1])
import grpc
from parameter_server_distributed_b200.control import protos
stub = protos.Stub(grpc.insecure_channel(f"127.0.0.1:{sys.argv[2]}"), "parameter_server.ParameterServer")
r = stub.SaveCheckpoint(protos.msg("parameter_server.SaveCheckpointRequest")(epoch=0, path="scale.ckpt"), timeout=30)
print("checkpoint:", r.message, r.checkpoint_path)
PY
}

restart_ps() {  # $1 = new worker count
  if alive "$RUN_DIR/parameter_server.pid"; then save_checkpoint || true; fi
  stop_pid "$RUN_DIR/parameter_server.pid"
  PS_PORT="$PS_PORT" TOTAL_WORKERS="$1" CHECKPOINT_INTERVAL="$CHECKPOINT_INTERVAL" \
    BINARY_PATH="$BIN/parameter_server" LOG_FILE="$RUN_DIR/parameter_server.log" \
    PID_FILE="$RUN_DIR/parameter_server.pid" PS_WORKDIR="$RUN_DIR" PS_FLAGS="${PS_FLAGS:-}" \
    bash "$ROOT/scripts/start_parameter_server.sh"
  sleep 0.5
}

start_worker() {  # $1 = id, $2 = optional checkpoint
  WORKER_ID="$1" COORDINATOR_ADDR="127.0.0.1:$COORDINATOR_PORT" ITERATIONS="$ITERATIONS" \
    CHECKPOINT_PATH="${2:-}" BINARY_PATH="$BIN/worker_main" LOG_FILE="$RUN_DIR/worker_$1.log" \
    PID_FILE="$RUN_DIR/worker_$1.pid" WORKER_FLAGS="${WORKER_FLAGS:-}" \
    bash "$ROOT/scripts/start_worker.sh"
}

case "${1:-status}" in
  up|down)
    TARGET="${2:?target worker count}"
    for id in $(workers); do stop_pid "$RUN_DIR/worker_$id.pid"; done  # all restart at iteration 0
    restart_ps "$TARGET"
    CKPT=""; [ -f "$RUN_DIR/scale.ckpt" ] && CKPT="scale.ckpt"
    for i in $(seq 0 $((TARGET - 1))); do
      if [ "$i" -eq 0 ]; then start_worker "$i" "$CKPT"; sleep 0.5; else start_worker "$i"; fi
    done
    echo "scaled to $TARGET workers" ;;
  status)
    SERVERS="parameter_server"
    if ls "$RUN_DIR"/parameter_server_*.pid >/dev/null 2>&1; then  # PS_SHARDS > 1
      SERVERS="$(ls "$RUN_DIR"/parameter_server_*.pid | sed 's/.*\/\(parameter_server_[0-9]*\)\.pid/\1/' | sort)"
    fi
    for f in coordinator $SERVERS; do
      alive "$RUN_DIR/$f.pid" && echo "$f: running (pid $(cat "$RUN_DIR/$f.pid"))" || echo "$f: stopped"
    done
    for id in $(workers); do
      alive "$RUN_DIR/worker_$id.pid" && echo "worker $id: running" || echo "worker $id: finished"
    done ;;
  stop)
    for id in $(workers); do stop_pid "$RUN_DIR/worker_$id.pid"; done
    stop_pid "$RUN_DIR/parameter_server.pid"; stop_pid "$RUN_DIR/coordinator.pid"
    for f in "$RUN_DIR"/parameter_server_*.pid; do [ -f "$f" ] && stop_pid "$f"; done  # PS_SHARDS > 1
    echo "stopped" ;;
  *) echo "usage: $0 up N | down N | status | stop"; exit 2 ;;
esac
