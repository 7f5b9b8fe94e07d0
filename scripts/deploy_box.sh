#!/bin/bash
# Single-box deployment — the local analog of the reference's scripts/deploy.sh (which builds,
# runs `terraform apply`, copies binaries to EC2 instances over ssh and starts coordinator,
# parameter server and workers there).  AWS / Terraform are out of scope for the 8xB200 target,
# so the same sequence runs on this machine: build -> coordinator -> parameter server ->
# workers, each through its start_*.sh script with logs and pidfiles under $RUN_DIR.
#
#   WORKER_COUNT=3 ITERATIONS=100 CHECKPOINT_INTERVAL=10 scripts/deploy_box.sh
#   PS_SHARDS=2 WORKER_COUNT=4 scripts/deploy_box.sh   # model split over 2 parameter servers
#   PS_SHARDS=8 PS_ON_GPUS=1 scripts/deploy_box.sh     # ... shard s keeps its weights in the HBM of GPU s
#   MODE=gpu GPUS=8 scripts/deploy_box.sh [script.py args...]   # HBM data plane: torchrun, 1 rank/GPU
set -euo pipefail

ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
RUN_DIR="${RUN_DIR:-/tmp/psb_run}"
WORKER_COUNT="${WORKER_COUNT:-3}"
ITERATIONS="${ITERATIONS:-100}"
CHECKPOINT_INTERVAL="${CHECKPOINT_INTERVAL:-10}"
COORDINATOR_PORT="${COORDINATOR_PORT:-50052}"
PS_PORT="${PS_PORT:-50051}"
MODE="${MODE:-host}"
PS_SHARDS="${PS_SHARDS:-1}"
mkdir -p "$RUN_DIR"

echo "[1/4] building"
make -C "$ROOT" -j8 binaries >/dev/null
BIN="$ROOT/build/bin"

if [ "$MODE" = "gpu" ]; then
  GPUS="${GPUS:-$(nvidia-smi -L | wc -l)}"
  python -c "import sys; sys.path.insert(0, '$ROOT'); from parameter_server_distributed_b200 import _build; _build.build_kernels()"
  SCRIPT="${1:-$ROOT/bench.py}"
  [ $# -gt 0 ] && shift
  echo "[2/4] starting coordinator on :$COORDINATOR_PORT"
  COORDINATOR_PORT="$COORDINATOR_PORT" PS_ADDRESS="127.0.0.1:$PS_PORT" BINARY_PATH="$BIN/coordinator" \
    LOG_FILE="$RUN_DIR/coordinator.log" PID_FILE="$RUN_DIR/coordinator.pid" \
    bash "$ROOT/scripts/start_coordinator.sh"
  trap 'kill "$(cat "$RUN_DIR/coordinator.pid")" 2>/dev/null || true' EXIT
  echo "[3/4] launching ${GPUS} ranks (one worker + one PS shard per GPU): $SCRIPT $*"
  echo "      the ranks register, exchange heap handles and synchronise through the coordinator"
  # the launcher only sets RANK / LOCAL_RANK / WORLD_SIZE; membership, shard map, handle
  # exchange and host barriers go through the coordinator (PSB_COORDINATOR)
  PSB_COORDINATOR="127.0.0.1:$COORDINATOR_PORT" python -m torch.distributed.run --nnodes=1 \
    --nproc-per-node "$GPUS" --master-addr 127.0.0.1 --master-port "${MASTER_PORT:-29500}" \
    "$SCRIPT" --gpus "$GPUS" "$@"
  echo "[4/4] done; coordinator log: $RUN_DIR/coordinator.log"
  exit 0
fi

echo "[2/4] starting coordinator on :$COORDINATOR_PORT"
COORDINATOR_PORT="$COORDINATOR_PORT" PS_ADDRESS="127.0.0.1:$PS_PORT" BINARY_PATH="$BIN/coordinator" \
  LOG_FILE="$RUN_DIR/coordinator.log" PID_FILE="$RUN_DIR/coordinator.pid" \
  bash "$ROOT/scripts/start_coordinator.sh"
if [ "$PS_SHARDS" -gt 1 ]; then
  echo "[3/4] starting $PS_SHARDS parameter-server shards on :$PS_PORT, :$((PS_PORT + 10)), .. for $WORKER_COUNT workers"
  for s in $(seq 0 $((PS_SHARDS - 1))); do
    mkdir -p "$RUN_DIR/shard_$s"   # every shard checkpoints into its own directory
    # shard s listens on PS_PORT + 10 s (PS_PORT + 1 is the coordinator's default port)
    SHARD_PORT="$((PS_PORT + 10 * s))"
    PS_PORT="$SHARD_PORT" TOTAL_WORKERS="$WORKER_COUNT" CHECKPOINT_INTERVAL="$CHECKPOINT_INTERVAL" \
      COORDINATOR_ADDR="127.0.0.1:$COORDINATOR_PORT" SHARD_ID="$s" \
      ADVERTISE_ADDR="127.0.0.1:$SHARD_PORT" \
      BINARY_PATH="$BIN/parameter_server" LOG_FILE="$RUN_DIR/parameter_server_$s.log" \
      PID_FILE="$RUN_DIR/parameter_server_$s.pid" PS_WORKDIR="$RUN_DIR/shard_$s" \
      PS_FLAGS="${PS_FLAGS:-}$([ "${PS_ON_GPUS:-0}" = 1 ] && echo " --device=$s")" \
      bash "$ROOT/scripts/start_parameter_server.sh"
  done
  WORKER_FLAGS="${WORKER_FLAGS:-} --shards=$PS_SHARDS"
else
echo "[3/4] starting parameter server on :$PS_PORT for $WORKER_COUNT workers"
PS_PORT="$PS_PORT" TOTAL_WORKERS="$WORKER_COUNT" CHECKPOINT_INTERVAL="$CHECKPOINT_INTERVAL" \
  COORDINATOR_ADDR="127.0.0.1:$COORDINATOR_PORT" \
  BINARY_PATH="$BIN/parameter_server" LOG_FILE="$RUN_DIR/parameter_server.log" \
  PID_FILE="$RUN_DIR/parameter_server.pid" PS_WORKDIR="$RUN_DIR" PS_FLAGS="${PS_FLAGS:-}" \
  bash "$ROOT/scripts/start_parameter_server.sh"
fi
sleep 0.5
echo "[4/4] starting $WORKER_COUNT workers x $ITERATIONS iterations"
for i in $(seq 0 $((WORKER_COUNT - 1))); do
  WORKER_ID="$i" COORDINATOR_ADDR="127.0.0.1:$COORDINATOR_PORT" ITERATIONS="$ITERATIONS" \
    BINARY_PATH="$BIN/worker_main" LOG_FILE="$RUN_DIR/worker_$i.log" PID_FILE="$RUN_DIR/worker_$i.pid" \
    WORKER_FLAGS="${WORKER_FLAGS:-}" bash "$ROOT/scripts/start_worker.sh"
done
echo "deployed: logs and pidfiles in $RUN_DIR (stop with scripts/scale_workers.sh stop)"
