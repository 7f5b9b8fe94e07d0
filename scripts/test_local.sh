#!/bin/bash
# Local plumbing test — BASELINE.json config 1: coordinator + 1 parameter server + 2 workers on
# localhost, 5 iterations of the synthetic gradient push/pull, CPU only.  Same topology and log
# files as the reference's scripts/test_local.sh, but with ASSERTIONS (the reference only
# `cat`s the logs): 10 x done=true, final weight == -0.03 (golden trace) and a
# byte-exact 458-byte checkpoint (App. B).  Exit code 0 = pass.
set -euo pipefail

ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
BIN="${BIN_DIR:-$ROOT/build/bin}"
COORD_PORT="${COORD_PORT:-50052}"
PS_PORT="${PS_PORT:-50051}"
ITERS="${ITERS:-5}"
LOG_DIR="${LOG_DIR:-/tmp}"
WORK="$(mktemp -d /tmp/psb_local.XXXXXX)"

if [ ! -x "$BIN/coordinator" ] || [ ! -x "$BIN/parameter_server" ] || [ ! -x "$BIN/worker_main" ]; then
  echo "building binaries..."
  make -C "$ROOT" -j8 binaries >/dev/null
fi

cleanup() {
  kill "${COORD_PID:-}" "${PS_PID:-}" 2>/dev/null || true
  rm -rf "$WORK"
}
trap cleanup EXIT

"$BIN/coordinator" "localhost:${COORD_PORT}" "localhost:${PS_PORT}" >"$LOG_DIR/coordinator.log" 2>&1 &
COORD_PID=$!
(cd "$WORK" && exec "$BIN/parameter_server" "localhost:${PS_PORT}" 2 0) >"$LOG_DIR/ps.log" 2>&1 &
PS_PID=$!
sleep 0.5

"$BIN/worker_main" "localhost:${COORD_PORT}" 0 "$ITERS" >"$LOG_DIR/worker0.log" 2>&1 &
W0=$!
"$BIN/worker_main" "localhost:${COORD_PORT}" 1 "$ITERS" >"$LOG_DIR/worker1.log" 2>&1 &
W1=$!
wait "$W0" "$W1"

for f in coordinator ps worker0 worker1; do
  echo "== $LOG_DIR/$f.log"
  cat "$LOG_DIR/$f.log"
done

fail() { echo "FAIL: $*"; exit 1; }
DONE=$(cat "$LOG_DIR/worker0.log" "$LOG_DIR/worker1.log" | grep -c "done=true" || true)
[ "$DONE" -eq $((2 * ITERS)) ] || fail "expected $((2 * ITERS)) x done=true, saw $DONE"
grep -q "coordinator listening on" "$LOG_DIR/coordinator.log" || fail "coordinator banner missing"
grep -q "parameter server listening on" "$LOG_DIR/ps.log" || fail "parameter server banner missing"

# final parameters + checkpoint through the wire API (python is only the verifier)
python3 - "$ROOT" "$PS_PORT" "$WORK" "$ITERS" <<'PY'
import sys
root, port, work, iters = sys.argv[1], int(sys.argv[2]), sys.argv[3], int(sys.argv[4])
sys.path.insert(0, root)
import grpc
from parameter_server_distributed_b200.control import protos
M = protos.msg
stub = protos.Stub(grpc.insecure_channel(f"localhost:{port}"), "parameter_server.ParameterServer")
p = stub.ServeParameters(M("parameter_server.PullRequest")(worker_id=0, iteration=iters - 1), timeout=10)
assert p.ready and len(p.parameters) == 1 and p.parameters[0].name == "weight"
exp = 0.01 - 0.01 * (iters - 1)
assert all(abs(v - exp) < 1e-6 for v in p.parameters[0].data), p.parameters[0].data[:4]
r = stub.SaveCheckpoint(M("parameter_server.SaveCheckpointRequest")(epoch=0), timeout=10)
assert r.success, r.message
raw = open(f"{work}/checkpoint_epoch_0.ckpt", "rb").read()
assert len(raw) == 458, len(raw)
print(f"final weight = {p.parameters[0].data[0]:.4f}, checkpoint = {len(raw)} bytes")
PY
echo "PASS: test_local ($DONE x done=true)"
