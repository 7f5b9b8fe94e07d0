#!/bin/bash
# Fault-injection companion of test_local.sh: coordinator + parameter server (following the
# coordinator's registry) + 3 workers on localhost; worker 2 is killed mid-run and the other
# two must finish every iteration with done=true.  The reference stalls for ever here: its
# coordinator evicts the dead worker but the parameter server never hears of it
# (/root/reference/src/coordinator.cpp:52-67, src/parameter_main.cpp:13-15).
#   BIN=build/bin-thread scripts/test_eviction.sh     # same run under ThreadSanitizer binaries
set -uo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
BIN="${BIN:-$ROOT/build/bin}"
case "$BIN" in /*) ;; *) BIN="$ROOT/$BIN" ;; esac
ITERS="${ITERS:-1500}"
COORD_PORT="${COORD_PORT:-47052}"
PS_PORT="${PS_PORT:-47051}"
[ -x "$BIN/coordinator" ] || make -C "$ROOT" -j8 binaries >/dev/null
WORK="$(mktemp -d)"
cleanup() { kill "${CO:-}" "${PS:-}" "${W2:-}" 2>/dev/null || true; rm -rf "$WORK"; }
trap cleanup EXIT
cd "$WORK"
PSB_STALE_SCAN_MS=100 PSB_STALE_TIMEOUT_S=1 "$BIN/coordinator" "127.0.0.1:$COORD_PORT" "127.0.0.1:$PS_PORT" >co.log 2>&1 &
CO=$!
"$BIN/parameter_server" "127.0.0.1:$PS_PORT" 3 0 "--coordinator=127.0.0.1:$COORD_PORT" --membership-poll-ms=100 >ps.log 2>&1 &
PS=$!
sleep 1
"$BIN/worker_main" "127.0.0.1:$COORD_PORT" 0 "$ITERS" --heartbeat-ms=100 --sync-timeout-ms=8000 >w0.log 2>&1 &
W0=$!
"$BIN/worker_main" "127.0.0.1:$COORD_PORT" 1 "$ITERS" --heartbeat-ms=100 --sync-timeout-ms=8000 >w1.log 2>&1 &
W1=$!
"$BIN/worker_main" "127.0.0.1:$COORD_PORT" 2 100000000 --heartbeat-ms=100 >w2.log 2>&1 &
W2=$!
# let the job get going, then kill worker 2 without ceremony
for _ in $(seq 1 400); do
  [ "$(grep -c "done=true" w0.log 2>/dev/null || true)" -ge 20 ] && break
  sleep 0.05
done
KILLED_AT=$(grep -c "done=true" w0.log || true)
kill -9 "$W2"
wait "$W0" "$W1"
[ "$KILLED_AT" -lt "$ITERS" ] || { echo "FAIL: worker 2 was killed after the others had finished"; exit 1; }
fail() { echo "FAIL: $*"; tail -n 5 ps.log co.log; exit 1; }
for w in 0 1; do
  ok=$(grep -c "done=true" "w$w.log" || true)
  bad=$(grep -c "done=false" "w$w.log" || true)
  [ "$ok" -eq "$ITERS" ] && [ "$bad" -eq 0 ] || fail "worker $w: $ok x done=true, $bad x done=false (wanted $ITERS / 0)"
done
grep -q "membership: total_workers 3 -> 2" ps.log || fail "the parameter server never shrank its barrier"
grep -q "evicted stale worker 2" co.log || fail "the coordinator never evicted worker 2"
if grep -q "ThreadSanitizer" ./*.log; then fail "ThreadSanitizer reported a race"; fi
echo "PASS: worker 2 killed at iteration $KILLED_AT; workers 0 and 1 finished all $ITERS with done=true ($(grep -h membership ps.log | tail -1))"
