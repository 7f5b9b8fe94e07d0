#!/bin/bash
# 8-GPU A/B of the push variants (bench.py, 30 steps each) + bandwidth sweep with NVLS rows.
N=${1:-8}
OUT=gpurun_out
mkdir -p $OUT
ab() {  # name, env assignments..., "--", bench args...
  local name=$1; shift
  local envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done
  shift
  local port=$((29500 + RANDOM % 400))
  env "${envs[@]}" timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --steps 30 --warmup 5 "$@" 2>$OUT/ab${N}_$name.err | tail -1 > $OUT/ab${N}_$name.json
  python - "$OUT/ab${N}_$name.json" "$name" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read())
    print(sys.argv[2].ljust(18), 'value', round(d['value']), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']),
          'pp', round(d.get('push_pull_ms_per_step') or 0, 4), 'loss', round(d.get('final_loss'), 6), d['clocks'].get('sm_mhz'), d['clocks'].get('reasons'))
except Exception as e:
    print(sys.argv[2], 'FAILED', e)
PY
}
ab default X=1 --
ab nosplit PSB_SPLIT_MB=0 --
ab grid16 PSB_OVERLAP_GRID=16 --
ab nvls X=1 -- --push nvls
ab nvls_nosplit PSB_SPLIT_MB=0 -- --push nvls
ab nvls_grid16 PSB_OVERLAP_GRID=16 -- --push nvls
ab hybrid X=1 -- --pull hybrid
ab nccl X=1 -- --impl nccl
if [ -z "$NO_SWEEP" ]; then bash benchmarks/r2_sweep.sh $N 256; fi
