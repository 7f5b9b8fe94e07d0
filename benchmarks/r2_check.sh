#!/bin/bash
# Round-2 check on N GPUs: GPU tests, then bench.py A/Bs.  Usage: r2_check.sh N [skip_tests]
N=${1:-2}
OUT=gpurun_out
mkdir -p $OUT
run_bench() {  # name, env..., -- args
  local name=$1; shift
  local port=$((29500 + RANDOM % 400))
  if [ "$N" -gt 1 ]; then
    env "${ENVV[@]}" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --steps 30 --warmup 5 "$@" 2>$OUT/$name.err | tail -1 > $OUT/$name.json
  else
    env "${ENVV[@]}" timeout 300 python bench.py --gpus 1 --steps 30 --warmup 5 "$@" 2>$OUT/$name.err | tail -1 > $OUT/$name.json
  fi
  python - "$OUT/$name.json" "$name" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read())
    print(sys.argv[2], 'value', round(d['value']), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']),
          'pp', d.get('push_pull_ms_per_step'), 'loss', d.get('final_loss'), d['clocks'].get('sm_mhz'), d['clocks'].get('reasons'))
except Exception as e:
    print(sys.argv[2], 'FAILED', e)
PY
}
if [ -z "$2" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_${N}gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/pytest_gpu_${N}gpu.log
fi
ENVV=(PSB_X=1); run_bench bench_${N}gpu_default
ENVV=(PSB_SPLIT_MB=0); run_bench bench_${N}gpu_nosplit
ENVV=(PSB_OVERLAP_GRID=16); run_bench bench_${N}gpu_grid16
ENVV=(PSB_OVERLAP_GRID=24); run_bench bench_${N}gpu_grid24
