#!/bin/bash
# second 8-GPU A/B: the round-2 default against each of its ingredients switched off, the multi-GPU
# test suite, and a kernel timeline of the default.
N=${1:-8}
OUT=gpurun_out
mkdir -p $OUT
source_ab() { :; }
ab() {
  local name=$1; shift
  local envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done
  shift
  local port=$((29500 + RANDOM % 400))
  env "${envs[@]}" timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --steps 30 --warmup 5 "$@" 2>$OUT/abb${N}_$name.err | tail -1 > $OUT/abb${N}_$name.json
  python - "$OUT/abb${N}_$name.json" "$name" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read())
    print(sys.argv[2].ljust(18), 'value', round(d['value']), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']),
          'pp', round(d.get('push_pull_ms_per_step') or 0, 4), 'loss', round(d.get('final_loss'), 6), d['clocks'].get('sm_mhz'), d['clocks'].get('reasons'), d['config'].get('buckets'))
except Exception as e:
    print(sys.argv[2], 'FAILED', e)
PY
}
timeout 500 python -m pytest tests/test_multigpu.py tests/test_elastic_gpu.py -x -q > $OUT/pytest_multigpu_${N}gpu.log 2>&1; echo "multigpu pytest rc=$?"; tail -3 $OUT/pytest_multigpu_${N}gpu.log
ab default X=1 --
ab nosplitwait PSB_SPLIT_WAIT=0 --
ab nosplit PSB_SPLIT_MB=0 --
ab p2p PSB_PUSH_NVLS=0 --
ab grid16 PSB_OVERLAP_GRID=16 --
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) benchmarks/timeline.py --steps 3 > $OUT/timeline_${N}gpu_default.tsv 2>$OUT/timeline_${N}gpu_default.err; wc -l $OUT/timeline_${N}gpu_default.tsv
