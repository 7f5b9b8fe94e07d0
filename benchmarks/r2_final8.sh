#!/bin/bash
# Final 8-GPU evidence run of round 2 (N = first argument): multi-GPU tests, sustained-clock runs of
# both arms, one more split A/B, K2f bench, NVLS sweep up to 1 GiB, ResNet-50.
N=${1:-8}
OUT=gpurun_out
mkdir -p $OUT
run() {  # name steps env... -- args
  local name=$1 steps=$2; shift 2
  local envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done
  shift
  local port=$((29500 + RANDOM % 400))
  if [ "$N" -gt 1 ]; then
    env "${envs[@]}" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --steps $steps --warmup 5 "$@" 2>$OUT/f${N}_$name.err | tail -1 > $OUT/f${N}_$name.json
  else
    env "${envs[@]}" timeout 400 python bench.py --gpus 1 --steps $steps --warmup 5 "$@" 2>$OUT/f${N}_$name.err | tail -1 > $OUT/f${N}_$name.json
  fi
  python - "$OUT/f${N}_$name.json" "$name" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read())
    print(sys.argv[2].ljust(20), 'value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value'], 1),
          'pp', round(d.get('push_pull_ms_per_step') or 0, 4), d['clocks'].get('sm_mhz'), d['clocks'].get('reasons'))
except Exception as e:
    print(sys.argv[2], 'FAILED', e)
PY
}
if [ "$N" -gt 1 ]; then
  timeout 600 python -m pytest tests/test_multigpu.py tests/test_elastic_gpu.py -x -q > $OUT/pytest_multigpu_${N}gpu.log 2>&1; echo "multigpu pytest rc=$?"; tail -3 $OUT/pytest_multigpu_${N}gpu.log
fi
run ours 30 X=1 --
run ours_2000 2000 X=1 --
run nccl 30 X=1 -- --impl nccl
run nccl_2000 2000 X=1 -- --impl nccl
if [ "$N" -ge 4 ] && [ -z "$NO_SPLIT32" ]; then run split32 30 PSB_SPLIT_MB=32 --; fi
run resnet50 20 X=1 -- --config resnet50 --optimizer adamw
if [ -n "$GPT2" ]; then run gpt2m 10 X=1 -- --config gpt2m --optimizer adamw; fi
if [ "$N" -gt 1 ] && [ -z "$NO_FUSED" ]; then
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) benchmarks/fused_gemm_bench.py 2>$OUT/fused_${N}gpu.err | grep '^{' > $OUT/fused_gemm_${N}gpu_r2.jsonl; wc -l $OUT/fused_gemm_${N}gpu_r2.jsonl
fi
if [ "$N" -gt 1 ] && [ -z "$NO_SWEEP" ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) benchmarks/bw_sweep.py --min-mb ${SWEEP_MIN:-64} --max-mb ${SWEEP_MAX:-1024} --tma-grids 32 > $OUT/bw_sweep_${N}gpu_r2b.jsonl 2> $OUT/bw_sweep_${N}gpu_r2b.err; echo "sweep rc=$?"; grep -c variant $OUT/bw_sweep_${N}gpu_r2b.jsonl
fi
