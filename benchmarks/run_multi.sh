#!/bin/bash
# bench.py + kernel timeline at N GPUs (one box).  Usage: run_multi.sh N [extra bench args]
N=${1:-8}; shift
PORT=$((29500 + RANDOM % 500))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 20 --warmup 5 "$@" 2>gpurun_out/bench_${N}gpu.err | tail -1 > gpurun_out/bench_${N}gpu.json
cat gpurun_out/bench_${N}gpu.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('n=$N', round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), 'pp', d['push_pull_ms_per_step'], d['clocks'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((PORT+1)) benchmarks/timeline.py > gpurun_out/timeline_${N}gpu.tsv 2>gpurun_out/timeline_${N}gpu.err
wc -l gpurun_out/timeline_${N}gpu.tsv
if [ -n "$HYBRID" ]; then
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((PORT+2)) bench.py --gpus $N --steps 20 --warmup 5 --pull hybrid 2>gpurun_out/bench_${N}gpu_hybrid.err | tail -1 > gpurun_out/bench_${N}gpu_hybrid.json
cat gpurun_out/bench_${N}gpu_hybrid.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('hybrid n=$N', round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), 'pp', d['push_pull_ms_per_step'])"
fi
