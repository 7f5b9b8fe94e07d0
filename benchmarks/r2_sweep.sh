#!/bin/bash
# bandwidth sweep with NVLS rows at N GPUs.  Usage: r2_sweep.sh N [max_mb]
N=${1:-2}; MAX=${2:-256}
PORT=$((29500 + RANDOM % 400))
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT benchmarks/bw_sweep.py --min-mb 16 --max-mb $MAX --tma-grids 16,32 > gpurun_out/bw_sweep_${N}gpu_r2.jsonl 2> gpurun_out/bw_sweep_${N}gpu_r2.err
echo "sweep rc=$?"; grep -c variant gpurun_out/bw_sweep_${N}gpu_r2.jsonl; tail -3 gpurun_out/bw_sweep_${N}gpu_r2.err
