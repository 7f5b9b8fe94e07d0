#!/bin/bash
# fused all-gather GEMM (K2f) vs NCCL all-gather + cuBLAS at N GPUs, gather-CTA sweep
N=${1:-2}
for g in auto 8 16 32; do
  PORT=$((29500 + RANDOM % 400))
  if [ $g = auto ]; then unset PSB_GATHER_CTAS; else export PSB_GATHER_CTAS=$g; fi
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT benchmarks/fused_gemm_bench.py 2>gpurun_out/fused_${N}gpu_g$g.err | grep '^{' > gpurun_out/fused_gemm_${N}gpu_g$g.jsonl
  echo "gather_ctas=$g"; python - gpurun_out/fused_gemm_${N}gpu_g$g.jsonl <<'PY'
import json,sys
for l in open(sys.argv[1]):
    d=json.loads(l); print(' ',d['M'],d['N'],d['K'],'fused',d['fused_ms'],'nccl+cublas',d['nccl_allgather_cublas_ms'],'local',d['local_gemm_ms'],'cublas',d['local_cublas_ms'],'frac',d['fused_frac_of_roofline'],d['gathered_weight_exact'])
PY
done
