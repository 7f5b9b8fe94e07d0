#!/bin/bash
# A/B of the MLP step at N GPUs: own tcgen05 GEMMs (CLC / static tile walk) vs library GEMMs.
N=${1:-2}
run() { timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) bench.py --gpus $N --steps 20 --warmup 5 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', 'n=$N', round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), 'pp', d['push_pull_ms_per_step'])"; }
PSB_GEMM_SCHED=clc run clc
PSB_GEMM_SCHED=static run static
PSB_TCGEN05_LINEAR=0 run library
PSB_GEMM_SCHED=clc run clc
PSB_GEMM_SCHED=static run static
