#!/bin/bash
# A/B at N GPUs: alternating parameter replicas (early pushes) on / off
N=${1:-2}
run() { timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) bench.py --gpus $N --steps 15 --warmup 5 2>gpurun_out/ab_alt.err | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', 'n=$N', round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), 'loss', d['final_loss'], 'alt', d['config']['alt_param_replicas'])"; }
PSB_ALT_PARAMS=1 run alt
PSB_ALT_PARAMS=0 run noalt
