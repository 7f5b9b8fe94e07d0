for sched in static clc; do for blk in 512 256 128; do
  PSB_GEMM_SCHED=$sched PSB_PUSH_BLOCK=$blk timeout 100 python bench.py --steps 20 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$sched $blk', round(d['ms_per_step'],3), round(d['e2e']['ms_per_step'],3))"
done; done
for blk in 512 256; do PSB_TCGEN05_LINEAR=0 PSB_PUSH_BLOCK=$blk timeout 100 python bench.py --steps 20 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('library $blk', round(d['ms_per_step'],3), round(d['e2e']['ms_per_step'],3))"; done
