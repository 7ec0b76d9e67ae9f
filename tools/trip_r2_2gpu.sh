#!/bin/bash
# 2-GPU A/B: fused peer-memory all-reduce + Adam vs NCCL all-reduce + Adam
cd "$(dirname "$0")/.."
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus $1 --steps 50 --warmup 5 --no-render --no-sustained --no-cpu-baseline; }
for rep in 1 2; do
  NB_FUSED_AR=1 run 2 > gpurun_out/r2_2gpu_fused_$rep.json 2> gpurun_out/r2_2gpu_fused_$rep.err
  NB_NVLS=0 NB_FUSED_AR=1 run 2 > gpurun_out/r2_2gpu_fusedp2p_$rep.json 2> gpurun_out/r2_2gpu_fusedp2p_$rep.err
  NB_FUSED_AR=0 run 2 > gpurun_out/r2_2gpu_nccl_$rep.json 2> gpurun_out/r2_2gpu_nccl_$rep.err
done
python bench.py --steps 50 --warmup 5 --no-render --no-sustained --no-cpu-baseline > gpurun_out/r2_2gpu_single.json 2>/dev/null
for f in gpurun_out/r2_2gpu_*.json; do python -c "
import json,sys
try:
    d=json.load(open('$f')); print('$f', round(d['ms_per_step'],4), round(d['value']), 'e2e', round(d['e2e']['value']))
except Exception as e: print('$f ERR', e)"; done
