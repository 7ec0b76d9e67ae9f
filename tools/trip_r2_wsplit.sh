#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_train.py tests/test_gpu_mlp.py -m gpu -q -x 2>&1 | tail -3
for p in 0 15 25 40; do
  echo "== NB_WGRAD_POOL=$p"
  for i in 1 2; do
    NB_WGRAD_POOL=$p timeout 100 python tools/wgrad_skew.py 1024 192 2>&1 | grep "kernel span"
    NB_WGRAD_POOL=$p timeout 100 python tools/wgrad_skew.py 1024 64 2>&1 | grep "kernel span"
  done
done
for p in 0 25; do
NB_WGRAD_POOL=$p python bench.py --no-cpu-baseline --no-render --no-sustained --steps 100 --warmup 10 2>gpurun_out/b.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['launch_ms'])
"
done
