#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_train.py tests/test_gpu_mlp.py -m gpu -q -x 2>&1 | tail -3
for i in 1 2; do
  timeout 100 python tools/wgrad_skew.py 1024 192 2>&1 | grep "kernel span"
  timeout 100 python tools/wgrad_skew.py 1024 64 2>&1 | grep "kernel span"
done
python bench.py --no-cpu-baseline --no-render --no-sustained --steps 50 --warmup 5 2>gpurun_out/b.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['launch_ms'])
print({k:v['ms'] for k,v in d['roofline']['per_kernel'].items()})
"
