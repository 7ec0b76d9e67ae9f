#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_train.py tests/test_gpu_mlp.py tests/test_gpu_pose.py -m gpu -q -x 2>&1 | tail -3
for i in 1 2; do
python bench.py --no-cpu-baseline --no-render --no-sustained --steps 100 --warmup 10 2>gpurun_out/b.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'])
print({k:v['ms'] for k,v in d['roofline']['per_kernel'].items()})
"
done
