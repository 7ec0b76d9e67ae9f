#!/bin/bash
# A/B of the two-branch training graph (NB_OVERLAP=0/1) + the train tests
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_train.py -m gpu -q --tb=short -x > gpurun_out/t_train.log 2>&1; echo "train tests exit $?"; tail -5 gpurun_out/t_train.log
for ov in 0 1 0 1; do
  NB_OVERLAP=$ov timeout 600 python bench.py --no-cpu-baseline --no-render --steps 40 2>gpurun_out/bench_ov$ov.err | tail -n 1 > gpurun_out/bench_ov$ov.json; echo "bench ov=$ov exit $?"
  python -c "
import json;d=json.load(open('gpurun_out/bench_ov$ov.json'));print('overlap=$ov', d['value'],d['ms_per_step'],d['e2e']['value'], d['gpu_launches'], d['config']['launch'])" || tail -20 gpurun_out/bench_ov$ov.err
done
