#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_train.py tests/test_gpu_checkpoint.py -m gpu -q --tb=short -x > gpurun_out/t_train.log 2>&1; echo "train tests exit $?"; tail -2 gpurun_out/t_train.log
for i in 1 2 3; do
timeout 900 python bench.py --no-cpu-baseline --no-render --steps 40 2>gpurun_out/bench_q.err | tail -n 1 > gpurun_out/bench_q.json; echo "bench exit $?"
python -c "
import json;d=json.load(open('gpurun_out/bench_q.json'));print(d['value'],d['ms_per_step'],d['e2e']['value'])" || tail -30 gpurun_out/bench_q.err
done
timeout 300 python tools/timeline.py > gpurun_out/timeline.txt 2>gpurun_out/timeline.err; echo timeline exit $?; head -14 gpurun_out/timeline.txt | cut -c1-100; tail -1 gpurun_out/timeline.txt
