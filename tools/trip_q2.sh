#!/bin/bash
mkdir -p gpurun_out
for t in mlp train pose; do
timeout 600 python -m pytest tests/test_gpu_$t.py -m gpu -q --tb=short -x > gpurun_out/t_$t.log 2>&1; echo "$t tests exit $?"; tail -2 gpurun_out/t_$t.log
done
for i in 1 2; do
timeout 900 python bench.py --no-cpu-baseline --no-render --steps 40 2>gpurun_out/bench_q.err | tail -n 1 > gpurun_out/bench_q.json; echo "bench exit $?"
python -c "
import json;d=json.load(open('gpurun_out/bench_q.json'));print(d['value'],d['ms_per_step'],d['e2e']['value']);print({k:(v['ms'],v['tflops']) for k,v in d['roofline']['per_kernel'].items()})" || tail -30 gpurun_out/bench_q.err
done
