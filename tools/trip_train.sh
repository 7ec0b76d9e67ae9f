#!/bin/bash
mkdir -p gpurun_out
for t in mlp train; do
timeout 300 python -m pytest tests/test_gpu_$t.py -m gpu -q --tb=short -x > gpurun_out/t_$t.log 2>&1; echo "$t exit $?"; tail -4 gpurun_out/t_$t.log
done
timeout 600 python bench.py --no-cpu-baseline --no-render 2>gpurun_out/bench.err | tail -n 1 > gpurun_out/bench_quick.json; echo "bench exit $?"
python -c "
import json;d=json.load(open('gpurun_out/bench_quick.json'));print(d['value'],d['ms_per_step'],d['e2e'],{k:(v['ms'],v['tflops']) for k,v in d['roofline']['per_kernel'].items()})"
