#!/bin/bash
mkdir -p gpurun_out
for t in kernels render; do
timeout 600 python -m pytest tests/test_gpu_$t.py -m gpu -q --tb=short -x > gpurun_out/t_$t.log 2>&1; echo "$t tests exit $?"; tail -2 gpurun_out/t_$t.log
done
timeout 300 python tools/bench_stream.py 262144 > gpurun_out/bench_stream.log 2>&1; echo "stream exit $?"; cat gpurun_out/bench_stream.log | tail -14 | head -11
NB_SPDF_WARP=1 timeout 300 python tools/bench_stream.py 262144 2>&1 | grep sample_pdf_merge | head -2
timeout 600 python bench.py --no-cpu-baseline --no-render --steps 40 2>gpurun_out/bench_q.err | tail -n 1 > gpurun_out/bench_q.json; echo "bench exit $?"
python -c "
import json;d=json.load(open('gpurun_out/bench_q.json'));print(d['value'],d['ms_per_step'],d['e2e']['value'],d['clocks'])" || tail -20 gpurun_out/bench_q.err
